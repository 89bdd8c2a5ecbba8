// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
// extern "C" surface of the CPU oracle, loaded with ctypes from tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs. Parity status: see oracle/README.md.
//
// curve ids: 0 = Ristretto255, 1 = secp256k1, 2 = BLS12-381 G1.
// return codes mirror include/bpp.h: 0 ok, 1 verification failed, 2 format, 3 length/argument.
#include <map>
#include <mutex>
#include <thread>
#include <utility>
#include "range.hpp"

using namespace orc;

struct orc_params {  // same layout as bpp_params in include/bpp.h
  uint8_t y[32], z[32], e_final[32], alpha[32], r[32], s[32], delta[32], eta[32];
  uint32_t n_rounds;
  const uint8_t* e_round;
  const uint8_t* d_L;
  const uint8_t* d_R;
};

namespace {

template <class G> struct Ctx {
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;

  static const PublicKey<G>& pk(size_t len) {
    static std::map<size_t, PublicKey<G>> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(len);
    if (it == cache.end()) it = cache.emplace(len, PublicKey<G>::create(len)).first;
    return it->second;
  }
  static bool sc(const uint8_t* b, Sc& out) { return out.from_bytes_le(b, 32); }
  static bool load_params(const orc_params* p, size_t m, size_t rounds, Params<G>& out) {
    if (!p) { out = Params<G>::reference(m, rounds); return true; }
    bool ok = sc(p->y, out.y) && sc(p->z, out.z) && sc(p->e_final, out.e_final) && sc(p->alpha, out.alpha) &&
              sc(p->r, out.r) && sc(p->s, out.s) && sc(p->delta, out.delta) && sc(p->eta, out.eta);
    if (!ok || p->n_rounds < rounds) return false;
    out.e_round.resize(rounds); out.d_L.resize(rounds); out.d_R.resize(rounds);
    for (size_t i = 0; i < rounds; i++)
      if (!sc(p->e_round + 32 * i, out.e_round[i]) || !sc(p->d_L + 32 * i, out.d_L[i]) || !sc(p->d_R + 32 * i, out.d_R[i]))
        return false;
    return true;
  }
  static size_t wip_len(size_t rounds) { return (2 * rounds + 2) * G::PB + 96; }
  static void put_wip(const WipProof<G>& w, uint8_t* o) {
    for (auto& p : w.L_vec) { G::encode(p, o); o += G::PB; }
    for (auto& p : w.R_vec) { G::encode(p, o); o += G::PB; }
    G::encode(w.A, o); o += G::PB;
    G::encode(w.B, o); o += G::PB;
    w.r_prime.to_bytes_le(o, 32); w.s_prime.to_bytes_le(o + 32, 32); w.d_prime.to_bytes_le(o + 64, 32);
  }
  static int get_wip(const uint8_t* b, size_t rounds, WipProof<G>& w) {
    w.L_vec.resize(rounds); w.R_vec.resize(rounds);
    for (size_t i = 0; i < rounds; i++) { if (!G::decode(b, w.L_vec[i])) return 2; b += G::PB; }
    for (size_t i = 0; i < rounds; i++) { if (!G::decode(b, w.R_vec[i])) return 2; b += G::PB; }
    if (!G::decode(b, w.A)) return 2; b += G::PB;
    if (!G::decode(b, w.B)) return 2; b += G::PB;
    if (!sc(b, w.r_prime) || !sc(b + 32, w.s_prime) || !sc(b + 64, w.d_prime)) return 2;
    return 0;
  }
  static size_t log2_exact(size_t n) { size_t l = 0; while (((size_t)1 << l) < n) l++; return l; }

  static int pk_new(size_t len, uint8_t* g, uint8_t* h, uint8_t* Gs, uint8_t* Hs) {
    const PublicKey<G>& k = pk(len);
    G::encode(k.g, g); G::encode(k.h, h);
    for (size_t i = 0; i < len; i++) { G::encode(k.G_vec[i], Gs + i * G::PB); G::encode(k.H_vec[i], Hs + i * G::PB); }
    return 0;
  }
  static int commit(size_t count, const uint64_t* v, const uint8_t* gamma, uint8_t* out) {
    const PublicKey<G>& k = pk(0);
    for (size_t i = 0; i < count; i++) {
      Sc gm; if (!sc(gamma + 32 * i, gm)) return 2;
      G::encode(k.commit_u64(v[i], gm), out + i * G::PB);
    }
    return 0;
  }
  static int msm_naive(size_t k, const uint8_t* scalars, const uint8_t* points, uint8_t* out) {
    MulVec<G> mv;
    for (size_t i = 0; i < k; i++) {
      Sc s; Point p;
      if (!sc(scalars + 32 * i, s) || !G::decode(points + G::PB * i, p)) return 2;
      mv.add_scalar(s); mv.add_point(p);
    }
    G::encode(mv.calculate(), out);
    return 0;
  }
  static int scalar_mul_pt(const uint8_t* k32, const uint8_t* pt, uint8_t* out) {
    Sc s; Point p;
    if (!sc(k32, s)) return 2;
    if (pt) { if (!G::decode(pt, p)) return 2; } else p = G::base();
    G::encode(scalar_mul<G>(p, s), out);
    return 0;
  }
  static int point_add(const uint8_t* a, const uint8_t* b, uint8_t* out) {
    Point p, q;
    if (!G::decode(a, p) || !G::decode(b, q)) return 2;
    G::encode(G::add(p, q), out);
    return 0;
  }
  static int point_neg(const uint8_t* a, uint8_t* out) {
    Point p;
    if (!G::decode(a, p)) return 2;
    G::encode(G::neg(p), out);
    return 0;
  }
  static int sc_op(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
    Sc x, y;
    if (!sc(a, x) || (b && !sc(b, y))) return 2;
    Sc r;
    switch (op) {
      case 0: r = x + y; break;
      case 1: r = x - y; break;
      case 2: r = x * y; break;
      case 3: r = x.inv(); break;
      case 4: r = -x; break;
      default: return 3;
    }
    r.to_bytes_le(out, 32);
    return 0;
  }
  static int fq_op(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
    typename G::Fq x, y;
    const size_t FB = 8 * G::Fq::N;
    if (!x.from_bytes_le(a, FB) || (b && !y.from_bytes_le(b, FB))) return 2;
    typename G::Fq r;
    switch (op) {
      case 0: r = x + y; break;
      case 1: r = x - y; break;
      case 2: r = x * y; break;
      case 3: r = x.inv(); break;
      case 4: r = -x; break;
      default: return 3;
    }
    r.to_bytes_le(out, FB);
    return 0;
  }
  static int sc_reduce_wide(const uint8_t* in, size_t nbytes, uint8_t* out) {
    Sc::from_bytes_reduce(in, nbytes).to_bytes_le(out, 32);
    return 0;
  }

  static int range_prove(size_t n, size_t m, const uint64_t* v, const uint8_t* gamma, const orc_params* prm,
                         uint8_t* proof, size_t* proof_len) {
    size_t mn = n * m;
    if (mn == 0 || (mn & (mn - 1))) return 3;
    size_t rounds = log2_exact(mn);
    size_t need = G::PB + wip_len(rounds);
    if (*proof_len < need) { *proof_len = need; return 3; }
    Params<G> P;
    if (!load_params(prm, m, rounds, P)) return 2;
    const PublicKey<G>& k = pk(mn);
    RangeProver<G> prover;
    for (size_t j = 0; j < m; j++) {
      Sc gm; if (!sc(gamma + 32 * j, gm)) return 2;
      prover.commit(k, v[j], gm);
    }
    RangeProof<G> rp = RangeProof<G>::prove(k, n, prover, P);
    G::encode(rp.A, proof);
    put_wip(rp.proof, proof + G::PB);
    *proof_len = need;
    return 0;
  }
  static int range_verify(size_t n, size_t m, const uint8_t* V, const uint8_t* proof, size_t proof_len, const orc_params* prm) {
    size_t mn = n * m;
    if (mn == 0 || (mn & (mn - 1))) return 3;
    size_t rounds = log2_exact(mn);
    if (proof_len != G::PB + wip_len(rounds)) return 2;
    Params<G> P;
    if (!load_params(prm, m, rounds, P)) return 2;
    RangeProof<G> rp;
    if (!G::decode(proof, rp.A)) return 2;
    int rc = get_wip(proof + G::PB, rounds, rp.proof);
    if (rc) return rc;
    std::vector<Point> Vs(m);
    for (size_t j = 0; j < m; j++) if (!G::decode(V + j * G::PB, Vs[j])) return 2;
    return rp.verify(pk(mn), n, Vs, P) ? 0 : 1;
  }
  static int wip_prove(size_t n, const uint8_t* a, const uint8_t* b, const uint8_t* ypow, const uint8_t* gamma,
                       const orc_params* prm, uint8_t* proof, size_t* proof_len) {
    if (n == 0 || (n & (n - 1))) return 3;
    size_t rounds = log2_exact(n);
    size_t need = wip_len(rounds);
    if (*proof_len < need) { *proof_len = need; return 3; }
    Params<G> P;
    if (!load_params(prm, 2, rounds, P)) return 2;
    std::vector<Sc> av(n), bv(n), yv(n);
    Sc gm;
    if (!sc(gamma, gm)) return 2;
    for (size_t i = 0; i < n; i++)
      if (!sc(a + 32 * i, av[i]) || !sc(b + 32 * i, bv[i]) || !sc(ypow + 32 * i, yv[i])) return 2;
    WipProof<G> w = WipProof<G>::prove(pk(n), av, bv, yv, gm, P);
    put_wip(w, proof);
    *proof_len = need;
    return 0;
  }
  static int wip_verify(size_t n, const uint8_t* ypow, const uint8_t* G_exp_c, const uint8_t* H_exp_c, const uint8_t* g_exp_c,
                        size_t nV, const uint8_t* V_exp_c, const uint8_t* A_prime, const uint8_t* V, const uint8_t* proof,
                        size_t proof_len, const orc_params* prm) {
    if (n == 0 || (n & (n - 1))) return 3;
    size_t rounds = log2_exact(n);
    if (proof_len != wip_len(rounds)) return 2;
    Params<G> P;
    if (!load_params(prm, 2, rounds, P)) return 2;
    WipProof<G> w;
    int rc = get_wip(proof, rounds, w);
    if (rc) return rc;
    std::vector<Sc> yv(n), ge(n), he(n), ve(nV);
    std::vector<Point> Vs(nV);
    Sc gc; Point Ap;
    if (!sc(g_exp_c, gc) || !G::decode(A_prime, Ap)) return 2;
    for (size_t i = 0; i < n; i++)
      if (!sc(ypow + 32 * i, yv[i]) || !sc(G_exp_c + 32 * i, ge[i]) || !sc(H_exp_c + 32 * i, he[i])) return 2;
    for (size_t i = 0; i < nV; i++)
      if (!sc(V_exp_c + 32 * i, ve[i]) || !G::decode(V + i * G::PB, Vs[i])) return 2;
    return w.verify(pk(n), yv, ge, he, gc, ve, Ap, Vs, P) ? 0 : 1;
  }
};

// B2 hook for Ristretto255: oracle points -> 128-byte extended coordinates -> baseline_msm.cpp -> oracle point
extern "C" int b2_msm_ext(size_t k, const uint8_t* scalars, const uint8_t* points, int threads, uint8_t* out);
int g_b2_threads = 1;
Ristretto::Point b2_hook(const std::vector<Ristretto::Sc>& sc, const std::vector<Ristretto::Point>& pts) {
  const size_t k = sc.size();
  std::vector<uint8_t> sb(32 * k), pb(128 * k);
  for (size_t i = 0; i < k; i++) {
    sc[i].to_bytes_le(&sb[32 * i], 32);
    pts[i].X.to_bytes_le(&pb[128 * i], 32); pts[i].Y.to_bytes_le(&pb[128 * i + 32], 32);
    pts[i].Z.to_bytes_le(&pb[128 * i + 64], 32); pts[i].T.to_bytes_le(&pb[128 * i + 96], 32);
  }
  uint8_t out[128];
  b2_msm_ext(k, sb.data(), pb.data(), g_b2_threads, out);
  Ristretto::Point r;
  r.X.from_bytes_le(out, 32); r.Y.from_bytes_le(out + 32, 32); r.Z.from_bytes_le(out + 64, 32); r.T.from_bytes_le(out + 96, 32);
  return r;
}

template <class F> int guard(F f) {
  try { return f(); } catch (const std::exception&) { return 3; }
}

}  // namespace

#define DISPATCH(curve, call)                                  \
  switch (curve) {                                             \
    case 0: return guard([&] { return Ctx<Ristretto>::call; });  \
    case 1: return guard([&] { return Ctx<Secp256k1>::call; });  \
    case 2: return guard([&] { return Ctx<Bls12381G1>::call; }); \
    default: return 3;                                         \
  }

extern "C" {

// CPU baselines only (bench.py): backend 0 = the reference's per-term loop (B1, default), 1 = bucket MSM of baseline_msm.cpp
// (B2, Ristretto255 only) with `threads` host threads inside each MSM. Process-wide; not thread-safe against running calls.
void orc_set_msm_backend(int backend, int threads) {
  MsmHook<Ristretto>::fn() = b2_hook;
  g_b2_threads = threads < 1 ? 1 : threads;
  msm_backend() = backend;
}
// B2 MSM timing (Ristretto255): n points built from two oracle points, MSM alone timed inside baseline_msm.cpp
extern "C" double b2_bench_msm(size_t n, const uint8_t* p0, const uint8_t* q, int threads, uint8_t* out);
double orc_b2_bench_msm(size_t n, int threads) {
  auto ext = [](const Ristretto::Point& p, uint8_t* b) {
    p.X.to_bytes_le(b, 32); p.Y.to_bytes_le(b + 32, 32); p.Z.to_bytes_le(b + 64, 32); p.T.to_bytes_le(b + 96, 32);
  };
  uint8_t p0[128], q[128], out[128];
  ext(scalar_mul<Ristretto>(Ristretto::base(), Ristretto::Sc::from_i32(1234567)), p0);
  ext(scalar_mul<Ristretto>(Ristretto::base(), Ristretto::Sc::from_i32(7654321)), q);
  return b2_bench_msm(n, p0, q, threads, out);
}
int orc_point_bytes(int curve) { return curve == 0 ? 32 : curve == 1 ? 64 : curve == 2 ? 96 : -1; }
int orc_pk_new(int curve, size_t len, uint8_t* g, uint8_t* h, uint8_t* G, uint8_t* H) { DISPATCH(curve, pk_new(len, g, h, G, H)) }
int orc_commit(int curve, size_t count, const uint64_t* v, const uint8_t* gamma, uint8_t* out) { DISPATCH(curve, commit(count, v, gamma, out)) }
int orc_msm_naive(int curve, size_t k, const uint8_t* scalars, const uint8_t* points, uint8_t* out) { DISPATCH(curve, msm_naive(k, scalars, points, out)) }
// point == NULL -> base point
int orc_scalar_mul(int curve, const uint8_t* k32, const uint8_t* point, uint8_t* out) { DISPATCH(curve, scalar_mul_pt(k32, point, out)) }
int orc_point_add(int curve, const uint8_t* a, const uint8_t* b, uint8_t* out) { DISPATCH(curve, point_add(a, b, out)) }
int orc_point_neg(int curve, const uint8_t* a, uint8_t* out) { DISPATCH(curve, point_neg(a, out)) }
// op: 0 add, 1 sub, 2 mul, 3 inv(a), 4 neg(a)
int orc_sc_op(int curve, int op, const uint8_t* a, const uint8_t* b, uint8_t* out) { DISPATCH(curve, sc_op(op, a, b, out)) }
int orc_fq_op(int curve, int op, const uint8_t* a, const uint8_t* b, uint8_t* out) { DISPATCH(curve, fq_op(op, a, b, out)) }
int orc_sc_reduce_wide(int curve, const uint8_t* in, size_t nbytes, uint8_t* out) { DISPATCH(curve, sc_reduce_wide(in, nbytes, out)) }
int orc_range_prove(int curve, size_t n, size_t m, const uint64_t* v, const uint8_t* gamma, const orc_params* prm,
                    uint8_t* proof, size_t* proof_len) { DISPATCH(curve, range_prove(n, m, v, gamma, prm, proof, proof_len)) }
int orc_range_verify(int curve, size_t n, size_t m, const uint8_t* V, const uint8_t* proof, size_t proof_len,
                     const orc_params* prm) { DISPATCH(curve, range_verify(n, m, V, proof, proof_len, prm)) }
int orc_wip_prove(int curve, size_t n, const uint8_t* a, const uint8_t* b, const uint8_t* ypow, const uint8_t* gamma,
                  const orc_params* prm, uint8_t* proof, size_t* proof_len) { DISPATCH(curve, wip_prove(n, a, b, ypow, gamma, prm, proof, proof_len)) }
int orc_wip_verify(int curve, size_t n, const uint8_t* ypow, const uint8_t* G_exp_c, const uint8_t* H_exp_c,
                   const uint8_t* g_exp_c, size_t nV, const uint8_t* V_exp_c, const uint8_t* A_prime, const uint8_t* V,
                   const uint8_t* proof, size_t proof_len, const orc_params* prm) {
  DISPATCH(curve, wip_verify(n, ypow, G_exp_c, H_exp_c, g_exp_c, nV, V_exp_c, A_prime, V, proof, proof_len, prm))
}

}  // extern "C"
