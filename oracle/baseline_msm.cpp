// ORACLE / BASELINE — TEST AND MEASUREMENT INFRASTRUCTURE ONLY. Nothing under bulletproofsplus_b200/ links or loads this.
//
// B2, the "fair CPU" baseline of SURVEY.md §8(d): the SAME algorithm family as the GPU path — signed-digit bucket
// (Pippenger) multi-scalar multiplication over edwards25519 in extended coordinates — written the way a tuned CPU
// library does it: GF(2^255-19) in five 51-bit limbs with unsigned __int128 products, one MSM per host thread
// (batch verification) or one point range per thread (a single large MSM), std::thread over all cores.
// It stands beside B1 (protocol.hpp `MulVec::calculate`, the literal restatement of the reference's per-term
// double-and-add loop, src/bls12_381/building_block/mulvec.rs:20-33) so that a GPU/CPU ratio is not inflated by the
// reference's naivety. It is NOT the reference's algorithm and is never used as a parity oracle; its own results are
// checked against B1 / the exponent-domain value (tests/test_oracle_kat.py).
//
// Points cross this file's boundary as 128 bytes: affine (x, y) are not needed — the callers hand over extended
// coordinates X|Y|Z|T, 32 bytes little-endian each, canonical (< p), exactly what oracle_api.cpp can export from its own
// Ristretto point type, and results go back the same way (the caller encodes them with the oracle's RFC 9496 encoder).
#include <cstdint>
#include <cstring>
#include <chrono>
#include <thread>
#include <vector>

namespace b2 {

typedef unsigned __int128 u128;
static const uint64_t MASK51 = (1ULL << 51) - 1;

struct fe { uint64_t v[5]; };

static inline fe fe_zero() { fe r = {{0, 0, 0, 0, 0}}; return r; }
static inline fe fe_one() { fe r = {{1, 0, 0, 0, 0}}; return r; }
static inline fe fe_add(const fe& a, const fe& b) {
  fe r;
  for (int i = 0; i < 5; i++) r.v[i] = a.v[i] + b.v[i];
  return r;
}
// a - b with a bias of 4p so that limbs stay non-negative (inputs: limbs < 2^53)
static inline fe fe_sub(const fe& a, const fe& b) {
  fe r;
  r.v[0] = a.v[0] + 0x1FFFFFFFFFFFB4ULL - b.v[0];
  for (int i = 1; i < 5; i++) r.v[i] = a.v[i] + 0x1FFFFFFFFFFFFCULL - b.v[i];
  return r;
}
static inline fe fe_carry(u128 t0, u128 t1, u128 t2, u128 t3, u128 t4) {
  fe r;
  t1 += (uint64_t)(t0 >> 51); r.v[0] = (uint64_t)t0 & MASK51;
  t2 += (uint64_t)(t1 >> 51); r.v[1] = (uint64_t)t1 & MASK51;
  t3 += (uint64_t)(t2 >> 51); r.v[2] = (uint64_t)t2 & MASK51;
  t4 += (uint64_t)(t3 >> 51); r.v[3] = (uint64_t)t3 & MASK51;
  uint64_t c = (uint64_t)(t4 >> 51); r.v[4] = (uint64_t)t4 & MASK51;
  r.v[0] += c * 19;
  c = r.v[0] >> 51; r.v[0] &= MASK51;
  r.v[1] += c;
  return r;
}
// inputs with limbs < 2^55 (sums / biased differences of reduced elements); output limbs < 2^51 + small
static inline fe fe_mul(const fe& a, const fe& b) {
  const uint64_t a0 = a.v[0], a1 = a.v[1], a2 = a.v[2], a3 = a.v[3], a4 = a.v[4];
  const uint64_t b0 = b.v[0], b1 = b.v[1], b2 = b.v[2], b3 = b.v[3], b4 = b.v[4];
  const uint64_t b1_19 = b1 * 19, b2_19 = b2 * 19, b3_19 = b3 * 19, b4_19 = b4 * 19;
  u128 t0 = (u128)a0 * b0 + (u128)a1 * b4_19 + (u128)a2 * b3_19 + (u128)a3 * b2_19 + (u128)a4 * b1_19;
  u128 t1 = (u128)a0 * b1 + (u128)a1 * b0 + (u128)a2 * b4_19 + (u128)a3 * b3_19 + (u128)a4 * b2_19;
  u128 t2 = (u128)a0 * b2 + (u128)a1 * b1 + (u128)a2 * b0 + (u128)a3 * b4_19 + (u128)a4 * b3_19;
  u128 t3 = (u128)a0 * b3 + (u128)a1 * b2 + (u128)a2 * b1 + (u128)a3 * b0 + (u128)a4 * b4_19;
  u128 t4 = (u128)a0 * b4 + (u128)a1 * b3 + (u128)a2 * b2 + (u128)a3 * b1 + (u128)a4 * b0;
  return fe_carry(t0, t1, t2, t3, t4);
}
static inline fe fe_sqr(const fe& a) {
  const uint64_t a0 = a.v[0], a1 = a.v[1], a2 = a.v[2], a3 = a.v[3], a4 = a.v[4];
  const uint64_t d0 = 2 * a0, d1 = 2 * a1, d2 = 2 * a2, a3_19 = 19 * a3, a4_19 = 19 * a4;
  u128 t0 = (u128)a0 * a0 + (u128)d1 * a4_19 + (u128)d2 * a3_19;
  u128 t1 = (u128)d0 * a1 + (u128)d2 * a4_19 + (u128)a3 * a3_19;
  u128 t2 = (u128)d0 * a2 + (u128)a1 * a1 + (u128)(2 * a3) * a4_19;
  u128 t3 = (u128)d0 * a3 + (u128)d1 * a2 + (u128)a4 * a4_19;
  u128 t4 = (u128)d0 * a4 + (u128)d1 * a3 + (u128)a2 * a2;
  return fe_carry(t0, t1, t2, t3, t4);
}
static inline fe fe_weak(const fe& a) { return fe_carry(a.v[0], a.v[1], a.v[2], a.v[3], a.v[4]); }

static fe fe_from_bytes(const uint8_t* s) {  // 32 bytes little-endian, value < 2^255
  uint64_t w[4];
  memcpy(w, s, 32);
  fe r;
  r.v[0] = w[0] & MASK51;
  r.v[1] = ((w[0] >> 51) | (w[1] << 13)) & MASK51;
  r.v[2] = ((w[1] >> 38) | (w[2] << 26)) & MASK51;
  r.v[3] = ((w[2] >> 25) | (w[3] << 39)) & MASK51;
  r.v[4] = (w[3] >> 12) & MASK51;
  return r;
}
static void fe_to_bytes(uint8_t* s, const fe& a) {  // canonical
  fe t = fe_weak(fe_weak(a));
  uint64_t q = (t.v[0] + 19) >> 51;  // q = 1 iff t >= p
  for (int i = 1; i < 5; i++) q = (t.v[i] + q) >> 51;
  t.v[0] += 19 * q;
  uint64_t c = 0;
  for (int i = 0; i < 5; i++) { t.v[i] += c; c = t.v[i] >> 51; t.v[i] &= MASK51; }
  uint64_t w[4];
  w[0] = t.v[0] | (t.v[1] << 51);
  w[1] = (t.v[1] >> 13) | (t.v[2] << 38);
  w[2] = (t.v[2] >> 26) | (t.v[3] << 25);
  w[3] = (t.v[3] >> 39) | (t.v[4] << 12);
  memcpy(s, w, 32);
}

// ---- edwards25519, a = -1, extended coordinates ---------------------------------------------------------------------
struct ge { fe X, Y, Z, T; };          // T = XY/Z
struct ge_cached { fe ypx, ymx, Z, t2d; };  // (Y+X, Y-X, Z, 2dT)

static fe fe_2d() {  // 2d, d = -121665/121666 (little-endian bytes of the canonical value)
  static const uint8_t b[32] = {0x59, 0xf1, 0xb2, 0x26, 0x94, 0x9b, 0xd6, 0xeb, 0x56, 0xb1, 0x83, 0x82, 0x9a, 0x14, 0xe0, 0x00,
                                0x30, 0xd1, 0xf3, 0xee, 0xf2, 0x80, 0x8e, 0x19, 0xe7, 0xfc, 0xdf, 0x56, 0xdc, 0xd9, 0x06, 0x24};
  return fe_from_bytes(b);
}
static const fe K2D = fe_2d();

static inline ge ge_identity() { ge r; r.X = fe_zero(); r.Y = fe_one(); r.Z = fe_one(); r.T = fe_zero(); return r; }
static inline ge_cached ge_to_cached(const ge& p) {
  ge_cached c;
  c.ypx = fe_weak(fe_add(p.Y, p.X)); c.ymx = fe_weak(fe_sub(p.Y, p.X)); c.Z = p.Z; c.t2d = fe_mul(p.T, K2D);
  return c;
}
// p + q (neg: p - q); add-2008-hwcd-3: 8M
static inline ge ge_add_cached(const ge& p, const ge_cached& q, bool neg) {
  const fe& qp = neg ? q.ymx : q.ypx;
  const fe& qm = neg ? q.ypx : q.ymx;
  fe A = fe_mul(fe_sub(p.Y, p.X), qm);
  fe B = fe_mul(fe_add(p.Y, p.X), qp);
  fe C = fe_mul(p.T, q.t2d);
  fe D = fe_mul(p.Z, q.Z); D = fe_add(D, D);
  fe E = fe_sub(B, A), H = fe_add(B, A);
  fe F = neg ? fe_add(D, C) : fe_sub(D, C);
  fe G = neg ? fe_sub(D, C) : fe_add(D, C);
  ge r;
  r.X = fe_mul(E, F); r.Y = fe_mul(G, H); r.T = fe_mul(E, H); r.Z = fe_mul(F, G);
  return r;
}
static inline ge ge_add(const ge& p, const ge& q) { return ge_add_cached(p, ge_to_cached(q), false); }  // 9M
// dbl-2008-hwcd with a = -1: 4S + 4M
static inline ge ge_dbl(const ge& p) {
  fe A = fe_sqr(p.X), B = fe_sqr(p.Y), C = fe_sqr(p.Z);
  C = fe_add(C, C);
  fe xy = fe_add(p.X, p.Y);
  fe E = fe_sub(fe_sub(fe_sqr(xy), A), B);
  fe G = fe_sub(B, A);           // D + B with D = -A
  fe F = fe_sub(G, C);
  fe H = fe_sub(fe_zero(), fe_add(A, B));  // D - B
  H = fe_weak(H); F = fe_weak(F);
  ge r;
  r.X = fe_mul(E, F); r.Y = fe_mul(G, H); r.T = fe_mul(E, H); r.Z = fe_mul(F, G);
  return r;
}
static inline ge ge_neg(const ge& p) {
  ge r = p;
  r.X = fe_weak(fe_sub(fe_zero(), p.X)); r.T = fe_weak(fe_sub(fe_zero(), p.T));
  return r;
}
static ge ge_from_bytes(const uint8_t* b) {
  ge p;
  p.X = fe_from_bytes(b); p.Y = fe_from_bytes(b + 32); p.Z = fe_from_bytes(b + 64); p.T = fe_from_bytes(b + 96);
  return p;
}
static void ge_to_bytes(uint8_t* b, const ge& p) {
  fe_to_bytes(b, p.X); fe_to_bytes(b + 32, p.Y); fe_to_bytes(b + 64, p.Z); fe_to_bytes(b + 96, p.T);
}

// ---- signed-digit Pippenger over one range of terms -----------------------------------------------------------------
static inline uint32_t bits_at(const uint8_t* s, unsigned off, unsigned c) {  // c <= 24 bits at bit offset off of a 256-bit LE value
  if (off >= 256) return 0;
  uint64_t w = 0;
  const unsigned byte = off >> 3;
  for (unsigned i = 0; i < 5 && byte + i < 32; i++) w |= (uint64_t)s[byte + i] << (8 * i);
  return (uint32_t)(w >> (off & 7)) & ((1u << c) - 1u);
}
static unsigned choose_c(size_t k) {
  unsigned best = 1;
  double best_cost = 1e300;
  for (unsigned c = 1; c <= 20; c++) {
    const double W = (254 + c - 1) / c;
    const double cost = W * ((double)k * 8 + 2.0 * (double)(1u << (c - 1)) * 9) + 254.0 * 8;
    if (cost < best_cost) { best_cost = cost; best = c; }
  }
  return best;
}
static ge pippenger(size_t k, const uint8_t* scalars, const ge* pts, const ge_cached* cached) {
  if (k == 0) return ge_identity();
  const unsigned c = choose_c(k), W = (254 + c - 1) / c;
  const uint32_t half = 1u << (c - 1);
  // signed digits, window-major so that each window's pass streams its digits
  std::vector<int32_t> dig((size_t)W * k);
  for (size_t i = 0; i < k; i++) {
    uint32_t carry = 0;
    for (unsigned w = 0; w < W; w++) {
      uint32_t raw = bits_at(scalars + 32 * i, w * c, c) + carry;
      carry = raw > half;
      dig[(size_t)w * k + i] = carry ? (int32_t)raw - (int32_t)(1u << c) : (int32_t)raw;
    }
  }
  std::vector<ge> bucket(half);
  std::vector<uint8_t> live(half);
  ge acc = ge_identity();
  bool acc_live = false;
  for (int w = (int)W - 1; w >= 0; w--) {
    if (acc_live) for (unsigned j = 0; j < c; j++) acc = ge_dbl(acc);
    std::fill(live.begin(), live.end(), 0);
    const int32_t* d = &dig[(size_t)w * k];
    for (size_t i = 0; i < k; i++) {
      const int32_t x = d[i];
      if (!x) continue;
      const bool neg = x < 0;
      const uint32_t b = (uint32_t)(neg ? -x : x) - 1;
      if (live[b]) bucket[b] = ge_add_cached(bucket[b], cached[i], neg);
      else { bucket[b] = neg ? ge_neg(pts[i]) : pts[i]; live[b] = 1; }
    }
    ge run = ge_identity(), sum = ge_identity();
    bool run_live = false, sum_live = false;
    for (int b = (int)half - 1; b >= 0; b--) {
      if (live[b]) { run = run_live ? ge_add(run, bucket[b]) : bucket[b]; run_live = true; }
      if (run_live) { sum = sum_live ? ge_add(sum, run) : run; sum_live = true; }
    }
    if (sum_live) { acc = acc_live ? ge_add(acc, sum) : sum; acc_live = true; }
  }
  return acc;
}

}  // namespace b2

extern "C" {

// One MSM, `threads` host threads, each over a contiguous range of the terms (point-range split, partial sums added at
// the end). points: k x 128 B extended coordinates X|Y|Z|T, canonical little-endian; out: the same format.
int b2_msm_ext(size_t k, const uint8_t* scalars, const uint8_t* points, int threads, uint8_t* out) {
  using namespace b2;
  if (threads < 1) threads = 1;
  if ((size_t)threads > k / 64 + 1) threads = (int)(k / 64 + 1);
  std::vector<ge> pts(k);
  std::vector<ge_cached> cached(k);
  std::vector<ge> part(threads);
  auto work = [&](int t) {
    const size_t lo = k * t / threads, hi = k * (t + 1) / threads;
    for (size_t i = lo; i < hi; i++) { pts[i] = ge_from_bytes(points + 128 * i); cached[i] = ge_to_cached(pts[i]); }
    part[t] = pippenger(hi - lo, scalars + 32 * lo, pts.data() + lo, cached.data() + lo);
  };
  if (threads == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  ge acc = part[0];
  for (int t = 1; t < threads; t++) acc = ge_add(acc, part[t]);
  ge_to_bytes(out, acc);
  return 0;
}


// Timing aid for bench.py's cpu_baseline leg: an MSM over n points P_i = P_0 + i * Q (built here by repeated addition from
// the caller's two extended points, so that no decoding or conversion is inside the timed region) with xorshift scalars
// below 2^252. Returns the seconds spent in the MSM alone (conversion to the cached form included, as a CPU library would
// pay it per call); out = the result, so the work cannot be optimised away.
double b2_bench_msm(size_t n, const uint8_t* p0, const uint8_t* q, int threads, uint8_t* out) {
  using namespace b2;
  std::vector<uint8_t> pts(128 * n), sc(32 * n);
  ge cur = ge_from_bytes(p0);
  const ge_cached step = ge_to_cached(ge_from_bytes(q));
  uint64_t x = 0x9E3779B97F4A7C15ULL;
  for (size_t i = 0; i < n; i++) {
    ge_to_bytes(&pts[128 * i], cur);
    cur = ge_add_cached(cur, step, false);
    for (int w = 0; w < 4; w++) {
      x ^= x << 13; x ^= x >> 7; x ^= x << 17;
      memcpy(&sc[32 * i + 8 * w], &x, 8);
    }
    sc[32 * i + 31] &= 0x0F;
  }
  const auto t0 = std::chrono::steady_clock::now();
  b2_msm_ext(n, sc.data(), pts.data(), threads, out);
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // extern "C"
