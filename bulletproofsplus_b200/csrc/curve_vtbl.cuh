// One table of host entry points per curve. Each curve is its own translation unit (curve_*.cu instantiates
// Api<C, M> once), so the three build in parallel and bpp.cu — the extern "C" surface — only dispatches.
#pragma once
#include "api_impl.cuh"
#include "debug_ops.cuh"

namespace bpp {

struct CurveVtbl {
  int curve, point_bytes, raw_bytes, field_bytes;
  int (*pk_new)(bpp_ctx*, size_t, bpp_pk**);
  int (*pk_new_hashed)(bpp_ctx*, size_t, const uint8_t*, size_t, bpp_pk**);
  int (*pk_precompute)(bpp_ctx*, bpp_pk*, size_t, int*);
  int (*pk_msm)(bpp_ctx*, const bpp_pk*, const uint8_t*, size_t, uint8_t*);
  int (*pk_export)(bpp_ctx*, const bpp_pk*, uint8_t*, uint8_t*, uint8_t*, uint8_t*);
  int (*commit_batch)(bpp_ctx*, const bpp_pk*, const uint64_t*, const uint8_t*, size_t, uint8_t*);
  int (*commit_batch_u64)(bpp_ctx*, const bpp_pk*, const uint64_t*, const uint8_t*, size_t, uint8_t*);
  int (*msm)(bpp_ctx*, const uint8_t*, const uint8_t*, size_t, uint8_t*);
  int (*points_upload)(bpp_ctx*, const uint8_t*, size_t, bool, bpp_points**);
  int (*points_precompute)(bpp_ctx*, bpp_points*);
  int (*points_precompute_direct)(bpp_ctx*, bpp_points*, size_t, int*);
  int (*msm_prepared)(bpp_ctx*, const bpp_points*, const uint8_t*, size_t, size_t, uint8_t*, bool);
  int (*msm_prepared_dev)(bpp_ctx*, const bpp_points*, const void*, size_t, size_t, void*);
  int (*msm_prepared_raw_dev)(bpp_ctx*, const bpp_points*, const void*, size_t, void*);
  int (*points_sum_raw)(bpp_ctx*, const uint8_t*, size_t, uint8_t*);
  int (*points_sum_raw_dev)(bpp_ctx*, const void*, size_t, void*);
  int (*range_prove_batch)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const uint64_t*, const uint8_t*, const bpp_params*, uint8_t*);
  int (*range_verify_batch)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const uint8_t*, const uint8_t*, const bpp_params*, uint8_t*);
  int (*range_prove_batch_dev)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const void*, const void*, const bpp_params*, void*);
  int (*range_verify_batch_dev)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const void*, const void*, const bpp_params*, void*);
  int (*range_verify_batch_fs)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const uint8_t*, const uint8_t*, const uint8_t*, uint8_t*);
  int (*range_verify_batch_fs_dev)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const void*, const void*, const void*, void*);
  int (*range_verify_batch_rlc)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const uint8_t*, const uint8_t*, const bpp_params*, const uint8_t*,
                                const uint8_t*, uint8_t*, int*);
  int (*range_verify_batch_rlc_dev)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const void*, const void*, const bpp_params*, const void*,
                                    const void*, void*, int*);
  int (*prove_begin)(bpp_ctx*, const bpp_pk*, size_t, size_t, size_t, const uint64_t*, const uint8_t*, const uint8_t*, bpp_prove_session**, uint8_t*);
  int (*prove_step)(bpp_ctx*, bpp_prove_session*, const uint8_t*, uint8_t*);
  int (*prove_finish)(bpp_ctx*, bpp_prove_session*, const uint8_t*, uint8_t*);
  int (*wip_prove)(bpp_ctx*, const bpp_pk*, const uint8_t*, const uint8_t*, const uint8_t*, const uint8_t*, const bpp_params*, uint8_t*, size_t*);
  int (*wip_verify)(bpp_ctx*, const bpp_pk*, const uint8_t*, const uint8_t*, const uint8_t*, const uint8_t*, size_t, const uint8_t*,
                    const uint8_t*, const uint8_t*, const uint8_t*, size_t, const bpp_params*);
  int (*debug_op)(bpp_ctx*, int, size_t, const uint8_t*, const uint8_t*, uint8_t*, uint8_t*);
};

// test surface: element-wise device arithmetic (see debug_ops.cuh)
template <class C, class M>
int debug_op_host(bpp_ctx* ctx, int op, size_t n, const uint8_t* a, const uint8_t* b, uint8_t* out, uint8_t* ok) {
  const size_t eb = op < 10 ? (size_t)C::FIELD_BYTES : op < 20 ? (size_t)C::POINT_BYTES : 32;
  BPP_CK(ctx->buf[B_IN0].ensure(n * eb));
  BPP_CK(ctx->buf[B_IN1].ensure(n * eb));
  BPP_CK(ctx->buf[B_OUT0].ensure(n * eb));
  BPP_CK(ctx->buf[B_DYN_OK].ensure(n));
  BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, a, n * eb, cudaMemcpyHostToDevice, ctx->stream));
  BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, b, n * eb, cudaMemcpyHostToDevice, ctx->stream));
  auto k = debug_op_kernel<C, M>;
  BPP_LAUNCH(k, ceil_div(n, 64), 64, 0, op, (uint32_t)n, ctx->buf[B_IN0].as<uint8_t>(), ctx->buf[B_IN1].as<uint8_t>(),
             ctx->buf[B_OUT0].as<uint8_t>(), ctx->buf[B_DYN_OK].as<uint8_t>());
  BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_OUT0].p, n * eb, cudaMemcpyDeviceToHost, ctx->stream));
  BPP_CK(cudaMemcpyAsync(ok, ctx->buf[B_DYN_OK].p, n, cudaMemcpyDeviceToHost, ctx->stream));
  BPP_CK(cudaStreamSynchronize(ctx->stream));
  return BPP_OK;
}

// ORD::le() = the group order, 32 bytes little-endian (host side canonicality checks of bpp_params).
// The table is filled in three parts so that every curve compiles as three translation units in parallel (the prover, the
// verifier and the key / MSM entry points instantiate disjoint sets of protocol kernels; the build is dominated by the 12-limb
// BLS12-381 instantiation otherwise).
template <class C, class M, class ORD>
void fill_vtbl_core(CurveVtbl& v) {
  typedef Api<C, M> A;
  v.curve = C::CURVE_ID; v.point_bytes = C::POINT_BYTES; v.raw_bytes = C::EXT_WORDS * 4; v.field_bytes = C::FIELD_BYTES;
  v.pk_new = A::pk_new;
  v.pk_new_hashed = A::pk_new_hashed;
  v.pk_precompute = A::pk_precompute;
  v.pk_msm = A::pk_msm;
  v.pk_export = A::pk_export;
  v.commit_batch = A::commit_batch_i32;
  v.commit_batch_u64 = A::commit_batch_u64;
  v.msm = A::msm;
  v.points_upload = A::points_upload;
  v.points_precompute = A::points_precompute;
  v.points_precompute_direct = A::points_precompute_direct;
  v.msm_prepared = A::msm_prepared;
  v.msm_prepared_dev = A::msm_prepared_dev;
  v.msm_prepared_raw_dev = A::msm_prepared_raw_dev;
  v.points_sum_raw = A::points_sum_raw;
  v.points_sum_raw_dev = A::points_sum_raw_dev;
  v.debug_op = debug_op_host<C, M>;
}
template <class C, class M, class ORD>
void fill_vtbl_verify(CurveVtbl& v) {
  typedef Api<C, M> A;
  v.range_verify_batch = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const uint8_t* V, const uint8_t* pr,
                            const bpp_params* p, uint8_t* o) { return A::range_verify_batch(c, pk, n, m, cnt, V, pr, p, o, ORD::le()); };
  v.range_verify_batch_dev = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const void* V, const void* pr,
                                const bpp_params* p, void* o) { return A::range_verify_batch_dev(c, pk, n, m, cnt, V, pr, p, o, ORD::le()); };
  v.range_verify_batch_fs = A::range_verify_batch_fs;
  v.range_verify_batch_fs_dev = A::range_verify_batch_fs_dev;
  v.range_verify_batch_rlc = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const uint8_t* V, const uint8_t* pr, const bpp_params* p,
                                const uint8_t* ch, const uint8_t* w, uint8_t* o, int* fb) {
    return A::range_verify_batch_rlc(c, pk, n, m, cnt, V, pr, p, ch, w, o, ORD::le(), fb);
  };
  v.range_verify_batch_rlc_dev = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const void* V, const void* pr, const bpp_params* p,
                                    const void* ch, const void* w, void* o, int* fb) {
    return A::range_verify_batch_rlc_dev(c, pk, n, m, cnt, V, pr, p, ch, w, o, ORD::le(), fb);
  };
  v.wip_verify = [](bpp_ctx* c, const bpp_pk* pk, const uint8_t* y, const uint8_t* Gc, const uint8_t* Hc, const uint8_t* gc, size_t nV,
                    const uint8_t* Vc, const uint8_t* Ap, const uint8_t* V, const uint8_t* pr, size_t l, const bpp_params* p) {
    return A::wip_verify(c, pk, y, Gc, Hc, gc, nV, Vc, Ap, V, pr, l, p, ORD::le());
  };
}
template <class C, class M, class ORD>
void fill_vtbl_prove(CurveVtbl& v) {
  typedef Api<C, M> A;
  v.range_prove_batch = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const uint64_t* vv, const uint8_t* g,
                           const bpp_params* p, uint8_t* o) { return A::range_prove_batch(c, pk, n, m, cnt, vv, g, p, o, ORD::le()); };
  v.range_prove_batch_dev = [](bpp_ctx* c, const bpp_pk* pk, size_t n, size_t m, size_t cnt, const void* vv, const void* g,
                               const bpp_params* p, void* o) { return A::range_prove_batch_dev(c, pk, n, m, cnt, vv, g, p, o, ORD::le()); };
  v.prove_begin = A::prove_begin;
  v.prove_step = A::prove_step;
  v.prove_finish = A::prove_finish;
  v.wip_prove = [](bpp_ctx* c, const bpp_pk* pk, const uint8_t* a, const uint8_t* b, const uint8_t* y, const uint8_t* g,
                   const bpp_params* p, uint8_t* o, size_t* l) { return A::wip_prove(c, pk, a, b, y, g, p, o, l, ORD::le()); };
}

// A curve's translation unit is compiled three times, with BPP_VTBL_PART = 0, 1, 2 (Makefile): part 0 defines the accessor
// `ACCESSOR()` and fills the key / MSM entries, parts 1 and 2 define the two helpers it calls.
#ifndef BPP_VTBL_PART
#define BPP_VTBL_PART 0
#endif
#if BPP_VTBL_PART == 0
#define BPP_CURVE_TU(ACCESSOR, TAG, C, M, ORD)                 \
  void fill_verify_##TAG(CurveVtbl& v);                        \
  void fill_prove_##TAG(CurveVtbl& v);                         \
  const CurveVtbl* ACCESSOR() {                                \
    static const CurveVtbl v = [] {                            \
      CurveVtbl t{};                                           \
      fill_vtbl_core<C, M, ORD>(t);                            \
      fill_verify_##TAG(t);                                    \
      fill_prove_##TAG(t);                                     \
      return t;                                                \
    }();                                                       \
    return &v;                                                 \
  }
#elif BPP_VTBL_PART == 1
#define BPP_CURVE_TU(ACCESSOR, TAG, C, M, ORD) \
  void fill_verify_##TAG(CurveVtbl& v) { fill_vtbl_verify<C, M, ORD>(v); }
#else
#define BPP_CURVE_TU(ACCESSOR, TAG, C, M, ORD) \
  void fill_prove_##TAG(CurveVtbl& v) { fill_vtbl_prove<C, M, ORD>(v); }
#endif

}  // namespace bpp
