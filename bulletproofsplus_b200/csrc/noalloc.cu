// bpp_points_sum_raw_noalloc_dev — the one entry point of the library that allocates nothing (include/bpp.h).
//
// It has no counterpart in the reference; it was added for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). It is a variant of
// bpp_points_sum_raw_dev — the epilogue of a point-range-split MSM: sum `count` raw partial points and encode the total —
// whose intermediate results live in a file-scope __device__ array instead of the context's grow-only cudaMalloc buffers.
//
// The array below is held by every GPU that loads the library until the process exits: BPP_NOALLOC_SLOTS raw points of
// at most 192 B (BLS12-381 G1, XYZZ over 12 limbs) = 12 480 B. Every call shares it; it keeps the previous call's contents.
// Nothing in it has to start at zero: phase 1 writes every slot phase 2 reads (idle blocks write the identity), so no
// clearing is needed. Two calls that overlap on the GPU (different streams / contexts on one device) RACE on it.
//
// In host code the NAME of the array is a host-side stand-in, not a device address: handing it to a kernel faults (or,
// where the GPU can reach host memory, silently uses the wrong memory). Its address is resolved once per device, with
// that device current, by cudaGetSymbolAddress, and that pointer is what the launches below receive.
#include <mutex>
#include "curve_params.cuh"
#include "ed25519.cuh"
#include "engine.cuh"
#include "proto_misc.cuh"

namespace bpp {

constexpr uint32_t NOALLOC_MAX_WORDS = 48;                    // words of the largest raw point (BLS12-381 G1)
constexpr uint32_t NOALLOC_PER_BLOCK = 1024;                  // partial points summed by one block of phase 1
constexpr uint32_t NOALLOC_MAX_BLOCKS = BPP_NOALLOC_MAX_POINTS / NOALLOC_PER_BLOCK;  // 64
constexpr uint32_t NOALLOC_SLOTS = NOALLOC_MAX_BLOCKS + 1;    // + the total

static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_SLOTS * NOALLOC_MAX_WORDS];

// phase 1: block b sums points [b * PER_BLOCK, (b + 1) * PER_BLOCK) into slot b (the identity if its range is empty)
template <class C>
__global__ void __launch_bounds__(128) noalloc_partial_sum_kernel(uint32_t count, const uint32_t* __restrict__ pts, uint32_t* __restrict__ scratch) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  typename C::Ext acc, p;
  C::set_identity(acc);
  const uint32_t lo = blockIdx.x * NOALLOC_PER_BLOCK;
  const uint32_t hi = lo + NOALLOC_PER_BLOCK < count ? lo + NOALLOC_PER_BLOCK : count;
  for (uint32_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    load_words<C::EXT_WORDS>(p, pts + (size_t)i * C::EXT_WORDS);
    C::add(acc, p);
  }
  uint32_t* mine = sh + threadIdx.x * C::EXT_WORDS;
  *reinterpret_cast<typename C::Ext*>(mine) = acc;
  __syncthreads();
  for (uint32_t s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (threadIdx.x + s) * C::EXT_WORDS);
      C::add(acc, o);
      *reinterpret_cast<typename C::Ext*>(mine) = acc;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) store_words<C::EXT_WORDS>(scratch + (size_t)blockIdx.x * C::EXT_WORDS, acc);
}

// device address of the scratch array, per device (resolved on first use with that device current; no allocation)
static int noalloc_scratch_ptr(bpp_ctx* ctx, uint32_t** out) {
  static std::mutex mu;
  static uint32_t* ptr[64] = {nullptr};
  if (ctx->device < 0 || ctx->device >= 64) { ctx->err = "noalloc: device index out of range"; return BPP_ERR_UNSUPPORTED; }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[ctx->device]) {
    void* p = nullptr;
    BPP_CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    ptr[ctx->device] = reinterpret_cast<uint32_t*>(p);
  }
  *out = ptr[ctx->device];
  return BPP_OK;
}

template <class C>
static int points_sum_raw_noalloc(bpp_ctx* ctx, const void* d_raw, size_t count, void* d_out) {
  static_assert(C::EXT_WORDS <= (int)NOALLOC_MAX_WORDS, "raw point larger than a scratch slot");
  uint32_t* scratch = nullptr;
  BPP_TRY(noalloc_scratch_ptr(ctx, &scratch));
  const uint32_t n = (uint32_t)count, blocks = ceil_div(n, NOALLOC_PER_BLOCK);
  uint32_t* total = scratch + (size_t)NOALLOC_MAX_BLOCKS * C::EXT_WORDS;
  // plain launches (BPP_LAUNCH would create events when the launch tracer is on)
  noalloc_partial_sum_kernel<C><<<blocks, 128, 0, ctx->stream>>>(n, reinterpret_cast<const uint32_t*>(d_raw), scratch);
  ext_sum_kernel<C><<<1, 128, 0, ctx->stream>>>(blocks, scratch, total);
  points_encode_kernel<C><<<1, 32, 0, ctx->stream>>>(1u, total, reinterpret_cast<uint8_t*>(d_out));
  ctx->launches += 3;
  BPP_CK(cudaGetLastError());
  return BPP_OK;
}

}  // namespace bpp

extern "C" int bpp_points_sum_raw_noalloc_dev(bpp_ctx* ctx, int curve, const void* d_raw_points, size_t count, void* d_out_point) {
  using namespace bpp;
  if (!ctx) return BPP_ERR_NOT_INIT;
  BPP_CK(cudaSetDevice(ctx->device));
  if (!d_raw_points || !d_out_point || count == 0) { ctx->err = "points_sum_raw_noalloc: null argument or count == 0"; return BPP_ERR_LENGTH; }
  if (count > BPP_NOALLOC_MAX_POINTS) {
    ctx->err = "points_sum_raw_noalloc: count exceeds BPP_NOALLOC_MAX_POINTS (65536): the static device scratch holds 64 block partials; "
               "use bpp_points_sum_raw_dev for larger inputs";
    return BPP_ERR_LENGTH;
  }
  switch (curve) {
    case BPP_RISTRETTO255: return points_sum_raw_noalloc<Ed25519>(ctx, d_raw_points, count, d_out_point);
    case BPP_SECP256K1: return points_sum_raw_noalloc<Secp256k1>(ctx, d_raw_points, count, d_out_point);
    case BPP_BLS12_381_G1: return points_sum_raw_noalloc<Bls12381G1>(ctx, d_raw_points, count, d_out_point);
    default: ctx->err = "unknown curve id"; return BPP_ERR_UNSUPPORTED;
  }
}
