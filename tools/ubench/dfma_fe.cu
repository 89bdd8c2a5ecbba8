// tools/ubench/dfma_fe.cu — can the FP64 pipe of sm_100a multiply in GF(2^255-19) faster than the integer pipe?
//
// The library's field multiply is 72 IMAD.WIDE per product and sits at the ceiling of the integer-multiply pipe
// (~330 issue cycles per warp-multiply, 112 G multiplies/s per chip). B200 — unlike B300 — keeps a
// full-rate FP64 pipe (64 DFMA lanes/clk/SM), so this probe measures the classic double-precision alternative:
//   * 5 limbs of 51 bits held in doubles, balanced (|limb| <= 2^50 + small), lazy sums up to 2^51.5 allowed;
//   * each 51x51-bit partial product split exactly by two DFMAs and one DADD:
//         hi = fma_rz(a, b, 1.5*2^104)            = 1.5*2^104 + floor(ab / 2^52) * 2^52
//         lo = fma_rz(a, b, (1.5*2^104 + 2^52) - hi) = 2^52 + (ab mod 2^52)
//     whose BIT PATTERNS are integers plus a constant, so the 25 products are summed per column with 64-bit integer adds;
//   * 2^255 = 19 fold in int64, one parallel (carry-save) balanced carry, conversion back by the 1.5*2^52 magic constant.
// Prints: raw DFMA / DADD / 64-bit-add issue rates (loop-produced operands), the multiply's rate by ILP and occupancy,
// and checks the result of a 1000-step recurrence against a 128-bit host implementation.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o /tmp/dfma_fe tools/ubench/dfma_fe.cu && /tmp/dfma_fe
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "../../bulletproofsplus_b200/csrc/fe25519.cuh"  // the library's 8 x 32-bit IMAD.WIDE multiply, for the mixed-pipe run

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

struct F5 { double v[5]; };

__device__ __forceinline__ void f5_mul(F5& r, const F5& a, const F5& b) {
  const double C1 = 0x1.8p104;               // 1.5 * 2^104: ulp 2^52 over [2^104, 2^105)
  const double C2 = 0x1.8p104 + 0x1p52;      // C1 + 2^52 (exact)
  long long lo[9], hi[9];
#pragma unroll
  for (int k = 0; k < 9; k++) { lo[k] = 0; hi[k] = 0; }
#pragma unroll
  for (int i = 0; i < 5; i++) {
#pragma unroll
    for (int j = 0; j < 5; j++) {
      const double h = __fma_rz(a.v[i], b.v[j], C1);
      const double d = __dsub_rn(C2, h);
      const double l = __fma_rz(a.v[i], b.v[j], d);
      hi[i + j] += __double_as_longlong(h);
      lo[i + j] += __double_as_longlong(l);
    }
  }
  // remove the exponent patterns: n_k products per column
  long long c[10];
#pragma unroll
  for (int k = 0; k < 10; k++) {
    long long x = 0;
    if (k < 9) { const int n = (k < 4 ? k : 8 - k) + 1; x = lo[k] - (long long)n * 0x4330000000000000ll; }
    if (k > 0) { const int n = (k - 1 < 4 ? k - 1 : 9 - k) + 1; x += 2 * (hi[k - 1] - (long long)n * 0x4678000000000000ll); }
    c[k] = x;
  }
  // 2^255 = 19
#pragma unroll
  for (int k = 0; k < 5; k++) c[k] += 19 * c[k + 5];
  // parallel balanced carry: t = c + 2^50; carry = t >> 51; rem = (t & (2^51 - 1)) - 2^50
  long long cr[5], rem[5];
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const long long t = c[k] + (1ll << 50);
    cr[k] = t >> 51;
    rem[k] = t & ((1ll << 51) - 1);
  }
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const long long in = k ? cr[k - 1] : 19 * cr[4];
    const long long bits = rem[k] + in + (0x4338000000000000ll - (1ll << 50));
    r.v[k] = __dsub_rn(__longlong_as_double(bits), 0x1.8p52);
  }
}

// ---- raw issue rates ----------------------------------------------------------------------------------------------
template <int OP>
__global__ void rate_kernel(int iters, double* out, double seed) {
  double x0 = seed + threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  long long y0 = threadIdx.x, y1 = y0 + 1, y2 = y0 + 2, y3 = y0 + 3, y4 = y0 + 4, y5 = y0 + 5, y6 = y0 + 6, y7 = y0 + 7;
  for (int i = 0; i < iters; i++) {
    if (OP == 0) {  // DFMA
      x0 = __fma_rz(x0, x1, x2); x1 = __fma_rz(x1, x2, x3); x2 = __fma_rz(x2, x3, x4); x3 = __fma_rz(x3, x4, x5);
      x4 = __fma_rz(x4, x5, x6); x5 = __fma_rz(x5, x6, x7); x6 = __fma_rz(x6, x7, x0); x7 = __fma_rz(x7, x0, x1);
    } else if (OP == 1) {  // DADD
      x0 = __dadd_rn(x0, x1); x1 = __dadd_rn(x1, x2); x2 = __dadd_rn(x2, x3); x3 = __dadd_rn(x3, x4);
      x4 = __dadd_rn(x4, x5); x5 = __dadd_rn(x5, x6); x6 = __dadd_rn(x6, x7); x7 = __dadd_rn(x7, x0);
    } else {  // 64-bit integer add
      y0 += y1; y1 += y2; y2 += y3; y3 += y4; y4 += y5; y5 += y6; y6 += y7; y7 += y0;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + (double)(y0 ^ y1 ^ y2 ^ y3 ^ y4 ^ y5 ^ y6 ^ y7);
}

// ---- the multiply, ILP independent chains per thread ----------------------------------------------------------------
template <int ILP>
__global__ void mul_kernel(int iters, const F5* __restrict__ in, F5* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  F5 x[ILP], y[ILP];
#pragma unroll
  for (int q = 0; q < ILP; q++) { x[q] = in[2 * q]; y[q] = in[2 * q + 1]; x[q].v[0] += (double)(t & 7); }
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int q = 0; q < ILP; q++) {
      F5 z;
      f5_mul(z, x[q], y[q]);
      y[q] = x[q];
      x[q] = z;
    }
  }
#pragma unroll
  for (int q = 0; q < ILP; q++) out[(size_t)t * ILP + q] = x[q];
}

// ---- both multiplies side by side: warps whose (index mod 4) bit is set in `pattern` run the FP64 multiply, the others the
// integer one. The two forms load different pipes (FP64 + ALU vs FMA-heavy), so a mix can beat either alone.
// by_block = 1: the role is chosen per BLOCK (index mod 4): a block's four warps sit on the four sub-partitions of an SM, so
// every sub-partition sees both kinds of warp. by_block = 0: per warp index within the block, which pins each role to
// fixed sub-partitions (no mixing inside a sub-partition).
__global__ void mix_kernel(int iters, uint32_t pattern, int by_block, const F5* __restrict__ in, F5* __restrict__ out, uint32_t* __restrict__ out_i) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const bool f64 = (pattern >> ((by_block ? blockIdx.x : (threadIdx.x >> 5)) & 3)) & 1u;
  if (f64) {
    F5 x = in[0], y = in[1];
    x.v[0] += (double)(t & 7);
    for (int i = 0; i < iters; i++) {
      F5 z;
      f5_mul(z, x, y);
      y = x;
      x = z;
    }
    out[t] = x;
  } else {
    bpp::Fe25519 x, y, z;
#pragma unroll
    for (int k = 0; k < 8; k++) { x.v[k] = 0x9e3779b9u * (k + 1) + t; y.v[k] = 0x85ebca6bu * (k + 3) ^ t; }
    for (int i = 0; i < iters; i++) {
      bpp::fe_mul(z, x, y);
      y = x;
      x = z;
    }
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) acc ^= x.v[k];
    out_i[t] = acc;
  }
}

// ---- host reference: 5 x 51-bit limbs, unsigned __int128 -------------------------------------------------------------
typedef unsigned __int128 u128;
static const uint64_t M51 = (1ull << 51) - 1;
static void h_mul(uint64_t h[5], const uint64_t f[5], const uint64_t g[5]) {
  u128 t[5];
  for (int k = 0; k < 5; k++) {
    t[k] = 0;
    for (int i = 0; i < 5; i++) {
      int j = k - i;
      if (j >= 0) t[k] += (u128)f[i] * g[j];
      else t[k] += (u128)f[i] * (19 * g[j + 5]);
    }
  }
  uint64_t c = 0;
  for (int k = 0; k < 5; k++) { t[k] += c; h[k] = (uint64_t)t[k] & M51; c = (uint64_t)(t[k] >> 51); }
  h[0] += 19 * c;
  c = h[0] >> 51; h[0] &= M51; h[1] += c;
}
static void h_canon(uint64_t h[5]) {
  for (int rep = 0; rep < 3; rep++) {
    uint64_t c = 0;
    for (int k = 0; k < 5; k++) { h[k] += c; c = h[k] >> 51; h[k] &= M51; }
    h[0] += 19 * c;
  }
  // h < 2^255 now (possibly >= p): subtract p if h >= p
  uint64_t t[5]; uint64_t c = 19;
  for (int k = 0; k < 5; k++) { t[k] = h[k] + c; c = t[k] >> 51; t[k] &= M51; }
  if (c) for (int k = 0; k < 5; k++) h[k] = t[k];
}
// signed double limbs -> canonical
static void from_dev(uint64_t h[5], const F5& d) {
  // add 4p in limb form so that every limb is positive: 4p = (2^53 - 76, 2^53 - 4, ...)
  long long l[5];
  for (int k = 0; k < 5; k++) l[k] = (long long)d.v[k] + (k == 0 ? (1ll << 53) - 76 : (1ll << 53) - 4);
  for (int k = 0; k < 5; k++) h[k] = (uint64_t)l[k];
  h_canon(h);
}

int main() {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
  printf("%s, %d SMs, clock attr %d MHz\n", prop.name, prop.multiProcessorCount, clk_khz / 1000);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int nsm = prop.multiProcessorCount;

  // raw rates
  {
    double* out;
    CK(cudaMalloc(&out, sizeof(double) * nsm * 8 * 256));
    const char* names[3] = {"DFMA (fma.rz.f64)", "DADD", "64-bit integer add"};
    for (int op = 0; op < 3; op++) {
      const int iters = 20000, blocks = nsm * 8, threads = 256;
      for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(e0));
        if (op == 0) rate_kernel<0><<<blocks, threads>>>(iters, out, 1.0000001);
        if (op == 1) rate_kernel<1><<<blocks, threads>>>(iters, out, 1.0000001);
        if (op == 2) rate_kernel<2><<<blocks, threads>>>(iters, out, 1.0000001);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
      }
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double ops = (double)blocks * threads * iters * 8;
      const double per_s = ops / (ms * 1e-3);
      printf("%-22s %8.2f T lane-ops/s  = %6.1f lanes/clk/SM at 1.9 GHz  (%.2f clk per warp instruction per SM sub-partition)\n", names[op],
             per_s / 1e12, per_s / nsm / 1.9e9, 32.0 / (per_s / nsm / 1.9e9 / 4));
    }
    CK(cudaFree(out));
  }

  // correctness + throughput of the multiply
  uint64_t hx[5] = {0x1234567890abcull & M51, 0x7ffffffffffffull & M51, 0x2468ace013579ull & M51, 0x5555555555555ull & M51, 0x3fedcba987654ull & M51};
  uint64_t hy[5] = {0x7ffffffffffecull & M51, 0x0000000000001ull, 0x7edcba9876543ull & M51, 0x1111111111111ull, 0x2aaaaaaaaaaaaull & M51};
  F5 hin[8];
  for (int q = 0; q < 4; q++)
    for (int k = 0; k < 5; k++) {
      // balanced start values: subtract 2^50 from every limb of x (value changes, fine: the host mirrors it below)
      hin[2 * q].v[k] = (double)(long long)(hx[k] >> 1) - (double)(q * 3);
      hin[2 * q + 1].v[k] = (double)(long long)(hy[k] >> 1) + (double)q;
    }
  F5 *din, *dout;
  const int max_threads = nsm * 2048;
  CK(cudaMalloc(&din, sizeof(hin)));
  CK(cudaMalloc(&dout, sizeof(F5) * (size_t)max_threads * 4));
  CK(cudaMemcpy(din, hin, sizeof(hin), cudaMemcpyHostToDevice));
  {
    const int iters = 1000;
    mul_kernel<1><<<1, 32>>>(iters, din, dout);
    CK(cudaDeviceSynchronize());
    F5 res[32];
    CK(cudaMemcpy(res, dout, sizeof(res), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int t = 0; t < 32; t++) {
      // host: same recurrence from the same start (thread t adds (t & 7) to limb 0 of x)
      uint64_t x[5], y[5], z[5];
      F5 sx = hin[0], sy = hin[1];
      sx.v[0] += (double)(t & 7);
      from_dev(x, sx); from_dev(y, sy);
      for (int i = 0; i < iters; i++) { h_mul(z, x, y); memcpy(y, x, sizeof(x)); memcpy(x, z, sizeof(z)); }
      h_canon(x);
      uint64_t g[5];
      from_dev(g, res[t]);
      if (memcmp(g, x, sizeof(x)) != 0) bad++;
      for (int k = 0; k < 5; k++)
        if (res[t].v[k] > 0x1p50 + 0x1p14 || res[t].v[k] < -0x1p50 - 0x1p14) bad += 100;
    }
    printf("correctness: 32 threads x %d dependent multiplies vs 128-bit host arithmetic: %s\n", iters, bad ? "MISMATCH" : "identical");
    if (bad) return 1;
  }
  for (int threads_per_sm : {256, 512, 1024, 2048}) {
    for (int ilp : {1, 2, 4}) {
      const int iters = 2000, threads = 128, blocks = nsm * threads_per_sm / threads;
      for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(e0));
        if (ilp == 1) mul_kernel<1><<<blocks, threads>>>(iters, din, dout);
        if (ilp == 2) mul_kernel<2><<<blocks, threads>>>(iters, din, dout);
        if (ilp == 4) mul_kernel<4><<<blocks, threads>>>(iters, din, dout);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
      }
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double muls = (double)blocks * threads * iters * ilp;
      printf("f5_mul (FP64 pipe): %4d threads/SM, ILP %d: %7.1f G multiplies/s  (%.0f clk per warp-multiply per sub-partition at 1.9 GHz)\n",
             threads_per_sm, ilp, muls / (ms * 1e-3) / 1e9, 1.9e9 * nsm * 4 * 32 / (muls / (ms * 1e-3)));
    }
  }
  {
    uint32_t* out_i;
    CK(cudaMalloc(&out_i, sizeof(uint32_t) * (size_t)max_threads));
    const struct { uint32_t pat; const char* name; } pats[] = {{0x0, "all warps integer multiply (8x32 IMAD.WIDE)"}, {0xf, "all warps FP64 multiply (5x51 DFMA)"},
      {0x5, "1/2 FP64 + 1/2 integer"}, {0x1, "1/4 FP64 + 3/4 integer"}, {0x7, "3/4 FP64 + 1/4 integer"}};
    for (int by_block = 1; by_block >= 0; by_block--)
    for (int threads_per_sm : {512, 1024, 2048}) {
      for (auto& pt : pats) {
        if (!by_block && (pt.pat == 0 || pt.pat == 0xf)) continue;
        const int iters = 2000, threads = 128, blocks = nsm * threads_per_sm / threads;
        for (int rep = 0; rep < 2; rep++) {
          CK(cudaEventRecord(e0));
          mix_kernel<<<blocks, threads>>>(iters, pt.pat, by_block, din, dout, out_i);
          CK(cudaEventRecord(e1));
          CK(cudaEventSynchronize(e1));
          CK(cudaGetLastError());
        }
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double muls = (double)blocks * threads * iters;
        printf("mix (%s): %4d threads/SM, %-46s %7.1f G multiplies/s\n", by_block ? "role per block" : "role per warp ", threads_per_sm, pt.name, muls / (ms * 1e-3) / 1e9);
      }
    }
  }
  printf("reference: the library's 8x32-bit IMAD.WIDE multiply runs at 108-113 G multiplies/s (profiles/r01c_fe_probe.txt)\n");
  return 0;
}
