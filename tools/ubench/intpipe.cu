// Development micro-benchmark (not part of libbpp_b200.so): which integer instruction mixes the sm_100a SM sustains, and what
// that means for a GF(2^255-19) multiply. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o intpipe intpipe.cu
// Every kernel is pure register arithmetic: 148 x 8 blocks of 256 threads, ILP-8 independent chains.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define DEV __device__ __forceinline__

// ---------------------------------------------------------------------------------------------------------------
// pipe probes
// ---------------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) pipe_kernel(uint32_t iters, uint32_t seed, uint64_t* __restrict__ sink) {
  uint32_t a = seed + threadIdx.x * 2654435761u + 1u, b = a ^ 0x9e3779b9u;
  uint64_t acc[8];
  uint32_t lo[8], hi[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { acc[i] = a + i; lo[i] = b + i; hi[i] = a * (i + 3); }
  for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        // the multiplicand is the accumulator's own low word: ptxas cannot hoist the product out of the loop
        if (MODE == 0 || MODE == 2 || MODE == 7 || MODE == 8)
          acc[i] += (uint64_t)(uint32_t)acc[i] * b;
        if (MODE == 1 || MODE == 2) asm volatile("add.u32 %0, %0, %1;" : "+r"(lo[i]) : "r"(hi[i]));
        if (MODE == 8) {  // two ALU ops per wide multiply-add
          asm volatile("add.u32 %0, %0, %1;" : "+r"(lo[i]) : "r"(hi[i]));
          asm volatile("add.u32 %0, %0, %1;" : "+r"(hi[i]) : "r"(lo[i]));
        }
        if (MODE == 11) asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(acc[i]) : "r"((uint32_t)acc[i]), "r"(b));
        if (MODE == 12) asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.u32 %1, %2, %3, %1;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(lo[i]), "r"(b));
        if (MODE == 13) asm volatile("lop3.b32 %0, %0, %1, %2, 0x6a;" : "+r"(lo[i]) : "r"(hi[i]), "r"(a));
        if (MODE == 14) {  // wide multiply-add + 1 logic op
          acc[i] += (uint64_t)(uint32_t)acc[i] * b;
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x6a;" : "+r"(lo[i]) : "r"(hi[i]), "r"(a));
        }
        if (MODE == 15) {  // 64-bit shift right + mask + 64-bit add: one step of a radix-2^26 carry chain
          uint64_t c = acc[i] >> 26;
          acc[i] &= 0x3ffffffull;
          acc[(i + 1) & 7] += c + b;
        }
        if (MODE == 3) {  // unfused: product then a 64-bit add with carry
          uint32_t pl, ph;
          asm volatile("mul.lo.u32 %0, %2, %3; mul.hi.u32 %1, %2, %3;" : "=r"(pl), "=r"(ph) : "r"(a), "r"(lo[i]));
          asm volatile("add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(pl), "r"(ph));
        }
        if (MODE == 4) {  // carry-chained wide multiply-add (IMAD.WIDE.U32.X form): row of 4 per chain pair
          if ((i & 3) == 0) asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(a), "r"(b));
          else asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(a), "r"(b));
        }
        if (MODE == 5 || MODE == 7) asm volatile("xor.b32 %0, %0, %1;" : "+r"(lo[i]) : "r"(hi[i] + i));
        if (MODE == 6) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(lo[i]) : "r"(hi[i]));
        if (MODE == 9) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo[i]) : "r"(a), "r"(b));
        if (MODE == 10) {  // add with carry chains only (IADD3 / IADD3.X)
          if ((i & 3) == 0) asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(lo[i]) : "r"(a));
          else asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(lo[i]) : "r"(a));
        }
      }
      if (MODE == 4 || MODE == 10) asm volatile("addc.u32 %0, %0, 0;" : "+r"(hi[7]));
    }
  }
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i] + lo[i] + hi[i];
  if (s == 0x1234567ull) sink[0] = s;
}

// ---------------------------------------------------------------------------------------------------------------
// field multiply candidates. Each: struct with LIMBS, init from seed, mul(r, a, b).
// ---------------------------------------------------------------------------------------------------------------
template <int NL>
DEV void mul_wide_n(uint32_t* r, const uint32_t* a, const uint32_t* b) {
  uint32_t E[2 * NL], O[2 * NL];
#pragma unroll
  for (int i = 0; i < 2 * NL; i++) { E[i] = 0; O[i] = 0; }
#pragma unroll
  for (int i = 0; i < NL; i++) {
    const int je = i & 1, jo = 1 - je;
    {
      const int c = i + je;
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(E[c]), "+r"(E[c + 1]) : "r"(a[je]), "r"(b[i]));
#pragma unroll
      for (int j = je + 2; j < NL; j += 2)
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(E[i + j]), "+r"(E[i + j + 1]) : "r"(a[j]), "r"(b[i]));
      if (c + NL < 2 * NL) asm volatile("addc.u32 %0, %0, 0;" : "+r"(E[c + NL]));
    }
    {
      const int c = i + jo - 1;
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(O[c]), "+r"(O[c + 1]) : "r"(a[jo]), "r"(b[i]));
#pragma unroll
      for (int j = jo + 2; j < NL; j += 2)
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(O[i + j - 1]), "+r"(O[i + j]) : "r"(a[j]), "r"(b[i]));
      if (c + NL < 2 * NL) asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[c + NL]));
    }
  }
  r[0] = E[0];
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r[1]) : "r"(E[1]), "r"(O[0]));
#pragma unroll
  for (int i = 2; i < 2 * NL - 1; i++) asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[i]) : "r"(E[i]), "r"(O[i - 1]));
  asm volatile("addc.u32 %0, %1, %2;" : "=r"(r[2 * NL - 1]) : "r"(E[2 * NL - 1]), "r"(O[2 * NL - 2]));
}

// V0: 8 x 32 saturated limbs, loose mod 2^256 - 38 (the committed multiply)
struct F8 {
  static constexpr int NL = 8;
  static constexpr uint32_t MASK = 0xffffffffu;
  DEV static void reduce_wide(uint32_t* o, const uint32_t* t) {
    uint32_t r[8], q[8], top;
#pragma unroll
    for (int i = 0; i < 8; i++) r[i] = t[i];
    const uint32_t k38 = 38;
    asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[0]), "+r"(r[1]) : "r"(t[8]), "r"(k38));
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[2]), "+r"(r[3]) : "r"(t[10]), "r"(k38));
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[4]), "+r"(r[5]) : "r"(t[12]), "r"(k38));
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[6]), "+r"(r[7]) : "r"(t[14]), "r"(k38));
    asm volatile("addc.u32 %0, 0, 0;" : "=r"(top));
#pragma unroll
    for (int j = 0; j < 4; j++)
      asm volatile("mul.lo.u32 %0, %2, %3; mul.hi.u32 %1, %2, %3;" : "=r"(q[2 * j]), "=r"(q[2 * j + 1]) : "r"(t[9 + 2 * j]), "r"(k38));
    asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(r[1]) : "r"(q[0]));
#pragma unroll
    for (int i = 2; i < 8; i++) asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(q[i - 1]));
    asm volatile("addc.u32 %0, %0, %1;" : "+r"(top) : "r"(q[7]));
    uint32_t f = top * 38u, c2;
    asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(r[0]) : "r"(f));
#pragma unroll
    for (int i = 1; i < 8; i++) asm volatile("addc.cc.u32 %0, %0, 0;" : "+r"(r[i]));
    asm volatile("addc.u32 %0, 0, 0;" : "=r"(c2));
    r[0] += c2 * 38u;
#pragma unroll
    for (int i = 0; i < 8; i++) o[i] = r[i];
  }
  DEV static void mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t t[16];
    mul_wide_n<8>(t, a, b);
    reduce_wide(r, t);
  }
};

constexpr uint32_t M26 = 0x3ffffffu, M25 = 0x1ffffffu;

DEV void carry10(uint32_t* r, uint64_t* h) {
  h[1] += h[0] >> 26; h[0] &= M26;
  h[5] += h[4] >> 26; h[4] &= M26;
  h[2] += h[1] >> 25; h[1] &= M25;
  h[6] += h[5] >> 25; h[5] &= M25;
  h[3] += h[2] >> 26; h[2] &= M26;
  h[7] += h[6] >> 26; h[6] &= M26;
  h[4] += h[3] >> 25; h[3] &= M25;
  h[8] += h[7] >> 25; h[7] &= M25;
  h[5] += h[4] >> 26; h[4] &= M26;
  h[9] += h[8] >> 26; h[8] &= M26;
  h[0] += 19ull * (h[9] >> 25); h[9] &= M25;
  h[1] += h[0] >> 26; h[0] &= M26;
#pragma unroll
  for (int k = 0; k < 10; k++) r[k] = (uint32_t)h[k];
}

// V1: 10 x 25.5, plain C accumulation (row order)
struct F10c {
  static constexpr int NL = 10;
  static constexpr uint32_t MASK = M25;
  DEV static void mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t f2[10], g19[10];
#pragma unroll
    for (int i = 0; i < 10; i++) { f2[i] = a[i] << 1; g19[i] = b[i] * 19u; }
    uint64_t h[10];
#pragma unroll
    for (int k = 0; k < 10; k++) h[k] = 0;
#pragma unroll
    for (int i = 0; i < 10; i++)
#pragma unroll
      for (int j = 0; j < 10; j++) {
        const int k = i + j;
        const uint32_t fi = ((i & 1) && (j & 1)) ? f2[i] : a[i];
        const uint32_t gj = (k >= 10) ? g19[j] : b[j];
        h[k >= 10 ? k - 10 : k] += (uint64_t)fi * gj;
      }
    carry10(r, h);
  }
};

// V2: 10 x 25.5, asm volatile mad.wide chains, column order (each column one dependent chain of 10)
struct F10col {
  static constexpr int NL = 10;
  static constexpr uint32_t MASK = M25;
  DEV static void mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t f2[10], g19[10];
#pragma unroll
    for (int i = 0; i < 10; i++) { f2[i] = a[i] << 1; g19[i] = b[i] * 19u; }
    uint64_t h[10];
#pragma unroll
    for (int k = 0; k < 10; k++) {
      h[k] = 0;
#pragma unroll
      for (int i = 0; i < 10; i++) {
        const int j = (k - i + 10) % 10;
        const uint32_t fi = ((i & 1) && (j & 1)) ? f2[i] : a[i];
        const uint32_t gj = (i + j >= 10) ? g19[j] : b[j];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(h[k]) : "r"(fi), "r"(gj));
      }
    }
    carry10(r, h);
  }
};

// V3: 10 x 25.5, asm volatile mad.wide chains, row order
struct F10row {
  static constexpr int NL = 10;
  static constexpr uint32_t MASK = M25;
  DEV static void mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t f2[10], g19[10];
#pragma unroll
    for (int i = 0; i < 10; i++) { f2[i] = a[i] << 1; g19[i] = b[i] * 19u; }
    uint64_t h[10];
#pragma unroll
    for (int k = 0; k < 10; k++) h[k] = 0;
#pragma unroll
    for (int i = 0; i < 10; i++)
#pragma unroll
      for (int j = 0; j < 10; j++) {
        const int k = i + j;
        const uint32_t fi = ((i & 1) && (j & 1)) ? f2[i] : a[i];
        const uint32_t gj = (k >= 10) ? g19[j] : b[j];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(h[k >= 10 ? k - 10 : k]) : "r"(fi), "r"(gj));
      }
    carry10(r, h);
  }
};

// V4: 10 x 25.5 with the accumulators as explicit (lo, hi) register pairs packed by mov.b64 so that the multiply-add
// is the only 64-bit instruction ptxas sees.
struct F10pair {
  static constexpr int NL = 10;
  static constexpr uint32_t MASK = M25;
  DEV static void mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t f2[10], g19[10];
#pragma unroll
    for (int i = 0; i < 10; i++) { f2[i] = a[i] << 1; g19[i] = b[i] * 19u; }
    uint64_t h[10];
    // first product of every column initialises it (mul.wide), the other nine accumulate
#pragma unroll
    for (int k = 0; k < 10; k++) {
      const int i0 = 0, j0 = k;
      asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(h[k]) : "r"(a[i0]), "r"(b[j0]));
    }
#pragma unroll
    for (int i = 1; i < 10; i++)
#pragma unroll
      for (int j = 0; j < 10; j++) {
        const int k = i + j;
        const uint32_t fi = ((i & 1) && (j & 1)) ? f2[i] : a[i];
        const uint32_t gj = (k >= 10) ? g19[j] : b[j];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(h[k >= 10 ? k - 10 : k]) : "r"(fi), "r"(gj));
      }
    carry10(r, h);
  }
};

template <class F, int ILP>
__global__ void __launch_bounds__(256) fe_kernel(uint32_t iters, uint32_t seed, uint32_t* __restrict__ sink) {
  uint32_t x[ILP][F::NL], y[F::NL];
  for (int k = 0; k < ILP; k++)
    for (int i = 0; i < F::NL; i++) x[k][i] = (seed * (k + 3) + threadIdx.x * 2654435761u + i * 40503u) & F::MASK;
  for (int i = 0; i < F::NL; i++) y[i] = ((seed + threadIdx.x) ^ (0x9e3779b9u * (i + 1))) & F::MASK;
  for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < ILP; k++) F::mul(x[k], x[k], y);
  }
  uint32_t acc = 0;
  for (int k = 0; k < ILP; k++)
    for (int i = 0; i < F::NL; i++) acc ^= x[k][i];
  if (acc == 0x12345u) sink[0] = acc;
}

// ---------------------------------------------------------------------------------------------------------------
static float time_launch(void (*launch)(int, int, uint32_t, void*), int blocks, int threads, uint32_t iters, void* sink) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    launch(blocks, threads, iters, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return best;
}

template <int MODE>
static void launch_pipe(int blocks, int threads, uint32_t iters, void* sink) { pipe_kernel<MODE><<<blocks, threads>>>(iters, 12345u, (uint64_t*)sink); }
template <class F, int ILP>
static void launch_fe(int blocks, int threads, uint32_t iters, void* sink) { fe_kernel<F, ILP><<<blocks, threads>>>(iters, 77u, (uint32_t*)sink); }

int main() {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no device\n"); return 1; }
  const int sms = prop.multiProcessorCount;
  void* sink;
  cudaMalloc(&sink, 64);
  const uint32_t iters = 4096;
  const double threads = (double)sms * 8 * 256;
  struct { const char* name; void (*fn)(int, int, uint32_t, void*); double ops_per_iter; } pipes[] = {
      {"0  mad.wide fused (IMAD.WIDE)          ", launch_pipe<0>, 32},
      {"1  add.u32 (IADD3)                     ", launch_pipe<1>, 32},
      {"2  mad.wide + add 1:1  [per pair]      ", launch_pipe<2>, 32},
      {"8  mad.wide + 2 add    [per triple]    ", launch_pipe<8>, 32},
      {"3  mul lo/hi + add.cc/addc [per group] ", launch_pipe<3>, 32},
      {"4  mad.lo.cc/madc.hi.cc (IMAD.WIDE.X)  ", launch_pipe<4>, 32},
      {"5  xor (LOP3)                          ", launch_pipe<5>, 32},
      {"7  mad.wide + xor 1:1  [per pair]      ", launch_pipe<7>, 32},
      {"6  shf.r.wrap (SHF)                    ", launch_pipe<6>, 32},
      {"9  mad.lo.u32 (IMAD)                   ", launch_pipe<9>, 32},
      {"10 add.cc/addc.cc chains (IADD3.X)     ", launch_pipe<10>, 32},
      {"11 mul.wide.u32 (IMAD.WIDE, RZ addend) ", launch_pipe<11>, 32},
      {"12 mad.lo.cc + madc.hi (carry-out only)", launch_pipe<12>, 32},
      {"13 lop3                                ", launch_pipe<13>, 32},
      {"14 mad.wide + lop3 1:1 [per pair]      ", launch_pipe<14>, 32},
      {"15 carry step shr/and/add64 [per step] ", launch_pipe<15>, 32},
  };
  double clk = 0;
  for (auto& p : pipes) {
    float ms = time_launch(p.fn, sms * 8, 256, iters, sink);
    double rate = threads * iters * p.ops_per_iter / (ms * 1e-3);
    if (clk == 0) clk = rate / (sms * 64.0);  // mode 0 is known to run 64 lanes/clk/SM: gives the SM clock under load
    printf("pipe %s %8.2f T/s  = %6.2f lanes/clk/SM (clock taken as %.3f GHz)\n", p.name, rate / 1e12, rate / (sms * clk), clk / 1e9);
  }
  struct { const char* name; void (*fn)(int, int, uint32_t, void*); int ilp; int blocks_per_sm; int tpb; } fes[] = {
      {"F8      8x32 carry chain  ilp1", launch_fe<F8, 1>, 1, 4, 128},
      {"F8      8x32 carry chain  ilp2", launch_fe<F8, 2>, 2, 4, 128},
      {"F10c    10x25.5 C         ilp1", launch_fe<F10c, 1>, 1, 4, 128},
      {"F10c    10x25.5 C         ilp2", launch_fe<F10c, 2>, 2, 4, 128},
      {"F10col  10x25.5 asm col   ilp1", launch_fe<F10col, 1>, 1, 4, 128},
      {"F10col  10x25.5 asm col   ilp2", launch_fe<F10col, 2>, 2, 4, 128},
      {"F10row  10x25.5 asm row   ilp1", launch_fe<F10row, 1>, 1, 4, 128},
      {"F10row  10x25.5 asm row   ilp2", launch_fe<F10row, 2>, 2, 4, 128},
      {"F10pair 10x25.5 mul+mad   ilp1", launch_fe<F10pair, 1>, 1, 4, 128},
      {"F10pair 10x25.5 mul+mad   ilp2", launch_fe<F10pair, 2>, 2, 4, 128},
  };
  for (auto& f : fes) {
    for (int bps : {4, 8}) {
      float ms = time_launch(f.fn, sms * bps, f.tpb, 1024, sink);
      double rate = (double)sms * bps * f.tpb * 1024 * f.ilp / (ms * 1e-3);
      printf("fe %s threads/SM=%4d: %7.1f G mul/s, %6.1f clk per warp-mul per SMSP\n", f.name, bps * f.tpb, rate / 1e9, clk / (rate / 32 / (sms * 4)));
    }
  }
  return 0;
}
