// Straus window sums for the per-call ("generic") points of many small jobs — the 16 to 39 points a range proof brings with
// it (A, A', B', L_j, R_j, V_j: range/mod.rs:480-503, weighted_inner_product_proof.rs:297-320) next to its ~130..2050
// generator terms.
//
// The bucket method serves those few points badly: with D = 16..39 points per job it picks 3- or 4-bit windows, i.e. 64-85
// bucket sets of 4-8 buckets per job that hold one point or none, and runs them through the whole pipeline — per-job sort,
// size sort, three accumulation kernels, chunk and set reduction: twelve dependent launches before the Horner pass, which on
// the auxiliary stream are the critical path of every batch of small proofs (config 4: 4.6 ms against a 1.5 ms direct kernel).
// Here ONE kernel does all of it, one block of 64 threads per job:
//   phase 0  the job's terms are scanned for per-call points (point index >= n_fixed); at most STRAUS_MAX_D of them
//   phase 1  thread i builds 1P_i .. 8P_i (seven mixed additions) in shared memory and recodes scalar i into 64 signed 4-bit
//            digits
//   phase 2  thread w adds up window w: S_w = sum_i +-(|d_iw| P_i), one complete addition per nonzero digit
// and the 64 window sums per job go to the same `windows` array the bucket pipeline fills, so the existing Horner kernel
// (msm_combine4_kernel, c = 4, 64 windows) finishes the job. Operation count per job: 7 D mixed + ~60 D complete additions
// + Horner — about the bucket method's, but in one launch with no global intermediate but the window sums.
// Same group element as the bucket path, hence the same encoding: the parity tests cover both (BPP_NO_STRAUS=1 switches back).
#pragma once
#include "msm.cuh"

namespace bpp {

constexpr uint32_t STRAUS_MAX_D = 48;      // per-call points per job this path accepts
constexpr uint32_t STRAUS_WINDOWS = 64;    // 4-bit signed windows: 256 bits >= SCALAR_BITS + 1
constexpr uint32_t STRAUS_ROW = 33;        // words per table entry in shared memory (32 + 1: entries k and k' on distinct banks)

inline size_t straus_smem_bytes(uint32_t d) { return (size_t)d * 8 * STRAUS_ROW * 4 + (size_t)d * STRAUS_WINDOWS + (size_t)d * 4; }

template <class C>
__global__ void __launch_bounds__(64)
msm_straus_kernel(MsmPlan pl, uint32_t d_max, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                  const uint32_t* __restrict__ job_start, const uint32_t* __restrict__ dyn, uint32_t* __restrict__ windows) {
  static_assert(C::EXT_WORDS == 32, "table rows are written for 32-word extended points");
  extern __shared__ uint32_t sm[];
  uint32_t* tab = sm;                                                       // [d_max][8][STRAUS_ROW]
  uint8_t* dig = reinterpret_cast<uint8_t*>(tab + (size_t)d_max * 8 * STRAUS_ROW);  // [d_max][64]: bit 7 = negative, bits 0..3 = magnitude
  uint32_t* list = reinterpret_cast<uint32_t*>(dig + (size_t)d_max * STRAUS_WINDOWS);  // [d_max] term indices
  __shared__ uint32_t n_found;
  const uint32_t job = blockIdx.x, tid = threadIdx.x;
  const uint32_t t0 = pl.uniform ? job * pl.uniform : job_start[job];
  const uint32_t t1 = pl.uniform ? t0 + pl.uniform : job_start[job + 1];
  if (tid == 0) n_found = 0;
  __syncthreads();
  for (uint32_t t = t0 + tid; t < t1; t += 64)
    if (pidx[t] >= pl.n_fixed) {
      const uint32_t pos = atomicAdd(&n_found, 1u);
      if (pos < d_max) list[pos] = t;
    }
  __syncthreads();
  const uint32_t nd = n_found < d_max ? n_found : d_max;
  // phase 1: multiples and digits of point / scalar i
  for (uint32_t i = tid; i < nd; i += 64) {
    const uint32_t t = list[i];
    const uint32_t* q = dyn + (size_t)(pidx[t] - pl.n_fixed) * C::AFF_WORDS;
    typename C::Ext e;
    C::init_from_ptr(e, q, 0u);
    uint32_t* row = tab + (size_t)i * 8 * STRAUS_ROW;
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      if (k) C::madd_ptr(e, q, 0u);  // (k + 1) P: the addition law is complete, so P + P needs no doubling formula
      uint32_t* r = row + k * STRAUS_ROW;
#pragma unroll
      for (int j = 0; j < 8; j++) { r[j] = e.X.v[j]; r[8 + j] = e.Y.v[j]; r[16 + j] = e.Z.v[j]; r[24 + j] = e.T.v[j]; }
    }
    const uint4* sp = reinterpret_cast<const uint4*>(scalars + 8ull * t);
    const uint4 a = __ldg(sp), b = __ldg(sp + 1);
    const uint32_t s[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t carry = 0;
#pragma unroll
    for (int l = 0; l < 8; l++)
#pragma unroll
      for (int n = 0; n < 8; n++) {
        const uint32_t raw = ((s[l] >> (4 * n)) & 15u) + carry;
        const uint32_t neg = raw > 8u;
        const uint32_t mag = neg ? 16u - raw : raw;  // 0..8
        carry = neg;
        dig[i * STRAUS_WINDOWS + l * 8 + n] = (uint8_t)(mag | (neg << 7));
      }
    // scalars are canonical (< order < 2^253): the top window is at most 1 + carry, so no carry leaves it
  }
  __syncthreads();
  // phase 2: window sums
  typename C::Ext acc;
  C::set_identity(acc);
  for (uint32_t i = 0; i < nd; i++) {
    const uint32_t d = dig[i * STRAUS_WINDOWS + tid];
    const uint32_t mag = d & 15u;
    if (mag) {
      typename C::Ext e;
      const uint32_t* r = tab + ((size_t)i * 8 + (mag - 1)) * STRAUS_ROW;
#pragma unroll
      for (int j = 0; j < 8; j++) { e.X.v[j] = r[j]; e.Y.v[j] = r[8 + j]; e.Z.v[j] = r[16 + j]; e.T.v[j] = r[24 + j]; }
      if (d >> 7) C::neg(e);
      C::add(acc, e);
    }
  }
  store_words<C::EXT_WORDS>(windows + ((size_t)job * STRAUS_WINDOWS + tid) * C::EXT_WORDS, acc);
}

}  // namespace bpp
