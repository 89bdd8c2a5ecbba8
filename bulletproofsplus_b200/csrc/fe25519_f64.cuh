// GF(2^255 - 19) on the FP64 pipe of sm_100a — the representation used INSIDE the bucket-accumulation kernels only.
//
// The 8 x 32-bit multiply of fe25519.cuh holds the integer-multiply pipe ~330 cycles per warp-multiply (72 IMAD.WIDE at
// ~4.6 clk each); the FP64 pipe of B200 idles beside it. Here an element is 5 limbs of 51 bits in doubles, BALANCED
// (|limb| <= 2^50 + 2^13: "tight"), and sums / differences of two tight elements ("lazy", < 2^51.01) may be multiplied
// without a carry pass: every partial product a*b (|ab| < 2^103) is split exactly by two DFMAs and one DADD,
//     hi = fma_rz(a, b, 1.5 * 2^104)                 = 1.5 * 2^104 + floor(ab / 2^52) * 2^52
//     lo = fma_rz(a, b, (1.5 * 2^104 + 2^52) - hi)   = 2^52 + (ab mod 2^52)
// and the BIT PATTERNS of hi / lo are (constant + integer), so the 25 partial products are summed per column with
// 64-bit integer additions (IADD3 pairs on the ALU pipe, in the shadow of the DFMAs). Then 2^255 = 19 in int64, one
// parallel balanced carry and the 1.5 * 2^52 magic constant bring the result back to 5 tight doubles.
// Measured against the integer multiply by tools/ubench/dfma_fe.cu (profiles/r01e_dfma_fe.txt).
//
// Stand-in for the reference's field layer (mcl Fp via mcl_rust) like fe25519.cuh; exactness is checked
// by the GPU parity tests (every MSM result is compared byte for byte with an independent CPU computation).
#pragma once
#include "fe25519.cuh"

namespace bpp {

struct F5 {
  double v[5];
};

BPP_DEV void f5_add(F5& r, const F5& a, const F5& b) {
#pragma unroll
  for (int i = 0; i < 5; i++) r.v[i] = __dadd_rn(a.v[i], b.v[i]);
}
BPP_DEV void f5_sub(F5& r, const F5& a, const F5& b) {
#pragma unroll
  for (int i = 0; i < 5; i++) r.v[i] = __dsub_rn(a.v[i], b.v[i]);
}
// r = neg ? -a : a (sign bit of the high word)
BPP_DEV void f5_cneg(F5& r, const F5& a, uint32_t neg) {
#pragma unroll
  for (int i = 0; i < 5; i++) {
    const uint32_t hi = (uint32_t)__double2hiint(a.v[i]) ^ (neg << 31);
    r.v[i] = __hiloint2double((int)hi, __double2loint(a.v[i]));
  }
}

// r = a * b; |a_i * b_j| < 2^103 for all i, j (tight x tight, lazy x lazy); r tight
BPP_DEV void f5_mul(F5& r, const F5& a, const F5& b) {
  const double C1 = 0x1.8p104;           // ulp 2^52 over [2^104, 2^105)
  const double C2 = 0x1.8p104 + 0x1p52;  // exact
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
  // strip the exponent patterns (n products per column); the high parts sit one 51-bit limb up, times two
  long long c[10];
#pragma unroll
  for (int k = 0; k < 10; k++) {
    long long x = 0;
    if (k < 9) { const int n = (k < 4 ? k : 8 - k) + 1; x = lo[k] - (long long)n * 0x4330000000000000ll; }
    if (k > 0) { const int n = (k - 1 < 4 ? k - 1 : 9 - k) + 1; x += 2 * (hi[k - 1] - (long long)n * 0x4678000000000000ll); }
    c[k] = x;
  }
#pragma unroll
  for (int k = 0; k < 5; k++) c[k] += 19 * c[k + 5];  // |c| < 2^60
  long long cr[5], rem[5];
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const long long t = c[k] + (1ll << 50);
    cr[k] = t >> 51;
    rem[k] = t & ((1ll << 51) - 1);  // = balanced remainder + 2^50
  }
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const long long in = k ? cr[k - 1] : 19 * cr[4];
    const long long bits = rem[k] + in + (0x4338000000000000ll - (1ll << 50));  // pattern of 1.5 * 2^52 + limb
    r.v[k] = __dsub_rn(__longlong_as_double(bits), 0x1.8p52);
  }
}

// 8 x 32 loose -> 5 tight balanced doubles (canonicalises first)
BPP_DEV void f5_from_fe(F5& r, const Fe25519& a) {
  Fe25519 c;
  fe_canon(c, a);
  unsigned long long w[4];
#pragma unroll
  for (int i = 0; i < 4; i++) w[i] = (unsigned long long)c.v[2 * i] | ((unsigned long long)c.v[2 * i + 1] << 32);
  const unsigned long long M = (1ull << 51) - 1;
  long long l[5];
  l[0] = (long long)(w[0] & M);
  l[1] = (long long)(((w[0] >> 51) | (w[1] << 13)) & M);
  l[2] = (long long)(((w[1] >> 38) | (w[2] << 26)) & M);
  l[3] = (long long)(((w[2] >> 25) | (w[3] << 39)) & M);
  l[4] = (long long)(w[3] >> 12);
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (l[k] > (1ll << 50)) { l[k] -= (1ll << 51); l[k + 1] += 1; }
  if (l[4] > (1ll << 50)) { l[4] -= (1ll << 51); l[0] += 19; }  // 2^255 = 19
#pragma unroll
  for (int k = 0; k < 5; k++) r.v[k] = (double)l[k];
}

// 5 doubles (tight or lazy) -> 8 x 32 loose
BPP_DEV void f5_to_fe(Fe25519& r, const F5& a) {
  // + 8p in limb form: every limb positive (|limb| < 2^52 < 2^54 - 152)
  unsigned long long l[5];
#pragma unroll
  for (int k = 0; k < 5; k++) l[k] = (unsigned long long)((long long)a.v[k] + (k == 0 ? (1ll << 54) - 152 : (1ll << 54) - 8));
  const unsigned long long M = (1ull << 51) - 1;
#pragma unroll
  for (int k = 0; k < 4; k++) { l[k + 1] += l[k] >> 51; l[k] &= M; }
  const unsigned long long top = l[4] >> 51;  // < 2^4
  l[4] &= M;
  unsigned long long w[4];
  w[0] = l[0] | (l[1] << 51);
  w[1] = (l[1] >> 13) | (l[2] << 38);
  w[2] = (l[2] >> 26) | (l[3] << 25);
  w[3] = (l[3] >> 39) | (l[4] << 12);
  // + 19 * top (value < 2^255 + 2^9: stays below 2^256, the loose range)
  const unsigned long long add = 19ull * top;
  unsigned long long s = w[0] + add;
  unsigned long long cy = s < w[0];
  w[0] = s;
#pragma unroll
  for (int i = 1; i < 4; i++) { s = w[i] + cy; cy = s < w[i]; w[i] = s; }
#pragma unroll
  for (int i = 0; i < 4; i++) { r.v[2 * i] = (uint32_t)w[i]; r.v[2 * i + 1] = (uint32_t)(w[i] >> 32); }
}

BPP_DEV void f5_load(F5& r, const double* p) {
#pragma unroll
  for (int i = 0; i < 5; i++) r.v[i] = __ldg(p + i);
}

}  // namespace bpp
