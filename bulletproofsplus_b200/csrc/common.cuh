// Shared device helpers: the NL x NL limb product with IMAD.WIDE carry chains (used by the Montgomery fields: scalar
// fields and the Weierstrass base fields) and 128-bit vectorised loads/stores of whole AoS objects.
#pragma once
#include <cstdint>

namespace bpp {

#define BPP_DEV __device__ __forceinline__

// r[0..2NL) = a * b
template <int NL>
BPP_DEV void mul_wide_n(uint32_t* r, const uint32_t* a, const uint32_t* b) {
  uint32_t E[2 * NL], O[2 * NL];
#pragma unroll
  for (int i = 0; i < 2 * NL; i++) { E[i] = 0; O[i] = 0; }
#pragma unroll
  for (int i = 0; i < NL; i++) {
    const int je = i & 1, jo = 1 - je;
    {  // products with i+j even, 64-bit aligned on even columns
      const int c = i + je;
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                   : "+r"(E[c]), "+r"(E[c + 1]) : "r"(a[je]), "r"(b[i]));
#pragma unroll
      for (int j = je + 2; j < NL; j += 2)
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                     : "+r"(E[i + j]), "+r"(E[i + j + 1]) : "r"(a[j]), "r"(b[i]));
      if (c + NL < 2 * NL) asm volatile("addc.u32 %0, %0, 0;" : "+r"(E[c + NL]));
    }
    {  // products with i+j odd, kept one column down in O
      const int c = i + jo - 1;
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                   : "+r"(O[c]), "+r"(O[c + 1]) : "r"(a[jo]), "r"(b[i]));
#pragma unroll
      for (int j = jo + 2; j < NL; j += 2)
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                     : "+r"(O[i + j - 1]), "+r"(O[i + j]) : "r"(a[j]), "r"(b[i]));
      if (c + NL < 2 * NL) asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[c + NL]));
    }
  }
  r[0] = E[0];
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r[1]) : "r"(E[1]), "r"(O[0]));
#pragma unroll
  for (int i = 2; i < 2 * NL - 1; i++) asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[i]) : "r"(E[i]), "r"(O[i - 1]));
  asm volatile("addc.u32 %0, %1, %2;" : "=r"(r[2 * NL - 1]) : "r"(E[2 * NL - 1]), "r"(O[2 * NL - 2]));
}

// r[0..2NL) = a * a with NL(NL+1)/2 wide products instead of NL^2: the off-diagonal products a_i * a_j (i < j) are summed
// once in the same even/odd-column carry chains as mul_wide_n, the sum is doubled by a one-bit funnel shift and the diagonal
// squares are added in a single 2NL-limb carry chain.
template <int NL>
BPP_DEV void sqr_wide_n(uint32_t* r, const uint32_t* a) {
  uint32_t E[2 * NL], O[2 * NL], t[2 * NL];
#pragma unroll
  for (int i = 0; i < 2 * NL; i++) { E[i] = 0; O[i] = 0; }
#pragma unroll
  for (int i = 0; i < NL - 1; i++) {
    // partial sums after row i stay below 2^(32 * (i + 1 + NL)): a chain whose last product ends at limb i + NL has no carry out
    if (i + 2 < NL) {  // i + j even: j = i + 2, i + 4, ...; product at even column i + j -> E[i + j], E[i + j + 1]
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                   : "+r"(E[2 * i + 2]), "+r"(E[2 * i + 3]) : "r"(a[i + 2]), "r"(a[i]));
      int last = i + 2;
#pragma unroll
      for (int j = i + 4; j < NL; j += 2) {
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                     : "+r"(E[i + j]), "+r"(E[i + j + 1]) : "r"(a[j]), "r"(a[i]));
        last = j;
      }
      if (last < NL - 1) asm volatile("addc.u32 %0, %0, 0;" : "+r"(E[i + last + 2]));
    }
    {  // i + j odd: j = i + 1, i + 3, ...; product at odd column i + j -> O[i + j - 1], O[i + j]
      asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                   : "+r"(O[2 * i]), "+r"(O[2 * i + 1]) : "r"(a[i + 1]), "r"(a[i]));
      int last = i + 1;
#pragma unroll
      for (int j = i + 3; j < NL; j += 2) {
        asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                     : "+r"(O[i + j - 1]), "+r"(O[i + j]) : "r"(a[j]), "r"(a[i]));
        last = j;
      }
      if (last < NL - 1) asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[i + last + 1]));
    }
  }
  // t = E + (O << 32): the off-diagonal sum, below 2^(64 NL - 1)
  t[0] = 0;  // E[0] = E[1] = 0: the lowest off-diagonal product sits at column 1
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(t[1]) : "r"(E[1]), "r"(O[0]));
#pragma unroll
  for (int i = 2; i < 2 * NL - 1; i++) asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(t[i]) : "r"(E[i]), "r"(O[i - 1]));
  asm volatile("addc.u32 %0, %1, %2;" : "=r"(t[2 * NL - 1]) : "r"(E[2 * NL - 1]), "r"(O[2 * NL - 2]));
  // r = 2 t
#pragma unroll
  for (int i = 2 * NL - 1; i > 0; i--) r[i] = __funnelshift_l(t[i - 1], t[i], 1);
  r[0] = 0;
  // r += sum_i a_i^2 * 2^(64 i)
  asm volatile("mad.lo.cc.u32 %0, %2, %2, %0; madc.hi.cc.u32 %1, %2, %2, %1;" : "+r"(r[0]), "+r"(r[1]) : "r"(a[0]));
#pragma unroll
  for (int i = 1; i < NL - 1; i++)
    asm volatile("madc.lo.cc.u32 %0, %2, %2, %0; madc.hi.cc.u32 %1, %2, %2, %1;" : "+r"(r[2 * i]), "+r"(r[2 * i + 1]) : "r"(a[i]));
  asm volatile("madc.lo.cc.u32 %0, %2, %2, %0; madc.hi.u32 %1, %2, %2, %1;" : "+r"(r[2 * NL - 2]), "+r"(r[2 * NL - 1]) : "r"(a[NL - 1]));
}

// r[0..16) = a * b for 8-limb operands with one level of Karatsuba: three 4 x 4 products (48 wide multiplies) instead of 64.
// The 16 multiplies saved cost ~60 extra integer additions — on sm_100a a gain, because IMAD.WIDE holds the FMA-heavy pipe
// ~4.2-4.6 cycles per warp instruction (the pipe the MSM kernels saturate: 88 % busy) while IADD3 / LOP3 issue to the ALU
// pipe, which those kernels leave mostly idle (ncu: issue slots 44 %).
//   a = a0 + 2^128 a1, b = b0 + 2^128 b1:  a b = z0 + 2^128 ((a0 + a1)(b0 + b1) - z0 - z2) + 2^256 z2
BPP_DEV void mul_wide_8_karatsuba(uint32_t* r, const uint32_t* a, const uint32_t* b) {
  uint32_t z0[8], z2[8], zm[8], sa[4], sb[4], m[9], ca, cb;
  mul_wide_n<4>(z0, a, b);
  mul_wide_n<4>(z2, a + 4, b + 4);
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(sa[0]) : "r"(a[0]), "r"(a[4]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sa[1]) : "r"(a[1]), "r"(a[5]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sa[2]) : "r"(a[2]), "r"(a[6]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sa[3]) : "r"(a[3]), "r"(a[7]));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(ca));
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(sb[0]) : "r"(b[0]), "r"(b[4]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sb[1]) : "r"(b[1]), "r"(b[5]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sb[2]) : "r"(b[2]), "r"(b[6]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(sb[3]) : "r"(b[3]), "r"(b[7]));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(cb));
  mul_wide_n<4>(zm, sa, sb);
  // m = (sa + 2^128 ca)(sb + 2^128 cb) = zm + 2^128 (ca ? sb : 0) + 2^128 (cb ? sa : 0) + 2^256 (ca & cb)   (< 2^258: 9 limbs)
  const uint32_t ma = 0u - ca, mb = 0u - cb;
#pragma unroll
  for (int i = 0; i < 4; i++) m[i] = zm[i];
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(m[4]) : "r"(zm[4]), "r"(sb[0] & ma));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(m[5]) : "r"(zm[5]), "r"(sb[1] & ma));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(m[6]) : "r"(zm[6]), "r"(sb[2] & ma));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(m[7]) : "r"(zm[7]), "r"(sb[3] & ma));
  asm volatile("addc.u32 %0, %1, 0;" : "=r"(m[8]) : "r"(ca & cb));
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(m[4]) : "r"(sa[0] & mb));
  asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(m[5]) : "r"(sa[1] & mb));
  asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(m[6]) : "r"(sa[2] & mb));
  asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(m[7]) : "r"(sa[3] & mb));
  asm volatile("addc.u32 %0, %0, 0;" : "+r"(m[8]));
  // m -= z0; m -= z2   (the middle term a0 b1 + a1 b0 is non-negative: no borrow out)
  asm volatile("sub.cc.u32 %0, %0, %1;" : "+r"(m[0]) : "r"(z0[0]));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("subc.cc.u32 %0, %0, %1;" : "+r"(m[i]) : "r"(z0[i]));
  asm volatile("subc.u32 %0, %0, 0;" : "+r"(m[8]));
  asm volatile("sub.cc.u32 %0, %0, %1;" : "+r"(m[0]) : "r"(z2[0]));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("subc.cc.u32 %0, %0, %1;" : "+r"(m[i]) : "r"(z2[i]));
  asm volatile("subc.u32 %0, %0, 0;" : "+r"(m[8]));
  // r = z0 + 2^256 z2 + 2^128 m
#pragma unroll
  for (int i = 0; i < 4; i++) r[i] = z0[i];
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r[4]) : "r"(z0[4]), "r"(m[0]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[5]) : "r"(z0[5]), "r"(m[1]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[6]) : "r"(z0[6]), "r"(m[2]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[7]) : "r"(z0[7]), "r"(m[3]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[8]) : "r"(z2[0]), "r"(m[4]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[9]) : "r"(z2[1]), "r"(m[5]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[10]) : "r"(z2[2]), "r"(m[6]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[11]) : "r"(z2[3]), "r"(m[7]));
  asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[12]) : "r"(z2[4]), "r"(m[8]));
  asm volatile("addc.cc.u32 %0, %1, 0;" : "=r"(r[13]) : "r"(z2[5]));
  asm volatile("addc.cc.u32 %0, %1, 0;" : "=r"(r[14]) : "r"(z2[6]));
  asm volatile("addc.u32 %0, %1, 0;" : "=r"(r[15]) : "r"(z2[7]));
}

// 128-bit vectorised global loads/stores of whole points (AoS, 32-byte aligned)
template <int WORDS, class T>
BPP_DEV void load_words(T& dst, const void* src) {
  const uint4* s = reinterpret_cast<const uint4*>(src);
  uint32_t* d = reinterpret_cast<uint32_t*>(&dst);
#pragma unroll
  for (int i = 0; i < WORDS / 4; i++) {
    uint4 q = __ldg(s + i);
    d[4 * i] = q.x; d[4 * i + 1] = q.y; d[4 * i + 2] = q.z; d[4 * i + 3] = q.w;
  }
}
template <int WORDS, class T>
BPP_DEV void store_words(void* dst, const T& src) {
  uint4* d = reinterpret_cast<uint4*>(dst);
  const uint32_t* s = reinterpret_cast<const uint32_t*>(&src);
#pragma unroll
  for (int i = 0; i < WORDS / 4; i++) d[i] = make_uint4(s[4 * i], s[4 * i + 1], s[4 * i + 2], s[4 * i + 3]);
}

}  // namespace bpp
