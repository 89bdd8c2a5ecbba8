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
