// bpp_debug_op: unit-level access to the device arithmetic (field, group law, scalar field) so that the
// parity tests can localise a failure below the MSM / protocol level. Test surface only.
#pragma once
#include "proto_misc.cuh"

namespace bpp {

// ops 0..4 (field): C::FIELD_BYTES per element; curve 0: raw 256-bit values (any value < 2^256), curves 1, 2: canonical; out canonical
// ops 10..14 (group): a, b are encoded points; out = encoded result
// ops 20..23 (scalar field): canonical scalars in, canonical out
template <class C, class M>
__global__ void debug_op_kernel(int op, uint32_t n, const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, uint8_t* __restrict__ out,
                                uint8_t* __restrict__ ok) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  if (op < 10) {
    C::dbg_field(op, a + (size_t)t * C::FIELD_BYTES, b + (size_t)t * C::FIELD_BYTES, out + (size_t)t * C::FIELD_BYTES);
    ok[t] = 1;
  } else if (op < 20) {
    typename C::Aff qa, qb;
    bool good = C::decode(qa, a + (size_t)t * C::POINT_BYTES);
    good = C::decode(qb, b + (size_t)t * C::POINT_BYTES) && good;
    typename C::Ext pa, pb;
    C::aff_to_ext(pa, qa);
    C::aff_to_ext(pb, qb);
    switch (op) {
      case 10: break;                         // decode -> encode
      case 11: C::madd(pa, qb, false); break;  // a + b (mixed)
      case 12: C::dbl(pa); break;              // 2a
      case 13: C::add(pa, pb); break;          // a + b (extended)
      default: C::madd(pa, qb, true); break;   // a - b
    }
    C::encode(out + (size_t)t * C::POINT_BYTES, pa);
    ok[t] = good ? 1 : 0;
  } else {
    const uint32_t* x = reinterpret_cast<const uint32_t*>(a) + 8 * (size_t)t;
    const uint32_t* y = reinterpret_cast<const uint32_t*>(b) + 8 * (size_t)t;
    Sc<M> sx = sc_load<M>(x), sy = sc_load<M>(y), r;
    switch (op) {
      case 20: r = sx * sy; break;
      case 21: r = sx.inv(); break;
      case 22: r = sx + sy; break;
      default: r = sx - sy; break;
    }
    sc_store<M>(reinterpret_cast<uint32_t*>(out) + 8 * (size_t)t, r);
    ok[t] = 1;
  }
}

}  // namespace bpp
