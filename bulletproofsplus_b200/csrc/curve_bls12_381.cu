// BLS12-381 G1 (C-ABI curve 2): the group the reference snapshot actually runs on (`Point(G1)` over mcl,
// src/bls12_381/building_block/point/point.rs:12; Fr = PrimeFieldElem, scalar/prime_field_elem.rs:12).
#include "curve_params.cuh"
#include "curve_vtbl.cuh"

namespace bpp {
struct OrderBlsR {
  static const uint8_t* le() {
    static const uint8_t o[32] = {0x01, 0x00, 0x00, 0x00, 0xff, 0xff, 0xff, 0xff, 0xfe, 0x5b, 0xfe, 0xff, 0x02, 0xa4, 0xbd, 0x53,
                                  0x05, 0xd8, 0xa1, 0x09, 0x08, 0xd8, 0x39, 0x33, 0x48, 0x7d, 0x9d, 0x29, 0x53, 0xa7, 0xed, 0x73};
    return o;
  }
};
BPP_CURVE_TU(vtbl_bls12_381_g1, bls, Bls12381G1, ModBlsR, OrderBlsR)
}  // namespace bpp
