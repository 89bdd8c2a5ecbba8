// secp256k1 (C-ABI curve 1): the reference's previous backend (src/secp256k1/, constants secp256k1.rs:20-53); secondary path.
#include "curve_params.cuh"
#include "curve_vtbl.cuh"

namespace bpp {
struct OrderSecpN {
  static const uint8_t* le() {
    static const uint8_t o[32] = {0x41, 0x41, 0x36, 0xd0, 0x8c, 0x5e, 0xd2, 0xbf, 0x3b, 0xa0, 0x48, 0xaf, 0xe6, 0xdc, 0xae, 0xba,
                                  0xfe, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff};
    return o;
  }
};
const CurveVtbl* vtbl_secp256k1() {
  static const CurveVtbl v = make_vtbl<Secp256k1, ModSecpN, OrderSecpN>();
  return &v;
}
}  // namespace bpp
