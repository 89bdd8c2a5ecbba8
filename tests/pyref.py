"""Small pure-Python group arithmetic used ONLY by tests to cross-check the C++ oracle
(independent second implementation; Python ints, affine formulas, slow)."""
from __future__ import annotations

import hashlib

# ---------------------------------------------------------------- edwards25519 / ristretto255
P25519 = 2**255 - 19
ELL = 2**252 + 27742317777372353535851937790883648493
ED_D = (-121665 * pow(121666, -1, P25519)) % P25519
SQRT_M1 = pow(2, (P25519 - 1) // 4, P25519)


def _is_neg(x):
    return x & 1


def sqrt_ratio_m1(u, v):
    p = P25519
    v3 = v * v % p * v % p
    v7 = v3 * v3 % p * v % p
    r = u * v3 % p * pow(u * v7 % p, (p - 5) // 8, p) % p
    check = v * r % p * r % p
    correct = check == u % p
    flipped = check == (-u) % p
    flipped_i = check == (-u) * SQRT_M1 % p
    if flipped or flipped_i:
        r = r * SQRT_M1 % p
    if _is_neg(r):
        r = p - r
    return (correct or flipped), r


INVSQRT_A_MINUS_D = sqrt_ratio_m1(1, (-1 - ED_D) % P25519)[1]


def ed_add(P, Q):
    """Affine twisted Edwards addition, a = -1 (complete)."""
    p = P25519
    x1, y1 = P
    x2, y2 = Q
    k = ED_D * x1 * x2 % p * y1 % p * y2 % p
    x3 = (x1 * y2 + x2 * y1) * pow(1 + k, -1, p) % p
    y3 = (y1 * y2 + x1 * x2) * pow(1 - k, -1, p) % p
    return x3, y3


def ed_base():
    y = 4 * pow(5, -1, P25519) % P25519
    ok, x = sqrt_ratio_m1((y * y - 1) % P25519, (ED_D * y * y + 1) % P25519)
    assert ok
    return x, y


def ed_mul(k, P):
    acc = (0, 1)
    while k:
        if k & 1:
            acc = ed_add(acc, P)
        P = ed_add(P, P)
        k >>= 1
    return acc


def ed25519_encode(P):
    x, y = P
    return int(y | ((x & 1) << 255)).to_bytes(32, "little")


def ristretto_encode(P):
    """RFC 9496 §4.3.2 from an affine point (z = 1, t = x*y)."""
    p = P25519
    x0, y0 = P
    z0, t0 = 1, x0 * y0 % p
    u1 = (z0 + y0) * (z0 - y0) % p
    u2 = x0 * y0 % p
    _, invsqrt = sqrt_ratio_m1(1, u1 * u2 % p * u2 % p)
    den1 = invsqrt * u1 % p
    den2 = invsqrt * u2 % p
    z_inv = den1 * den2 % p * t0 % p
    ix0 = x0 * SQRT_M1 % p
    iy0 = y0 * SQRT_M1 % p
    ench = den1 * INVSQRT_A_MINUS_D % p
    rotate = _is_neg(t0 * z_inv % p)
    x, y, den_inv = (iy0, ix0, ench) if rotate else (x0, y0, den2)
    if _is_neg(x * z_inv % p):
        y = (-y) % p
    s = den_inv * (z0 - y) % p
    if _is_neg(s):
        s = p - s
    return s.to_bytes(32, "little")


def ristretto_decode(b):
    p = P25519
    s = int.from_bytes(b, "little")
    if s >= p or _is_neg(s):
        return None
    ss = s * s % p
    u1 = (1 - ss) % p
    u2 = (1 + ss) % p
    u2s = u2 * u2 % p
    v = (-(ED_D * u1 % p * u1) - u2s) % p
    ok, invsqrt = sqrt_ratio_m1(1, v * u2s % p)
    den_x = invsqrt * u2 % p
    den_y = invsqrt * den_x % p * v % p
    x = 2 * s * den_x % p
    if _is_neg(x):
        x = p - x
    y = u1 * den_y % p
    t = x * y % p
    if not ok or _is_neg(t) or y == 0:
        return None
    return x, y



# ---- RFC 9496 §4.3.4: element derivation (hash-to-ristretto255), for the hash-derived generators of bpp_pk_new_hashed
ONE_MINUS_D_SQ = (1 - ED_D * ED_D) % P25519
D_MINUS_ONE_SQ = (ED_D - 1) * (ED_D - 1) % P25519
SQRT_AD_MINUS_ONE = 25063068953384623474111414158702152701244531502492656460079210482610430750235
assert SQRT_AD_MINUS_ONE * SQRT_AD_MINUS_ONE % P25519 == (-ED_D - 1) % P25519


def ristretto_map(t):
    """MAP of RFC 9496 §4.3.4 -> affine (x, y) on edwards25519."""
    p = P25519
    r = SQRT_M1 * t % p * t % p
    u = (r + 1) * ONE_MINUS_D_SQ % p
    v = (-1 - r * ED_D) * (r + ED_D) % p
    was_square, s = sqrt_ratio_m1(u, v)
    s_prime = s * t % p
    if not _is_neg(s_prime):
        s_prime = (-s_prime) % p
    if not was_square:
        s = s_prime
    c = (p - 1) if was_square else r
    N = (c * (r - 1) % p * D_MINUS_ONE_SQ - v) % p
    w0 = 2 * s * v % p
    w1 = N * SQRT_AD_MINUS_ONE % p
    w2 = (1 - s * s) % p
    w3 = (1 + s * s) % p
    X, Y, Z = w0 * w3 % p, w2 * w1 % p, w1 * w3 % p
    zi = pow(Z, -1, p)
    return X * zi % p, Y * zi % p


def ristretto_from_uniform_bytes(b: bytes):
    """64 uniform bytes -> group element (affine edwards25519 representative)."""
    assert len(b) == 64
    t1 = (int.from_bytes(b[:32], "little") & ((1 << 255) - 1)) % P25519
    t2 = (int.from_bytes(b[32:], "little") & ((1 << 255) - 1)) % P25519
    return ed_add(ristretto_map(t1), ristretto_map(t2))


def hashed_generator_seed(domain_sep: bytes, label: bytes, index: int) -> bytes:
    """The 64 uniform bytes bpp_pk_new_hashed feeds to the element derivation for one generator (include/bpp.h)."""
    return hashlib.sha512(b"bpp-b200 generators v1" + len(domain_sep).to_bytes(4, "little") + domain_sep + label + index.to_bytes(8, "little")).digest()


def hashed_generator(domain_sep: bytes, label: bytes, index: int) -> bytes:
    return ristretto_encode(ristretto_from_uniform_bytes(hashed_generator_seed(domain_sep, label, index)))

# ---------------------------------------------------------------- short Weierstrass a = 0
SECP_P = 2**256 - 2**32 - 977
SECP_N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
SECP_G = (0x79BE667EF9DCBBAC55A06295CE870B07029BFCDB2DCE28D959F2815B16F81798,
          0x483ADA7726A3C4655DA4FBFC0E1108A8FD17B448A68554199C47D08FFB10D4B8)
BLS_P = 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB
BLS_R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
BLS_G = (3685416753713387016781088315183077757961620795782546409894578378688607592378376318836054947676345821548104185464507,
         1339506544944476473020471379941921221584933875938349620426543736416511423956333506472724655353366534992391756441569)


def sw_add(P, Q, p):
    if P is None:
        return Q
    if Q is None:
        return P
    x1, y1 = P
    x2, y2 = Q
    if x1 == x2:
        if (y1 + y2) % p == 0:
            return None
        lam = 3 * x1 * x1 * pow(2 * y1, -1, p) % p
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, p) % p
    x3 = (lam * lam - x1 - x2) % p
    return x3, (lam * (x1 - x3) - y1) % p


def sw_mul(k, P, p):
    acc = None
    while k:
        if k & 1:
            acc = sw_add(acc, P, p)
        P = sw_add(P, P, p)
        k >>= 1
    return acc


def sw_encode(P, fb):
    if P is None:
        return bytes(2 * fb)
    return P[0].to_bytes(fb, "little") + P[1].to_bytes(fb, "little")


CURVES = {
    0: dict(order=ELL, pb=32),
    1: dict(order=SECP_N, pb=64),
    2: dict(order=BLS_R, pb=96),
}


def base_mul_encoded(curve, k):
    """encoding of k*g computed by this file's arithmetic."""
    k %= CURVES[curve]["order"]
    if curve == 0:
        return ristretto_encode(ed_mul(k, ed_base()))
    if curve == 1:
        return sw_encode(sw_mul(k, SECP_G, SECP_P), 32)
    return sw_encode(sw_mul(k, BLS_G, BLS_P), 48)


# ---------------------------------------------------------------- synthetic input stream
def prng_bytes(seed: int, index: int, n: int = 64) -> bytes:
    """Counter-based stream SHA-256(seed_le64 || index_le64 || block_le32), widened to n bytes."""
    out = b""
    blk = 0
    while len(out) < n:
        out += hashlib.sha256(seed.to_bytes(8, "little") + index.to_bytes(8, "little") + blk.to_bytes(4, "little")).digest()
        blk += 1
    return out[:n]


def prng_scalar(seed: int, index: int, order: int) -> int:
    return int.from_bytes(prng_bytes(seed, index, 64), "little") % order
