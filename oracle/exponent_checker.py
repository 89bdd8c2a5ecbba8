"""ORACLE — TEST INFRASTRUCTURE ONLY.

Exponent-domain checker: because the reference's generators are known
multiples of the base point (src/publickey.rs:21-48: g=1, h=2, G_i=3(i+1), H_i=5(i+1)),
every point on the path is k*g for a k computable with modular integers alone, and
MulVec::calculate(s, k*g) = (sum s_i k_i)*g.  This file re-derives prove/verify with Python
ints only — independently of the C++ restatement in range.hpp/protocol.hpp — and is used to
(a) pin the C++ oracle against SURVEY's KAT set 2 and (b) give large-N MSM expectations with a
single scalar multiplication.

Citations are to /root/reference/src.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass, field

ORDERS = {
    0: 2**252 + 27742317777372353535851937790883648493,  # Ristretto255, ell
    1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,  # secp256k1 n (secp256k1.rs:25)
    2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001,  # BLS12-381 r
}


@dataclass
class Params:
    """Challenges / nonces. `reference(m, rounds)` = the constants hard-coded in the snapshot."""
    y: int
    z: int
    e_final: int
    alpha: int
    r: int
    s: int
    delta: int
    eta: int
    e_round: list = field(default_factory=list)
    d_L: list = field(default_factory=list)
    d_R: list = field(default_factory=list)

    @staticmethod
    def reference(m: int, rounds: int) -> "Params":
        single = m == 1  # range/mod.rs:94,109-110 vs :256,278-279
        return Params(y=7 if single else 12, z=7 if single else 23, e_final=99, alpha=7 if single else 33,
                      r=33, s=44, delta=88, eta=123, e_round=[7] * rounds, d_L=[4] * rounds, d_R=[5] * rounds)


def as_i32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


@dataclass
class ExpProof:
    """Discrete logs (w.r.t. g) of the proof points, and the three scalars."""
    A: int
    L: list
    R: list
    wipA: int
    wipB: int
    r_prime: int
    s_prime: int
    d_prime: int

    def kat_digest(self) -> str:
        """SHA-256 over the values, 32 B big-endian each, order A, L[..], R[..], proof.A, proof.B, r', s', d' (SURVEY KAT set 2)."""
        h = hashlib.sha256()
        for x in [self.A] + self.L + self.R + [self.wipA, self.wipB, self.r_prime, self.s_prime, self.d_prime]:
            h.update(int(x).to_bytes(32, "big"))
        return h.hexdigest()


class Checker:
    def __init__(self, order: int):
        self.q = order

    # generators as discrete logs (publickey.rs:21-48)
    def pk(self, length: int):
        return 1, 2, [3 * (i + 1) for i in range(length)], [5 * (i + 1) for i in range(length)]

    def commit(self, v: int, gamma: int) -> int:  # prover.rs:36-39 + publickey.rs:50-52
        return (as_i32(v) + 2 * gamma) % self.q

    def inv(self, x: int) -> int:
        return pow(x, -1, self.q)

    def wip_prove(self, N, a, b, ypow, gamma, prm: Params, Gd=None, Hd=None):
        """weighted_inner_product_proof.rs:36-226 in the exponent domain."""
        q = self.q
        g, h, G0, H0 = self.pk(N)
        G = list(Gd if Gd is not None else G0)
        H = list(Hd if Hd is not None else H0)
        a, b, py = list(a), list(b), list(ypow)
        alpha = gamma % q
        n = N
        Ls, Rs = [], []
        rnd = 0
        while n != 1:
            n //= 2
            a1, a2, b1, b2 = a[:n], a[n:2 * n], b[:n], b[n:2 * n]
            y1, y2 = py[:n], py[n:2 * n]
            G1, G2, H1, H2 = G[:n], G[n:2 * n], H[:n], H[n:2 * n]
            c_L = sum(a1[i] * b2[i] * y1[i] for i in range(n)) % q
            c_R = sum(a2[i] * b1[i] * y2[i] for i in range(n)) % q
            d_L, d_R = prm.d_L[rnd], prm.d_R[rnd]
            yh = y1[n - 1]
            yhi = self.inv(yh)
            L = (sum(yhi * a1[i] * G2[i] + b2[i] * H1[i] for i in range(n)) + c_L * g + d_L * h) % q
            R = (sum(yh * a2[i] * G1[i] + b1[i] * H2[i] for i in range(n)) + c_R * g + d_R * h) % q
            Ls.append(L)
            Rs.append(R)
            e = prm.e_round[rnd]
            ei = self.inv(e)
            a = [(a1[i] * e + a2[i] * yh * ei) % q for i in range(n)]
            b = [(b1[i] * ei + b2[i] * e) % q for i in range(n)]
            G = [(ei * G1[i] + yhi * e * G2[i]) % q for i in range(n)]
            H = [(e * H1[i] + ei * H2[i]) % q for i in range(n)]
            py = y1
            alpha = (alpha + e * e * d_L + ei * ei * d_R) % q
            rnd += 1
        r, s, delta, eta = prm.r, prm.s, prm.delta, prm.eta
        rcbsca = (r * py[0] * b[0] + s * py[0] * a[0]) % q
        rcs = r * py[0] * s % q
        A = (r * G[0] + s * H[0] + rcbsca * g + delta * h) % q
        B = (rcs * g + eta * h) % q
        e = prm.e_final
        return Ls, Rs, A, B, (r + a[0] * e) % q, (s + b[0] * e) % q, (eta + delta * e + alpha * e * e) % q

    def range_prove(self, n, m, v, gamma, prm: Params | None = None) -> ExpProof:
        """range/mod.rs:80-187 (m == 1) and :240-403 (m >= 2)."""
        q = self.q
        N = n * m
        rounds = N.bit_length() - 1
        prm = prm or Params.reference(m, rounds)
        g, h, G, H = self.pk(N)
        y, z, alpha = prm.y, prm.z, prm.alpha
        bits = [((v[i // n] >> (i % n)) & 1) if (i % n) < 64 else 0 for i in range(N)]
        A = (alpha * h + sum(G[i] if bits[i] else -H[i] for i in range(N))) % q
        pow2 = [pow(2, i, q) for i in range(n)]
        ypow = [pow(y, i + 1, q) for i in range(N)]
        if m == 1:
            d = pow2
            V_exp = [pow(y, n + 1, q)]
            gam = gamma[0] * V_exp[0]
        else:
            zpow = [pow(z * z, j + 1, q) for j in range(m)]
            d = [pow2[i] * zpow[j] % q for j in range(m) for i in range(n)]
            y1 = pow(y, N + 1, q)
            V_exp = [zp * y1 % q for zp in zpow]
            gam = sum(zpow[j] * gamma[j] for j in range(m)) * y1
        H_exp = [(d[i] * ypow[N - 1 - i] + z) % q for i in range(N)]
        a = [(1 - z) % q if bits[i] else (-z) % q for i in range(N)]
        b = [H_exp[i] if bits[i] else (H_exp[i] - 1) % q for i in range(N)]
        alpha_hat = (alpha + gam) % q
        Ls, Rs, wA, wB, rp, sp, dp = self.wip_prove(N, a, b, ypow, alpha_hat, prm)
        return ExpProof(A, Ls, Rs, wA, wB, rp, sp, dp)

    def verification_scalars(self, N, rounds, prm: Params):
        """weighted_inner_product_proof.rs:330-382."""
        q = self.q
        ch = list(prm.e_round[:rounds])
        chi = [self.inv(c) for c in ch]
        allinv = 1
        for c in chi:
            allinv = allinv * c % q
        sq = [c * c % q for c in ch]
        isq = [c * c % q for c in chi]
        s = [allinv]
        for i in range(1, N):
            lg = i.bit_length() - 1
            s.append(s[i - (1 << lg)] * sq[rounds - 1 - lg] % q)
        return sq, isq, s

    def range_verify(self, n, m, V, p: ExpProof, prm: Params | None = None) -> bool:
        """range/mod.rs:189-238 -> wip.rs:238-328 (m == 1); range/mod.rs:405-510 (m >= 2)."""
        q = self.q
        N = n * m
        rounds = len(p.L)
        if N != 1 << rounds:
            return False
        prm = prm or Params.reference(m, rounds)
        g, h, G, H = self.pk(N)
        y, z = prm.y, prm.z
        sq, isq, s = self.verification_scalars(N, rounds, prm)
        e = prm.e_final
        pow2 = [pow(2, i, q) for i in range(n)]
        ypow = [pow(y, i + 1, q) for i in range(N)]
        yinv = self.inv(y)
        if m == 1:
            Gc = [(-z) % q] * n
            Hc = [(pow2[i] * ypow[n - 1 - i] + z) % q for i in range(n)]
            Vc = pow(y, n + 1, q)
            gc = (sum(ypow) * (z - z * z) - (pow(2, n, q) - 1) * Vc * z) % q
            e2 = e * e % q
            tot = p.wipB + e * p.wipA + e2 * p.A
            tot += ((-p.r_prime * y * p.s_prime) + gc * e2) * g + (-p.d_prime) * h
            tot += sum(sq[k] * e2 * p.L[k] + isq[k] * e2 * p.R[k] for k in range(rounds))
            tot += sum((-s[i] * pow(yinv, i + 1, q) * p.r_prime * e * y + Gc[i] * e2) * G[i] for i in range(n))
            tot += sum((-s[n - 1 - i] * p.s_prime * e + Hc[i] * e2) * H[i] for i in range(n))
            tot += Vc * e2 * V[0]
            return tot % q == 0
        zz = z * z % q
        zpow = [pow(zz, j + 1, q) for j in range(m)]
        d = [pow2[i] * zpow[j] % q for j in range(m) for i in range(n)]
        y1 = pow(y, N + 1, q)
        ei = self.inv(e)
        e2i = self.inv(e * e % q)
        tot = p.A + ei * p.wipA + e2i * p.wipB
        sum_y = sum(ypow) % q
        sum_2 = (pow(2, n, q) - 1) % q
        sum_z = sum(zpow) % q
        g_exp = (-p.r_prime * p.s_prime * y * e2i + (sum_y * (z - zz) - y1 * z * sum_2 * sum_z)) % q
        tot += g_exp * g + (-p.d_prime * e2i) * h
        tot += sum(sq[k] * p.L[k] + isq[k] * p.R[k] for k in range(rounds))
        tot += sum(((-z) - s[i] * pow(yinv, i + 1, q) * p.r_prime * ei * y) * G[i] for i in range(N))
        tot += sum((-p.s_prime * ei * s[N - 1 - i] + d[i] * ypow[N - 1 - i] + z) * H[i] for i in range(N))
        tot += sum(zpow[j] * y1 * V[j] for j in range(m))
        return tot % q == 0


def main_rs_kat(curve: int) -> ExpProof:
    """The src/main.rs configuration: n=64, m=2, v={2,5}, gamma={3,7}."""
    return Checker(ORDERS[curve]).range_prove(64, 2, [2, 5], [3, 7])


if __name__ == "__main__":
    for c in (2, 0, 1):
        pr = main_rs_kat(c)
        ck = Checker(ORDERS[c])
        ok = ck.range_verify(64, 2, [ck.commit(2, 3), ck.commit(5, 7)], pr)
        print(c, pr.kat_digest(), "verify:", ok)
