//! The reference crate's API (gogoex/BulletProofsPlus `src/lib.rs:11-13`) over the B200 backend.
//!
//! Same names, argument meaning and error behaviour as the reference: caller errors panic (`mulvec.rs:23-25`,
//! `weighted_inner_product_proof.rs:61-67`, `range/mod.rs:90-91,252-253`, `arith.rs:12-14`), verification failure is
//! `Err(ProofError::VerificationError)` (`errors.rs:17`). Transcript and Fiat-Shamir logic stay here on the host; the library
//! only ever sees challenges (`transcript.rs` below follows the label protocol of `bulletproofsplus_b200/merlin.py` and
//! `include/bpp_merlin.hpp`, which are compiled and tested against each other and against merlin's own test vector).
//!
//! NOT COMPILED in the build image (no Rust toolchain there). `bpp-sys` is generated from `include/bpp.h`.
use bpp_sys::*;
use std::ptr;

pub const POINT_BYTES: usize = 32; // Ristretto255, RFC 9496

/// `point/point.rs:12` — here a canonical compressed Ristretto255 point.
#[derive(Clone, Copy, PartialEq, Eq, Debug)]
pub struct Point(pub [u8; 32]);
/// `scalar/prime_field_elem.rs:12` — here 32 bytes little-endian, canonical (< ell).
#[derive(Clone, Copy, PartialEq, Eq, Debug)]
pub struct PrimeFieldElem(pub [u8; 32]);

/// `errors.rs:14-50`; the path only ever produces `VerificationError` (and `FormatError` for undecodable input).
#[derive(Clone, Copy, PartialEq, Eq, Debug)]
pub enum ProofError {
    VerificationError,
    FormatError,
}

/// ell = 2^252 + 27742317777372353535851937790883648493, little-endian 64-bit limbs
const ELL: [u64; 4] = [0x5812631a5cf5d3ed, 0x14def9dea2f79cd6, 0, 0x1000000000000000];

impl PrimeFieldElem {
    /// `prime_field_elem.rs:191-195`: negative n -> modulus - |n|
    pub fn new(n: i32) -> Self {
        let mut le = [0u8; 32];
        le[..8].copy_from_slice(&(n.unsigned_abs() as u64).to_le_bytes());
        let e = PrimeFieldElem(le);
        if n < 0 { e.neg() } else { e }
    }
    pub fn neg(&self) -> Self {
        if self.0 == [0u8; 32] { return *self; }
        let mut out = [0u8; 32];
        let mut borrow = 0u128;
        for i in 0..4 {
            let mut w = [0u8; 8];
            w.copy_from_slice(&self.0[8 * i..8 * i + 8]);
            let d = (ELL[i] as u128).wrapping_sub(u64::from_le_bytes(w) as u128).wrapping_sub(borrow);
            out[8 * i..8 * i + 8].copy_from_slice(&(d as u64).to_le_bytes());
            borrow = (d >> 64) & 1;
        }
        PrimeFieldElem(out)
    }
    /// 64 uniform bytes reduced mod ell (bit-serial; off the hot path). None when the result is zero.
    pub fn from_bytes_wide(wide: &[u8; 64]) -> Option<Self> {
        let mut r = [0u64; 5];
        for bit in (0..512).rev() {
            for i in (1..5).rev() { r[i] = (r[i] << 1) | (r[i - 1] >> 63); }
            r[0] = (r[0] << 1) | ((wide[bit >> 3] >> (bit & 7)) & 1) as u64;
            let mut ge = r[4] != 0;
            if !ge {
                ge = true;
                for i in (0..4).rev() { if r[i] != ELL[i] { ge = r[i] > ELL[i]; break; } }
            }
            if ge {
                let mut borrow = 0u128;
                for i in 0..4 {
                    let d = (r[i] as u128).wrapping_sub(ELL[i] as u128).wrapping_sub(borrow);
                    r[i] = d as u64;
                    borrow = (d >> 64) & 1;
                }
                r[4] = r[4].wrapping_sub(borrow as u64);
            }
        }
        let mut le = [0u8; 32];
        for i in 0..4 { le[8 * i..8 * i + 8].copy_from_slice(&r[i].to_le_bytes()); }
        if le == [0u8; 32] { None } else { Some(PrimeFieldElem(le)) }
    }
}

/// `arith.rs:6-17`: the reference initialises libmcl once per process behind a `Once`; a `bpp_ctx` is per thread (one CUDA
/// stream + scratch arena each; not thread-safe), so it lives in a thread-local and is created on first use.
pub struct Arith;
struct Ctx(*mut bpp_ctx);
impl Drop for Ctx { fn drop(&mut self) { unsafe { bpp_destroy(self.0) } } }
thread_local! {
    static CTX: Ctx = unsafe {
        let mut c = ptr::null_mut();
        // arith.rs:12-14 panics when init fails; so does this (no CUDA device = no backend: there is no CPU fallback)
        assert_eq!(bpp_init(0, ptr::null_mut(), &mut c), BPP_OK, "bpp_init failed: no usable CUDA device");
        Ctx(c)
    };
}
impl Arith {
    pub fn init() { CTX.with(|_| ()); }
}
fn ctx() -> *mut bpp_ctx { CTX.with(|c| c.0) }
fn check(rc: i32, what: &str) {
    if rc != BPP_OK {
        let msg = unsafe { std::ffi::CStr::from_ptr(bpp_last_error(ctx())) }.to_string_lossy().into_owned();
        panic!("{}: bpp error {} ({})", what, rc, msg); // the reference panics on caller errors
    }
}
fn flat<T: Copy, const N: usize>(items: &[T], f: impl Fn(&T) -> [u8; N]) -> Vec<u8> {
    items.iter().flat_map(|x| f(x).to_vec()).collect()
}

/// `publickey.rs:13-52`. The generator table lives in HBM; `g`, `h`, `G_vec`, `H_vec` are fetched on demand.
pub struct PublicKey { h: *mut bpp_pk, len: usize }
unsafe impl Send for PublicKey {} // immutable after creation, shareable across contexts on one device (include/bpp.h)
impl PublicKey {
    pub fn new(len: usize) -> Self {
        let mut h = ptr::null_mut();
        check(unsafe { bpp_pk_new(ctx(), BPP_RISTRETTO255, len, &mut h) }, "PublicKey::new");
        PublicKey { h, len }
    }
    /// Hash-derived generators (`bpp_pk_new_hashed`): no known discrete-log relations — use this outside parity tests.
    pub fn new_hashed(len: usize, domain_sep: &[u8]) -> Self {
        let mut h = ptr::null_mut();
        check(unsafe { bpp_pk_new_hashed(ctx(), BPP_RISTRETTO255, len, domain_sep.as_ptr(), domain_sep.len(), &mut h) }, "PublicKey::new_hashed");
        PublicKey { h, len }
    }
    pub fn len(&self) -> usize { self.len }
    /// Optional, once per key (`bpp_pk_precompute`): direct fixed-base tables within `max_bytes` of HBM; every later prove /
    /// verify against the key, one proof or a batch, takes them. Returns the window width. No counterpart in the reference.
    pub fn precompute(&mut self, max_bytes: usize) -> i32 {
        let mut c: std::os::raw::c_int = 0;
        check(unsafe { bpp_pk_precompute(ctx(), self.h, max_bytes, &mut c) }, "PublicKey::precompute");
        c as i32
    }
    /// (g, h, G_vec, H_vec) — the pub fields of the reference struct
    pub fn generators(&self) -> (Point, Point, Vec<Point>, Vec<Point>) {
        let (mut g, mut h) = ([0u8; 32], [0u8; 32]);
        let (mut gv, mut hv) = (vec![0u8; 32 * self.len + 1], vec![0u8; 32 * self.len + 1]);
        check(unsafe { bpp_pk_export(ctx(), self.h, g.as_mut_ptr(), h.as_mut_ptr(), gv.as_mut_ptr(), hv.as_mut_ptr()) }, "pk_export");
        let split = |v: &[u8]| (0..self.len).map(|i| { let mut p = [0u8; 32]; p.copy_from_slice(&v[32 * i..32 * i + 32]); Point(p) }).collect();
        (Point(g), Point(h), split(&gv), split(&hv))
    }
    /// `publickey.rs:50-52`, with the `v as i32` of `range/prover.rs:37` inside (parity with the reference)
    pub fn commitment(&self, v: u64, gamma: &PrimeFieldElem) -> Point {
        let mut out = [0u8; 32];
        check(unsafe { bpp_commit_batch(ctx(), self.h, &v, gamma.0.as_ptr(), 1, out.as_mut_ptr()) }, "commitment");
        Point(out)
    }
    /// all 64 bits of v: what a proof with n = 64 is about
    pub fn commitment_u64(&self, v: u64, gamma: &PrimeFieldElem) -> Point {
        let mut out = [0u8; 32];
        check(unsafe { bpp_commit_batch_u64(ctx(), self.h, &v, gamma.0.as_ptr(), 1, out.as_mut_ptr()) }, "commitment_u64");
        Point(out)
    }
}
impl Drop for PublicKey { fn drop(&mut self) { unsafe { bpp_pk_free(self.h) } } }

/// `range/prover.rs:13-42`
pub struct RangeProver { pub v_vec: Vec<u64>, pub gamma_vec: Vec<PrimeFieldElem>, pub commitment_vec: Vec<Point> }
impl RangeProver {
    pub fn new() -> Self { RangeProver { v_vec: vec![], gamma_vec: vec![], commitment_vec: vec![] } }
    pub fn commit(&mut self, pk: &PublicKey, v: u64, gamma: PrimeFieldElem) {
        self.commitment_vec.push(pk.commitment(v, &gamma));
        self.v_vec.push(v);
        self.gamma_vec.push(gamma);
    }
    pub fn commit_u64(&mut self, pk: &PublicKey, v: u64, gamma: PrimeFieldElem) {
        self.commitment_vec.push(pk.commitment_u64(v, &gamma));
        self.v_vec.push(v);
        self.gamma_vec.push(gamma);
    }
}

/// Named by the task statement; the snapshot passes a bare slice to `verify` (`range/mod.rs:57-62`).
pub struct RangeVerifier { pub commitment_vec: Vec<Point> }
impl RangeVerifier {
    pub fn new() -> Self { RangeVerifier { commitment_vec: vec![] } }
    pub fn allocate(&mut self, commitments: &[Point]) { self.commitment_vec.extend_from_slice(commitments); }
}

/// `bls12_381/building_block/mulvec.rs:13-53`
pub struct MulVec { scalars: Vec<PrimeFieldElem>, points: Vec<Point> }
impl MulVec {
    pub fn new() -> Self { MulVec { scalars: vec![], points: vec![] } }
    pub fn add_scalar(&mut self, s: &PrimeFieldElem) { self.scalars.push(*s); }
    pub fn add_scalars(&mut self, s: &[PrimeFieldElem]) { self.scalars.extend_from_slice(s); }
    pub fn add_point(&mut self, p: &Point) { self.points.push(*p); }
    pub fn add_points(&mut self, p: &[Point]) { self.points.extend_from_slice(p); }
    pub fn calculate(&self) -> Point {
        assert_eq!(self.scalars.len(), self.points.len(), "lengths of scalars and points must match"); // mulvec.rs:23-25
        let (s, p) = (flat(&self.scalars, |x| x.0), flat(&self.points, |x| x.0));
        let mut out = [0u8; 32];
        check(unsafe { bpp_msm(ctx(), BPP_RISTRETTO255, s.as_ptr(), p.as_ptr(), self.scalars.len(), out.as_mut_ptr()) }, "MulVec::calculate");
        Point(out)
    }
}

/// `range/mod.rs:24-28` + `weighted_inner_product_proof.rs:25-33`, in the wire layout
/// `A | L[0..L) | R[0..L) | wipA | wipB | r' | s' | delta'` (include/bpp.h).
#[derive(Clone, PartialEq, Eq, Debug)]
pub struct RangeProof(pub Vec<u8>);

fn verdict(rc: i32) -> Result<(), ProofError> {
    match rc {
        BPP_OK => Ok(()),
        BPP_ERR_VERIFICATION => Err(ProofError::VerificationError), // errors.rs:17
        BPP_ERR_FORMAT => Err(ProofError::FormatError),             // errors.rs:18-20
        rc => panic!("bpp error {}", rc),
    }
}

impl RangeProof {
    /// `range/mod.rs:31-55` — the reference's constants for every challenge and nonce (parity mode; NOT zero-knowledge)
    pub fn prove(pk: &PublicKey, n: usize, prover: &RangeProver) -> RangeProof {
        let m = prover.v_vec.len();
        assert_eq!(pk.len(), n * m, "pk.G_vec.len() != n*m"); // range/mod.rs:90-91, :252-253
        let mut len = unsafe { bpp_range_proof_len(BPP_RISTRETTO255, n, m) };
        let mut buf = vec![0u8; len];
        let gam = flat(&prover.gamma_vec, |g| g.0);
        check(unsafe { bpp_range_prove(ctx(), pk.h, n, m, prover.v_vec.as_ptr(), gam.as_ptr(), ptr::null(), buf.as_mut_ptr(), &mut len) }, "prove");
        RangeProof(buf)
    }
    /// `range/mod.rs:57-78`
    pub fn verify(&self, pk: &PublicKey, n: usize, commitment_vec: &[Point]) -> Result<(), ProofError> {
        let v = flat(commitment_vec, |p| p.0);
        verdict(unsafe { bpp_range_verify(ctx(), pk.h, n, commitment_vec.len(), v.as_ptr(), self.0.as_ptr(), self.0.len(), ptr::null()) })
    }
    /// Batch of independent proofs, one verdict each (`bpp_range_verify_batch`).
    pub fn verify_batch(proofs: &[RangeProof], pk: &PublicKey, n: usize, m: usize, commitments: &[Point]) -> Vec<Result<(), ProofError>> {
        let (p, v) = (proofs.iter().flat_map(|x| x.0.clone()).collect::<Vec<u8>>(), flat(commitments, |c| c.0));
        let mut out = vec![0u8; proofs.len()];
        check(unsafe { bpp_range_verify_batch(ctx(), pk.h, n, m, proofs.len(), v.as_ptr(), p.as_ptr(), ptr::null(), out.as_mut_ptr()) }, "verify_batch");
        out.into_iter().map(|rc| verdict(rc as i32)).collect()
    }
}

/// `weighted_inner_product_proof.rs:25-33`: `L | R | A | B | r' | s' | delta'`
#[derive(Clone, PartialEq, Eq, Debug)]
pub struct WeightedInnerProductProof(pub Vec<u8>);
impl WeightedInnerProductProof {
    /// `weighted_inner_product_proof.rs:36-43`. The reference's `commitment` argument only feeds a dead update (`:137-142`).
    pub fn prove(pk: &PublicKey, a_vec: &[PrimeFieldElem], b_vec: &[PrimeFieldElem], power_of_y_vec: &[PrimeFieldElem],
                 gamma: PrimeFieldElem, _commitment: Point) -> Self {
        let n = pk.len();
        assert!(a_vec.len() == n && b_vec.len() == n && power_of_y_vec.len() == n); // :61-64
        assert!(n.is_power_of_two());                                               // :67
        let (a, b, y) = (flat(a_vec, |x| x.0), flat(b_vec, |x| x.0), flat(power_of_y_vec, |x| x.0));
        let mut len = unsafe { bpp_wip_proof_len(BPP_RISTRETTO255, n) };
        let mut buf = vec![0u8; len];
        check(unsafe { bpp_wip_prove(ctx(), pk.h, a.as_ptr(), b.as_ptr(), y.as_ptr(), gamma.0.as_ptr(), ptr::null(), buf.as_mut_ptr(), &mut len) }, "wip prove");
        WeightedInnerProductProof(buf)
    }
    /// `weighted_inner_product_proof.rs:238-248`
    #[allow(clippy::too_many_arguments)]
    pub fn verify(&self, pk: &PublicKey, power_of_y_vec: &[PrimeFieldElem], g_vec_exp_of_commitment: &[PrimeFieldElem],
                  h_vec_exp_of_commitment: &[PrimeFieldElem], g_exp_of_commitment: PrimeFieldElem, v_exp_of_commitment: &[PrimeFieldElem],
                  a_prime: Point, v: &[Point]) -> Result<(), ProofError> {
        let (y, ge, he) = (flat(power_of_y_vec, |x| x.0), flat(g_vec_exp_of_commitment, |x| x.0), flat(h_vec_exp_of_commitment, |x| x.0));
        let (ve, vp) = (flat(v_exp_of_commitment, |x| x.0), flat(v, |x| x.0));
        verdict(unsafe { bpp_wip_verify(ctx(), pk.h, y.as_ptr(), ge.as_ptr(), he.as_ptr(), g_exp_of_commitment.0.as_ptr(), v.len(),
                                        if v.is_empty() { ptr::null() } else { ve.as_ptr() }, a_prime.0.as_ptr(),
                                        if v.is_empty() { ptr::null() } else { vp.as_ptr() }, self.0.as_ptr(), self.0.len(), ptr::null()) })
    }
}

/// Host-owned Fiat-Shamir: protocol "rangeproof v1" (label set and order: `bulletproofsplus_b200/merlin.py`; compiled twins
/// `include/bpp_merlin.hpp` and that file are pinned to each other byte for byte by `tests/test_gpu_cpp_host.py`).
pub mod transcript {
    use super::*;
    use merlin::Transcript;

    fn challenge_scalar(t: &mut Transcript, label: &'static [u8]) -> PrimeFieldElem {
        loop {
            let mut wide = [0u8; 64];
            t.challenge_bytes(label, &mut wide);
            if let Some(e) = PrimeFieldElem::from_bytes_wide(&wide) { return e; } // a zero challenge is drawn again
        }
    }
    fn start(t: &mut Transcript, n: usize, m: usize, commitments: &[Point]) {
        t.append_message(b"dom-sep", b"rangeproof v1");
        t.append_u64(b"n", n as u64);
        t.append_u64(b"m", m as u64);
        for v in commitments { t.append_message(b"V", &v.0); }
    }
    fn after_a(t: &mut Transcript, a: &[u8], big_n: usize) -> (PrimeFieldElem, PrimeFieldElem) {
        t.append_message(b"A", a);
        let (y, z) = (challenge_scalar(t, b"y"), challenge_scalar(t, b"z"));
        t.append_message(b"dom-sep", b"wipp v1");
        t.append_u64(b"n", big_n as u64);
        (y, z)
    }
    fn after_pair(t: &mut Transcript, l0: &'static [u8], p0: &[u8], l1: &'static [u8], p1: &[u8]) -> PrimeFieldElem {
        t.append_message(l0, p0);
        t.append_message(l1, p1);
        challenge_scalar(t, b"e")
    }

    /// alpha, r, s, delta, eta, d_L[L], d_R[L]: fresh uniform blinding scalars from the OS CSPRNG — a proof is zero-knowledge
    /// only with these secret; the reference's constants (NULL at the ABI) are for parity tests.
    fn fresh_nonces(rounds: usize) -> Vec<u8> {
        use rand_core::{OsRng, RngCore};
        let mut out = Vec::with_capacity((5 + 2 * rounds) * 32);
        for _ in 0..5 + 2 * rounds {
            let e = loop {
                let mut wide = [0u8; 64];
                OsRng.fill_bytes(&mut wide);
                if let Some(e) = PrimeFieldElem::from_bytes_wide(&wide) { break e; }
            };
            out.extend_from_slice(&e.0);
        }
        out
    }

    impl RangeProof {
        /// README.md:39-46 of the reference: `RangeProof::prove(&pk, n, &prover, &mut transcript)`
        pub fn prove_with_transcript(pk: &PublicKey, n: usize, prover: &RangeProver, t: &mut Transcript) -> RangeProof {
            let m = prover.v_vec.len();
            assert_eq!(pk.len(), n * m, "pk.G_vec.len() != n*m");
            let rounds = (n * m).trailing_zeros() as usize;
            start(t, n, m, &prover.commitment_vec);
            let gam = flat(&prover.gamma_vec, |g| g.0);
            let nonces = fresh_nonces(rounds);
            let mut s: *mut bpp_prove_session = ptr::null_mut();
            let mut a = [0u8; 32];
            check(unsafe { bpp_range_prove_begin(ctx(), pk.h, n, m, 1, prover.v_vec.as_ptr(), gam.as_ptr(), nonces.as_ptr(), &mut s, a.as_mut_ptr()) }, "prove_begin");
            struct Guard(*mut bpp_prove_session);
            impl Drop for Guard { fn drop(&mut self) { unsafe { bpp_prove_session_free(self.0) } } }
            let _guard = Guard(s);
            let (y, z) = after_a(t, &a, n * m);
            let mut ch: Vec<u8> = [y.0, z.0].concat();
            let mut out = [0u8; 64];
            for k in 0..=rounds {
                check(unsafe { bpp_range_prove_step(ctx(), s, ch.as_ptr(), out.as_mut_ptr()) }, "prove_step");
                let e = if k < rounds { after_pair(t, b"L", &out[..32], b"R", &out[32..]) } else { after_pair(t, b"A", &out[..32], b"B", &out[32..]) };
                ch = e.0.to_vec();
            }
            let mut proof = vec![0u8; unsafe { bpp_range_proof_len(BPP_RISTRETTO255, n, m) }];
            check(unsafe { bpp_range_prove_finish(ctx(), s, ch.as_ptr(), proof.as_mut_ptr()) }, "prove_finish");
            RangeProof(proof)
        }

        /// README.md:47-52 of the reference: `proof.verify(&pk, n, &verifier.commitment_vec, &mut transcript)`
        pub fn verify_with_transcript(&self, pk: &PublicKey, n: usize, commitment_vec: &[Point], t: &mut Transcript) -> Result<(), ProofError> {
            let m = commitment_vec.len();
            let rounds = (n * m).trailing_zeros() as usize;
            if self.0.len() != (2 * rounds + 3) * 32 + 96 { return Err(ProofError::FormatError); }
            let pt = |i: usize| &self.0[32 * i..32 * i + 32];
            start(t, n, m, commitment_vec);
            let (y, z) = after_a(t, pt(0), n * m);
            let mut ch = vec![0u8; (3 + rounds) * 32];
            ch[..32].copy_from_slice(&y.0);
            ch[32..64].copy_from_slice(&z.0);
            for k in 0..rounds {
                let e = after_pair(t, b"L", pt(1 + k), b"R", pt(1 + rounds + k));
                ch[(3 + k) * 32..(4 + k) * 32].copy_from_slice(&e.0);
            }
            let e = after_pair(t, b"A", pt(1 + 2 * rounds), b"B", pt(2 + 2 * rounds));
            ch[64..96].copy_from_slice(&e.0);
            let v = flat(commitment_vec, |p| p.0);
            let mut out = [BPP_ERR_CUDA as u8];
            check(unsafe { bpp_range_verify_batch_fs(ctx(), pk.h, n, m, 1, v.as_ptr(), self.0.as_ptr(), ch.as_ptr(), out.as_mut_ptr()) }, "verify_batch_fs");
            verdict(out[0] as i32)
        }
    }
}

#[cfg(test)]
mod tests {
    use super::*;
    /// `src/main.rs:6-57` of the reference
    #[test]
    fn main_rs_flow() {
        Arith::init();
        let (n, m) = (64, 2);
        let pk = PublicKey::new(n * m);
        let mut prover = RangeProver::new();
        prover.commit(&pk, 2, PrimeFieldElem::new(3));
        prover.commit(&pk, 5, PrimeFieldElem::new(7));
        let proof = RangeProof::prove(&pk, n, &prover);
        assert_eq!(proof.verify(&pk, n, &prover.commitment_vec), Ok(()));
        let mut bad = proof.clone();
        let k = bad.0.len() - 96;
        bad.0[k] ^= 1;
        assert_eq!(bad.verify(&pk, n, &prover.commitment_vec), Err(ProofError::VerificationError));
    }
    #[test]
    fn transcript_flow() {
        let (n, m) = (32, 4);
        let pk = PublicKey::new(n * m);
        let mut prover = RangeProver::new();
        for j in 0..m as u64 { prover.commit_u64(&pk, 31 + j, PrimeFieldElem::new(11 + j as i32)); }
        let proof = RangeProof::prove_with_transcript(&pk, n, &prover, &mut merlin::Transcript::new(b"RangeProof Test"));
        assert_eq!(proof.verify_with_transcript(&pk, n, &prover.commitment_vec, &mut merlin::Transcript::new(b"RangeProof Test")), Ok(()));
        assert_eq!(proof.verify_with_transcript(&pk, n, &prover.commitment_vec, &mut merlin::Transcript::new(b"other")), Err(ProofError::VerificationError));
    }
}
