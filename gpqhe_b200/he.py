"""Host-side mirror of the reference interface for the RNS polynomial hot path.

Same names, argument meaning and error behaviour as the reference's C functions
(file:line cited per function, GPQHE commit 5e70953c), on numpy arrays instead of
libgcrypt MPIs.  Everything computes on the GPU through the C ABI
(include/gpqhe_b200.h); there is no CPU path here.

Representation
  limb-poly / poly_rns : np.uint64 [n] / [dim, n]
  big poly (poly_mpi)  : np.uint64 [W, n], word-major two's complement
  HeCt                 : l, nu, B, c0, c1   (struct he_ct, src/gpqhe.h:84-91)
  HeEvk                : device-resident key handle built from (p0, p1) uint64 [dimevk, n]
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

from . import capi
from .capi import GpqheError

LOGP = 59  # GPQHE_LOGP, src/params.h:27
SIGMA = 3.1915382432114616  # GPQHE_SIGMA, src/params.h:56
BLKSIZ = 64  # GPQHE_BLKSIZ, src/params.h:48


def int_to_words(x: int, nwords: int | None = None) -> np.ndarray:
    nwords = nwords or max(1, (x.bit_length() + 63) // 64)
    return np.frombuffer(x.to_bytes(8 * nwords, "little"), dtype="<u8").astype(np.uint64)


def ints_to_big(a, W: int) -> np.ndarray:
    """list of Python ints -> word-major two's complement [W, n]."""
    mask = (1 << (64 * W)) - 1
    buf = b"".join((int(x) & mask).to_bytes(8 * W, "little") for x in a)
    return np.ascontiguousarray(np.frombuffer(buf, dtype="<u8").reshape(len(a), W).T).astype(np.uint64)


def big_to_ints(w: np.ndarray) -> list[int]:
    W, n = w.shape
    raw = np.ascontiguousarray(w.T).astype("<u8").tobytes()
    half, full = 1 << (64 * W - 1), 1 << (64 * W)
    out = []
    for i in range(n):
        x = int.from_bytes(raw[i * 8 * W:(i + 1) * 8 * W], "little")
        out.append(x - full if x >= half else x)
    return out


def _ptr(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"], "need contiguous uint64"
    return a.ctypes.data


def _c(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint64)


@dataclass
class HeCt:
    """struct he_ct (src/gpqhe.h:84-91)."""
    l: int
    c0: np.ndarray
    c1: np.ndarray
    nu: float = 1.0
    B: float = 0.0

    def copy(self) -> "HeCt":   # he_copy_ct, src/he-mem.c:88-97
        return HeCt(self.l, self.c0.copy(), self.c1.copy(), self.nu, self.B)


@dataclass
class HePt:
    """struct he_pt (src/gpqhe.h:93-97)."""
    m: np.ndarray
    nu: float = 1.0


class HeEvk:
    """he_evk_t (src/gpqhe.h:78-82) uploaded once and kept on the device."""

    def __init__(self, ctx: "HeContext", p0: np.ndarray, p1: np.ndarray):
        p0, p1 = _c(p0).reshape(-1), _c(p1).reshape(-1)
        want = ctx.dimevk * ctx.n
        if p0.size != want or p1.size != want:
            raise GpqheError(f"evk must be 2 x [dimevk={ctx.dimevk}, n={ctx.n}] uint64")
        self._lib, self._h = ctx._lib, C.c_void_p()
        capi.check(self._lib, self._lib.gq_evk_create(ctx._h, C.byref(self._h), _ptr(p0), _ptr(p1)))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.gq_evk_destroy(self._h)
            self._h = C.c_void_p()


class HeContext:
    """hectx_init(logn, q, slots, Delta) (src/precomp.c:431-452); with Delta=None only
    polyctx_init(logn, q) (src/precomp.c:328-384).  Errors the reference reports by
    abort() are raised as GpqheError with the same message."""

    def __init__(self, logn: int, q: int, slots: int | None = None, Delta: int | None = None, *,
                 device: int = 0, logp: int = LOGP, logqub: int = 0, lib_path: str | None = None,
                 ref_tables=None):
        self._lib = capi.load(lib_path)
        self._h = C.c_void_p()
        qw = int_to_words(q)
        primes = zet = zin = None
        if ref_tables is not None:   # (primes[K], zetas[K][n], zetas_inv[K][n]) in Montgomery form
            pr, z, zi = ref_tables
            self._keep = (_c(pr), [_c(x) for x in z], [_c(x) for x in zi])
            K = len(self._keep[1])
            primes = self._keep[0].ctypes.data_as(capi.u64p)
            zet = (capi.u64p * K)(*[x.ctypes.data_as(capi.u64p) for x in self._keep[1]])
            zin = (capi.u64p * K)(*[x.ctypes.data_as(capi.u64p) for x in self._keep[2]])
        capi.check(self._lib, self._lib.gq_ctx_create(
            C.byref(self._h), logn, qw.ctypes.data_as(capi.u64p), qw.size, slots or 0, Delta or 0,
            logp, logqub, device, primes, zet, zin))
        info = capi.GqInfo()
        capi.check(self._lib, self._lib.gq_ctx_info(self._h, C.byref(info)))
        self.info = info
        self.logn, self.n, self.logp, self.dimub = info.logn, info.n, info.logp, info.dimub
        self.L, self.dim, self.dimevk, self.W = info.L, info.dim, info.dimevk, info.W
        self.q, self.slots, self.Delta = q, slots, Delta
        pr = np.zeros(self.dimub, dtype=np.uint64)
        capi.check(self._lib, self._lib.gq_ctx_primes(self._h, pr.ctypes.data_as(capi.u64p)))
        self.primes = [int(x) for x in pr]
        if Delta:
            self.ql = [0] * (self.L + 1)
            t = q
            for l in range(self.L, -1, -1):   # qtable_init, src/precomp.c:394-400
                self.ql[l] = t
                t //= Delta
            self._bounds_init()

    def close(self):
        if self._h.value:
            self._lib.gq_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- metadata helpers -------------------------------------------------------
    def _bounds_init(self):
        """bounds_init, src/precomp.c:411-429 (host metadata only; plain doubles here)."""
        n, h, s = self.n, BLKSIZ, SIGMA
        self.Bclean = 8 * math.sqrt(2) * s * n + 6 * s * math.sqrt(n) + 16 * s * math.sqrt(h * n)
        self.Brs = math.sqrt(n / 3.0) * (3 + 8 * math.sqrt(h))
        self.Bks = 8 * s * n / math.sqrt(3)
        pinv = 1.0
        for p in self.primes:
            pinv *= 1.0 / p
        x = pinv * float(self.ql[0])
        self.Bmult = []
        for l in range(self.L + 1):
            self.Bmult.append(x * self.Bks + self.Brs)
            x *= float(self.Delta)

    def dims(self, l: int) -> tuple[int, int]:
        """(tensor limbs, key-switch limbs) at level l: he-mult.c:99 and :51."""
        a, b = C.c_uint(), C.c_uint()
        capi.check(self._lib, self._lib.gq_ctx_dims(self._h, l, C.byref(a), C.byref(b)))
        return a.value, b.value

    def prime_consts(self, limb: int):
        out = np.zeros(4, dtype=np.uint64)
        capi.check(self._lib, self._lib.gq_ctx_prime_consts(self._h, limb, out.ctypes.data_as(capi.u64p)))
        return [int(x) for x in out]

    @property
    def launch_count(self) -> int:
        return int(self._lib.gq_ctx_launch_count(self._h))

    @property
    def crt_fixups(self) -> int:
        """coefficients the fast CRT kernels passed on to the exact full-width kernel"""
        return int(self._lib.gq_ctx_crt_fixups(self._h))

    def sync(self):
        capi.check(self._lib, self._lib.gq_ctx_sync(self._h))

    # ---- compute lanes (gq_ctx_set_lanes: independent operations overlap on one GPU) ----
    def set_lanes(self, lanes: int):
        capi.check(self._lib, self._lib.gq_ctx_set_lanes(self._h, lanes))

    @property
    def lanes(self) -> int:
        return int(self._lib.gq_ctx_lane_count(self._h))

    def lane(self, i: int) -> C.c_void_p:
        """handle of lane i for the *_dev entry points (lane 0 is the context itself)"""
        h = self._lib.gq_ctx_lane(self._h, i)
        if not h:
            raise GpqheError(f"lane {i} does not exist (lanes={self.lanes})")
        return C.c_void_p(h)

    # ---- limb layer ---------------------------------------------------------------
    def ntt(self, a: np.ndarray, limb: int) -> np.ndarray:
        """ntt(a, rns) (src/ntt.c:37-52); `limb` = index of the rns list node."""
        a = _c(a).copy()
        capi.check(self._lib, self._lib.gq_ntt(self._h, limb, _ptr(a)))
        return a

    def invntt(self, a: np.ndarray, limb: int) -> np.ndarray:
        """invntt(a, rns) (src/ntt.c:54-73)."""
        a = _c(a).copy()
        capi.check(self._lib, self._lib.gq_invntt(self._h, limb, _ptr(a)))
        return a

    def ntt_batch(self, data: np.ndarray, limb0: int = 0, inverse: bool = False) -> np.ndarray:
        """data[npolys, nlimbs, n]: every limb-poly transformed with its own prime."""
        d = _c(data).copy()
        npolys, nlimbs, n = d.shape
        capi.check(self._lib, self._lib.gq_ntt_batch(self._h, limb0, nlimbs, npolys, _ptr(d), int(inverse)))
        return d

    def poly_rns_add(self, a, b, limb: int) -> np.ndarray:
        """poly_rns_add (src/poly.c:71-76)."""
        a, b = _c(a), _c(b)
        r = np.empty_like(a)
        capi.check(self._lib, self._lib.gq_poly_rns_add(self._h, limb, _ptr(r), _ptr(a), _ptr(b)))
        return r

    def poly_rns_mul(self, a, b, limb: int) -> np.ndarray:
        """poly_rns_mul (src/poly.c:77-82)."""
        a, b = _c(a), _c(b)
        r = np.empty_like(a)
        capi.check(self._lib, self._lib.gq_poly_rns_mul(self._h, limb, _ptr(r), _ptr(a), _ptr(b)))
        return r

    # ---- MPI <-> RNS ----------------------------------------------------------------
    def rns_decompose(self, a: np.ndarray, limb0: int = 0, nlimbs: int = 1) -> np.ndarray:
        """rns_decompose (src/rns.c:37-48) for limbs limb0..limb0+nlimbs-1; a is [W, n]."""
        a = _c(a)
        out = np.empty((nlimbs, self.n), dtype=np.uint64)
        capi.check(self._lib, self._lib.gq_rns_decompose(self._h, limb0, nlimbs, _ptr(out), _ptr(a), a.shape[0]))
        return out

    def poly_rns2mpi(self, rhat: np.ndarray, dim: int, q: int, W: int | None = None) -> np.ndarray:
        """poly_rns2mpi(r, rhat, rns[dim-1], q) (src/poly.c:109-120)."""
        rhat = _c(rhat)
        W = W or max(1, (q.bit_length() + 63) // 64)
        out = np.empty((W, self.n), dtype=np.uint64)
        qw = int_to_words(q)
        capi.check(self._lib, self._lib.gq_poly_rns2mpi(self._h, dim, _ptr(out), W, _ptr(rhat),
                                                        qw.ctypes.data_as(capi.u64p), qw.size))
        return out

    def poly_mul(self, a: np.ndarray, b: np.ndarray, dim: int, q: int, W: int | None = None) -> np.ndarray:
        """poly_mul(r, a, b, dim, q) (src/poly.c:84-107)."""
        a, b = _c(a), _c(b)
        W = W or max(1, (q.bit_length() + 63) // 64)
        out = np.empty((W, self.n), dtype=np.uint64)
        qw = int_to_words(q)
        capi.check(self._lib, self._lib.gq_poly_mul(self._h, _ptr(out), W, _ptr(a), _ptr(b), a.shape[0], dim,
                                                    qw.ctypes.data_as(capi.u64p), qw.size))
        return out

    def poly_rot(self, a: np.ndarray, rot: int) -> np.ndarray:
        """poly_rot (src/poly.c:263-276)."""
        a = _c(a)
        out = np.empty_like(a)
        capi.check(self._lib, self._lib.gq_poly_rot(self._h, _ptr(out), _ptr(a), a.shape[0], rot))
        return out

    def poly_conj(self, a: np.ndarray) -> np.ndarray:
        """poly_conj (src/poly.c:278-283)."""
        a = _c(a)
        out = np.empty_like(a)
        capi.check(self._lib, self._lib.gq_poly_conj(self._h, _ptr(out), _ptr(a), a.shape[0]))
        return out

    # ---- HE primitives ----------------------------------------------------------------
    def genevk(self, p0: np.ndarray, p1: np.ndarray) -> HeEvk:
        return HeEvk(self, p0, p1)

    def he_mul(self, ct1: HeCt, ct2: HeCt, rlk: HeEvk) -> HeCt:
        """he_mul(ct, ct1, ct2, rlk) (src/he-mult.c:88-156)."""
        if ct1.l != ct2.l:
            raise GpqheError("he_mul: ct1->l != ct2->l")   # assert at he-mult.c:90
        l = ct1.l
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        capi.check(self._lib, self._lib.gq_he_mul(self._h, l, _ptr(c0), _ptr(c1), _ptr(_c(ct1.c0)),
                                                  _ptr(_c(ct1.c1)), _ptr(_c(ct2.c0)), _ptr(_c(ct2.c1)), rlk._h))
        nu = ct1.nu * ct2.nu                                  # he-mult.c:93-95
        B = ct1.nu * ct2.B + ct2.nu * ct1.B + ct1.B * ct2.B + self.Bmult[l]
        return HeCt(l, c0, c1, nu, B)

    def _linear(self, name, a: HeCt, b0, b1):
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        args = [self._h, a.l, _ptr(c0), _ptr(c1), _ptr(_c(a.c0)), _ptr(_c(a.c1))]
        args += [_ptr(_c(x)) for x in (b0, b1) if x is not None]
        capi.check(self._lib, getattr(self._lib, name)(*args))
        return c0, c1

    def he_add(self, ct1: HeCt, ct2: HeCt) -> HeCt:
        """he_add(ct, ct1, ct2) (src/he-add.c:32-54): nu = max, B = sum."""
        if ct1.l != ct2.l:
            raise GpqheError("he_add: ct1->l != ct2->l")
        c0, c1 = self._linear("gq_he_add", ct1, ct2.c0, ct2.c1)
        return HeCt(ct1.l, c0, c1, max(ct1.nu, ct2.nu), ct1.B + ct2.B)

    def he_sub(self, ct1: HeCt, ct2: HeCt) -> HeCt:
        """he_sub(ct, ct1, ct2) (src/he-add.c:57-79)."""
        if ct1.l != ct2.l:
            raise GpqheError("he_sub: ct1->l != ct2->l")
        c0, c1 = self._linear("gq_he_sub", ct1, ct2.c0, ct2.c1)
        return HeCt(ct1.l, c0, c1, max(ct1.nu, ct2.nu), ct1.B + ct2.B)

    def he_addpt(self, ct: HeCt, pt: HePt) -> HeCt:
        """he_addpt(dest, src, pt) (src/he-add.c:82-103)."""
        c0, c1 = self._linear("gq_he_addpt", ct, pt.m, None)
        return HeCt(ct.l, c0, c1, max(ct.nu, pt.nu), ct.B)

    def he_subpt(self, ct: HeCt, pt: HePt) -> HeCt:
        """he_subpt(dest, src, pt) (src/he-add.c:106-127)."""
        c0, c1 = self._linear("gq_he_subpt", ct, pt.m, None)
        return HeCt(ct.l, c0, c1, max(ct.nu, pt.nu), ct.B)

    def he_neg(self, ct: HeCt) -> HeCt:
        """he_neg(ct) (src/he-add.c:130-146); returns the negated ciphertext."""
        c0, c1 = _c(ct.c0).copy(), _c(ct.c1).copy()
        capi.check(self._lib, self._lib.gq_he_neg(self._h, ct.l, _ptr(c0), _ptr(c1)))
        return HeCt(ct.l, c0, c1, ct.nu, ct.B)

    def he_rot_hoisted(self, ct: HeCt, rots, rk) -> list:
        """[he_rot(copy of ct, r, rk) for r in rots] (src/he-automorphism.c:102-115) with the
        decomposition and forward NTT of c1 shared between the rotations (gq_he_rot_hoisted)."""
        rots = list(rots)
        k = len(rots)
        outs = [(np.empty((self.W, self.n), dtype=np.uint64), np.empty((self.W, self.n), dtype=np.uint64))
                for _ in range(k)]
        c0, c1 = _c(ct.c0), _c(ct.c1)
        arr = lambda xs: (C.c_void_p * k)(*xs)
        capi.check(self._lib, self._lib.gq_he_rot_hoisted(
            self._h, ct.l, k, (C.c_uint * k)(*rots), arr([o[0].ctypes.data for o in outs]),
            arr([o[1].ctypes.data for o in outs]), _ptr(c0), _ptr(c1), arr([rk[r]._h for r in rots])))
        return [HeCt(ct.l, o0, o1, ct.nu, ct.B) for o0, o1 in outs]

    def set_plaintext_bits(self, bits: int):
        """gq_ctx_set_plaintext_bits: |m| < 2^bits for the plaintexts of the next he_gemv-type calls (0 = unknown)."""
        capi.check(self._lib, self._lib.gq_ctx_set_plaintext_bits(self._h, int(bits)))

    @staticmethod
    def plaintext_bits(pts) -> int:
        """smallest b with |m| < 2^b over every coefficient of the word-major plaintexts"""
        best = 1
        for p in pts:
            m = np.asarray(p.m, dtype=np.uint64)
            neg = (m[-1] >> np.uint64(63)).astype(bool)
            mag = np.where(neg[None, :], ~m, m)          # |m| - 1 for negatives: same bit length bound
            for w in range(m.shape[0] - 1, -1, -1):
                top = int(mag[w].max())
                if top:
                    best = max(best, 64 * w + top.bit_length() + 1)
                    break
        return best

    def he_gemv_encoded(self, ct: HeCt, pts, n1: int, n2: int, rk) -> HeCt:
        """The ciphertext arithmetic of he_gemv(ct_dest, A, ct, rk) (src/he-algo.c:47-90) with the
        diagonals already encoded: pts[i*n1+j] is the HePt of zrotdiag(A, i*n1+j, -i*n1).
        Metadata follows the reference's sequence of he_mulpt / he_add / he_rs updates."""
        pts = list(pts)
        if len(pts) != n1 * n2:
            raise GpqheError("he_gemv: need n1*n2 encoded diagonals")
        dim = self.mulpt_dim(ct.l, pts[0].nu)
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        hold = [_c(p.m) for p in pts]
        nk = n1 * n2
        keys = (C.c_void_p * nk)(*[(rk[r]._h if (r < n1 or r % n1 == 0) else None) for r in range(nk)])
        capi.check(self._lib, self._lib.gq_he_gemv(
            self._h, ct.l, n1, n2, dim, _ptr(c0), _ptr(c1), _ptr(_c(ct.c0)), _ptr(_c(ct.c1)),
            (C.c_void_p * nk)(*[h.ctypes.data for h in hold]), keys))
        nu_o = B_o = 0.0
        for i in range(n2):
            nu_i = B_i = 0.0
            for j in range(n1):
                pt = pts[i * n1 + j]
                nu_t, B_t = ct.nu * pt.nu, ct.B * pt.nu                     # he-mult.c:162-164
                nu_i, B_i = (nu_t, B_t) if j == 0 else (max(nu_i, nu_t), B_i + B_t)   # he-add.c:36-38
            nu_o, B_o = (nu_i, B_i) if i == 0 else (max(nu_o, nu_i), B_o + B_i)
        return HeCt(ct.l - 1, c0, c1, nu_o / float(self.Delta), B_o / float(self.Delta) + self.Brs)

    def he_coeff2slot_encoded(self, ct: HeCt, pts4, cpt: HePt, n1: int, n2: int, ck: HeEvk, rk):
        """The ciphertext arithmetic of he_coeff2slot(ct_real, ct_imag, ct, ck, rk) (src/he-bootstrap.c:359-409)
        with the plaintexts already encoded: pts4 = four lists of n1*n2 HePt (diagonals of U0_conjT, U0_T,
        U1_conjT, U1_T as he_gemv_encoded wants them), cpt = the constant plaintext.  Returns (real, imag)."""
        m = n1 * n2
        flat = [p for part in pts4 for p in part]
        if len(flat) != 4 * m:
            raise GpqheError("he_coeff2slot: need 4 x n1*n2 encoded diagonals")
        l = ct.l
        dim_gemv = self.mulpt_dim(l, flat[0].nu)
        outs = [np.empty((self.W, self.n), dtype=np.uint64) for _ in range(4)]
        hold = [_c(p.m) for p in flat]
        keys = (C.c_void_p * m)(*[(rk[r]._h if (r < n1 or r % n1 == 0) else None) for r in range(m)])
        # metadata of the two he_gemv + he_add, before the constant multiply (same for both parts)
        g = HeCt(l, ct.c0, ct.c1, ct.nu, ct.B)
        nu_o = B_o = 0.0
        for i in range(n2):
            nu_i = B_i = 0.0
            for j in range(n1):
                pt = flat[i * n1 + j]
                nu_t, B_t = g.nu * pt.nu, g.B * pt.nu
                nu_i, B_i = (nu_t, B_t) if j == 0 else (max(nu_i, nu_t), B_i + B_t)
            nu_o, B_o = (nu_i, B_i) if i == 0 else (max(nu_o, nu_i), B_o + B_i)
        nu_g, B_g = nu_o / float(self.Delta), B_o / float(self.Delta) + self.Brs
        nu_a, B_a = nu_g, 2 * B_g                                              # he-add.c:36-38
        dim_pt = self.mulpt_dim(l - 1, cpt.nu)
        capi.check(self._lib, self._lib.gq_he_coeff2slot(
            self._h, l, n1, n2, dim_gemv, dim_pt, _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2]), _ptr(outs[3]),
            _ptr(_c(ct.c0)), _ptr(_c(ct.c1)), (C.c_void_p * (4 * m))(*[h.ctypes.data for h in hold]),
            _ptr(_c(cpt.m)), ck._h, keys))
        nu_f = nu_a * cpt.nu / float(self.Delta)                               # he-mult.c:162-164, he-rescale.c:36-38
        B_f = B_a * cpt.nu / float(self.Delta) + self.Brs
        return (HeCt(l - 2, outs[0], outs[1], nu_f, B_f), HeCt(l - 2, outs[2], outs[3], nu_f, B_f))

    # ---- metadata replay: the (nu, B) updates of the reference, one per primitive ---------------------
    def _m_mul(self, l, a, b):      # he-mult.c:93-95
        return (a[0] * b[0], a[0] * b[1] + b[0] * a[1] + a[1] * b[1] + self.Bmult[l])

    def _m_mul_d2(self, l, a, b):   # he_mul(&x, &y, &x): B reads the nu just written (he-mult.c:94-95)
        nu = a[0] * b[0]
        return (nu, a[0] * b[1] + nu * a[1] + a[1] * b[1] + self.Bmult[l])

    def _m_sqr_inplace(self, l, a):  # he_mul(&x, &x, &x)
        nu = a[0] * a[0]
        return (nu, nu * a[1] + nu * a[1] + a[1] * a[1] + self.Bmult[l])

    def _m_rs(self, a):             # he-rescale.c:36-38
        return (a[0] / float(self.Delta), a[1] / float(self.Delta) + self.Brs)

    @staticmethod
    def _m_mulpt(a, ptnu):          # he-mult.c:162-164
        return (a[0] * ptnu, a[1] * ptnu)

    @staticmethod
    def _m_add(a, b):               # he-add.c:36-38
        return (max(a[0], b[0]), a[1] + b[1])

    @staticmethod
    def _m_addpt(a, ptnu):          # he-add.c:85-87
        return (max(a[0], ptnu), a[1])

    def _exp_meta(self, l, meta, ptnu, iters):
        """(nu, B) of he_exp (src/he-algo.c:364-458) by replaying its primitives' updates in order, including
        the aliased he_mul calls whose B statement reads the nu it has just overwritten."""
        act = self._m_rs(self._m_mulpt(meta, ptnu))
        k = l - 1
        ct2 = self._m_rs(self._m_mul(k, act, act))
        ct4 = self._m_rs(self._m_mul(k - 1, ct2, ct2))
        lin = lambda: self._m_rs(self._m_mulpt(self._m_addpt(act, ptnu), ptnu))
        ct01 = lin()
        ct23 = self._m_rs(self._m_mul_d2(k - 1, ct2, lin()))
        ct0123 = self._m_add(ct01, ct23)
        ct45 = lin()
        ct67 = self._m_rs(self._m_mul_d2(k - 1, ct2, lin()))
        ct4567 = self._m_rs(self._m_mul_d2(k - 2, ct4, self._m_add(ct45, ct67)))
        out = self._m_add(ct0123, ct4567)
        for t in range(iters):
            out = self._m_rs(self._m_sqr_inplace(k - 3 - t, out))
        return out

    def he_exp_encoded(self, ct: HeCt, pts, iters: int, rlk: HeEvk) -> HeCt:
        """The ciphertext arithmetic of he_exp(ct_exp, a, ct, rlk, iter) (src/he-algo.c:435-458 over
        he_exp_taylor, :364-433) with the plaintexts already encoded: pts[0] = he_ecd of a/2^iter in every
        slot, pts[1..7] = he_const_pt of 1, 3, 1/6, 5, 1/120, 7, 1/5040.  One crossing (gq_he_exp)."""
        pts = list(pts)
        if len(pts) != 8:
            raise GpqheError("he_exp: need 8 encoded plaintexts")
        l = ct.l
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        hold = [_c(p.m) for p in pts]
        capi.check(self._lib, self._lib.gq_he_exp(
            self._h, l, iters, self.mulpt_dim(l, pts[0].nu), self.mulpt_dim(l - 1, pts[1].nu), _ptr(c0), _ptr(c1),
            _ptr(_c(ct.c0)), _ptr(_c(ct.c1)), (C.c_void_p * 8)(*[h.ctypes.data for h in hold]), rlk._h))
        nu, B = self._exp_meta(l, (ct.nu, ct.B), pts[0].nu, iters)
        return HeCt(l - 4 - iters, c0, c1, nu, B)

    def he_rlsin_encoded(self, ct: HeCt, pts, iters: int, rlk: HeEvk, ck: HeEvk) -> HeCt:
        """The ciphertext arithmetic of he_rlsin(ct_dest, a, ct, rlk, ck, iter) (src/he-bootstrap.c:444-475):
        pts[0..7] as he_exp_encoded (pts[0] encodes i*a/2^iter), pts[8] = he_ecd of 1/(2*i*a)."""
        pts = list(pts)
        if len(pts) != 9:
            raise GpqheError("he_rlsin: need 9 encoded plaintexts")
        l = ct.l
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        hold = [_c(p.m) for p in pts]
        m = l - 4 - iters
        capi.check(self._lib, self._lib.gq_he_rlsin(
            self._h, l, iters, self.mulpt_dim(l, pts[0].nu), self.mulpt_dim(l - 1, pts[1].nu),
            self.mulpt_dim(max(m, 0), pts[8].nu), _ptr(c0), _ptr(c1), _ptr(_c(ct.c0)), _ptr(_c(ct.c1)),
            (C.c_void_p * 9)(*[h.ctypes.data for h in hold]), rlk._h, ck._h))
        e = self._exp_meta(l, (ct.nu, ct.B), pts[0].nu, iters)
        nu, B = self._m_rs(self._m_mulpt(self._m_add(e, e), pts[8].nu))      # he_conj keeps nu and B; he_sub as he_add
        return HeCt(m - 1, c0, c1, nu, B)

    def he_slot2coeff_encoded(self, ct0: HeCt, ct1: HeCt, pts2, n1: int, n2: int, rk) -> HeCt:
        """The ciphertext arithmetic of he_slot2coeff(ct, ct0, ct1, rk) (src/he-bootstrap.c:414-438):
        he_gemv(U0, ct0) + he_gemv(U1, ct1) with pts2 = the encoded diagonals of U0 and of U1."""
        m = n1 * n2
        flat = [p for part in pts2 for p in part]
        if len(flat) != 2 * m or ct0.l != ct1.l:
            raise GpqheError("he_slot2coeff: need 2 x n1*n2 encoded diagonals and equal levels")
        l = ct0.l
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        hold = [_c(p.m) for p in flat]
        keys = (C.c_void_p * m)(*[(rk[r]._h if (r < n1 or r % n1 == 0) else None) for r in range(m)])
        capi.check(self._lib, self._lib.gq_he_slot2coeff(
            self._h, l, n1, n2, self.mulpt_dim(l, flat[0].nu), _ptr(c0), _ptr(c1), _ptr(_c(ct0.c0)), _ptr(_c(ct0.c1)),
            _ptr(_c(ct1.c0)), _ptr(_c(ct1.c1)), (C.c_void_p * (2 * m))(*[h.ctypes.data for h in hold]), keys))
        metas = []
        for part, src in enumerate((ct0, ct1)):
            nu_o = B_o = 0.0
            for i in range(n2):
                nu_i = B_i = 0.0
                for j in range(n1):
                    pt = flat[part * m + i * n1 + j]
                    nu_t, B_t = src.nu * pt.nu, src.B * pt.nu
                    nu_i, B_i = (nu_t, B_t) if j == 0 else (max(nu_i, nu_t), B_i + B_t)
                nu_o, B_o = (nu_i, B_i) if i == 0 else (max(nu_o, nu_i), B_o + B_i)
            metas.append(self._m_rs((nu_o, B_o)))
        nu, B = self._m_add(*metas)
        return HeCt(l - 1, c0, c1, nu, B)

    def he_mul_rs(self, ct1: HeCt, ct2: HeCt, rlk: HeEvk) -> HeCt:
        """he_mul then he_rs on the product, one crossing (gq_he_mul_rs)."""
        return self.he_mul_batch([(ct1, ct2)], rlk, rescale=True, fused_single=True)[0]

    def he_mul_batch(self, pairs, rlk: HeEvk, rescale: bool = True, fused_single: bool = False):
        """[he_mul(a, b) (+ he_rs) for a, b in pairs] through the pipelined batch entry point
        (gq_he_mul_batch); metadata follows he-mult.c:93-95 and he-rescale.c:40-44."""
        if not pairs:
            return []
        l = pairs[0][0].l
        for a, b in pairs:
            if a.l != l or b.l != l:
                raise GpqheError("he_mul_batch: every ciphertext must be at the same level")
        k = len(pairs)
        hold = [[_c(a.c0), _c(a.c1), _c(b.c0), _c(b.c1)] for a, b in pairs]
        outs = [(np.empty((self.W, self.n), dtype=np.uint64), np.empty((self.W, self.n), dtype=np.uint64))
                for _ in range(k)]
        if fused_single and k == 1:
            h = hold[0]
            capi.check(self._lib, self._lib.gq_he_mul_rs(self._h, l, _ptr(outs[0][0]), _ptr(outs[0][1]),
                                                         _ptr(h[0]), _ptr(h[1]), _ptr(h[2]), _ptr(h[3]), rlk._h))
        else:
            arr = lambda xs: (C.c_void_p * k)(*[x.ctypes.data for x in xs])
            capi.check(self._lib, self._lib.gq_he_mul_batch(
                self._h, l, k, arr([o[0] for o in outs]), arr([o[1] for o in outs]),
                arr([h[0] for h in hold]), arr([h[1] for h in hold]), arr([h[2] for h in hold]),
                arr([h[3] for h in hold]), rlk._h, 1 if rescale else 0))
        res = []
        for (a, b), (c0, c1) in zip(pairs, outs):
            nu = a.nu * b.nu
            B = a.nu * b.B + b.nu * a.B + a.B * b.B + self.Bmult[l]
            if rescale:
                res.append(HeCt(l - 1, c0, c1, nu / float(self.Delta), B / float(self.Delta) + self.Brs))
            else:
                res.append(HeCt(l, c0, c1, nu, B))
        return res

    def mulpt_dim(self, l: int, nu: float) -> int:
        """limb count of he_mulpt, double arithmetic truncated (src/he-mult.c:168)."""
        return int((self.ql[l].bit_length() + math.log2(nu) + self.logn) / self.logp + 1)

    def he_mulpt(self, ct: HeCt, pt: HePt) -> HeCt:
        """he_mulpt(dest, src, pt) (src/he-mult.c:159-196)."""
        dim = self.mulpt_dim(ct.l, pt.nu)
        c0 = np.empty((self.W, self.n), dtype=np.uint64)
        c1 = np.empty_like(c0)
        capi.check(self._lib, self._lib.gq_he_mulpt(self._h, ct.l, dim, _ptr(c0), _ptr(c1), _ptr(_c(ct.c0)),
                                                    _ptr(_c(ct.c1)), _ptr(_c(pt.m))))
        return HeCt(ct.l, c0, c1, ct.nu * pt.nu, ct.B * pt.nu)

    def he_rs(self, ct: HeCt) -> HeCt:
        """he_rs(ct) (src/he-rescale.c:33-54); returns the rescaled ciphertext."""
        c0, c1 = _c(ct.c0).copy(), _c(ct.c1).copy()
        capi.check(self._lib, self._lib.gq_he_rs(self._h, ct.l, _ptr(c0), _ptr(c1)))
        return HeCt(ct.l - 1, c0, c1, ct.nu / float(self.Delta), ct.B / float(self.Delta) + self.Brs)

    def he_moddown(self, ct: HeCt) -> HeCt:
        """he_moddown(ct) (src/he-rescale.c:56-70)."""
        c0, c1 = _c(ct.c0).copy(), _c(ct.c1).copy()
        capi.check(self._lib, self._lib.gq_he_moddown(self._h, ct.l, _ptr(c0), _ptr(c1)))
        return HeCt(ct.l - 1, c0, c1, ct.nu, ct.B)

    def he_rot(self, ct: HeCt, rot: int, rk) -> HeCt:
        """he_rot(ct, rot, rk) (src/he-automorphism.c:102-115); `rk` is the key array
        indexed by rot (a dict or list of HeEvk)."""
        c0, c1 = _c(ct.c0).copy(), _c(ct.c1).copy()
        capi.check(self._lib, self._lib.gq_he_rot(self._h, ct.l, rot, _ptr(c0), _ptr(c1), rk[rot]._h))
        return HeCt(ct.l, c0, c1, ct.nu, ct.B)

    def he_conj(self, ct: HeCt, ck: HeEvk) -> HeCt:
        """he_conj(ct, ck) (src/he-automorphism.c:87-100)."""
        c0, c1 = _c(ct.c0).copy(), _c(ct.c1).copy()
        capi.check(self._lib, self._lib.gq_he_conj(self._h, ct.l, _ptr(c0), _ptr(c1), ck._h))
        return HeCt(ct.l, c0, c1, ct.nu, ct.B)


def hectx_init(logn: int, q: int, slots: int, Delta: int, **kw) -> HeContext:
    """hectx_init (src/precomp.c:431)."""
    return HeContext(logn, q, slots, Delta, **kw)


def polyctx_init(logn: int, q: int, **kw) -> HeContext:
    """polyctx_init (src/precomp.c:328)."""
    return HeContext(logn, q, None, None, **kw)
