"""CPU restatement of GPQHE's RNS polynomial hot path (big-integer side).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, as the checker.  The product
(gpqhe_b200/, include/, dropin/) never imports it.

Big integers are Python ints (exact, independent of libgcrypt and of the device
multiword code); the 64-bit limb layer is oracle.c through ctypes
(oracle/liboracle.so, `make -C oracle port`).  Every function cites the reference
file:line (commit 5e70953c) it restates.

Parity status: PINNED.  tests/test_oracle.py compares this module with the
unmodified reference built in oracle/_ref (when present) and with the committed
fixtures in tests/golden/ that tests/golden/make_golden.py generated from that
reference build.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LOGP = 59   # src/params.h:27  GPQHE_LOGP
ROT = 5     # src/params.h:54  GPQHE_ROT
_u64p = ctypes.POINTER(ctypes.c_uint64)


def _load():
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.c")
    if (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    lib = ctypes.CDLL(so)
    u64 = ctypes.c_uint64
    lib.orc_montgomery_inv.restype = u64
    lib.orc_montgomery_inv.argtypes = [u64]
    lib.orc_barrett_inv.restype = u64
    lib.orc_barrett_inv.argtypes = [u64]
    lib.orc_barrett_reduce128.restype = u64
    lib.orc_barrett_reduce128.argtypes = [u64, u64, u64, u64, ctypes.POINTER(ctypes.c_int)]
    lib.orc_ntt.restype = None
    lib.orc_ntt.argtypes = [_u64p, ctypes.c_uint, u64, u64, _u64p]
    lib.orc_invntt.restype = None
    lib.orc_invntt.argtypes = [_u64p, ctypes.c_uint, u64, u64, u64, _u64p]
    for f in (lib.orc_poly_rns_add, lib.orc_poly_rns_mul):
        f.restype = None
        f.argtypes = [_u64p, _u64p, _u64p, ctypes.c_uint, u64, u64]
    lib.orc_next_prime.restype = u64
    lib.orc_next_prime.argtypes = [u64, ctypes.c_uint]
    lib.orc_root_of_unity.restype = u64
    lib.orc_root_of_unity.argtypes = [u64, u64]
    lib.orc_ntt_tables.restype = None
    lib.orc_ntt_tables.argtypes = [u64, ctypes.c_uint, _u64p, _u64p, _u64p]
    lib.orc_powm.restype = u64
    lib.orc_powm.argtypes = [u64, u64, u64]
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u64p)


# --------------------------------------------------------------------------- #
# scalar big-integer helpers (src/types.c)
# --------------------------------------------------------------------------- #
def mpi_smod(r: int, q: int, qh: int) -> int:
    """src/types.c:108-113: r mod q in [0,q), minus q iff >= qh."""
    r %= q
    return r - q if r >= qh else r


def mpi_rdiv(a: int, m: int) -> int:
    """src/types.c:115-128: floor division, +1 iff remainder > floor(m/2)."""
    assert m > 0
    q, r = divmod(a, m)          # Python divmod == libgcrypt floor division (round=-1)
    return q + 1 if r > m // 2 else q


def nbits(a: int) -> int:
    """gcry_mpi_get_nbits: bit length of |a| (0 for 0)."""
    return abs(a).bit_length()


# --------------------------------------------------------------------------- #
# context (src/precomp.c)
# --------------------------------------------------------------------------- #
class RnsNode:
    """One node of the reference's linked list, struct rns_ctx (src/poly.h:28-41)."""
    __slots__ = ("dim", "p", "pinv_mont", "pinv_barr", "ninv", "zetas", "zetas_inv",
                 "P", "P_2", "phat", "phat_invmp")


class Ctx:
    """polyctx_init (src/precomp.c:328-384) + hectx_init/qtable_init (:386-452).

    `logp` is GPQHE_LOGP (59; 9 under -DTEST_CRT, src/params.h:30-33)."""

    def __init__(self, logn: int, q: int, slots: int | None = None, Delta: int | None = None,
                 logp: int = LOGP, logqub: int | None = None):
        L = lib()
        self.logn, self.n, self.m = logn, 1 << logn, 2 << logn
        self.logp = logp
        self.q = q
        self.logq = nbits(q) - 1
        if logqub is None:
            std = {10: 27, 11: 54, 12: 109, 13: 218, 14: 438, 15: 881}  # precomp.c:54-61
            logqub = std.get(logn, self.logq)
            if logn < 10 or logn > 15:
                logqub = self.logq
        if self.logq > logqub:
            raise ValueError("The input modulus q is too large.")       # precomp.c:344-350
        self.logqub = logqub
        self.dimub = (1 + logn + 4 * logqub) // logp + 1                 # precomp.c:357
        self.rns: list[RnsNode] = []
        p = (1 << logp) + 1
        P = 1
        for dim in range(1, self.dimub + 1):
            p = int(L.orc_next_prime(p, self.n))
            nd = RnsNode()
            nd.dim, nd.p = dim, p
            nd.pinv_mont = int(L.orc_montgomery_inv(p))
            nd.pinv_barr = int(L.orc_barrett_inv(p))
            nd.zetas = np.empty(self.n, dtype=np.uint64)
            nd.zetas_inv = np.empty(self.n, dtype=np.uint64)
            ninv = ctypes.c_uint64()
            L.orc_ntt_tables(p, logn, _p(nd.zetas), _p(nd.zetas_inv), ctypes.byref(ninv))
            nd.ninv = int(ninv.value)
            P *= p                                                       # precomp.c:276-279
            nd.P, nd.P_2 = P, P // 2
            primes = [x.p for x in self.rns] + [p]
            nd.phat = [P // pd for pd in primes]                          # precomp.c:287
            nd.phat_invmp = [pow((P // pd) % pd, pd - 2, pd) for pd in primes]  # :289
            self.rns.append(nd)
        self.primes = [x.p for x in self.rns]
        if Delta is not None:
            self._he_init(slots, Delta)

    def _he_init(self, slots, Delta):
        if slots & (slots - 1):
            raise ValueError("The slots must be the power of 2.")
        if slots > self.n // 2:
            raise ValueError("Must guarantee slots<=(n/2).")
        self.slots, self.Delta = slots, Delta
        logq = nbits(self.q) - 1
        logDelta = nbits(Delta) - 1
        self.L = logq // logDelta                                        # precomp.c:391 (integer division)
        self.ql = [0] * (self.L + 1)
        self.qh = [0] * (self.L + 1)
        q = self.q
        for l in range(self.L, -1, -1):                                  # precomp.c:394-400
            self.ql[l], self.qh[l] = q, q // 2
            q //= Delta
        self.dim = (nbits(self.ql[self.L]) + self.logn) // self.logp + 1  # :401
        self.P = self.rns[self.dim - 1].P                                # :402-404
        self.PqL = self.P * self.q
        self.dimevk = (nbits(self.ql[self.L]) + nbits(self.PqL) + self.logn) // self.logp + 1  # :407

    # limb counts used by the HE ops
    def dim_tensor(self, l):   # src/he-mult.c:99
        return (nbits(self.ql[l]) * 2 + self.logn) // self.logp + 1

    def dim_ks(self, l):       # src/he-mult.c:51, src/he-automorphism.c:52
        return (nbits(self.ql[l]) + nbits(self.PqL) + self.logn) // self.logp + 1


# --------------------------------------------------------------------------- #
# limb layer (src/ntt.c, src/poly.c:71-82) - thin wrappers over oracle.c
# --------------------------------------------------------------------------- #
def ntt(a: np.ndarray, nd: RnsNode) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_ntt(_p(a), a.size, nd.p, nd.pinv_mont, _p(nd.zetas))
    return a


def invntt(a: np.ndarray, nd: RnsNode) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_invntt(_p(a), a.size, nd.p, nd.pinv_mont, nd.ninv, _p(nd.zetas_inv))
    return a


def poly_rns_mul(a, b, nd: RnsNode) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = np.ascontiguousarray(b, dtype=np.uint64)
    r = np.empty_like(a)
    lib().orc_poly_rns_mul(_p(r), _p(a), _p(b), a.size, nd.p, nd.pinv_barr)
    return r


def poly_rns_add(a, b, nd: RnsNode) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = np.ascontiguousarray(b, dtype=np.uint64)
    r = np.empty_like(a)
    lib().orc_poly_rns_add(_p(r), _p(a), _p(b), a.size, nd.p, nd.pinv_barr)
    return r


# --------------------------------------------------------------------------- #
# MPI <-> RNS (src/rns.c) and polynomial ops (src/poly.c)
# --------------------------------------------------------------------------- #
def rns_decompose(a: list[int], nd: RnsNode) -> np.ndarray:
    """src/rns.c:37-48: ahat[i] = a[i] mod p, non-negative for negative a[i]."""
    p = nd.p
    return np.array([x % p for x in a], dtype=np.uint64)


def rns_reconstruct(ahat: np.ndarray, i: int, nd: RnsNode, n: int) -> int:
    """src/rns.c:60-75: sum_d ahat[d*n+i] * ((P/p_d * phat_invmp_d) mod P) mod P.

    `nd` is the node whose `dim` defines the basis {p_0..p_{dim-1}}."""
    a = 0
    for d in range(nd.dim):
        c = (nd.phat[d] * nd.phat_invmp[d]) % nd.P
        b = (int(ahat[d * n + i]) * c) % nd.P
        a = (a + b) % nd.P
    return a


def poly_rns2mpi(rhat: np.ndarray, nd: RnsNode, q: int, n: int) -> list[int]:
    """src/poly.c:109-120: CRT, centre mod P, centre mod q.  rhat is limb-major."""
    qh = q // 2
    C = [(nd.phat[d] * nd.phat_invmp[d]) % nd.P for d in range(nd.dim)]
    cols = [rhat[d * n:(d + 1) * n].tolist() for d in range(nd.dim)]
    P, P_2 = nd.P, nd.P_2
    out = []
    for i in range(n):
        a = 0
        for d in range(nd.dim):
            a += cols[d][i] * C[d]
        a %= P
        a = mpi_smod(a, P, P_2)
        out.append(mpi_smod(a, q, qh))
    return out


def poly_mul(ctx: Ctx, a: list[int], b: list[int], dim: int, q: int) -> list[int]:
    """src/poly.c:84-107."""
    n = ctx.n
    rhat = np.empty(dim * n, dtype=np.uint64)
    for d in range(dim):
        nd = ctx.rns[d]
        ah = ntt(rns_decompose(a, nd), nd)
        bh = ntt(rns_decompose(b, nd), nd)
        rhat[d * n:(d + 1) * n] = invntt(poly_rns_mul(ah, bh, nd), nd)
    return poly_rns2mpi(rhat, ctx.rns[dim - 1], q, n)


def poly_rot(a: list[int], rot: int, n: int) -> list[int]:
    """src/poly.c:263-276 (power wraps mod 2^64 as size_t; harmless mod 2n)."""
    power = pow(ROT, rot, 1 << 64)
    m = 2 * n
    r = [0] * n
    for i in range(n):
        k = ((i * power) % (1 << 64)) % m
        if k < n:
            r[k] = a[i]
        else:
            r[k - n] = -a[i]
    return r


def poly_conj(a: list[int], n: int) -> list[int]:
    """src/poly.c:278-283."""
    return [a[0]] + [-a[n - i] for i in range(1, n)]


# --------------------------------------------------------------------------- #
# HE primitives (src/he-mult.c, src/he-rescale.c, src/he-automorphism.c)
# A ciphertext is a dict {l, c0, c1}; an evk is (p0, p1) limb-major uint64 arrays
# over dimevk limbs, already in ntt() output order (src/he-kem.c:103-110).
# --------------------------------------------------------------------------- #
def _keyswitch(ctx: Ctx, l: int, d: list[int], evk) -> tuple[list[int], list[int]]:
    """Common body of he_relin (src/he-mult.c:40-77) and he_swk
    (src/he-automorphism.c:40-76): round(d*evk.p{0,1} / P) before the additive term."""
    n = ctx.n
    ql = ctx.ql[l]
    Pql = ctx.P * ql
    dim = ctx.dim_ks(l)
    c0hat = np.empty(dim * n, dtype=np.uint64)
    c1hat = np.empty(dim * n, dtype=np.uint64)
    for k in range(dim):
        nd = ctx.rns[k]
        dh = ntt(rns_decompose(d, nd), nd)
        c0hat[k * n:(k + 1) * n] = invntt(poly_rns_mul(dh, evk[0][k * n:(k + 1) * n], nd), nd)
        c1hat[k * n:(k + 1) * n] = invntt(poly_rns_mul(dh, evk[1][k * n:(k + 1) * n], nd), nd)
    nd = ctx.rns[dim - 1]
    c0 = poly_rns2mpi(c0hat, nd, Pql, n)
    c1 = poly_rns2mpi(c1hat, nd, Pql, n)
    return [mpi_rdiv(x, ctx.P) for x in c0], [mpi_rdiv(x, ctx.P) for x in c1]


def he_relin(ctx: Ctx, l: int, d0, d1, d2, rlk):
    """src/he-mult.c:40-85."""
    ql, qlh = ctx.ql[l], ctx.qh[l]
    k0, k1 = _keyswitch(ctx, l, d2, rlk)
    c0 = [mpi_smod((x + y) % ql, ql, qlh) for x, y in zip(k0, d0)]
    c1 = [mpi_smod((x + y) % ql, ql, qlh) for x, y in zip(k1, d1)]
    return c0, c1


def he_mul_tensor(ctx: Ctx, ct1: dict, ct2: dict):
    """src/he-mult.c:88-141: the three tensor-product polynomials d0,d1,d2 mod q_l."""
    assert ct1["l"] == ct2["l"]
    l = ct1["l"]
    q, n = ctx.ql[l], ctx.n
    dim = ctx.dim_tensor(l)
    d0h = np.empty(dim * n, dtype=np.uint64)
    d1h = np.empty(dim * n, dtype=np.uint64)
    d2h = np.empty(dim * n, dtype=np.uint64)
    for k in range(dim):
        nd = ctx.rns[k]
        a0 = ntt(rns_decompose(ct1["c0"], nd), nd)
        a1 = ntt(rns_decompose(ct1["c1"], nd), nd)
        b0 = ntt(rns_decompose(ct2["c0"], nd), nd)
        b1 = ntt(rns_decompose(ct2["c1"], nd), nd)
        s = slice(k * n, (k + 1) * n)
        d0h[s] = invntt(poly_rns_mul(a0, b0, nd), nd)
        d2h[s] = invntt(poly_rns_mul(a1, b1, nd), nd)
        x = invntt(poly_rns_mul(a0, b1, nd), nd)
        y = invntt(poly_rns_mul(a1, b0, nd), nd)
        d1h[s] = poly_rns_add(x, y, nd)          # added AFTER the inverse NTTs (:136)
    nd = ctx.rns[dim - 1]
    return (poly_rns2mpi(d0h, nd, q, n), poly_rns2mpi(d1h, nd, q, n),
            poly_rns2mpi(d2h, nd, q, n))


def he_mul(ctx: Ctx, ct1: dict, ct2: dict, rlk) -> dict:
    """src/he-mult.c:88-156."""
    d0, d1, d2 = he_mul_tensor(ctx, ct1, ct2)
    c0, c1 = he_relin(ctx, ct1["l"], d0, d1, d2, rlk)
    return {"l": ct1["l"], "c0": c0, "c1": c1}


def he_mulpt(ctx: Ctx, ct: dict, m: list[int], dim: int) -> dict:
    """src/he-mult.c:159-196.  `dim` is computed by the caller from pt->nu in double
    arithmetic (:168); it is host metadata, passed in explicitly here."""
    l, n = ct["l"], ctx.n
    q = ctx.ql[l]
    c0h = np.empty(dim * n, dtype=np.uint64)
    c1h = np.empty(dim * n, dtype=np.uint64)
    for k in range(dim):
        nd = ctx.rns[k]
        ph = ntt(rns_decompose(m, nd), nd)
        s = slice(k * n, (k + 1) * n)
        c0h[s] = invntt(poly_rns_mul(ntt(rns_decompose(ct["c0"], nd), nd), ph, nd), nd)
        c1h[s] = invntt(poly_rns_mul(ntt(rns_decompose(ct["c1"], nd), nd), ph, nd), nd)
    nd = ctx.rns[dim - 1]
    return {"l": l, "c0": poly_rns2mpi(c0h, nd, q, n), "c1": poly_rns2mpi(c1h, nd, q, n)}


def he_rs(ctx: Ctx, ct: dict) -> dict:
    """src/he-rescale.c:33-54."""
    l = ct["l"] - 1
    q, qh, p = ctx.ql[l], ctx.qh[l], ctx.Delta
    return {"l": l,
            "c0": [mpi_smod(mpi_rdiv(x, p), q, qh) for x in ct["c0"]],
            "c1": [mpi_smod(mpi_rdiv(x, p), q, qh) for x in ct["c1"]]}


def he_moddown(ctx: Ctx, ct: dict) -> dict:
    """src/he-rescale.c:56-70."""
    l = ct["l"] - 1
    q, qh = ctx.ql[l], ctx.qh[l]
    return {"l": l, "c0": [mpi_smod(x, q, qh) for x in ct["c0"]],
            "c1": [mpi_smod(x, q, qh) for x in ct["c1"]]}


def he_swk(ctx: Ctx, l: int, d0, d1, swk) -> dict:
    """src/he-automorphism.c:40-85: only c0 receives the additive term (:73)."""
    ql, qlh = ctx.ql[l], ctx.qh[l]
    k0, k1 = _keyswitch(ctx, l, d1, swk)
    c0 = [mpi_smod((x + y) % ql, ql, qlh) for x, y in zip(k0, d0)]
    c1 = [mpi_smod(x, ql, qlh) for x in k1]
    return {"l": l, "c0": c0, "c1": c1}


def he_rot(ctx: Ctx, ct: dict, rot: int, rk_rot) -> dict:
    """src/he-automorphism.c:102-115 (`rk_rot` is rk[rot])."""
    n = ctx.n
    return he_swk(ctx, ct["l"], poly_rot(ct["c0"], rot, n), poly_rot(ct["c1"], rot, n), rk_rot)


def he_conj(ctx: Ctx, ct: dict, ck) -> dict:
    """src/he-automorphism.c:87-100."""
    n = ctx.n
    return he_swk(ctx, ct["l"], poly_conj(ct["c0"], n), poly_conj(ct["c1"], n), ck)


def he_add(ctx: Ctx, ct1: dict, ct2: dict) -> dict:
    """src/he-add.c:32-54: mpi_addm then mpi_smod per coefficient."""
    l = ct1["l"]
    q, qh = ctx.ql[l], ctx.qh[l]
    return {"l": l, "c0": [mpi_smod((x + y) % q, q, qh) for x, y in zip(ct1["c0"], ct2["c0"])],
            "c1": [mpi_smod((x + y) % q, q, qh) for x, y in zip(ct1["c1"], ct2["c1"])]}


def he_sub(ctx: Ctx, ct1: dict, ct2: dict) -> dict:
    """src/he-add.c:57-79: mpi_subm then mpi_smod."""
    l = ct1["l"]
    q, qh = ctx.ql[l], ctx.qh[l]
    return {"l": l, "c0": [mpi_smod((x - y) % q, q, qh) for x, y in zip(ct1["c0"], ct2["c0"])],
            "c1": [mpi_smod((x - y) % q, q, qh) for x, y in zip(ct1["c1"], ct2["c1"])]}


def he_addpt(ctx: Ctx, ct: dict, m: list[int], sign: int = 1) -> dict:
    """src/he-add.c:82-103 (sign=+1) and :106-127 (he_subpt, sign=-1): c0 +- m, c1 only reduced."""
    l = ct["l"]
    q, qh = ctx.ql[l], ctx.qh[l]
    return {"l": l, "c0": [mpi_smod((x + sign * y) % q, q, qh) for x, y in zip(ct["c0"], m)],
            "c1": [mpi_smod(x % q, q, qh) for x in ct["c1"]]}


def he_neg(ctx: Ctx, ct: dict) -> dict:
    """src/he-add.c:130-146."""
    l = ct["l"]
    q, qh = ctx.ql[l], ctx.qh[l]
    return {"l": l, "c0": [mpi_smod(-x, q, qh) for x in ct["c0"]], "c1": [mpi_smod(-x, q, qh) for x in ct["c1"]]}


def he_gemv_encoded(ctx: Ctx, ct: dict, pts: list, dim: int, n1: int, n2: int, rk) -> dict:
    """src/he-algo.c:47-90 with the encoded diagonals given: pts[i*n1+j] = coefficients of
    he_ecd(zrotdiag(A, i*n1+j, -i*n1)); rk[r] = (p0, p1) of rotation r.  The same loop, the same
    order of he_rot / he_mulpt / he_add / he_rs as the reference."""
    outer = None
    for i in range(n2):
        shift = i * n1
        inner = None
        for j in range(n1):
            ct_rot = he_rot(ctx, ct, j, rk[j])
            ct_rot = he_mulpt(ctx, ct_rot, pts[shift + j], dim)
            inner = ct_rot if inner is None else he_add(ctx, inner, ct_rot)
        inner = he_rot(ctx, inner, shift, rk[shift])
        outer = inner if outer is None else he_add(ctx, outer, inner)
    return he_rs(ctx, outer)


def he_coeff2slot_encoded(ctx: Ctx, ct: dict, pts4: list, cpt: list, dim_gemv: int, dim_pt: int,
                          n1: int, n2: int, ck, rk) -> tuple:
    """src/he-bootstrap.c:359-409 with the plaintexts given: pts4 = encoded diagonals of U0_conjT, U0_T,
    U1_conjT, U1_T, cpt = coefficients of the constant plaintext.  Returns (ct_real, ct_imag)."""
    ct_conj = he_conj(ctx, ct, ck)
    out = []
    for part in range(2):
        ct0 = he_gemv_encoded(ctx, ct, pts4[2 * part], dim_gemv, n1, n2, rk)
        ct1 = he_gemv_encoded(ctx, ct_conj, pts4[2 * part + 1], dim_gemv, n1, n2, rk)
        r = he_add(ctx, ct0, ct1)
        r = he_mulpt(ctx, r, cpt, dim_pt)
        out.append(he_rs(ctx, r))
    return tuple(out)


def he_exp_encoded(ctx: Ctx, ct: dict, pts: list, dim_a: int, dim_t: int, iters: int, rlk) -> dict:
    """src/he-algo.c:435-458 (he_exp) over he_exp_taylor (:364-433) with the plaintexts given: pts[0] =
    coefficients of he_ecd(a/2^iter), pts[1..7] = he_const_pt of 1, 3, 1/6, 5, 1/120, 7, 1/5040.  The same
    primitives in the same order as the reference."""
    act = he_rs(ctx, he_mulpt(ctx, ct, pts[0], dim_a))                       # :448-449
    ct2 = he_rs(ctx, he_mul(ctx, act, act, rlk))                             # :378-379
    ct4 = he_rs(ctx, he_mul(ctx, ct2, ct2, rlk))                             # :381-382
    lin = lambda ka, km: he_rs(ctx, he_mulpt(ctx, he_addpt(ctx, act, pts[ka]), pts[km], dim_t))
    ct01 = he_moddown(ctx, lin(1, 1))                                        # :384-388
    ct23 = he_rs(ctx, he_mul(ctx, ct2, lin(2, 3), rlk))                      # :390-396
    ct0123 = he_moddown(ctx, he_add(ctx, ct01, ct23))                        # :398-399
    ct45 = he_moddown(ctx, lin(4, 5))                                        # :401-406
    ct67 = he_rs(ctx, he_mul(ctx, ct2, lin(6, 7), rlk))                      # :408-414
    ct4567 = he_rs(ctx, he_mul(ctx, ct4, he_add(ctx, ct45, ct67), rlk))      # :416-418
    out = he_add(ctx, ct0123, ct4567)                                        # :420
    for _ in range(iters):                                                   # :451-454
        out = he_rs(ctx, he_mul(ctx, out, out, rlk))
    return out


def he_rlsin_encoded(ctx: Ctx, ct: dict, pts: list, dim_a: int, dim_t: int, dim_f: int, iters: int,
                     rlk, ck) -> dict:
    """src/he-bootstrap.c:444-475: pts[0..7] as he_exp_encoded (pts[0] encodes i*a/2^iter), pts[8] =
    he_ecd(1/(2*i*a))."""
    e = he_exp_encoded(ctx, ct, pts, dim_a, dim_t, iters, rlk)
    en = he_conj(ctx, e, ck)
    return he_rs(ctx, he_mulpt(ctx, he_sub(ctx, e, en), pts[8], dim_f))


def he_slot2coeff_encoded(ctx: Ctx, ct0: dict, ct1: dict, pts2: list, dim: int, n1: int, n2: int, rk) -> dict:
    """src/he-bootstrap.c:414-438: he_gemv(U0, ct0) + he_gemv(U1, ct1)."""
    return he_add(ctx, he_gemv_encoded(ctx, ct0, pts2[0], dim, n1, n2, rk),
                  he_gemv_encoded(ctx, ct1, pts2[1], dim, n1, n2, rk))


# --------------------------------------------------------------------------- #
# packed-word interchange format used at the C-ABI (include/gpqhe_b200.h):
# word-major two's complement, words[w, i] = bits 64w..64w+63 of coefficient i
# --------------------------------------------------------------------------- #
def to_words(a: list[int], W: int) -> np.ndarray:
    n = len(a)
    mask = (1 << (64 * W)) - 1
    buf = b"".join((x & mask).to_bytes(8 * W, "little") for x in a)
    return np.ascontiguousarray(np.frombuffer(buf, dtype="<u8").reshape(n, W).T)


def from_words(w: np.ndarray) -> list[int]:
    W, n = w.shape
    rows = np.ascontiguousarray(w.T).astype("<u8")
    half, full = 1 << (64 * W - 1), 1 << (64 * W)
    out = []
    raw = rows.tobytes()
    for i in range(n):
        x = int.from_bytes(raw[i * 8 * W:(i + 1) * 8 * W], "little")
        out.append(x - full if x >= half else x)
    return out
