"""ctypes binding to the UNMODIFIED reference library built in oracle/_ref.

TEST INFRASTRUCTURE ONLY (see oracle/oracle.c header).  Used to (a) pin the
restatement in oracle.c / pyoracle.py, (b) generate tests/golden fixtures, and
(c) time the reference's own CPU path for bench.py's cpu_baseline when
`oracle/_ref/libgpqhe.so` and the system libgcrypt.so.20 load on the box.

The reference keeps all state in process globals (`polyctx`, `hectx`,
src/precomp.c:38-47), so one process holds one context at a time.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
MPI = C.c_void_p
u64 = C.c_uint64
u64p = C.POINTER(C.c_uint64)


class RnsCtx(C.Structure):
    """struct rns_ctx, src/poly.h:28-41."""


RnsCtx._fields_ = [
    ("dim", C.c_uint), ("p", u64), ("pinv_mont", u64), ("pinv_barr", u64), ("ninv", u64),
    ("zetas", u64p), ("zetas_inv", u64p), ("P", MPI), ("P_2", MPI),
    ("phat", C.POINTER(MPI)), ("phat_invmp", u64p), ("next", C.POINTER(RnsCtx)),
]


class PolyCtx(C.Structure):
    """struct poly_ctx, src/poly.h:49-65 (u128 members are 16-byte aligned)."""
    _fields_ = [
        ("logn", C.c_uint), ("n", C.c_uint), ("m", C.c_uint), ("logqub", C.c_uint),
        ("logq", C.c_uint), ("_pad0", C.c_uint), ("q", MPI), ("logR", C.c_uint),
        ("_pad1", C.c_uint * 3), ("R", u64 * 2), ("Rsub1", u64 * 2), ("dimub", C.c_uint),
        ("_pad2", C.c_uint), ("rns", C.POINTER(RnsCtx)), ("cyc_group", C.c_void_p),
        ("ring_zetas", C.c_void_p), ("cM", C.c_double), ("_pad3", u64),
    ]


class HeCtx(C.Structure):
    """struct he_ctx, src/gpqhe.h:36-50."""
    _fields_ = [
        ("P", MPI), ("PqL", MPI), ("q", C.POINTER(MPI)), ("qh", C.POINTER(MPI)), ("p", MPI),
        ("slots", C.c_uint), ("Delta", C.c_double), ("L", C.c_uint), ("dim", C.c_uint),
        ("dimevk", C.c_uint), ("Bclean", C.c_double), ("Brs", C.c_double), ("Bks", C.c_double),
        ("bnd_Bmult", C.POINTER(C.c_double)), ("Bmult", C.c_double),
    ]


class PolyMpi(C.Structure):
    _fields_ = [("coeffs", C.POINTER(MPI))]


class PolyRns(C.Structure):
    _fields_ = [("coeffs", u64p)]


class HeCt(C.Structure):
    """struct he_ct, src/gpqhe.h:84-91."""
    _fields_ = [("l", C.c_uint), ("nu", C.c_double), ("B", C.c_double),
                ("c0", PolyMpi), ("c1", PolyMpi)]


class HeEvk(C.Structure):
    _fields_ = [("p0", PolyRns), ("p1", PolyRns)]


class HePt(C.Structure):
    _fields_ = [("nu", C.c_double), ("m", PolyMpi)]


def available(variant: str = "libgpqhe.so") -> bool:
    if not os.path.exists(os.path.join(_HERE, "_ref", variant)):
        return False
    try:
        C.CDLL("libgcrypt.so.20")
    except OSError:
        return False
    return True


class Ref:
    """One loaded copy of the reference library."""

    def __init__(self, variant: str = "libgpqhe.so"):
        self.gcry = C.CDLL("libgcrypt.so.20", mode=C.RTLD_GLOBAL)
        self.lib = C.CDLL(os.path.join(_HERE, "_ref", variant))
        g, L = self.gcry, self.lib
        g.gcry_mpi_new.restype = MPI
        g.gcry_mpi_new.argtypes = [C.c_uint]
        g.gcry_mpi_release.argtypes = [MPI]
        g.gcry_mpi_scan.argtypes = [C.POINTER(MPI), C.c_int, C.c_char_p, C.c_size_t, C.c_void_p]
        g.gcry_mpi_print.argtypes = [C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), MPI]
        g.gcry_mpi_set.argtypes = [MPI, MPI]
        g.gcry_mpi_set.restype = MPI
        g.gcry_mpi_is_neg.argtypes = [MPI]
        g.gcry_mpi_neg.argtypes = [MPI, MPI]
        self.polyctx = PolyCtx.in_dll(L, "polyctx")
        self.hectx = HeCtx.in_dll(L, "hectx")
        L.hectx_init.argtypes = [C.c_uint, MPI, C.c_uint, u64]
        L.polyctx_init.argtypes = [C.c_uint, MPI]
        for f in ("ntt", "invntt"):
            getattr(L, f).argtypes = [u64p, C.POINTER(RnsCtx)]
        for f in ("poly_rns_add", "poly_rns_mul"):
            getattr(L, f).argtypes = [u64p, u64p, u64p, C.POINTER(RnsCtx)]
        L.rns_decompose.argtypes = [u64p, C.POINTER(MPI), C.POINTER(RnsCtx)]
        L.rns_reconstruct.argtypes = [MPI, u64p, C.c_uint, C.POINTER(RnsCtx)]
        L.poly_rns2mpi.argtypes = [C.POINTER(PolyMpi), C.POINTER(PolyRns), C.POINTER(RnsCtx), MPI]
        L.poly_mul.argtypes = [C.POINTER(PolyMpi)] * 3 + [C.c_uint, MPI]
        L.poly_rot.argtypes = [C.POINTER(PolyMpi), C.POINTER(PolyMpi), C.c_size_t]
        L.poly_conj.argtypes = [C.POINTER(PolyMpi), C.POINTER(PolyMpi)]
        L.poly_mpi_alloc.argtypes = [C.POINTER(PolyMpi)]
        L.poly_mpi_free.argtypes = [C.POINTER(PolyMpi)]
        L.he_alloc_ct.argtypes = [C.POINTER(HeCt)]
        L.he_free_ct.argtypes = [C.POINTER(HeCt)]
        L.he_alloc_evk.argtypes = [C.POINTER(HeEvk)]
        L.he_alloc_pt.argtypes = [C.POINTER(HePt)]
        L.he_mul.argtypes = [C.POINTER(HeCt)] * 3 + [C.POINTER(HeEvk)]
        L.he_mulpt.argtypes = [C.POINTER(HeCt), C.POINTER(HeCt), C.POINTER(HePt)]
        L.he_rs.argtypes = [C.POINTER(HeCt)]
        L.he_moddown.argtypes = [C.POINTER(HeCt)]
        L.he_rot.argtypes = [C.POINTER(HeCt), C.c_int, C.POINTER(HeEvk)]
        L.he_conj.argtypes = [C.POINTER(HeCt), C.POINTER(HeEvk)]
        L.mpi_smod.argtypes = [MPI, MPI, MPI]
        L.mpi_rdiv.argtypes = [MPI, MPI, MPI]
        L.montgomery_inv.restype = u64
        L.montgomery_inv.argtypes = [u64]
        L.barrett_inv.restype = u64
        L.barrett_inv.argtypes = [u64]
        self.n = 0

    # ---- MPI <-> int ------------------------------------------------------
    def mpi(self, x: int) -> MPI:
        m = MPI()
        s = (("-" if x < 0 else "") + "0" + format(abs(x), "X")).encode()
        err = self.gcry.gcry_mpi_scan(C.byref(m), 4, s, 0, None)   # GCRYMPI_FMT_HEX
        assert err == 0
        return m

    def set_mpi(self, dst: MPI, x: int):
        t = self.mpi(x)
        self.gcry.gcry_mpi_set(dst, t)
        self.gcry.gcry_mpi_release(t)

    def to_int(self, m) -> int:
        buf = C.create_string_buffer(1024)
        nw = C.c_size_t()
        err = self.gcry.gcry_mpi_print(4, buf, 1024, C.byref(nw), m)
        assert err == 0
        return int(buf.value.decode(), 16)

    # ---- context ----------------------------------------------------------
    def hectx_init(self, logn: int, q: int, slots: int, Delta: int):
        self.lib.hectx_init(logn, self.mpi(q), slots, Delta)
        self.n = self.polyctx.n

    def polyctx_init(self, logn: int, q: int):
        self.lib.polyctx_init(logn, self.mpi(q))
        self.n = self.polyctx.n

    def node(self, d: int):
        """pointer to the d-th list node (0-based)."""
        nd = self.polyctx.rns
        for _ in range(d):
            nd = nd.contents.next
        return nd

    def primes(self):
        return [self.node(d).contents.p for d in range(self.polyctx.dimub)]

    # ---- containers -------------------------------------------------------
    def new_poly(self, vals=None) -> PolyMpi:
        p = PolyMpi()
        self.lib.poly_mpi_alloc(C.byref(p))
        if vals is not None:
            for i, x in enumerate(vals):
                self.set_mpi(p.coeffs[i], x)
        return p

    def poly_ints(self, p: PolyMpi):
        return [self.to_int(p.coeffs[i]) for i in range(self.n)]

    def new_ct(self, l: int, c0, c1, nu=1.0, B=0.0) -> HeCt:
        ct = HeCt()
        self.lib.he_alloc_ct(C.byref(ct))
        ct.l, ct.nu, ct.B = l, nu, B
        for i in range(self.n):
            self.set_mpi(ct.c0.coeffs[i], c0[i])
            self.set_mpi(ct.c1.coeffs[i], c1[i])
        return ct

    def ct_dict(self, ct: HeCt):
        return {"l": ct.l, "c0": self.poly_ints(ct.c0), "c1": self.poly_ints(ct.c1),
                "nu": ct.nu, "B": ct.B}

    def new_evk(self, p0: np.ndarray, p1: np.ndarray) -> HeEvk:
        evk = HeEvk()
        self.lib.he_alloc_evk(C.byref(evk))
        nw = self.hectx.dimevk * self.n
        assert p0.size == nw and p1.size == nw
        C.memmove(evk.p0.coeffs, p0.ctypes.data, nw * 8)
        C.memmove(evk.p1.coeffs, p1.ctypes.data, nw * 8)
        return evk


def ptr(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)
