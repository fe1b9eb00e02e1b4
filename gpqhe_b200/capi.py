"""ctypes binding of the C ABI in include/gpqhe_b200.h.

The product library is gpqhe_b200/libgpqhe_b200.so (hand-written CUDA, sm_100a, built
by gpqhe_b200/csrc/Makefile).  There is no CPU fallback: if the library is missing,
or no CUDA device is present when a context is created, this module raises.  Tests
that run without a GPU pass the path of the emulated build (tests/emul/) explicitly;
nothing in the package ever selects it.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(_HERE, "libgpqhe_b200.so")

u64 = C.c_uint64
u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p


class GqInfo(C.Structure):
    _fields_ = [(k, C.c_uint) for k in
                ("logn", "n", "logp", "dimub", "has_he", "L", "dim", "dimevk", "W", "q_is_pow2")] + \
               [("device", C.c_int)]


class GqP2pXchg(C.Structure):
    """gq_p2p_xchg: one exchange of the sharded pipeline, producer (`send`) or consumer (`wait`) side"""
    _fields_ = [("G", C.c_uint), ("me", C.c_uint), ("slot", C.c_uint), ("D", C.c_uint),
                ("dst", C.c_void_p), ("off", C.c_void_p), ("flags", C.c_void_p), ("flags_local", C.c_void_p),
                ("epoch_dev", C.c_void_p), ("err_dev", C.c_void_p), ("ticket_dev", C.c_void_p),
                ("pre", C.c_uint)]


class GpqheError(RuntimeError):
    pass


# name -> (restype, argtypes); every symbol include/gpqhe_b200.h declares
SIGNATURES = {
    "gq_last_error": (C.c_char_p, []),
    "gq_abi_version": (C.c_int, []),
    "gq_ctx_create": (C.c_int, [C.POINTER(vp), C.c_uint, u64p, C.c_uint, C.c_uint, u64, C.c_uint,
                                C.c_uint, C.c_int, u64p, C.POINTER(u64p), C.POINTER(u64p)]),
    "gq_ctx_destroy": (None, [vp]),
    "gq_ctx_info": (C.c_int, [vp, C.POINTER(GqInfo)]),
    "gq_ctx_primes": (C.c_int, [vp, u64p]),
    "gq_ctx_prime_consts": (C.c_int, [vp, C.c_uint, u64p]),
    "gq_ctx_dims": (C.c_int, [vp, C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint)]),
    "gq_ctx_sync": (C.c_int, [vp]),
    "gq_ctx_launch_count": (C.c_ulonglong, [vp]),
    "gq_ctx_stream": (vp, [vp]),
    "gq_ctx_crt_fixups": (C.c_longlong, [vp]),
    "gq_ctx_set_lanes": (C.c_int, [vp, C.c_uint]),
    "gq_ctx_lane_count": (C.c_uint, [vp]),
    "gq_ctx_lane": (vp, [vp, C.c_uint]),
    "gq_ctx_lanes_fork": (C.c_int, [vp]),
    "gq_ctx_lanes_join": (C.c_int, [vp]),
    "gq_host_alloc": (C.c_int, [C.POINTER(vp), C.c_size_t]),
    "gq_host_free": (None, [vp]),
    "gq_ntt": (C.c_int, [vp, C.c_uint, vp]),
    "gq_invntt": (C.c_int, [vp, C.c_uint, vp]),
    "gq_poly_rns_add": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_poly_rns_mul": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_ntt_batch": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.c_int]),
    "gq_rns_decompose": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, C.c_uint]),
    "gq_poly_rns2mpi": (C.c_int, [vp, C.c_uint, vp, C.c_uint, vp, u64p, C.c_uint]),
    "gq_poly_mul": (C.c_int, [vp, vp, C.c_uint, vp, vp, C.c_uint, C.c_uint, u64p, C.c_uint]),
    "gq_poly_rot": (C.c_int, [vp, vp, vp, C.c_uint, u64]),
    "gq_poly_conj": (C.c_int, [vp, vp, vp, C.c_uint]),
    "gq_evk_create": (C.c_int, [vp, C.POINTER(vp), vp, vp]),
    "gq_evk_create_dev": (C.c_int, [vp, C.POINTER(vp), vp, vp]),
    "gq_evk_export": (C.c_int, [vp, vp, vp]),
    "gq_evk_destroy": (None, [vp]),
    "gq_he_mul": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_add": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_sub": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_addpt": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp]),
    "gq_he_subpt": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp]),
    "gq_he_neg": (C.c_int, [vp, C.c_uint, vp, vp]),
    "gq_he_add_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_sub_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_addpt_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_subpt_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_neg_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_he_rot_hoisted_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_rot_hoisted": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_gemv": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_gemv_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_coeff2slot": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_coeff2slot_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_exp": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_exp_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_rlsin": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_rlsin_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_slot2coeff": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_slot2coeff_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_mul_rs": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_mul_batch": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp, vp, C.c_int]),
    "gq_he_mulpt": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp]),
    "gq_he_rs": (C.c_int, [vp, C.c_uint, vp, vp]),
    "gq_he_moddown": (C.c_int, [vp, C.c_uint, vp, vp]),
    "gq_he_rot": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp]),
    "gq_he_conj": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_ntt_batch_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.c_int, vp]),
    "gq_poly_rns_mul_batch_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp]),
    "gq_he_mul_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_ctx_set_plaintext_bits": (C.c_int, [vp, C.c_uint]),
    "gq_dev_alloc": (C.c_int, [vp, C.c_size_t, C.POINTER(vp)]),
    "gq_dev_free": (None, [vp, vp]),
    "gq_dev_upload": (C.c_int, [vp, vp, vp, C.c_size_t, vp]),
    "gq_dev_download": (C.c_int, [vp, vp, vp, C.c_size_t, vp]),
    "gq_dev_copy": (C.c_int, [vp, vp, vp, C.c_size_t, vp]),
    "gq_he_mul_rs_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gq_he_mulpt_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_mulpt_rs_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp, vp, vp]),
    "gq_he_rs_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_he_moddown_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp]),
    "gq_he_rot_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp]),
    "gq_he_conj_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp]),
    "gq_stage_decompose_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.POINTER(vp), C.c_uint, C.c_uint, vp, vp]),
    "gq_stage_ntt_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.c_int, vp]),
    "gq_stage_tensor_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp]),
    "gq_stage_tensor_inv_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp]),
    "gq_stage_mul_evk_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp]),
    "gq_stage_mul_evk_inv_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, vp]),
    "gq_stage_reconstruct_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.POINTER(vp), C.c_uint, vp]),
    "gq_stage_reconstruct_noalloc_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.POINTER(vp), C.c_uint, vp]),
    "gq_stage_ks_epilogue_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, C.c_uint, vp]),
    "gq_stage_ks_epilogue_rs_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, C.c_uint, vp]),
    "gq_stage_rescale_dev": (C.c_int, [vp, C.c_uint, vp, vp, C.c_int, C.c_uint, vp]),
    "gq_set_pdl": (C.c_int, [C.c_int]),
    "gq_stage_prescale_ok": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_int]),
    "gq_stage_tensor_inv_pre_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp]),
    "gq_stage_mul_evk_inv_pre_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, vp, vp, vp]),
    "gq_stage_reconstruct_pre_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.POINTER(vp), C.c_uint, vp]),
    "gq_stage_ks_epilogue_pre_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, C.c_uint, C.c_int, vp]),
    "gq_p2p_alloc": (C.c_int, [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]),
    "gq_p2p_open": (C.c_int, [vp, C.c_char_p, C.POINTER(vp)]),
    "gq_p2p_close": (C.c_int, [vp, vp]),
    "gq_p2p_free": (C.c_int, [vp, vp]),
    "gq_p2p_scatter_c2l_dev": (C.c_int, [vp, vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.POINTER(vp), C.POINTER(C.c_int), vp]),
    "gq_p2p_scatter_l2c_dev": (C.c_int, [vp, vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.POINTER(vp), C.POINTER(C.c_int), vp]),
    "gq_p2p_bump_dev": (C.c_int, [vp, vp, vp]),
    "gq_p2p_fused_ok": (C.c_int, [vp, C.c_uint]),
    "gq_p2p_decompose_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.POINTER(vp), C.c_uint, C.c_uint, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_ntt_fwd_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_tensor_inv_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_mul_evk_inv_dev": (C.c_int, [vp, C.c_uint, C.c_uint, vp, vp, vp, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_reconstruct_dev": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, vp, C.POINTER(vp), C.c_uint, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_ks_epilogue_dev": (C.c_int, [vp, C.c_uint, vp, vp, vp, vp, vp, C.c_uint, C.c_int, C.POINTER(GqP2pXchg), vp]),
    "gq_p2p_signal_dev": (C.c_int, [vp, C.POINTER(vp), C.c_uint, C.c_uint, C.c_uint, vp, vp, vp]),
    "gq_p2p_wait_dev": (C.c_int, [vp, vp, C.c_uint, C.c_uint, vp, vp, vp]),
}

_loaded: dict[str, C.CDLL] = {}


def load(path: str | None = None) -> C.CDLL:
    """Load the C-ABI library and attach prototypes.  Raises if it is not built."""
    path = os.path.abspath(path or DEFAULT_LIB)
    if path in _loaded:
        return _loaded[path]
    if not os.path.exists(path):
        raise GpqheError(
            f"{path} not found: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()' or make -C gpqhe_b200/csrc). "
            "gpqhe_b200 has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library misses a symbol
        fn.restype, fn.argtypes = res, args
    _loaded[path] = lib
    return lib


def check(lib: C.CDLL, rc: int):
    if rc != 0:
        raise GpqheError(lib.gq_last_error().decode(errors="replace"))
