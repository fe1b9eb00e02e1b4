"""gpqhe_b200 - B200-native RNS polynomial hot path of GPQHE (NTT, MPI<->RNS CRT,
relinearisation, rescale, Galois key switch) behind the reference's C API.

csrc/   hand-written CUDA for sm_100a + the C ABI (include/gpqhe_b200.h)
capi.py ctypes binding of that ABI
he.py   host-side mirror of the reference interface for this path
wire.py packed dump of ciphertexts / plaintexts / evaluation keys (checkpoints, fixtures)
"""
from .capi import GpqheError, load  # noqa: F401
from .he import (HeContext, HeCt, HeEvk, HePt, big_to_ints, hectx_init, ints_to_big,  # noqa: F401
                 polyctx_init)
