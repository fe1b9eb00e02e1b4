#!/usr/bin/env python
"""profiles/probe: register-resident butterfly rates (exact and approximate quotient) and the IMAD.WIDE peak."""
import ctypes as C
import os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = C.CDLL(os.path.join(ROOT, "profiles", "probe", "libgqprobe.so"))
lib.gqp_last_butterfly_approx_rate.restype = C.c_double
p = 576460752308273153
w = p // 3
bf, mw = C.c_double(0), C.c_double(0)
assert lib.gqp_bench_intpipe(0, C.c_uint64(p), C.c_uint64(w), C.c_uint64((w << 64) // p), C.byref(bf), C.byref(mw)) == 0
ap = lib.gqp_last_butterfly_approx_rate()
print(f"butterfly exact {bf.value/1e9:.1f} G/s  approx-quotient {ap/1e9:.1f} G/s  ratio {ap/bf.value:.3f}  madwide {mw.value/1e9:.1f} G/s")
