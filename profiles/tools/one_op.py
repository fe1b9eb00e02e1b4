#!/usr/bin/env python
"""One he_mul+he_rs at N=2^16 for ncu: 3 warm-up ops, then ONE op between cudaProfilerStart/Stop
(ncu --profile-from-start off ... python profiles/tools/one_op.py [--rot] [--gemv])."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gpqhe_b200 as g  # noqa: E402

ctx = g.HeContext(bench.LOGN, 1 << bench.LOGQ, bench.SLOTS, 1 << bench.LOGDELTA, device=0)
lib, h = ctx._lib, ctx._h
n, W, L = ctx.n, ctx.W, ctx.L
rng = np.random.default_rng(1)
evk = ctx.genevk(bench.synth_evk(rng, ctx.primes, n, ctx.dimevk), bench.synth_evk(rng, ctx.primes, n, ctx.dimevk))
dev = torch.device("cuda", 0)
inp = torch.from_numpy(np.stack([bench.synth_ct(rng, n, W, bench.LOGQ) for _ in range(4)]).view(np.int64)).to(dev)
out = torch.empty((2, W, n), dtype=torch.int64, device=dev)
poly = W * n * 8


def op():
    b = inp.data_ptr()
    if "--rot" in sys.argv:
        rc = lib.gq_he_rot_dev(h, L, 1, b, b + poly, evk._h, None)
    else:
        rc = lib.gq_he_mul_rs_dev(h, L, out.data_ptr(), out.data_ptr() + poly, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h, None)
    assert rc == 0, lib.gq_last_error()


if "--gemv" in sys.argv:         # configs[3]: he_gemv with slots = 16 (4 x 4 baby-step / giant-step), synthetic diagonals
    import ctypes as C
    import math
    NROT = 16
    rks = [ctx.genevk(bench.synth_evk(rng, ctx.primes, n, ctx.dimevk), bench.synth_evk(rng, ctx.primes, n, ctx.dimevk))
           for _ in range(NROT)]
    gdim = int((bench.LOGQ + 1 + math.log2(float(1 << bench.LOGDELTA)) + bench.LOGN) / 59 + 1)
    gpts = torch.zeros((NROT, W, n), dtype=torch.int64, device=dev)
    gpts[:, 0, :].random_(-(1 << bench.LOGDELTA), 1 << bench.LOGDELTA)
    gpts[:, 1:, :] = (gpts[:, :1, :] >> 63)
    VP = C.c_void_p * NROT
    ptsv = VP(*[gpts.data_ptr() + k * poly for k in range(NROT)])
    keys = VP(*[rk._h for rk in rks])
    assert lib.gq_ctx_set_plaintext_bits(h, bench.LOGDELTA + 1) == 0
    ctx.set_lanes(3)

    def op():
        b = inp.data_ptr()
        rc = lib.gq_he_gemv_dev(h, L, 4, 4, gdim, out.data_ptr(), out.data_ptr() + poly, b, b + poly, ptsv, keys, None)
        assert rc == 0, lib.gq_last_error()

if "--p2p" in sys.argv:          # the peer-to-peer sharded plan with ONE virtual rank (exchanges fused into the kernels
    from gpqhe_b200.sharded import P2PPlan      # unless GPQHE_B200_P2P_UNFUSED=1): the same launches a rank of N issues
    ext = torch.cuda.ExternalStream(lib.gq_ctx_stream(h), device=dev)
    with torch.cuda.stream(ext):
        GV = int(os.environ.get("GQ_P2P_RANKS", "1"))        # virtual ranks on this one GPU: the grids of a GV-GPU run
        plans = P2PPlan.loopback(ctx, GV, L, evk, rescale=True, device=dev)
        nl = n // GV
        for r, pl in enumerate(plans):
            for k in range(4):
                pl.inp[k].copy_(inp[k][:, r * nl:(r + 1) * nl])

    def op():
        with torch.cuda.stream(ext):
            P2PPlan.run_lockstep(plans)

for _ in range(3):
    op()
ctx.sync()
torch.cuda.profiler.start()
op()
ctx.sync()
torch.cuda.profiler.stop()
print("one_op done, launches so far:", ctx.launch_count)
ctx.close()
