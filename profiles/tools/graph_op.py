#!/usr/bin/env python
"""he_mul+he_rs at N=2^16 issued eagerly vs replayed from ONE CUDA graph (stream capture of gq_he_mul_rs_dev on the
caller's stream): how much of the single-stream latency is launch gaps.  python profiles/tools/graph_op.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gpqhe_b200 as g  # noqa: E402

ctx = g.HeContext(bench.LOGN, 1 << bench.LOGQ, bench.SLOTS, 1 << bench.LOGDELTA, device=0,
                  lib_path=os.environ.get("GQ_LIB_VARIANT"))      # A/B of a differently built library
lib, h = ctx._lib, ctx._h
n, W, L = ctx.n, ctx.W, ctx.L
rng = np.random.default_rng(1)
evk = ctx.genevk(bench.synth_evk(rng, ctx.primes, n, ctx.dimevk), bench.synth_evk(rng, ctx.primes, n, ctx.dimevk))
dev = torch.device("cuda", 0)
POOL = 12
inp = torch.from_numpy(np.stack([bench.synth_ct(rng, n, W, bench.LOGQ) for _ in range(4 * POOL)]).view(np.int64)).to(dev)
out = torch.empty((2, W, n), dtype=torch.int64, device=dev)
poly = W * n * 8
st = torch.cuda.Stream(device=dev)


def op(i, stream):
    b = inp.data_ptr() + (i % POOL) * 4 * poly
    rc = lib.gq_he_mul_rs_dev(h, L, out.data_ptr(), out.data_ptr() + poly, b, b + poly, b + 2 * poly, b + 3 * poly,
                              evk._h, stream.cuda_stream)
    assert rc == 0, lib.gq_last_error()


def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for i in range(5):
            fn(i)
        st.synchronize()
        e0.record(st)
        for i in range(reps):
            fn(i)
        e1.record(st)
        st.synchronize()
    return e0.elapsed_time(e1) / reps


eager = timed(lambda i: op(i, st), 400)
graphs = []
with torch.cuda.stream(st):
    for k in range(POOL):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st, capture_error_mode="thread_local"):
            op(k, st)
        graphs.append(gr)
replay = timed(lambda i: graphs[i % POOL].replay(), 400)
print(f"eager {eager * 1e3:.1f} us per op, one graph per op {replay * 1e3:.1f} us per op")
ctx.close()
