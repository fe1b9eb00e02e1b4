#!/usr/bin/env python
"""The forward NTT over bench.py's NTT batch (16 polys x 24 limbs = 384 limb-polys, 192 MiB) for ncu: 2 warm-up
transforms, then ONE transform (strided + tile kernel) between cudaProfilerStart/Stop.  Source of
profiles/traffic.json (dram bytes per launch of k_nttl_fwd_tile<12> in the shape roofline.achieved is quoted on)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gpqhe_b200 as g  # noqa: E402

ctx = g.HeContext(bench.LOGN, 1 << bench.LOGQ, bench.SLOTS, 1 << bench.LOGDELTA, device=0)
lib, h = ctx._lib, ctx._h
D2 = ctx.dims(ctx.L)[1]
dn = torch.empty((16, D2, ctx.n), dtype=torch.int64, device="cuda:0")
dn.random_(0, 1 << 59)
torch.cuda.synchronize()
for _ in range(2):
    assert lib.gq_stage_ntt_dev(h, 0, D2, 16, dn.data_ptr(), 0, None) == 0
ctx.sync()
dn.random_(0, 1 << 59)
torch.cuda.synchronize()
torch.cuda.profiler.start()
assert lib.gq_stage_ntt_dev(h, 0, D2, 16, dn.data_ptr(), 0, None) == 0
ctx.sync()
torch.cuda.profiler.stop()
print("ntt_tile done")
ctx.close()
