"""Where does the time of gq_he_mul_batch go?  Times the copy engines alone, the kernels alone and the
pipelined call at several batch sizes (run on the GPU box: python profiles/tools/probe_batch.py)."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import gpqhe_b200 as g  # noqa: E402
from gpqhe_b200 import capi  # noqa: E402

ctx = g.HeContext(16, 1 << 438, 4, 1 << 50)
lib, h = ctx._lib, ctx._h
n, W, L = ctx.n, ctx.W, ctx.L
poly = W * n * 8
POOL = 8
rng = np.random.default_rng(1)
host = torch.from_numpy(rng.integers(0, 1 << 62, size=(POOL, 4, W, n), dtype=np.int64)).pin_memory()
host[:, :, W - 1, :] &= (1 << (438 - 64 * (W - 1) - 1)) - 1
houts = torch.empty((POOL, 2, W, n), dtype=torch.int64).pin_memory()
primes = ctx.primes
ev = [np.concatenate([rng.integers(0, int(primes[d]), size=n, dtype=np.uint64) for d in range(ctx.dimevk)]) for _ in range(2)]
evk = ctx.genevk(ev[0], ev[1])
dev = torch.device("cuda", 0)
dbuf = torch.empty((POOL, 4, W, n), dtype=torch.int64, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def wall(fn, reps=1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def h2d(cnt):
    with torch.cuda.stream(s1):
        for i in range(cnt):
            dbuf[i % POOL].copy_(host[i % POOL], non_blocking=True)


def d2h(cnt):
    with torch.cuda.stream(s2):
        for i in range(cnt):
            houts[i % POOL].copy_(dbuf[i % POOL, :2], non_blocking=True)


def batch(cnt):
    VP = C.c_void_p * cnt
    ins = [VP(*[host.data_ptr() + (i % POOL) * 4 * poly + j * poly for i in range(cnt)]) for j in range(4)]
    outs = [VP(*[houts.data_ptr() + (i % POOL) * 2 * poly + j * poly for i in range(cnt)]) for j in range(2)]
    capi.check(lib, lib.gq_he_mul_batch(h, L, cnt, outs[0], outs[1], ins[0], ins[1], ins[2], ins[3], evk._h, 1))


def devonly(cnt):
    for i in range(cnt):
        b = dbuf.data_ptr() + (i % POOL) * 4 * poly
        capi.check(lib, lib.gq_he_mul_dev(h, L, b, b + poly, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h, None))
        capi.check(lib, lib.gq_he_rs_dev(h, L, b, b + poly, None))
    ctx.sync()


def devonly_host_time(cnt):
    t0 = time.perf_counter()
    for i in range(cnt):
        b = dbuf.data_ptr() + (i % POOL) * 4 * poly
        capi.check(lib, lib.gq_he_mul_dev(h, L, b, b + poly, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h, None))
        capi.check(lib, lib.gq_he_rs_dev(h, L, b, b + poly, None))
    t1 = time.perf_counter()
    ctx.sync()
    return (t1 - t0) / cnt


batch(4); devonly(4); h2d(4); d2h(4)
N = 100
print(f"h2d alone      : {wall(lambda: h2d(N)) / N * 1e3:.3f} ms/op  ({4 * poly / (wall(lambda: h2d(N)) / N) / 1e9:.1f} GB/s)")
print(f"d2h alone      : {wall(lambda: d2h(N)) / N * 1e3:.3f} ms/op  ({2 * poly / (wall(lambda: d2h(N)) / N) / 1e9:.1f} GB/s)")
print(f"h2d + d2h      : {wall(lambda: (h2d(N), d2h(N))) / N * 1e3:.3f} ms/op")
print(f"kernels alone  : {wall(lambda: devonly(N)) / N * 1e3:.3f} ms/op, host issue time {devonly_host_time(N) * 1e3:.3f} ms/op")
print(f"kernels + h2d + d2h (independent): {wall(lambda: (h2d(N), d2h(N), devonly(N))) / N * 1e3:.3f} ms/op")
for cnt in (5, 20, 100, 400):
    print(f"gq_he_mul_batch count={cnt:4d}: {wall(lambda: batch(cnt)) / cnt * 1e3:.3f} ms/op")
ctx.close()
