"""Does running independent he_mul+he_rs streams concurrently on ONE GPU fill the kernel tails?
Times LANES contexts (each its own stream + scratch) fed round-robin (run on the GPU box)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import gpqhe_b200 as g  # noqa: E402

POOL = 8
STEPS = 200
for LANES in (1, 2, 3, 4):
    ctxs = [g.HeContext(16, 1 << 438, 4, 1 << 50) for _ in range(LANES)]
    n, W, L = ctxs[0].n, ctxs[0].W, ctxs[0].L
    poly = W * n * 8
    rng = np.random.default_rng(1)
    primes = ctxs[0].primes
    ev = [np.concatenate([rng.integers(0, int(primes[d]), size=n, dtype=np.uint64) for d in range(ctxs[0].dimevk)])
          for _ in range(2)]
    evks = [c.genevk(ev[0], ev[1]) for c in ctxs]
    dev = torch.device("cuda", 0)
    dpool = torch.from_numpy(rng.integers(0, 1 << 62, size=(POOL, 4, W, n), dtype=np.int64))
    dpool[:, :, W - 1, :] &= (1 << (438 - 64 * (W - 1) - 1)) - 1
    dpool = dpool.to(dev)
    douts = [torch.empty((2, W, n), dtype=torch.int64, device=dev) for _ in range(LANES)]
    torch.cuda.synchronize()

    def step(i):
        k = i % LANES
        c = ctxs[k]
        b = dpool.data_ptr() + (i % POOL) * 4 * poly
        o0, o1 = douts[k].data_ptr(), douts[k].data_ptr() + poly
        assert c._lib.gq_he_mul_dev(c._h, L, o0, o1, b, b + poly, b + 2 * poly, b + 3 * poly, evks[k]._h, None) == 0
        assert c._lib.gq_he_rs_dev(c._h, L, o0, o1, None) == 0

    for i in range(4 * LANES):
        step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(STEPS):
        step(i)
    t_enq = time.perf_counter() - t0
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"lanes={LANES}: {dt / STEPS * 1e3:.4f} ms/op ({STEPS / dt:.0f} ops/s), host enqueue {t_enq / STEPS * 1e3:.4f} ms/op",
          flush=True)
    ref = douts[0].clone()
    for c in ctxs:
        c.close()
    del evks, ctxs
