"""gq_stage_reconstruct_noalloc_dev (gq_noalloc.cu): the CRT reconstruction stage with its fix-list scratch in a
file-scope __device__ array instead of cudaMalloc'd context memory - no counterpart in the reference, added
for callers that must not allocate.  Checked through the C ABI against the existing
entry point gq_stage_reconstruct_dev (same kernels: bit-exact is the tolerance), on two calls in a row, with
coefficients on the +P/2 edge so that the scratch list is really written and read.

This file sorts last on purpose: a failure here must not hide other tests."""
import contextlib
import ctypes as C

import numpy as np
import pytest
import torch

import gpqhe_b200 as g
from gpqhe_b200 import capi
from oracle import pyoracle as o


def _check(lib_path, device, logn, q, Delta, npolys, logqub=0, calls=2):
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub or None)
    try:
        lib, h = ctx._lib, ctx._h
        n, W, l = ctx.n, ctx.W, ctx.L
        D = ctx.dims(l)[0]
        P = oc.rns[D - 1].P
        Ph = P // 2
        edge = [Ph, Ph + 1, Ph - 1, 0, P - 1, 1]
        cm = contextlib.nullcontext()
        if torch.device(device).type == "cuda":
            cm = torch.cuda.stream(torch.cuda.ExternalStream(lib.gq_ctx_stream(h), device=device))
        rng = np.random.default_rng(logn)
        for call in range(calls):
            # fresh inputs per call: what the previous call left in the shared scratch must not matter
            r = np.stack([np.stack([rng.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(D)])
                          for _ in range(npolys)])
            for k in range(0, min(n, 48), 3 + call):          # a few coefficients on the +P/2 edge: they take the list path
                for b in range(npolys):
                    v = edge[(k + b + call) % len(edge)]
                    for d in range(D):
                        r[b, d, k] = v % ctx.primes[d]
            with cm:
                rhat = torch.from_numpy(r.view(np.int64)).to(device)
                want = torch.zeros((npolys, W, n), dtype=torch.int64, device=device)
                got = torch.full((npolys, W, n), -1, dtype=torch.int64, device=device)
                wp = (C.c_void_p * npolys)(*[want[k].data_ptr() for k in range(npolys)])
                gp = (C.c_void_p * npolys)(*[got[k].data_ptr() for k in range(npolys)])
                before = ctx.crt_fixups
                capi.check(lib, lib.gq_stage_reconstruct_dev(h, D, l, npolys, rhat.data_ptr(), wp, n, None))
                ctx.sync()
                assert ctx.crt_fixups > before, "the edge values must exercise the scratch list"
                capi.check(lib, lib.gq_stage_reconstruct_noalloc_dev(h, D, l, npolys, rhat.data_ptr(), gp, n, None))
                ctx.sync()
                assert torch.equal(got, want), f"call {call}: no-alloc reconstruct differs from gq_stage_reconstruct_dev"
        # a slice of the ring and a single polynomial work too
        with cm:
            got2 = torch.full((1, W, n // 2), -1, dtype=torch.int64, device=device)
            sl = rhat[0, :, : n // 2].contiguous()
            capi.check(lib, lib.gq_stage_reconstruct_noalloc_dev(h, D, l, 1, sl.data_ptr(),
                                                                 (C.c_void_p * 1)(got2[0].data_ptr()), n // 2, None))
            ctx.sync()
            assert torch.equal(got2[0], want[0][:, : n // 2])
        # bad arguments fail loudly
        assert lib.gq_stage_reconstruct_noalloc_dev(h, D, l, 5, rhat.data_ptr(), gp, n, None) != 0
        assert lib.gq_stage_reconstruct_noalloc_dev(h, D, l, npolys, rhat.data_ptr(), gp, n + 1, None) != 0
    finally:
        ctx.close()


def _check_limit(lib_path):
    """npolys * ncoef beyond the static array (262 144) is rejected with a clear error, nothing is launched"""
    ctx = g.HeContext(17, 1 << 120, 4, 1 << 40, lib_path=lib_path, logqub=120)
    try:
        lib, h = ctx._lib, ctx._h
        D = ctx.dims(ctx.L)[0]
        dummy = (C.c_void_p * 4)(*[8] * 4)               # never dereferenced: the size check comes first
        before = ctx.launch_count
        assert lib.gq_stage_reconstruct_noalloc_dev(h, D, ctx.L, 3, C.c_void_p(8), dummy, ctx.n, None) != 0
        assert b"exceeds the static scratch" in lib.gq_last_error()
        assert ctx.launch_count == before
    finally:
        ctx.close()


def test_no_alloc_reconstruct_emulated(emul_lib):
    """CPU tier: host logic and kernel indexing of the same entry point on the emulated build"""
    _check(emul_lib, "cpu", 6, 1 << 438, 1 << 50, 3, logqub=438)
    _check(emul_lib, "cpu", 5, 1 << 120, 1 << 40, 4, logqub=120)
    _check_limit(emul_lib)


@pytest.mark.gpu
def test_no_alloc_reconstruct_gpu(real_lib):
    """he_mul's shape at N=2^16 (3 polynomials over the 16 tensor limbs), the 4-polynomial maximum, and the limit"""
    dev = torch.device("cuda", 0)
    _check(real_lib, dev, 16, 1 << 438, 1 << 50, 3)
    _check(real_lib, dev, 16, 1 << 438, 1 << 50, 4)
    _check(real_lib, dev, 12, 1 << 109, 1 << 30, 2)
    _check_limit(real_lib)
