"""Limb-sharded multi-rank pipeline (gpqhe_b200/sharded.py).

CPU tier: world_size 2 and 4 over gloo, every rank driving the emulated build of the device
library; the sharded he_mul + he_rs must equal the single-rank C-ABI result and the oracle bit
for bit.  GPU tier: the same check on one GPU with world_size 1 (the exchange degenerates to
the identity but every gq_stage_* kernel path runs on the real library)."""
import os
import random
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _case(lib_path, logn, q, Delta, logqub, device):
    """runs on every rank: returns (sharded result gathered, single-rank result, oracle result)"""
    import gpqhe_b200 as g
    from gpqhe_b200.sharded import ShardedHe
    from oracle import pyoracle as o
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub or None)
    rng = random.Random(99)                       # same inputs on every rank
    n, W, l = ctx.n, ctx.W, ctx.L
    ql = oc.ql[l]
    polys = [[rng.randrange(ql) - ql // 2 for _ in range(n)] for _ in range(4)]
    nr = np.random.default_rng(5)
    p0 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
    p1 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
    evk = ctx.genevk(p0, p1)
    big = [g.ints_to_big(p, W) for p in polys]
    sh = ShardedHe(ctx, device=device)
    import contextlib
    on_ctx_stream = contextlib.nullcontext()
    if torch.device(device).type == "cuda":   # torch work must be ordered with the context's stream
        on_ctx_stream = torch.cuda.stream(torch.cuda.ExternalStream(ctx._lib.gq_ctx_stream(ctx._h), device=device))
    with on_ctx_stream:
        sl = [sh.shard(torch.from_numpy(b.view(np.int64)).to(device)) for b in big]
        c0, c1 = sh.he_mul(l, sl[0], sl[1], sl[2], sl[3], evk)
        c0, c1 = sh.he_rs(l, c0, c1)
        ctx.sync()
        got = [g.big_to_ints(sh.gather(c).cpu().numpy().view(np.uint64)) for c in (c0, c1)]
        # the planned variant (static buffers, CUDA graphs between the exchanges on a GPU)
        from gpqhe_b200.sharded import MulPlan
        plan = MulPlan(sh, l, evk, rescale=True)
        for k in range(4):
            plan.inp[k].copy_(sl[k])
        for _ in range(2):                       # twice: replays must not depend on stale state
            p0_, p1_ = plan.run()
        ctx.sync()
        got_plan = [g.big_to_ints(sh.gather(c).cpu().numpy().view(np.uint64)) for c in (p0_, p1_)]
        assert got_plan == got, "MulPlan differs from the eager sharded path"
    one = ctx.he_rs(ctx.he_mul(g.HeCt(l, big[0], big[1]), g.HeCt(l, big[2], big[3]), evk))
    single = [g.big_to_ints(one.c0), g.big_to_ints(one.c1)]
    e = o.he_rs(oc, o.he_mul(oc, {"l": l, "c0": polys[0], "c1": polys[1]}, {"l": l, "c0": polys[2], "c1": polys[3]}, (p0, p1)))
    stats = (sh.exchanges, sh.exchanged_bytes)
    ctx.close()
    return got, single, [e["c0"], e["c1"]], stats


def _worker(rank, world, port, lib_path, params, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        got, single, want, stats = _case(lib_path, *params, device="cpu")
        ok = got == want and single == want
        q_out.put((rank, ok, stats))
    except Exception as ex:  # pragma: no cover - reported to the parent
        q_out.put((rank, False, repr(ex)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,logn,logq,logDelta", [(2, 6, 438, 50), (4, 7, 438, 50), (2, 8, 120, 40)])
def test_sharded_he_mul_rs_gloo(emul_lib, world, logn, logq, logDelta):
    params = (logn, 1 << logq, 1 << logDelta, logq)
    ctxm = mp.get_context("spawn")
    q_out = ctxm.Queue()
    port = _free_port()
    procs = [ctxm.Process(target=_worker, args=(r, world, port, emul_lib, params, q_out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    # four all-to-alls per he_mul, none for he_rs
    assert all(st[0] == 12 for _, _, st in res), res    # eager op + two planned ops


@pytest.mark.gpu
def test_sharded_stage_entry_points_gpu(real_lib):
    """world_size 1 on the real library: the gq_stage_* kernels on limb ranges / coefficient
    slices give the same bits as the fused single-call path and the oracle at N=2^12."""
    dev = torch.device("cuda", 0)
    got, single, want, _ = _case(real_lib, 12, 1 << 109, 1 << 30, 0, dev)
    assert got == want and single == want


def _loopback_case(lib_path, logn, q, Delta, logqub, G, device, expect_fused=None):
    """G virtual ranks in ONE process through the peer-to-peer plan (scatter kernels + flags, or - for n > 2^12 -
    the exchanges fused into the producing / consuming kernels)."""
    import gpqhe_b200 as g
    from gpqhe_b200.sharded import P2PPlan
    from oracle import pyoracle as o
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub or None)
    rng = random.Random(7)
    n, W, l = ctx.n, ctx.W, ctx.L
    ql = oc.ql[l]
    if n > 4096:          # Python big integers are too slow up here: numpy words, checked against the single-GPU call
        polys = None
    else:
        polys = [[rng.randrange(ql) - ql // 2 for _ in range(n)] for _ in range(4)]
    nr = np.random.default_rng(11)
    p0 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
    p1 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
    evk = ctx.genevk(p0, p1)
    if polys is None:
        import parity_cases as pc
        words = [pc._rand_big(nr, n, W, ql.bit_length() - 1) for _ in range(4)]
        big = [torch.from_numpy(w.view(np.int64)).to(device) for w in words]
    else:
        big = [torch.from_numpy(g.ints_to_big(p, W).view(np.int64)).to(device) for p in polys]
    import contextlib
    cm = contextlib.nullcontext()
    if torch.device(device).type == "cuda":
        cm = torch.cuda.stream(torch.cuda.ExternalStream(ctx._lib.gq_ctx_stream(ctx._h), device=device))
    with cm:
        plans = P2PPlan.loopback(ctx, G, l, evk, rescale=True, device=device)
        if expect_fused is not None:
            assert all(pl.fused == expect_fused for pl in plans)
        nl = n // G
        for r, pl in enumerate(plans):
            for k in range(4):
                pl.inp[k].copy_(big[k][:, r * nl:(r + 1) * nl])
        for _ in range(2):                               # epochs must advance correctly
            P2PPlan.run_lockstep(plans)
        ctx.sync()
        assert not any(pl.failed() for pl in plans), "a flag wait timed out"
        got = [g.big_to_ints(torch.cat([pl.out[k] for pl in plans], dim=1).cpu().numpy().view(np.uint64)) for k in range(2)]
        for pl in plans:
            pl.close()
    if polys is None:
        want = ctx.he_mul_rs(g.HeCt(l, words[0], words[1]), g.HeCt(l, words[2], words[3]), evk)
        ctx.close()
        assert got == [g.big_to_ints(want.c0), g.big_to_ints(want.c1)]
        return
    e = o.he_rs(oc, o.he_mul(oc, {"l": l, "c0": polys[0], "c1": polys[1]}, {"l": l, "c0": polys[2], "c1": polys[3]}, (p0, p1)))
    ctx.close()
    assert got == [e["c0"], e["c1"]]


@pytest.mark.parametrize("G,logn,logq,logDelta", [(2, 6, 438, 50), (4, 7, 438, 50), (8, 8, 120, 40), (3, 6, 120, 40)])
def test_p2p_plan_loopback_emulated(emul_lib, G, logn, logq, logDelta):
    if (1 << logn) % G:
        pytest.skip("ring degree not divisible")
    _loopback_case(emul_lib, logn, 1 << logq, 1 << logDelta, logq, G, "cpu")


@pytest.mark.parametrize("G", [2])
def test_p2p_plan_fused_exchange_emulated(emul_lib, G, monkeypatch):
    """n = 2^13: the exchanges run inside the producing / consuming kernels (gq_p2p.h); same bits as one GPU, and
    the same bits with the fusion switched off"""
    _loopback_case(emul_lib, 13, 1 << 120, 1 << 40, 120, G, "cpu", expect_fused=True)
    monkeypatch.setenv("GPQHE_B200_SHARDED_PLAIN", "1")          # plain residues through the fused exchange
    _loopback_case(emul_lib, 13, 1 << 120, 1 << 40, 120, G, "cpu", expect_fused=True)
    monkeypatch.delenv("GPQHE_B200_SHARDED_PLAIN")
    monkeypatch.setenv("GPQHE_B200_P2P_UNFUSED", "1")
    _loopback_case(emul_lib, 13, 1 << 120, 1 << 40, 120, G, "cpu", expect_fused=False)


@pytest.mark.parametrize("G,logn,fused", [(2, 6, False)])
def test_p2p_plan_plain_and_prescaled_exchanges_emulated(emul_lib, G, logn, fused, monkeypatch):
    """the limb -> coefficient exchanges carry residues pre-scaled for the CRT (tensor-core CRT stages) by default and
    plain residues with GPQHE_B200_SHARDED_PLAIN=1: both forms give the bits of the single-GPU call / the oracle"""
    import gpqhe_b200 as g
    ctx = g.HeContext(logn, 1 << 120, 4, 1 << 40, lib_path=emul_lib, logqub=120)
    D1, D2 = ctx.dims(ctx.L)
    assert ctx._lib.gq_stage_prescale_ok(ctx._h, D1, ctx.L, 0) == 1 and ctx._lib.gq_stage_prescale_ok(ctx._h, D2, ctx.L, 1) == 1
    assert ctx._lib.gq_stage_prescale_ok(ctx._h, D2 - 1, ctx.L, 1) == 0      # not the key-switch basis of this level
    ctx.close()
    _loopback_case(emul_lib, logn, 1 << 120, 1 << 40, 120, G, "cpu", expect_fused=fused)
    monkeypatch.setenv("GPQHE_B200_SHARDED_PLAIN", "1")
    _loopback_case(emul_lib, logn, 1 << 120, 1 << 40, 120, G, "cpu", expect_fused=fused)


def test_p2p_timeout_is_sticky_and_poisons_peers(emul_lib):
    """A rank whose wait timed out (error flag set) must not let its peers consume its half-written slab: it signals
    a poison epoch, the peers' waits turn it into their own error flag, and run(check=True) raises on every rank."""
    import gpqhe_b200 as g
    from gpqhe_b200.sharded import P2PPlan
    ctx = g.HeContext(6, 1 << 120, 4, 1 << 40, lib_path=emul_lib, logqub=120)
    try:
        n = ctx.n
        nr = np.random.default_rng(3)
        key = lambda: np.concatenate([nr.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(ctx.dimevk)])
        evk = ctx.genevk(key(), key())
        plans = P2PPlan.loopback(ctx, 2, ctx.L, evk, rescale=True, device="cpu")
        for pl in plans:
            pl.inp.zero_()
        P2PPlan.run_lockstep(plans)
        assert not any(pl.failed() for pl in plans)
        # rank 1 skips its producer step of the next op: rank 0's wait times out (bounded spin) ...
        plans[0].seg(0)
        plans[0].seg(1)
        assert plans[0].failed() and not plans[1].failed()
        # ... and from then on rank 0 signals poison: rank 1 fails too instead of reading rank 0's stale slab
        plans[1].seg(0)
        plans[1].seg(1)
        plans[0].seg(2)                  # signals slot 2 with the poison value
        plans[1].seg(2)
        plans[1].seg(3)                  # waits on slot 2
        assert plans[1].failed()
        with pytest.raises(g.GpqheError):
            plans[1].run()
        for pl in plans:
            pl.close()
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("G", [1, 4, 8])
def test_p2p_plan_loopback_gpu(real_lib, G):
    _loopback_case(real_lib, 12, 1 << 109, 1 << 30, 0, G, torch.device("cuda", 0))


@pytest.mark.gpu
@pytest.mark.parametrize("G,logn", [(2, 16), (8, 16), (4, 14)])
def test_p2p_plan_fused_exchange_loopback_gpu(real_lib, G, logn):
    """the fused exchange at the headline ring (and a ragged limb split) on one real GPU, virtual ranks in lockstep"""
    if (1 << logn) % G:
        pytest.skip("ring degree not divisible")
    _loopback_case(real_lib, logn, 1 << 438, 1 << 50, 0, G, torch.device("cuda", 0), expect_fused=True)
