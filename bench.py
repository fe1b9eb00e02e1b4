#!/usr/bin/env python
"""bench.py - headline metric of BASELINE.json: HE mult + relinearise + rescale ops/s at
N=2^16 (logq=438, Delta=2^50: 16 tensor limbs, 24 key-switch limbs), one step = one
he_mul + he_rs on one synthetic ciphertext pair.  Also reports NTT limb-polys/s, the
roofline of the dominant kernel (the 4096-point NTT tile kernel), the reference's CPU
path timed on this box, clocks, and the end-to-end figure through the host C ABI.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  N>1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOGN, LOGQ, LOGDELTA, SLOTS = 16, 438, 50, 16
METRIC = "he_mul+he_rs ops/s (mult+relin+rescale, N=2^16, logq=438, Delta=2^50)"
WORKLOAD = "configs[2]: ciphertext mult + relinearisation key-switch + rescale at N=2^16, 16 tensor / 24 key-switch limbs"
POOL = 8


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 9 for i in range(4) if r[5 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synth_ct(rng, n, W, bits):
    """uniform on the centred range of q = 2^bits: random bits, sign-extended from bit bits-1."""
    import numpy as np
    w = rng.integers(0, 1 << 63, size=(W, n), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(W, n), dtype=np.uint64)
    sb = (bits - 1) % 64
    top = w[(bits - 1) // 64].astype(np.int64)
    top = (top << np.int64(63 - sb)) >> np.int64(63 - sb)
    w[(bits - 1) // 64] = top.astype(np.uint64)
    for k in range((bits - 1) // 64 + 1, W):
        w[k] = (top >> np.int64(63)).astype(np.uint64)
    return w


def synth_evk(rng, primes, n, dimevk):
    import numpy as np
    return np.stack([rng.integers(0, primes[d], size=n, dtype=np.uint64) for d in range(dimevk)])


# ------------------------------------------------------------------ reference CPU path
def cpu_reference_extras(budget_s=40.0, logn=LOGN):
    """the other CPU figures that stand beside ours - he_rot, one-limb ntt / invntt (the "NTT limbs/s" of
    BASELINE.json's metric) on the unmodified reference as shipped (-Og), and he_mul+he_rs / ntt on the same sources
    built -O3 (oracle/_ref/libgpqhe_O3.so, a courtesy build).  One thread: the reference has no threading."""
    import ctypes as C
    import random

    import numpy as np
    from oracle import refbind as rb
    out = {}
    q, Delta = 1 << LOGQ, 1 << LOGDELTA
    t_start = time.perf_counter()
    for variant, tag in (("libgpqhe.so", "Og"), ("libgpqhe_O3.so", "O3")):
        if not rb.available(variant) or time.perf_counter() - t_start > budget_s:
            continue
        R = rb.Ref(variant)
        R.hectx_init(logn, q, SLOTS, Delta)
        n, l = R.n, R.hectx.L
        rng, nrng = random.Random(2), np.random.default_rng(2)
        primes = R.primes()
        a = nrng.integers(0, primes[0], size=n, dtype=np.uint64)
        nd = R.node(0)
        ts = []
        for _ in range(5):
            b = a.copy(); t0 = time.perf_counter(); R.lib.ntt(rb.ptr(b), nd); ts.append(time.perf_counter() - t0)
        ti = []
        for _ in range(5):
            b = a.copy(); t0 = time.perf_counter(); R.lib.invntt(rb.ptr(b), nd); ti.append(time.perf_counter() - t0)
        out[f"ntt_limb_per_s_{tag}"] = 1.0 / statistics.median(ts)
        out[f"invntt_limb_per_s_{tag}"] = 1.0 / statistics.median(ti)
        ql = R.to_int(R.hectx.q[l])
        mk = lambda: [rng.randrange(ql) - ql // 2 for _ in range(n)]
        evk = R.new_evk(synth_evk(nrng, primes, n, R.hectx.dimevk).reshape(-1),
                        synth_evk(nrng, primes, n, R.hectx.dimevk).reshape(-1))
        if tag == "Og":
            ca = R.new_ct(l, mk(), mk())
            t0 = time.perf_counter(); R.lib.he_rot(C.byref(ca), 0, C.byref(evk)); out["he_rot_per_s_Og"] = 1.0 / (time.perf_counter() - t0)
        else:
            ca, cb = R.new_ct(l, mk(), mk()), R.new_ct(l, mk(), mk())
            cr = R.new_ct(l, [0] * n, [0] * n)
            t0 = time.perf_counter()
            R.lib.he_mul(C.byref(cr), C.byref(ca), C.byref(cb), C.byref(evk)); R.lib.he_rs(C.byref(cr))
            out["he_mul_rs_per_s_O3"] = 1.0 / (time.perf_counter() - t0)
    out["note"] = ("unmodified reference, 1 thread; Og = gcc -Og as shipped (Makefile:21), O3 = the same sources at -O3 "
                   "(courtesy); ntt/invntt: one limb-poly at N=2^%d (src/ntt.c:37,54), median of 5; he_rot: one call" % logn)
    return out


def ref_config(note):
    import math
    bits = LOGQ + 1
    return {"workload": WORKLOAD, "logn": LOGN, "logq": LOGQ, "logDelta": LOGDELTA, "level": LOGQ // LOGDELTA,
            "tensor_limbs": (2 * bits + LOGN) // 59 + 1, "keyswitch_limbs": 24, "words_per_coeff": (bits + 63) // 64,
            "note": note}


def cpu_reference(nmeas, budget_s=150.0, logn=LOGN):
    """Time he_mul + he_rs of the reference's own CPU implementation on this box.
    kind "reference": the unmodified library in oracle/_ref (gcc -Og as shipped, 1 thread:
    the reference has no threading).  kind "port": the oracle restatement, if the reference
    build or libgcrypt.so.20 is absent.  Returns (sec_per_op, ops_measured, kind, cores, sample)."""
    import ctypes as C
    import random

    import numpy as np
    q, Delta = 1 << LOGQ, 1 << LOGDELTA
    rng = random.Random(1)
    nrng = np.random.default_rng(1)
    from oracle import refbind as rb
    if rb.available():
        R = rb.Ref()
        R.hectx_init(logn, q, SLOTS, Delta)
        n, l = R.n, R.hectx.L
        ql = R.to_int(R.hectx.q[l])
        primes = R.primes()
        mk = lambda: [rng.randrange(ql) - ql // 2 for _ in range(n)]
        ca, cb = R.new_ct(l, mk(), mk()), R.new_ct(l, mk(), mk())
        cr = R.new_ct(l, [0] * n, [0] * n)
        evk = R.new_evk(synth_evk(nrng, primes, n, R.hectx.dimevk).reshape(-1),
                        synth_evk(nrng, primes, n, R.hectx.dimevk).reshape(-1))

        def op():
            R.lib.he_mul(C.byref(cr), C.byref(ca), C.byref(cb), C.byref(evk))
            R.lib.he_rs(C.byref(cr))
        kind = "reference"
        sample = "unmodified reference (oracle/_ref/libgpqhe.so, gcc -Og as shipped, libgcrypt 1.10), 1 thread"
    else:
        from oracle import pyoracle as o
        oc = o.Ctx(logn, q, SLOTS, Delta)
        n, l = oc.n, oc.L
        ql = oc.ql[l]
        mk = lambda: [rng.randrange(ql) - ql // 2 for _ in range(n)]
        c1, c2 = {"l": l, "c0": mk(), "c1": mk()}, {"l": l, "c0": mk(), "c1": mk()}
        ek = (synth_evk(nrng, oc.primes, n, oc.dimevk).reshape(-1), synth_evk(nrng, oc.primes, n, oc.dimevk).reshape(-1))

        def op():
            o.he_rs(oc, o.he_mul(oc, c1, c2, ek))
        kind = "port"
        sample = "oracle restatement (oracle.c limb layer + Python big integers), 1 thread"
    t0 = time.perf_counter(); op(); first = time.perf_counter() - t0
    times = [first]
    while len(times) < nmeas and sum(times) + first < budget_s:
        t0 = time.perf_counter(); op(); times.append(time.perf_counter() - t0)
    sec = statistics.median(times)
    return sec, len(times), kind, 1, f"{len(times)} x (he_mul+he_rs) at N=2^{logn}, logq={LOGQ}; {sample}"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sec, nm, kind, cores, sample = cpu_reference(max(1, args.steps))
    val = 1.0 / sec
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "ops/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": ref_config(f"CPU reference arm: {nm} op(s) measured (each ~{sec:.1f} s), median"),
            "cpu_baseline": {"value": val, "unit": "ops/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "ops/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- our arm
def run_ours(args):
    import ctypes as C
    # NCCL (and anything else native) may print banners on fd 1; the contract is ONE JSON line on
    # stdout, so everything but that line is sent to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    import numpy as np
    import torch
    import torch.distributed as dist

    import gpqhe_b200 as g
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gpqhe_b200 has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    ctx = g.HeContext(LOGN, 1 << LOGQ, SLOTS, 1 << LOGDELTA, device=local)
    lib, h = ctx._lib, ctx._h
    # measurement-only kernels (integer-pipe peaks, 32-bit NTT probe, tensor-core contraction probe): a separate library
    probe = C.CDLL(os.path.join(ROOT, "profiles", "probe", "libgqprobe.so"))
    probe.gqp_last_error.restype = C.c_char_p
    n, W, L = ctx.n, ctx.W, ctx.L
    D1, D2 = ctx.dims(L)
    rng = np.random.default_rng(1234 + rank)
    evk = ctx.genevk(synth_evk(rng, ctx.primes, n, ctx.dimevk), synth_evk(rng, ctx.primes, n, ctx.dimevk))

    host_pool = torch.empty((POOL, 4, W, n), dtype=torch.int64).pin_memory()
    hp = host_pool.numpy().view(np.uint64)
    for k in range(POOL):
        for j in range(4):
            hp[k, j] = synth_ct(rng, n, W, LOGQ)
    dpool = host_pool.to(dev, non_blocking=True)
    dout = torch.empty((2, W, n), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    poly = W * n * 8
    ext = torch.cuda.ExternalStream(lib.gq_ctx_stream(h), device=dev)

    def ck(rc):
        if rc != 0:
            raise RuntimeError(lib.gq_last_error().decode())

    def step(i):
        b = dpool.data_ptr() + (i % POOL) * 4 * poly
        o0, o1 = dout.data_ptr(), dout.data_ptr() + poly
        ck(lib.gq_he_mul_rs_dev(h, L, o0, o1, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h, None))

    # throughput mode: independent products round-robin over LANES compute lanes of this GPU (each lane =
    # stream + scratch, tables and key shared); the timed region forks from / joins into the context's stream
    LANES = max(1, args.lanes)
    ctx.set_lanes(LANES)
    lane_h = [ctx.lane(k) for k in range(LANES)]
    douts = [dout] + [torch.empty_like(dout) for _ in range(LANES - 1)]

    def lane_step(i):
        k = i % LANES
        b = dpool.data_ptr() + (i % POOL) * 4 * poly
        o0, o1 = douts[k].data_ptr(), douts[k].data_ptr() + poly
        ck(lib.gq_he_mul_rs_dev(lane_h[k], L, o0, o1, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h, None))

    def barrier():
        ctx.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps, warm, lanes=False):
        with torch.cuda.stream(ext):
            for i in range(warm):
                fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
            if lanes:
                ck(lib.gq_ctx_lanes_fork(h))
            for i in range(steps):
                fn(warm + i)
            if lanes:
                ck(lib.gq_ctx_lanes_join(h))
            e1.record()
        e1.synchronize()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed_regions(fn, steps, warm, lanes=False, min_total_s=0.5, max_regions=400):
        """the `steps`-sized timed region repeated until at least min_total_s of device time has been measured (at least
        3 regions): one region of 20 products is a few milliseconds.  Returns the per-region times in ms."""
        first = timed(fn, steps, warm, lanes)
        want = int(min(max_regions, max(3, min_total_s * 1e3 / max(first, 1e-3) + 1)))
        if world > 1:
            t = torch.tensor([want], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            want = int(t.item())
        return [first] + [timed(fn, steps, 0, lanes) for _ in range(want - 1)]

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launch_count
    regions = timed_regions(lane_step, args.steps, max(args.warmup, LANES), lanes=True)
    ms = statistics.median(regions)
    launches = (ctx.launch_count - l0) * args.steps // (args.steps * len(regions) + max(args.warmup, LANES))
    # every lane's last product equals the single-stream result of the same inputs
    lanes_ok = True
    for k in range(LANES):
        i_last = max(i for i in range(args.steps) if i % LANES == k)        # the last region started at step index 0
        got = douts[k].clone()
        with torch.cuda.stream(ext):
            step(i_last)
        barrier()
        lanes_ok = lanes_ok and bool(torch.equal(got, dout if k == 0 else douts[0]))
    ms_single = statistics.median(timed_regions(step, args.steps, args.warmup, min_total_s=0.2))
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms / args.steps
    value = world * args.steps / (ms * 1e-3)

    # ---- NTT sub-benchmark + roofline of the dominant kernel (4096-point tile kernel) ----
    NB = 16                                   # 16 polys x 24 limbs x 512 KiB = 192 MiB > L2
    dn = torch.empty((NB, D2, n), dtype=torch.int64, device=dev)
    dn.random_(0, 1 << 59)
    nlp = NB * D2

    def ntt_fn(flags, exact=False):
        if exact:
            return lambda i: ck(lib.gq_ntt_batch_dev(h, 0, D2, NB, dn.data_ptr(), flags, None))
        return lambda i: ck(lib.gq_stage_ntt_dev(h, 0, D2, NB, dn.data_ptr(), flags, None))
    reps = 10
    ms_f = timed(ntt_fn(0), reps, 3) / reps
    dn.random_(0, 1 << 59)                       # the inverse wants canonical input again
    ms_i = timed(ntt_fn(1), reps, 3) / reps
    dn.random_(0, 1 << 59)
    ms_tile = timed(ntt_fn(2), reps, 3) / reps   # profiling flag: forward tile kernel only
    dn.random_(0, 1 << 59)
    ms_fx = timed(ntt_fn(0, True), reps, 3) / reps   # exact-semantics kernels behind gq_ntt()
    ntt_bytes = 2 * n * 8 * nlp
    pk, pk_src = peaks()
    hbm = float(pk.get("hbm_gbs", 6650.0))
    achieved = ntt_bytes / (ms_tile * 1e-3) / 1e9
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tr.get("k_nttl_fwd_tile<12>", {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_nttl_fwd_tile<12> (last 12 of 16 forward NTT stages on 4096-point tiles)",
                "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": traffic,
                "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({pk_src})",
                "algorithmic_bytes_per_launch": ntt_bytes, "kernel_ms_per_launch": ms_tile,
                "note": "this kernel is bound by the integer-multiply pipe, not HBM: ncu shows sm__pipe_fmaheavy_cycles_active "
                        "61 % / alu 46 % / dram 13 % (profiles/r02_call2_full.md; SASS mix in profiles/r02_sass_fwd_tile.md); the HBM fraction is reported as the contract asks"}
    # integer-pipe roofline: what the tile kernel achieves in butterflies/s against the measured
    # register-resident butterfly rate of this GPU (gq_peak.cu)
    bf_peak, mw_peak = C.c_double(0), C.c_double(0)
    p0c = ctx.prime_consts(0)
    w0 = int(p0c[0]) // 3
    if probe.gqp_bench_intpipe(local, C.c_uint64(int(p0c[0])), C.c_uint64(w0), C.c_uint64((w0 << 64) // int(p0c[0])),
                               C.byref(bf_peak), C.byref(mw_peak)) != 0:
        raise RuntimeError(probe.gqp_last_error().decode())
    tile_bf = nlp * (n // 2) * 12 / (ms_tile * 1e-3)
    MUL_PER_BF = 8.0          # 6 IMAD.WIDE + 4 low-half IMAD per lazy Shoup butterfly, a low-half product counted 0.5
    roofline["int_pipe"] = {"bound": "integer multiply pipe (fmaheavy)", "unit": "G 32x32-multiply-equivalents/s",
                            "achieved": tile_bf * MUL_PER_BF / 1e9, "peak": mw_peak.value / 1e9,
                            "frac": tile_bf * MUL_PER_BF / mw_peak.value if mw_peak.value else None,
                            "mul32_per_butterfly": MUL_PER_BF,
                            "peak_source": "measured in this run: k_peak_madwide (dependent-free IMAD.WIDE.U32, profiles/probe)",
                            "algorithmic_butterflies_per_launch": nlp * (n // 2) * 12,
                            "butterfly_ceiling": {"achieved_Gbf": tile_bf / 1e9, "peak_Gbf": bf_peak.value / 1e9,
                                                  "frac": tile_bf / bf_peak.value if bf_peak.value else None,
                                                  "note": "secondary: against k_peak_butterfly, this kernel's own butterfly "
                                                          "instruction sequence run from registers (memory / barrier overhead only)"}}
    ntt = {"fwd_limb_ntt_per_s": world * nlp / (ms_f * 1e-3), "inv_limb_ntt_per_s": world * nlp / (ms_i * 1e-3),
           "fwd_ms_per_batch": ms_f, "inv_ms_per_batch": ms_i, "tile_kernel_ms_per_batch": ms_tile,
           "exact_fwd_limb_ntt_per_s": world * nlp / (ms_fx * 1e-3),
           "batch_limb_polys": nlp, "fwd_GBps_algorithmic": ntt_bytes / (ms_f * 1e-3) / 1e9,
           "kernels": "lazy (internal pipelines); exact_* = the ntt()-semantics kernels"}
    del dn
    # measurement only (gq_ntt32.cu, not on the product path): the same two-kernel forward transform over 30-bit
    # primes on 32-bit words, the building block of the internal 31-bit-class RNS basis planned in DESIGN.md
    # section 4; 720 limb-polys = 184 MB against the 192 MB of the 64-bit batch
    n32_rate, n32_bad = C.c_double(0), C.c_int(-1)
    ctx.sync()
    if probe.gqp_bench_ntt32(local, LOGN, 30, 24, 5, C.byref(n32_rate), C.byref(n32_bad)) != 0:
        raise RuntimeError(probe.gqp_last_error().decode())
    ntt["ntt32_probe"] = {"fwd_limb_ntt_per_s": world * n32_rate.value, "batch_limb_polys": 720, "prime_bits": 30,
                          "mismatches_vs_host": n32_bad.value,
                          "limbs_needed_vs_64bit": "30 (tensor basis) / 48 (key-switch basis) against 16 / 24",
                          "note": "not used by any product path yet; limb-NTT for limb-NTT it has to be compared with "
                                  "fwd_limb_ntt_per_s x 16/30"}
    # measurement only (profiles/probe/gqp_imma.cu): rns_decompose as a u8 tensor-core contraction (mma.sync m16n8k32,
    # exact int32 accumulation) in the shape of he_mul's first decomposition, next to the integer-pipe kernel's time
    imma = None
    try:
        raw, rate, msl, bad = C.c_double(0), C.c_double(0), C.c_double(0), C.c_int(-1)
        pr16 = (C.c_uint64 * 16)(*ctx.primes[:16])
        if probe.gqp_bench_imma_raw(local, C.byref(raw)) == 0 and \
           probe.gqp_bench_imma_decompose(local, LOGN, 16, pr16, 10, C.byref(rate), C.byref(msl), C.byref(bad)) == 0:
            imma = {"raw_u8_Tmac_per_s": raw.value / 1e12, "decompose_4x16_us": msl.value * 1e3,
                    "residues_per_s": rate.value, "mismatches_vs_host": bad.value,
                    "note": "prototype: A fragments straight from the word-major coefficients, B (2^(8j) mod p, byte split) "
                            "from shared memory, thread-local recombination of the 8 byte columns + pseudo-Mersenne fold; "
                            "compare with the 'decompose x4 -> D1' row of `stages`"}
        else:
            imma = {"error": probe.gqp_last_error().decode()}
    except Exception as ex:
        imma = {"error": repr(ex)}

    # ---- every stage kernel of one he_mul+he_rs, timed alone in the shapes the op uses (gq_stage_* = the
    # kernels the op launches), against both rooflines: algorithmic bytes / HBM peak and 32x32-multiply
    # work / measured IMAD.WIDE peak.  Multiplies per unit are the algorithmic counts of DESIGN.md section 4
    # (a 32x32->64 product = 1, a low-half 32x32 product = 0.5).
    def stage_table():
        VP = C.c_void_p
        b = dpool.data_ptr()
        xc = torch.empty((4, D1, n), dtype=torch.int64, device=dev)
        y = torch.empty((3, D1, n), dtype=torch.int64, device=dev)
        d = torch.empty((3, W, n), dtype=torch.int64, device=dev)
        x2 = torch.empty((1, D2, n), dtype=torch.int64, device=dev)
        y2 = torch.empty((2, D2, n), dtype=torch.int64, device=dev)
        o = torch.empty((2, W, n), dtype=torch.int64, device=dev)
        in4 = (VP * 4)(*[b + j * poly for j in range(4)])
        in1 = (VP * 1)(d[2].data_ptr())
        out3 = (VP * 3)(*[d[j].data_ptr() for j in range(3)])
        # the forms the op itself uses: inverse NTT ending in the CRT pre-multiplication, CRT stages on the tensor cores
        pre1 = bool(lib.gq_stage_prescale_ok(h, D1, L, 0))
        pre2 = bool(lib.gq_stage_prescale_ok(h, D2, L, 1))
        imma_peak = (imma or {}).get("raw_u8_Tmac_per_s")

        def tensor_inv(i):
            if pre1:
                ck(lib.gq_stage_tensor_inv_pre_dev(h, 0, D1, D1, y.data_ptr(), xc.data_ptr(), None))
            else:
                ck(lib.gq_stage_tensor_inv_dev(h, 0, D1, y.data_ptr(), xc.data_ptr(), None))

        def reconstruct(i):
            f = lib.gq_stage_reconstruct_pre_dev if pre1 else lib.gq_stage_reconstruct_dev
            ck(f(h, D1, L, 3, y.data_ptr(), out3, n, None))

        def evk_inv(i):
            if pre2:
                ck(lib.gq_stage_mul_evk_inv_pre_dev(h, 0, D2, D2, y2.data_ptr(), x2.data_ptr(), evk._h, None))
            else:
                ck(lib.gq_stage_mul_evk_inv_dev(h, 0, D2, y2.data_ptr(), x2.data_ptr(), evk._h, None))

        def ks_epilogue(i):
            if pre2:
                ck(lib.gq_stage_ks_epilogue_pre_dev(h, L, y2.data_ptr(), d[0].data_ptr(), d[1].data_ptr(), o[0].data_ptr(), o[1].data_ptr(), n, 1, None))
            else:
                ck(lib.gq_stage_ks_epilogue_rs_dev(h, L, y2.data_ptr(), d[0].data_ptr(), d[1].data_ptr(), o[0].data_ptr(), o[1].data_ptr(), n, None))
        # u8 multiply-accumulates per coefficient of the tensor-core CRT: 32 K-bytes per chunk of 4 limbs x 80 columns
        tc_mac1 = 32 * ((D1 + 3) // 4) * 80
        tc_mac2 = 32 * ((D2 + 3) // 4 + (ctx.dim + 3) // 4) * 80
        # decomposition: the tensor-core form for coefficients of at most 7 words over pseudo-Mersenne primes (64 K-bytes x
        # 8 byte columns per residue), else 28-bit digits on the integer pipe (15 digits x 2 multiply-adds + 4 in the fold)
        dec_tc = W <= 7 and pre1 and not os.environ.get("GPQHE_B200_NO_TC_CRT")
        dec_kern = "k_decompose_tc (u8 mma.sync contraction)" if dec_tc else "k_decompose28<7, pm>"
        dec_w = 512 if dec_tc else 34.0
        stages = [
            ("decompose x4 -> D1", dec_kern, lambda i: ck(lib.gq_stage_decompose_dev(h, 0, D1, 4, in4, W, n, xc.data_ptr(), None)),
             4 * D1 * n, dec_w, 4 * W * n * 8 + 4 * D1 * n * 8),
            ("forward NTT, 4 x D1 limb-polys", "k_nttl_fwd_strided<4> + k_nttl_fwd_tile<12>", lambda i: ck(lib.gq_stage_ntt_dev(h, 0, D1, 4, xc.data_ptr(), 0, None)),
             4 * D1 * (n // 2) * LOGN, 8.0, 2 * 2 * 4 * D1 * n * 8),
            ("tensor product + inverse NTT, 3 x D1", "k_nttl_inv_tile<12, TENSOR> + k_nttl_inv_strided<4>",
             tensor_inv,
             3 * D1 * (n // 2) * LOGN, 8.0 + 4 * 7.0 * 2 / (3 * LOGN), (4 + 3 + 2 * 3) * D1 * n * 8),
            ("reconstruct x3 over D1", "k_crt_tc<K5> (u8 mma.sync contraction)" if pre1 else "k_reconstruct_fast<7>", reconstruct,
             3 * n, (tc_mac1 if pre1 else D1 * 38.0 + 4 * W), 3 * D1 * n * 8 + 3 * W * n * 8),
            ("decompose d2 -> D2", dec_kern, lambda i: ck(lib.gq_stage_decompose_dev(h, 0, D2, 1, in1, W, n, x2.data_ptr(), None)),
             D2 * n, dec_w, W * n * 8 + D2 * n * 8),
            ("forward NTT, D2", "k_nttl_fwd_*", lambda i: ck(lib.gq_stage_ntt_dev(h, 0, D2, 1, x2.data_ptr(), 0, None)),
             D2 * (n // 2) * LOGN, 8.0, 2 * 2 * D2 * n * 8),
            ("multiply by rlk + inverse NTT, 2 x D2", "k_nttl_inv_tile<12, EVK> + k_nttl_inv_strided<4>",
             evk_inv,
             2 * D2 * (n // 2) * LOGN, 8.0 + 2 * 7.0 * 2 / (2 * LOGN), (3 + 2 + 2 * 2) * D2 * n * 8),
            ("key-switch epilogue + rescale", "k_crt_tc<K6> (u8 mma.sync contraction, he_rs fused)" if pre2 else "k_ks_epilogue_fast<8, 7> (he_rs fused)",
             ks_epilogue,
             2 * n, (tc_mac2 if pre2 else D2 * 42.0 + ctx.dim * 40.0 + 128 + 64), 2 * D2 * n * 8 + 4 * W * n * 8),
        ]
        with torch.cuda.stream(ext):
            for st in stages:          # one pass in order so that every stage sees real data
                st[2](0)
        barrier()
        rows, total = [], 0.0
        reps = 20
        for name, kern, fn, units, mulw, nbytes in stages:
            t = timed(fn, reps, 3) / reps
            total += t
            fh = nbytes / (t * 1e-3) / 1e9 / hbm
            if "_tc" in kern:            # tensor-core rows: the work is u8 MACs against the measured mma.sync u8 peak
                ft = units * mulw / (t * 1e-3) / (imma_peak * 1e12) if imma_peak else None
                rows.append({"stage": name, "kernel": kern, "us": t * 1e3, "units": units, "u8_mac_per_unit": mulw,
                             "algorithmic_bytes": nbytes, "frac_tensor_u8": ft, "frac_hbm": fh,
                             "bound": "latency (integer epilogue); nearest roof: " + ("hbm" if fh >= (ft or 0) else "tensor")})
                continue
            fi = units * mulw / (t * 1e-3) / mw_peak.value if mw_peak.value else None
            rows.append({"stage": name, "kernel": kern, "us": t * 1e3, "units": units, "mul32_per_unit": mulw,
                         "algorithmic_bytes": nbytes, "frac_int_pipe": fi, "frac_hbm": fh,
                         "bound": "int" if (fi or 0) >= fh else "hbm"})
        for r in rows:
            r["share"] = r["us"] / (total * 1e3)
        return {"rows": rows, "sum_us": total * 1e3,
                "peaks": {"madwide_Gops": mw_peak.value / 1e9, "hbm_GBps": hbm},
                "note": "the op's own launches (the stage-level entry points call the kernels gq_he_mul_rs_dev launches, in the "
                        "pre-scaled forms when gq_stage_prescale_ok: the inverse NTT folds (P/p_d)^-1 into its final scaling and the "
                        "CRT stages run as u8 tensor-core contractions, gq_crt_tc.cu); each stage timed alone on one stream (20 launches back to back, L2-warm as inside the op); "
                        "frac_int_pipe = 32x32-multiply work / measured IMAD.WIDE peak of this GPU (gq_bench_intpipe), "
                        "frac_hbm = algorithmic bytes / measured HBM copy peak; the NTT rows use 8 multiply-equivalents "
                        "per butterfly (6 wide + 4 low)"}
    stages = stage_table()

    # ---- configs[3]: rotation sweep, he_rot = Galois permutation + key switch, rot = 0..NROT-1 with
    # one switching key per rotation (the he_gemv workload: slots=16 rotations)
    def rotation_sweep(NROT=16):
        rks = [ctx.genevk(synth_evk(rng, ctx.primes, n, ctx.dimevk), synth_evk(rng, ctx.primes, n, ctx.dimevk))
               for _ in range(NROT)]
        drot = dpool[:, :2].clone()                   # [POOL][2][W][n] ciphertexts rotated in place

        def rot_step(i):
            b = drot.data_ptr() + (i % POOL) * 2 * poly
            ck(lib.gq_he_rot_dev(h, L, i % NROT, b, b + poly, rks[i % NROT]._h, None))

        def rot_lane_step(i):            # independent rotations round-robin over the lanes; a buffer stays on its lane
            k = i % LANES
            b = drot.data_ptr() + (k + LANES * ((i // LANES) % max(1, POOL // LANES))) * 2 * poly
            ck(lib.gq_he_rot_dev(lane_h[k], L, i % NROT, b, b + poly, rks[i % NROT]._h, None))
        nrot = max(args.steps, NROT)
        ms_rot1 = timed(rot_step, nrot, args.warmup)
        ms_rot = timed(rot_lane_step, nrot, max(args.warmup, LANES), lanes=True)
        # the same 16 rotations of ONE ciphertext with the decomposition + forward NTT of c1 shared
        hout_r = torch.empty((NROT, 2, W, n), dtype=torch.int64, device=dev)
        VP = C.c_void_p * NROT
        o0 = VP(*[hout_r.data_ptr() + k * 2 * poly for k in range(NROT)])
        o1 = VP(*[hout_r.data_ptr() + k * 2 * poly + poly for k in range(NROT)])
        keys = VP(*[rk._h for rk in rks])
        rots = (C.c_uint * NROT)(*range(NROT))
        src = dpool[0, :2].clone()

        def hoisted_step(i):
            ck(lib.gq_he_rot_hoisted_dev(h, L, NROT, rots, o0, o1, src.data_ptr(), src.data_ptr() + poly, keys, None))
        reps = max(2, args.steps // 4)
        ms_h = timed(hoisted_step, reps, 2)
        # spot check: hoisted rotation 3 == he_rot by 3
        chk = src.clone()
        with torch.cuda.stream(ext):
            ck(lib.gq_he_rot_dev(h, L, 3, chk.data_ptr(), chk.data_ptr() + poly, rks[3]._h, None))
            ctx.sync()
            same = bool(torch.equal(chk, hout_r[3]))
        # configs[3] as the reference composes it: he_gemv with slots = 16 (baby-step/giant-step 4 x 4:
        # 4 hoisted + 4 plain rotations, 16 he_mulpt, 15 he_add, 1 he_rs), device resident
        import math
        gdim = int((LOGQ + 1 + math.log2(float(1 << LOGDELTA)) + LOGN) / 59 + 1)       # he-mult.c:168
        gpts = torch.zeros((NROT, W, n), dtype=torch.int64, device=dev)
        gpts[:, 0, :].random_(-(1 << LOGDELTA), 1 << LOGDELTA)
        gpts[:, 1:, :] = (gpts[:, :1, :] >> 63)                                          # sign extension
        gout = torch.empty((2, W, n), dtype=torch.int64, device=dev)
        ptsv = VP(*[gpts.data_ptr() + k * poly for k in range(NROT)])

        # the synthetic diagonals are below 2^LOGDELTA in magnitude: with the bound he_gemv accumulates each giant step
        # in the NTT domain (one inverse NTT + CRT per giant step instead of one per he_mulpt)
        ck(lib.gq_ctx_set_plaintext_bits(h, LOGDELTA + 1))

        def gemv_step(i):
            ck(lib.gq_he_gemv_dev(h, L, 4, 4, gdim, gout.data_ptr(), gout.data_ptr() + poly, src.data_ptr(),
                                  src.data_ptr() + poly, ptsv, keys, None))
        ms_g = timed(gemv_step, reps, 2)
        gemv = {"gemv_per_s": world * reps / (ms_g * 1e-3), "ms_per_gemv": ms_g / reps, "slots": 16, "mulpt_limbs": gdim,
                "note": "gq_he_gemv_dev: the ciphertext arithmetic of he_gemv (src/he-algo.c:47-90) with pre-encoded "
                        f"diagonals, every intermediate ciphertext in HBM, giant steps over {LANES} lanes"}
        # configs[4]: the linear-transform half of bootstrapping as the reference composes it (he_coeff2slot,
        # src/he-bootstrap.c:359-409): he_conj + 4 he_gemv + 2 x (he_add, he_mulpt, he_rs), device resident
        cdim = int((LOGQ - LOGDELTA + 1 + math.log2(float(1 << LOGDELTA)) + LOGN) / 59 + 1)
        cpts = gpts.repeat(4, 1, 1)
        cptsv = (C.c_void_p * (4 * NROT))(*[cpts.data_ptr() + k * poly for k in range(4 * NROT)])
        cconst = gpts[0].clone()
        cout = torch.empty((4, W, n), dtype=torch.int64, device=dev)

        def c2s_step(i):
            ck(lib.gq_he_coeff2slot_dev(h, L, 4, 4, gdim, cdim, cout.data_ptr(), cout.data_ptr() + poly,
                                        cout.data_ptr() + 2 * poly, cout.data_ptr() + 3 * poly, src.data_ptr(),
                                        src.data_ptr() + poly, cptsv, cconst.data_ptr(), rks[1]._h, keys, None))
        ms_c = timed(c2s_step, reps, 2)
        c2s = {"coeff2slot_per_s": world * reps / (ms_c * 1e-3), "ms_per_coeff2slot": ms_c / reps, "slots": 16,
               "ops": "1 he_conj, 4 he_gemv (each 4 hoisted + 4 plain rotations, 16 he_mulpt, 15 he_add, 1 he_rs), 2 he_add, "
                      "2 he_mulpt, 2 he_rs",
               "note": "configs[4], CoeffToSlot (he_coeff2slot, src/he-bootstrap.c:359-409): gq_he_coeff2slot_dev, every "
                       "intermediate in HBM"}
        # configs[4], the rest of the chain the reference's he_bootstrap composes (src/he-bootstrap.c:477-545):
        # EvalMod = he_rlsin on both halves (he_exp: he_mulpt, he_rs, degree-7 Taylor polynomial with 7 he_mul,
        # then he_conj, he_sub, he_mulpt, he_rs), SlotToCoeff = 2 he_gemv + he_add.  L = 8 leaves iter = 0:
        # coeff2slot 8 -> 6, rlsin 6 -> 1, slot2coeff 1 -> 0.  Every intermediate ciphertext stays in HBM.
        dim_at = lambda lv: int((LOGQ - (L - lv) * LOGDELTA + 1 + math.log2(float(1 << LOGDELTA)) + LOGN) / 59 + 1)
        epts = gpts[:9].clone()
        epts[1:8, :, 1:] = 0                                                             # he_const_pt: coefficient 0 only
        eptsv = (C.c_void_p * 9)(*[epts.data_ptr() + k * poly for k in range(9)])
        sptsv = (C.c_void_p * (2 * NROT))(*[cpts.data_ptr() + k * poly for k in range(2 * NROT)])
        eout = torch.empty((4, W, n), dtype=torch.int64, device=dev)
        bout = torch.empty((2, W, n), dtype=torch.int64, device=dev)

        def evalmod_step(i, inp=None):
            x = cout if inp is None else inp
            for part in range(2):
                a = x.data_ptr() + 2 * part * poly
                o = eout.data_ptr() + 2 * part * poly
                ck(lib.gq_he_rlsin_dev(h, L - 2, 0, dim_at(L - 2), dim_at(L - 3), dim_at(L - 6), o, o + poly, a, a + poly,
                                       eptsv, evk._h, rks[1]._h, None))
        ms_e = timed(evalmod_step, reps, 2)

        def boot_step(i):
            c2s_step(i)
            evalmod_step(i)
            ck(lib.gq_he_slot2coeff_dev(h, L - 7, 4, 4, dim_at(L - 7), bout.data_ptr(), bout.data_ptr() + poly,
                                        eout.data_ptr(), eout.data_ptr() + poly, eout.data_ptr() + 2 * poly,
                                        eout.data_ptr() + 3 * poly, sptsv, keys, None))
        ms_b = timed(boot_step, reps, 2)
        evalmod = {"rlsin_per_s": world * reps * 2 / (ms_e * 1e-3), "ms_per_rlsin": ms_e / (reps * 2), "iter": 0,
                   "levels": f"{L - 2} -> {L - 7}",
                   "ops": "per he_rlsin: 7 he_mul, 6 he_mulpt, 13 he_rs/he_moddown, 4 he_addpt, 3 he_add/he_sub, 1 he_conj",
                   "note": "gq_he_rlsin_dev (src/he-bootstrap.c:444-475 over he_exp, src/he-algo.c:364-458): EvalMod of "
                           "he_bootstrap on one half; every intermediate ciphertext in HBM"}
        bootstrap = {"chains_per_s": world * reps / (ms_b * 1e-3), "ms_per_chain": ms_b / reps, "slots": 16,
                     "levels": f"{L} -> 0",
                     "note": "he_coeff2slot -> 2 x he_rlsin (iter = 0) -> he_slot2coeff, the ciphertext arithmetic of "
                             "he_bootstrap (src/he-bootstrap.c:477-545) as far as L = 8 levels reach, device resident "
                             "end to end (synthetic encoded plaintexts; the reference's own he_bootstrap is unfinished, "
                             "SURVEY fact 9)"}
        return {"he_rot_per_s": world * nrot / (ms_rot * 1e-3), "ms_per_rot": ms_rot / nrot, "rotations": NROT, "gemv": gemv,
                "coeff2slot": c2s, "evalmod": evalmod, "bootstrap_chain": bootstrap,
                "keyswitch_limbs": D2, "note": f"he_rot at l=L, 16 distinct rotation keys (384 MiB) cycled, independent rotations over {LANES} lanes",
                "single_stream": {"he_rot_per_s": world * nrot / (ms_rot1 * 1e-3), "ms_per_rot": ms_rot1 / nrot},
                "hoisted": {"he_rot_per_s": world * reps * NROT / (ms_h * 1e-3), "ms_per_rot": ms_h / (reps * NROT),
                            "bit_exact_vs_he_rot": same,
                            "note": "16 rotations of one ciphertext per call (gq_he_rot_hoisted_dev): rns_decompose + "
                                    "forward NTT of c1 once, Galois map applied as a permutation of the NTT slots"}}
    rot_sweep = None if args.no_rot else rotation_sweep()

    # ---- limb-sharded single-stream pipeline: all ranks cooperate on ONE he_mul+he_rs;
    # ciphertexts stay coefficient-sharded, four all-to-alls per op (gpqhe_b200/sharded.py)
    def all_ok(flag):
        if world == 1:
            return bool(flag)
        t = torch.tensor([1 if flag else 0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    sharded = None
    pdl_before_sharded = None
    if not args.no_sharded:
        pdl_before_sharded = lib.gq_set_pdl(1)          # the sharded drivers switch it off above two ranks (sharded.py)
        lib.gq_set_pdl(pdl_before_sharded)
        from gpqhe_b200.sharded import MulPlan, ShardedHe
        with torch.cuda.stream(ext):
            sh = ShardedHe(ctx, device=dev)
            srng = np.random.default_rng(4321)              # the same ciphertexts on every rank
            full = [torch.from_numpy(synth_ct(srng, n, W, LOGQ).view(np.int64)).to(dev) for _ in range(4)]
            sl = [sh.shard(t) for t in full]
            evk_s = evk
            if world > 1:                                   # the key must be the same on every rank too
                evk_s = ctx.genevk(synth_evk(srng, ctx.primes, n, ctx.dimevk), synth_evk(srng, ctx.primes, n, ctx.dimevk))

            def sstep(i):
                c0, c1 = sh.he_mul(L, sl[0], sl[1], sl[2], sl[3], evk_s)
                sh.he_rs(L, c0, c1)
            # the same op on ONE GPU (every rank holds the full inputs): what the sharded results must equal, bit for bit
            single = torch.empty((2, W, n), dtype=torch.int64, device=dev)
            ck(lib.gq_he_mul_rs_dev(h, L, single[0].data_ptr(), single[1].data_ptr(), full[0].data_ptr(), full[1].data_ptr(),
                                    full[2].data_ptr(), full[3].data_ptr(), evk_s._h, None))
            ctx.sync()
            want_sl = [sh.shard(single[0]), sh.shard(single[1])]
        ex0, by0 = sh.exchanges, sh.exchanged_bytes
        ms_eager = timed(sstep, args.steps, args.warmup)
        a2a_per_op = (sh.exchanges - ex0) // (args.steps + args.warmup)
        bytes_per_op = (sh.exchanged_bytes - by0) // (args.steps + args.warmup)
        # the eager pipeline is launch-bound (about 15 small launches per exchange from Python):
        # MulPlan records the work between two exchanges as one CUDA graph (5 graphs + 4 eager
        # all-to-alls per op, static buffers)
        ms_s, graphed, plan = ms_eager, False, None
        try:
            barrier()
            with torch.cuda.stream(ext):
                plan = MulPlan(sh, L, evk_s, rescale=True, graphs=True)
                for k in range(4):
                    plan.inp[k].copy_(sl[k])
            ms_s, graphed = timed(lambda i: plan.run(), args.steps, args.warmup), True
        except Exception as ex:   # keep the eager number and say why
            print(f"bench.py: graph-planned sharded op failed: {ex!r}", file=sys.stderr)
        nccl_exact = None
        if plan is not None:
            with torch.cuda.stream(ext):
                ctx.sync()
                nccl_exact = bool(torch.equal(plan.out[0], want_sl[0]) and torch.equal(plan.out[1], want_sl[1]))
        sharded = {"ops_per_s": args.steps / (ms_s * 1e-3), "ms_per_op": ms_s / args.steps, "ranks": world,
                   "bit_exact_vs_single_gpu": None if nccl_exact is None else all_ok(nccl_exact),
                   "scaling": "strong", "all_to_all_per_op": a2a_per_op, "bytes_sent_per_rank_per_op": bytes_per_op,
                   "cuda_graphs_between_exchanges": graphed, "eager_ms_per_op": ms_eager / args.steps,
                   "note": "one ciphertext stream sharded by RNS limb (NTT domain) and by coefficient (CRT domain); "
                           "NCCL all_to_all_single between the domains"}

        # peer-to-peer variant: the re-layouts are scatter kernels writing into the peers' slabs over
        # NVLink (gq_p2p.cu), epoch flags instead of NCCL, the whole op one CUDA graph per rank
        if not args.no_p2p and plan is not None:
            from gpqhe_b200.sharded import P2PPlan

            def p2p_leg(unfused):
                """one P2P plan: build, check against the single-GPU result, time eagerly and as one CUDA graph.
                unfused=True keeps the round-1 exchange (scatter + signal + wait launches) for comparison."""
                pp, ok, match = None, True, False
                if unfused:
                    os.environ["GPQHE_B200_P2P_UNFUSED"] = "1"
                try:
                    with torch.cuda.stream(ext):
                        pp = P2PPlan.build(sh, L, evk_s, rescale=True)
                        for k in range(4):
                            pp.inp[k].copy_(sl[k])
                        pp.run(check=False)
                        ctx.sync()
                        match = bool(torch.equal(pp.out[0], want_sl[0]) and torch.equal(pp.out[1], want_sl[1])) and not pp.failed()
                except Exception as ex:
                    print(f"bench.py: p2p plan failed: {ex!r}", file=sys.stderr)
                    ok = False
                finally:
                    os.environ.pop("GPQHE_B200_P2P_UNFUSED", None)
                if not all_ok(ok):
                    return {"error": "plan construction failed on some rank (see stderr)"}
                l0 = ctx.launch_count
                ms_pe = timed(lambda i: pp.run(check=False), args.steps, args.warmup)
                res = {"fused_exchange": bool(pp.fused), "bit_exact_vs_single_gpu": all_ok(match),
                       "eager_ms_per_op": ms_pe / args.steps,
                       "launches_per_op": (ctx.launch_count - l0) // (args.steps + args.warmup)}
                okc = True
                try:
                    with torch.cuda.stream(ext):
                        pp.capture()
                except Exception as ex:
                    print(f"bench.py: p2p graph capture failed: {ex!r}", file=sys.stderr)
                    okc = False
                if all_ok(okc):
                    ms_pg = statistics.median(timed(lambda i: pp.run(check=False), args.steps, args.warmup if r == 0 else 0)
                                              for r in range(9))
                    with torch.cuda.stream(ext):
                        ctx.sync()
                        same = bool(torch.equal(pp.out[0], want_sl[0]) and torch.equal(pp.out[1], want_sl[1]))
                    res.update({"ms_per_op": ms_pg / args.steps, "ops_per_s": args.steps / (ms_pg * 1e-3),
                                "one_graph_per_op": True, "bit_exact_after_replays": all_ok(same and not pp.failed())})
                else:
                    pp.graph = None
                if world > 1:
                    dist.barrier()
                pp.close()
                return res
            sharded["p2p"] = p2p_leg(False)
            if sharded["p2p"].get("fused_exchange"):
                sharded["p2p"]["unfused"] = p2p_leg(True)
    if pdl_before_sharded is not None:
        lib.gq_set_pdl(pdl_before_sharded)             # back to the default for the remaining legs

    # ---- end to end through the host C ABI: pinned host buffers, copies inside the timed region ----
    hout = torch.empty((2, W, n), dtype=torch.int64).pin_memory()
    houts = torch.empty((POOL, 2, W, n), dtype=torch.int64).pin_memory()

    def e2e_step(i):
        b = host_pool.data_ptr() + (i % POOL) * 4 * poly
        o0, o1 = hout.data_ptr(), hout.data_ptr() + poly
        ck(lib.gq_he_mul(h, L, o0, o1, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h))
        ck(lib.gq_he_rs(h, L, o0, o1))
    for i in range(args.warmup):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step(args.warmup + i)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    def wall(fn):
        barrier()
        t0 = time.perf_counter()
        fn()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt
    two_calls_s = e2e_s
    # one crossing per product (gq_he_mul_rs), still one synchronous call per op
    def fused_steps(k0, cnt):
        for i in range(k0, k0 + cnt):
            b = host_pool.data_ptr() + (i % POOL) * 4 * poly
            ck(lib.gq_he_mul_rs(h, L, hout.data_ptr(), hout.data_ptr() + poly, b, b + poly, b + 2 * poly, b + 3 * poly, evk._h))
    fused_steps(0, args.warmup)
    fused_s = wall(lambda: fused_steps(args.warmup, args.steps))
    # the batched entry point: `steps` products in one call, uploads / kernels / downloads pipelined
    def batch_args(k0, cnt):                     # the pointer tables a caller hands over
        VP = C.c_void_p * cnt
        hp, ho = host_pool.data_ptr(), houts.data_ptr()
        ins = [VP(*[hp + ((k0 + i) % POOL) * 4 * poly + j * poly for i in range(cnt)]) for j in range(4)]
        outs = [VP(*[ho + ((k0 + i) % POOL) * 2 * poly + j * poly for i in range(cnt)]) for j in range(2)]
        return outs, ins

    def batch_call(outs, ins, cnt):
        ck(lib.gq_he_mul_batch(h, L, cnt, outs[0], outs[1], ins[0], ins[1], ins[2], ins[3], evk._h, 1))
    # the host-buffer leg hides more of the copy / compute interplay with more lanes than the device-resident one
    # needs (B200, K = 20: 2 965 ops/s over 3 lanes, 3 060 over 6; `value` is flat to -2 % beyond 3): it gets its own count
    # (one process per node only: with 2 ranks the host side is the limit and six lanes cost 6 %)
    E2E_LANES = max(LANES, args.e2e_lanes) if world == 1 else LANES
    ctx.set_lanes(E2E_LANES)
    wcnt = max(args.warmup, 2 * E2E_LANES)
    batch_call(*batch_args(0, wcnt), wcnt)
    bo, bi = batch_args(args.warmup, args.steps)
    batch_runs = [wall(lambda: batch_call(bo, bi, args.steps)) for _ in range(3)]   # 3 calls of `steps` products each
    batch_s = sorted(batch_runs)[1]                                                 # the median call is reported
    ctx.sync()
    same = bool(torch.equal(houts[(args.warmup + args.steps - 1) % POOL].reshape(-1), hout.reshape(-1)))
    e2e = {"value": world * args.steps / batch_s, "unit": "ops/s", "h2d_bytes_per_step": 4 * poly,
           "d2h_bytes_per_step": 2 * poly, "ms_per_step": batch_s / args.steps * 1e3,
           "api": f"gq_he_mul_batch(count={args.steps}, rescale=1): host pointers (page-locked), every product's "
                  "4 input polys uploaded and 2 result polys downloaded inside the timed call; uploads, kernels and "
                  "downloads of neighbouring products overlap on the two copy engines, products round-robin over the compute "
                  "lanes; median of 3 consecutive calls",
           "batch_equals_single_call_result": same,
           "ms_per_step_of_each_call": [t / args.steps * 1e3 for t in batch_runs], "lanes": E2E_LANES,
           "single_call": {"gq_he_mul_rs": {"ops_per_s": world * args.steps / fused_s, "ms_per_step": fused_s / args.steps * 1e3,
                                            "h2d_bytes_per_step": 4 * poly, "d2h_bytes_per_step": 2 * poly},
                           "gq_he_mul+gq_he_rs": {"ops_per_s": world * args.steps / two_calls_s,
                                                  "ms_per_step": two_calls_s / args.steps * 1e3,
                                                  "h2d_bytes_per_step": 6 * poly, "d2h_bytes_per_step": 4 * poly,
                                                  "note": "one crossing per reference call, as the drop-in issues them"}}}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        sec, nm, kind, cores, sample = cpu_reference(1, budget_s=30.0)
        cpu = {"value": 1.0 / sec, "unit": "ops/s", "cores": cores, "kind": kind, "sample": sample,
               "host_cores_on_box": os.cpu_count()}
        try:
            if kind == "reference":
                cpu["more"] = cpu_reference_extras()
        except Exception as ex:
            cpu["more"] = {"error": repr(ex)}
    # the drop-in itself: he_mul + he_rs through the REFERENCE'S OWN SYMBOLS of dropin/_build/libgpqhe.so with gcry_mpi
    # ciphertexts (what a GPQHE caller links), timed by dropin/_build/dropin_bench in its own process
    dropin = None
    if rank == 0 and world == 1 and not args.no_dropin:
        exe = os.path.join(ROOT, "dropin", "_build", "dropin_bench")
        if os.path.exists(exe):
            ctx.sync()
            def run_tool(extra_env):
                out = subprocess.run([exe, str(LOGN), str(LOGQ), str(SLOTS), str(LOGDELTA), str(max(8, args.steps))],
                                     stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900,
                                     env=dict(os.environ, **extra_env))
                if out.returncode != 0:
                    return {"error": (out.stderr or out.stdout)[-400:]}
                return json.loads(out.stdout.strip().splitlines()[-1])
            try:
                dropin = run_tool({})
                if "chained" in dropin:
                    dropin["value"], dropin["unit"] = dropin["chained"]["ops_per_s"], "ops/s"
                    slow = run_tool({"GPQHE_DROPIN_SLOW_MPI": "1"})      # gcry_mpi_print / scan instead of the limb view
                    dropin["slow_mpi"] = {k: slow.get(k) for k in ("mpi_path", "isolated", "chained", "error") if k in slow}
            except Exception as ex:
                dropin = {"error": repr(ex)}
        else:
            dropin = {"error": "dropin/_build/dropin_bench is not built"}
    e2e["dropin"] = dropin

    if rank == 0:
        ws_mib = (4 + 3) * ctx.dimub * n * 8 / 2**20 + 2 * ctx.dimevk * n * 8 / 2**20
        line = {"metric": METRIC, "value": value, "unit": "ops/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": {"workload": WORKLOAD, "logn": LOGN, "logq": LOGQ, "logDelta": LOGDELTA, "level": L,
                           "tensor_limbs": D1, "keyswitch_limbs": D2, "words_per_coeff": W,
                           "parallelism": f"{world} GPU(s), each running independent he_mul+he_rs products round-robin over {LANES} "
                                          "compute lanes (streams) of its own context; no data-path collective",
                           "lanes_per_gpu": LANES, "timed_regions": len(regions),
                           "timed_region_ms": {"median": ms, "min": min(regions), "max": max(regions),
                                               "total_s": sum(regions) * 1e-3},
                           "sharded_ranks": (sharded or {}).get("ranks"),
                           "sharded_nccl_ms_per_op": (sharded or {}).get("ms_per_op"),
                           "sharded_p2p_ms_per_op": ((sharded or {}).get("p2p") or {}).get("ms_per_op"),
                           "sharded_p2p_fused_exchange": ((sharded or {}).get("p2p") or {}).get("fused_exchange"),
                           "sharded_p2p_unfused_ms_per_op": (((sharded or {}).get("p2p") or {}).get("unfused") or {}).get("ms_per_op"),
                           "sharded_bytes_sent_per_rank_per_op": (sharded or {}).get("bytes_sent_per_rank_per_op"),
                           "sharded_bit_exact_vs_single_gpu": (None if not sharded else bool(
                               sharded.get("bit_exact_vs_single_gpu") and
                               ((sharded.get("p2p") or {}).get("bit_exact_after_replays", (sharded.get("p2p") or {}).get("bit_exact_vs_single_gpu", True))))),
                           "l2": f"inputs rotate over a pool of {POOL} ciphertext pairs ({POOL * 4 * poly / 2**20:.0f} MiB) and the "
                                 f"per-op scratch + key is {ws_mib:.0f} MiB: working set > 126 MB L2"},
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "ntt": ntt,
                "stages": stages, "imma_probe": imma, "rot_sweep": rot_sweep, "sharded": sharded, "cpu_baseline": cpu,
                "single_stream": {"ops_per_s": world * args.steps / (ms_single * 1e-3), "ms_per_op": ms_single / args.steps,
                                  "note": "the same products issued back to back on ONE stream (latency of one op)"},
                "lanes_bit_exact_vs_single_stream": lanes_ok}
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--lanes", type=int, default=3, help="compute lanes per GPU for the throughput figure")
    ap.add_argument("--e2e-lanes", type=int, default=6, help="compute lanes per GPU for the host-buffer (e2e) leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-dropin", action="store_true", help="skip the drop-in (reference symbols, gcry_mpi) leg")
    ap.add_argument("--no-rot", action="store_true", help="skip the rotation sweep leg")
    ap.add_argument("--no-p2p", action="store_true", help="skip the peer-to-peer variant of the sharded leg")
    ap.add_argument("--no-sharded", action="store_true", help="skip the limb-sharded single-stream leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
