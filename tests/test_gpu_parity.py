"""GPU tier: the real sm_100a library through the C ABI vs golden fixtures and the oracle."""
import random

import numpy as np
import pytest

import gpqhe_b200 as g
import parity_cases as pc
from oracle import pyoracle as o

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", pc.GOLDEN_CASES)
def test_golden(real_lib, name):
    pc.check_golden(real_lib, name)


@pytest.mark.parametrize("logn,logq", [(3, 61), (4, 61), (6, 61), (7, 61), (8, 61), (9, 100), (10, 27), (11, 54),
                                       (12, 100), (13, 120), (14, 438), (15, 120), (16, 438), (17, 120)])
def test_limb_layer(real_lib, logn, logq):
    pc.check_limb_layer_vs_oracle(real_lib, logn, 1 << logq, seed=logn, limbs=(0, 1))


@pytest.mark.parametrize("logn,q,Delta,levels,logqub", [
    (8, 1 << 120, 1 << 40, None, 0),
    (10, 1 << 438, 1 << 50, [8, 5, 1], 438),      # logn in [10,15] consults the security table
    (12, 1 << 109, 1 << 30, None, 0),
    (7, (1 << 119) + 12345, (1 << 40) + 7, None, 0),
    (6, 3 ** 75, 1 << 40, None, 0),
])
def test_he_ops(real_lib, logn, q, Delta, levels, logqub):
    pc.check_he_vs_oracle(real_lib, logn, q, 4, Delta, seed=logn, levels=levels, logqub=logqub)


def test_kat_ring_n14_he_mul_rs(real_lib):
    """config[0] ring (logn=14, logq=438, Delta=2^50, tests/gpqhe.c:1296-1299): he_mul+he_rs and
    he_rot at the top level, bit-exact vs the oracle."""
    pc.check_he_vs_oracle(real_lib, 14, 1 << 438, 16, 1 << 50, seed=14, levels=[8], rots=(1,))


def test_batched_ntt_n16_all_limbs(real_lib):
    """config[1]: batched forward/inverse NTT + pointwise mul over 24 limbs at N=2^16."""
    rng = np.random.default_rng(16)
    ctx = g.HeContext(16, 1 << 438, 16, 1 << 50, lib_path=real_lib)
    oc = o.Ctx(16, 1 << 438, 16, 1 << 50)
    try:
        assert ctx.primes == oc.primes and ctx.dimevk == 24
        n, D = ctx.n, 24
        x = np.stack([rng.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(D)])[None]
        f = ctx.ntt_batch(x, 0, inverse=False)
        for d in range(D):
            pc.eq(f[0, d], o.ntt(x[0, d], oc.rns[d]), f"batched ntt limb {d}")
        pc.eq(ctx.ntt_batch(f, 0, inverse=True), x, "inverse of forward")
        y = np.stack([rng.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(D)])[None]
        iv = ctx.ntt_batch(y, 0, inverse=True)
        for d in (0, 7, 23):
            pc.eq(iv[0, d], o.invntt(y[0, d], oc.rns[d]), f"batched invntt limb {d}")
            pc.eq(ctx.poly_rns_mul(x[0, d], y[0, d], d), o.poly_rns_mul(x[0, d], y[0, d], oc.rns[d]), "mul")
    finally:
        ctx.close()


def test_full_size_n16_he_mul_rs_rot(real_lib):
    """config[2]/[3] at full size (N=2^16, 16 tensor limbs, 24 key-switch limbs): he_mul + he_rs
    and he_rot vs the oracle, plus size-independent properties."""
    rng = random.Random(2016)
    logn, q, Delta = 16, 1 << 438, 1 << 50
    ctx = g.HeContext(logn, q, 16, Delta, lib_path=real_lib)
    oc = o.Ctx(logn, q, 16, Delta)
    try:
        n, W, l = ctx.n, ctx.W, ctx.L
        nr = np.random.default_rng(7)
        p0 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
        p1 = np.concatenate([nr.integers(0, oc.primes[d], size=n, dtype=np.uint64) for d in range(oc.dimevk)])
        evk = ctx.genevk(p0, p1)
        ql = oc.ql[l]
        a0, a1, b0, b1 = (pc.rpoly(rng, ql, n) for _ in range(4))
        c1 = g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W))
        c2 = g.HeCt(l, g.ints_to_big(b0, W), g.ints_to_big(b1, W))
        r = ctx.he_mul(c1, c2, evk)
        e = o.he_mul(oc, {"l": l, "c0": a0, "c1": a1}, {"l": l, "c0": b0, "c1": b1}, (p0, p1))
        assert g.big_to_ints(r.c0) == e["c0"] and g.big_to_ints(r.c1) == e["c1"], "he_mul N=2^16"
        r2, e2 = ctx.he_rs(r), o.he_rs(oc, e)
        assert g.big_to_ints(r2.c0) == e2["c0"] and g.big_to_ints(r2.c1) == e2["c1"], "he_rs N=2^16"
        r3 = ctx.he_rot(c1, 5, {5: evk})
        e3 = o.he_rot(oc, {"l": l, "c0": a0, "c1": a1}, 5, (p0, p1))
        assert g.big_to_ints(r3.c0) == e3["c0"] and g.big_to_ints(r3.c1) == e3["c1"], "he_rot N=2^16"
        # properties: commutativity of he_mul; outputs centred mod q_l; rot then level check
        rs = ctx.he_mul(c2, c1, evk)
        pc.eq(rs.c0, r.c0, "he_mul commutes"); pc.eq(rs.c1, r.c1, "he_mul commutes")
        top = np.asarray(r.c0[W - 1], dtype=np.uint64).astype(np.int64) >> 53   # bits >= 437 are sign
        assert set(np.unique(top).tolist()) <= {0, -1}, "centred range of q_L"
    finally:
        ctx.close()


def test_crt_edge_values_take_exact_path(real_lib):
    pc.check_crt_edges(real_lib)
    pc.check_crt_edges(real_lib, logn=12, q=1 << 109, Delta=1 << 30, logqub=109)


def test_gemv_device_resident(real_lib):
    """he_gemv with slots = 16 (n1 = n2 = 4, the KAT's shape) at N=2^12."""
    pc.check_gemv(real_lib, 12, 1 << 109, 1 << 30, 4, 4)


def test_coeff2slot_device_resident(real_lib):
    """configs[4]: he_coeff2slot (he_conj + 4 he_gemv + he_add + he_mulpt + he_rs) at N=2^12, slots = 16,
    giant steps spread over 3 compute lanes."""
    pc.check_coeff2slot(real_lib, 12, 1 << 109, 1 << 30, 4, 4, lanes=3)
    pc.check_coeff2slot(real_lib, 10, 1 << 120, 1 << 40, 2, 2, logqub=120, lanes=1)


def test_rlsin_device_resident(real_lib):
    """configs[4], EvalMod: he_exp + he_rlsin (7+iter he_mul, 5 he_mulpt, he_conj, ...) at N=2^11 on the KAT modulus
    chain down to level 0, over compute lanes and on one stream."""
    pc.check_rlsin(real_lib, 11, 1 << 438, 1 << 50, 3, logqub=438, lanes=3)
    pc.check_rlsin(real_lib, 10, 1 << 438, 1 << 50, 1, logqub=438, lanes=1)


def test_slot2coeff_device_resident(real_lib):
    """configs[4], SlotToCoeff: two he_gemv + he_add at N=2^12, slots = 16."""
    pc.check_slot2coeff(real_lib, 12, 1 << 109, 1 << 30, 4, 4, lanes=3)


@pytest.mark.parametrize("logn,logq", [(10, 100), (12, 100), (13, 120), (15, 120), (16, 438), (17, 61)])
def test_lazy_ntt_vs_exact(real_lib, logn, logq):
    pc.check_lazy_ntt_vs_exact(real_lib, logn, 1 << logq, device="cuda:0",
                               limbs=(0, 5, 23) if logq == 438 else None)


def _probe():
    import ctypes as C
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "probe", "libgqprobe.so")
    if not os.path.exists(path):
        pytest.skip("profiles/probe/libgqprobe.so not built")
    lib = C.CDLL(path)
    lib.gqp_last_error.restype = C.c_char_p
    return lib


@pytest.mark.parametrize("logn", [13, 16])
def test_ntt32_probe_matches_host(logn):
    import ctypes as C
    lib = _probe()
    rate, bad = C.c_double(0), C.c_int(-1)
    assert lib.gqp_bench_ntt32(0, logn, 30, 2, 2, C.byref(rate), C.byref(bad)) == 0, lib.gqp_last_error()
    assert bad.value == 0 and rate.value > 0


def test_imma_decompose_probe_matches_host():
    """measurement-only tensor-core decomposition prototype at the op's shape: every residue it is checked on equals
    the plain modular reduction"""
    import ctypes as C
    lib = _probe()
    oc = o.Ctx(16, 1 << 438, 16, 1 << 50)
    pr = (C.c_uint64 * 16)(*oc.primes[:16])
    rate, ms, bad = C.c_double(0), C.c_double(0), C.c_int(-1)
    assert lib.gqp_bench_imma_decompose(0, 16, 16, pr, 3, C.byref(rate), C.byref(ms), C.byref(bad)) == 0, lib.gqp_last_error()
    assert bad.value == 0 and rate.value > 0


@pytest.mark.parametrize("logn", [16, 14])
def test_compositions_at_config_size(real_lib, logn):
    """configs[3] / configs[4] at the sizes they are quoted on (N=2^16; N=2^14 = the KAT ring), slots = 16:
    he_gemv, hoisted rotations, he_coeff2slot, he_rlsin (iter = 1), he_slot2coeff as ONE device-resident call
    each against the oracle's compositions run primitive by primitive through the C ABI."""
    pc.check_compositions_full_size(real_lib, logn, 1 << 438, 1 << 50, 4, 4, lanes=3, iters=1)


def test_fused_rescale_tie_rule(real_lib):
    pc.check_fused_rescale_ties(real_lib)
    pc.check_fused_rescale_ties(real_lib, logn=13, q=1 << 120, Delta=1 << 40, logqub=120)


@pytest.mark.parametrize("logn", [16, 14, 12])
def test_fused_inverse_equals_separate_kernels(real_lib, logn):
    pc.check_fused_vs_separate(real_lib, logn)


@pytest.mark.parametrize("logq", [500, 512, 65])
def test_tensor_core_crt_widths(real_lib, logq):
    """the tensor-core reconstruction (gq_crt_tc.cu) at 8 result words (every byte column in use), at the widest
    modulus it takes, and at 2 words (6 of 8 words are padding): product, rescale ties, edge values, fused vs separate"""
    q, Delta = 1 << logq, 1 << (50 if logq > 100 else 20)
    pc.check_he_vs_oracle(real_lib, 10, q, 4, Delta, seed=logq, levels=None if logq < 100 else [9, 5, 1], logqub=logq)
    pc.check_crt_edges(real_lib, logn=10, q=q, Delta=Delta, logqub=logq)
    pc.check_fused_rescale_ties(real_lib, logn=10, q=q, Delta=Delta, logqub=logq)
    pc.check_fused_vs_separate(real_lib, 10, q=q, Delta=Delta, logqub=logq)
