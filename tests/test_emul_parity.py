"""CPU tier: host logic + kernel index arithmetic.  The library sources are compiled
against the fibre-based CUDA stand-in in tests/emul (test infrastructure, not a product
path) and driven through the same C ABI and the same checks as the GPU tier."""
import pytest

import gpqhe_b200 as g
import parity_cases as pc


@pytest.mark.parametrize("name", pc.GOLDEN_CASES)
def test_golden(emul_lib, name):
    pc.check_golden(emul_lib, name)


@pytest.mark.parametrize("logn,logq", [(3, 61), (4, 61), (5, 61), (7, 61), (9, 100), (10, 27), (12, 100),
                                       (13, 120), (14, 120)])
def test_limb_layer(emul_lib, logn, logq):
    pc.check_limb_layer_vs_oracle(emul_lib, logn, 1 << logq, seed=logn)


@pytest.mark.parametrize("logn,q,Delta,levels", [
    (8, 1 << 120, 1 << 40, None),                 # power-of-two fast path, every level incl. q_0 = 1
    (5, 1 << 438, 1 << 50, [8, 4, 1]),            # the KAT modulus chain (24 key-switch limbs)
    (6, (1 << 119) + 12345, (1 << 40) + 7, None),  # arbitrary q and Delta: exact slow path
    (4, 3 ** 75, 1 << 40, None),                  # arbitrary q, power-of-two Delta
    (3, 1 << 881, 1 << 30, [29, 10, 1]),          # the reference's `cmp` ring: 47 key-switch limbs, a 2 774-bit basis
    (3, 1 << 438, 1 << 50, [8, 1]),               # n = 8: half a 16-row step of the tensor-core CRT kernels is padding
])
def test_he_ops(emul_lib, logn, q, Delta, levels):
    pc.check_he_vs_oracle(emul_lib, logn, q, 4, Delta, seed=logn, levels=levels, logqub=q.bit_length() - 1)


def test_error_behaviour(emul_lib):
    """The reference abort()s with these messages (src/precomp.c:344-350,434-445)."""
    with pytest.raises(g.GpqheError, match="too large"):
        g.HeContext(12, 1 << 120, 4, 1 << 40, lib_path=emul_lib)
    with pytest.raises(g.GpqheError, match="power of 2"):
        g.HeContext(8, 1 << 120, 3, 1 << 40, lib_path=emul_lib)
    with pytest.raises(g.GpqheError, match="slots"):
        g.HeContext(4, 1 << 120, 16, 1 << 40, lib_path=emul_lib)
    ctx = g.HeContext(8, 1 << 120, 4, 1 << 40, lib_path=emul_lib)
    import numpy as np
    z = np.zeros((ctx.W, ctx.n), dtype=np.uint64)
    evk = ctx.genevk(np.zeros((ctx.dimevk, ctx.n), dtype=np.uint64), np.zeros((ctx.dimevk, ctx.n), dtype=np.uint64))
    with pytest.raises(g.GpqheError):            # assert(ct1->l==ct2->l), src/he-mult.c:90
        ctx.he_mul(g.HeCt(3, z, z), g.HeCt(2, z, z), evk)
    with pytest.raises(g.GpqheError):
        ctx.he_rs(g.HeCt(0, z, z))
    ctx.close()


def test_small_primes_crt(emul_lib):
    """tests/crt.c: -DTEST_CRT rings (9-bit primes): decompose -> reconstruct round trip at
    every basis size; NTT entry points refuse (the reference aborts, src/reduce.c:94-100)."""
    import random
    import numpy as np
    from oracle import pyoracle as o
    ctx = g.HeContext(4, 1 << 10, lib_path=emul_lib, logp=9)
    oc = o.Ctx(4, 1 << 10, logp=9)
    assert ctx.primes == oc.primes
    rng = random.Random(5)
    n = ctx.n
    for dim in range(ctx.dimub, 1, -1):
        P = oc.rns[dim - 1].P
        a = [rng.randrange(P) - P // 2 for _ in range(n)]
        W = (P.bit_length() + 63) // 64
        dec = ctx.rns_decompose(g.ints_to_big(a, W), 0, dim)
        for d in range(dim):
            pc.eq(dec[d], o.rns_decompose(a, oc.rns[d]), "decompose")
        back = g.big_to_ints(ctx.poly_rns2mpi(dec.reshape(-1), dim, P, W))
        assert back == o.poly_rns2mpi(dec.reshape(-1), oc.rns[dim - 1], P, n)
        assert all((x - y) % P == 0 for x, y in zip(back, a))
    with pytest.raises(g.GpqheError):
        ctx.ntt(np.zeros(n, dtype=np.uint64), 0)
    ctx.close()


@pytest.mark.parametrize("logn", [9, 13, 14])
def test_poly_mul_lazy_ntt_two_kernel_path(emul_lib, logn):
    """poly_mul runs the lazy (Harvey) NTT kernels of gq_ntt_lazy.cu; above 2^12 that is the
    strided + tile pair.  Bit-exact vs the oracle, and identical to the exact-NTT pipeline
    (GPQHE_B200_EXACT_NTT=1 selects the kernels of gq_ntt.cu for the internal transforms)."""
    import os
    import random
    from oracle import pyoracle as o
    q = 1 << 100
    rng = random.Random(logn)
    oc = o.Ctx(logn, q, logqub=100)
    n = oc.n
    a = [rng.randrange(q) - q // 2 for _ in range(n)]
    b = [rng.randrange(q) - q // 2 for _ in range(n)]
    a[0], a[1], b[0] = 0, q // 2 - 1, -(q // 2)
    want = o.poly_mul(oc, a, b, 3, q)
    got = []
    for exact in (False, True):
        if exact:
            os.environ["GPQHE_B200_EXACT_NTT"] = "1"
        try:
            ctx = g.HeContext(logn, q, lib_path=emul_lib, logqub=100)
            W = ctx.W
            got.append(g.big_to_ints(ctx.poly_mul(g.ints_to_big(a, W), g.ints_to_big(b, W), 3, q)))
            ctx.close()
        finally:
            os.environ.pop("GPQHE_B200_EXACT_NTT", None)
    assert got[0] == want and got[1] == want


def test_crt_edge_values_take_exact_path(emul_lib):
    pc.check_crt_edges(emul_lib)


@pytest.mark.parametrize("logn,q,Delta,n1,n2", [(5, 1 << 120, 1 << 40, 2, 2), (5, 1 << 438, 1 << 50, 4, 2)])
def test_gemv_device_resident(emul_lib, logn, q, Delta, n1, n2):
    pc.check_gemv(emul_lib, logn, q, Delta, n1, n2, logqub=q.bit_length() - 1)


@pytest.mark.parametrize("logn,q,Delta,n1,n2,lanes", [(5, 1 << 120, 1 << 40, 2, 2, 1), (5, 1 << 438, 1 << 50, 2, 2, 3)])
def test_coeff2slot_device_resident(emul_lib, logn, q, Delta, n1, n2, lanes):
    pc.check_coeff2slot(emul_lib, logn, q, Delta, n1, n2, logqub=q.bit_length() - 1, lanes=lanes)


@pytest.mark.parametrize("logn,q,Delta,iters,lanes", [(4, 1 << 438, 1 << 50, 1, 1), (4, 1 << 438, 1 << 50, 3, 3),
                                                     (3, 3 ** 150, 1 << 40, 0, 1)])
def test_rlsin_device_resident(emul_lib, logn, q, Delta, iters, lanes):
    """he_exp / he_rlsin (EvalMod of he_bootstrap): pow2 chain down to level 0, and an arbitrary q."""
    pc.check_rlsin(emul_lib, logn, q, Delta, iters, logqub=q.bit_length() - 1, lanes=lanes)


@pytest.mark.parametrize("logn,q,Delta,n1,n2,lanes", [(5, 1 << 120, 1 << 40, 2, 2, 1), (5, 1 << 438, 1 << 50, 2, 2, 3)])
def test_slot2coeff_device_resident(emul_lib, logn, q, Delta, n1, n2, lanes):
    pc.check_slot2coeff(emul_lib, logn, q, Delta, n1, n2, logqub=q.bit_length() - 1, lanes=lanes)


@pytest.mark.parametrize("logn,logq", [(9, 100), (12, 100), (13, 120), (14, 120), (16, 61), (17, 61)])
def test_lazy_ntt_vs_exact(emul_lib, logn, logq):
    """single-tile (n <= 2^12) and two-kernel (n >= 2^13) lazy transforms on growth-maximising vectors"""
    pc.check_lazy_ntt_vs_exact(emul_lib, logn, 1 << logq)


def _probe(path):
    import ctypes as C
    lib = C.CDLL(path)
    lib.gqp_last_error.restype = C.c_char_p
    return lib


@pytest.mark.parametrize("logn,nlimbs,npolys", [(13, 2, 1), (14, 3, 2), (16, 2, 1)])
def test_ntt32_probe_matches_host(emul_lib, logn, nlimbs, npolys):
    """profiles/probe/gqp_ntt32.cu (measurement-only 32-bit transform): device output == plain host transform ==
    direct evaluation at psi^(2 brv(j) + 1)"""
    import ctypes as C
    import os
    lib = _probe(os.path.join(os.path.dirname(emul_lib), "libgqprobe_emul.so"))
    rate, bad = C.c_double(0), C.c_int(-1)
    assert lib.gqp_bench_ntt32(0, logn, nlimbs, npolys, 1, C.byref(rate), C.byref(bad)) == 0
    assert bad.value == 0 and rate.value > 0
    assert lib.gqp_bench_ntt32(0, 12, 1, 1, 1, C.byref(rate), C.byref(bad)) != 0   # outside 13..16


@pytest.mark.parametrize("logn,D", [(6, 8), (7, 4)])
def test_imma_decompose_probe_matches_host(emul_lib, logn, D):
    """profiles/probe/gqp_imma.cu (measurement only): rns_decompose as a u8 tensor-core contraction with thread-local
    byte-column recombination, under the emulated mma.sync fragment layouts, against plain modular reduction"""
    import ctypes as C
    import os
    from oracle import pyoracle as o
    lib = _probe(os.path.join(os.path.dirname(emul_lib), "libgqprobe_emul.so"))
    oc = o.Ctx(logn, 1 << 438, 4, 1 << 50, logqub=438)
    pr = (C.c_uint64 * D)(*oc.primes[:D])
    rate, ms, bad = C.c_double(0), C.c_double(0), C.c_int(-1)
    assert lib.gqp_bench_imma_decompose(0, logn, D, pr, 1, C.byref(rate), C.byref(ms), C.byref(bad)) == 0, lib.gqp_last_error()
    assert bad.value == 0


@pytest.mark.parametrize("logn,q,Delta,n1,n2,lanes", [(5, 1 << 438, 1 << 50, 2, 2, 1), (4, 1 << 438, 1 << 50, 2, 1, 3)])
def test_compositions_vs_primitives(emul_lib, logn, q, Delta, n1, n2, lanes):
    """the full-size composition check of the GPU tier, validated here at a toy ring"""
    pc.check_compositions_full_size(emul_lib, logn, q, Delta, n1, n2, lanes=lanes, logqub=438, iters=1)


def test_fused_rescale_tie_rule(emul_lib):
    pc.check_fused_rescale_ties(emul_lib)
    pc.check_fused_rescale_ties(emul_lib, logn=4, q=1 << 120, Delta=1 << 40, logqub=120)


@pytest.mark.parametrize("logn,level", [(13, 2), (9, None)])
def test_fused_inverse_equals_separate_kernels(emul_lib, logn, level):
    """two-kernel transforms (n = 2^13) and one-tile transforms: every loader mode of the fused inverse NTT"""
    pc.check_fused_vs_separate(emul_lib, logn, level=level, short=(logn > 12))


@pytest.mark.parametrize("logq", [500, 512, 65])
def test_tensor_core_crt_widths(emul_lib, logq):
    """the tensor-core reconstruction (gq_crt_tc.cu) at 8 result words (every byte column in use), at the widest
    modulus it takes, and at 2 words (6 of 8 words are padding): product, rescale ties, edge values, fused vs separate"""
    q, Delta = 1 << logq, 1 << (50 if logq > 100 else 20)
    pc.check_he_vs_oracle(emul_lib, 5, q, 4, Delta, seed=logq, levels=None if logq < 100 else [9, 5, 1], logqub=logq)
    pc.check_crt_edges(emul_lib, logn=5, q=q, Delta=Delta, logqub=logq)
    pc.check_fused_rescale_ties(emul_lib, logn=5, q=q, Delta=Delta, logqub=logq)
    pc.check_fused_vs_separate(emul_lib, 5, q=q, Delta=Delta, logqub=logq)
