"""CPU tier: pin the oracle (oracle/oracle.c + oracle/pyoracle.py) against the fixtures the
unmodified reference produced (tests/golden, made by tests/golden/make_golden.py) and,
when oracle/_ref exists in this container, against the reference build itself."""
import numpy as np
import pytest

from oracle import pyoracle as o
from parity_cases import GOLDEN_CASES, eq, load_golden, words_to_int


def ints(w):
    return o.from_words(np.asarray(w, dtype=np.uint64))


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_matches_reference_fixture(name):
    z, logn, q, slots, Delta = load_golden(name)
    oc = o.Ctx(logn, q, slots, Delta)
    n, W = oc.n, int(z["W"])
    assert (oc.L, oc.dim, oc.dimevk, oc.dimub) == tuple(int(z[k]) for k in ("L", "dim", "dimevk", "dimub"))
    eq(oc.primes, z["primes"], "primes")
    for d, nd in enumerate(oc.rns):
        eq([nd.p, nd.pinv_mont, nd.pinv_barr, nd.ninv, int(nd.zetas[1]), int(nd.zetas_inv[1])],
           z["prime_consts"][d], "per-prime constants")
    for tag in ("a", "b"):
        nd = oc.rns[int(z[f"limb_{tag}"])]
        x, y = z[f"x_{tag}"], z[f"y_{tag}"]
        eq(o.ntt(x, nd), z[f"ntt_{tag}"], "ntt")
        eq(o.invntt(y, nd), z[f"invntt_{tag}"], "invntt")
        eq(o.poly_rns_mul(x, y, nd), z[f"mul_{tag}"], "mul")
        eq(o.poly_rns_add(x, y, nd), z[f"add_{tag}"], "add")
        eq(o.ntt(z[f"hz_{tag}"], nd), z[f"hzntt_{tag}"], "ntt hazard")
        eq(o.invntt(z[f"hzntt_{tag}"], nd), z[f"hzinv_{tag}"], "invntt hazard")
    a, b = ints(z["big_a"]), ints(z["big_b"])
    for d in range(oc.dimub):
        eq(o.rns_decompose(a, oc.rns[d]), z["decompose_a"][d], "decompose")
    q2 = words_to_int(z["rns2mpi_q2"])
    assert o.poly_rns2mpi(z["rns2mpi_in"], oc.rns[oc.dimub - 1], q, n) == ints(z["rns2mpi_full"])
    assert o.poly_rns2mpi(z["rns2mpi_in"], oc.rns[oc.dim - 1], q2, n) == ints(z["rns2mpi_dim"])
    assert o.poly_mul(oc, a, b, oc.dimub, q) == ints(z["poly_mul"])
    for rot in (1, 5):
        assert o.poly_rot(a, rot, n) == ints(z[f"poly_rot{rot}"])
    assert o.poly_conj(a, n) == ints(z["poly_conj"])
    evk = (z["evk_p0"], z["evk_p1"])
    for l in [int(v) for v in z["levels"]]:
        c1 = {"l": l, "c0": ints(z[f"l{l}_a0"]), "c1": ints(z[f"l{l}_a1"])}
        c2 = {"l": l, "c0": ints(z[f"l{l}_b0"]), "c1": ints(z[f"l{l}_b1"])}
        r = o.he_mul(oc, c1, c2, evk)
        assert r["c0"] == ints(z[f"l{l}_mul_c0"]) and r["c1"] == ints(z[f"l{l}_mul_c1"])
        r = o.he_rs(oc, r)
        assert r["c0"] == ints(z[f"l{l}_mulrs_c0"]) and r["c1"] == ints(z[f"l{l}_mulrs_c1"])
        r = o.he_moddown(oc, c1)
        assert r["c0"] == ints(z[f"l{l}_moddown_c0"]) and r["c1"] == ints(z[f"l{l}_moddown_c1"])
        for rot in (0, 3):
            r = o.he_rot(oc, c1, rot, evk)
            assert r["c0"] == ints(z[f"l{l}_rot{rot}_c0"]) and r["c1"] == ints(z[f"l{l}_rot{rot}_c1"])
        r = o.he_conj(oc, c1, evk)
        assert r["c0"] == ints(z[f"l{l}_conj_c0"]) and r["c1"] == ints(z[f"l{l}_conj_c1"])
        nu = float(1 << 40)
        import math
        dim = int((oc.ql[l].bit_length() + math.log2(nu) + oc.logn) / oc.logp + 1)
        r = o.he_mulpt(oc, c1, ints(z[f"l{l}_pt_m"]), dim)
        assert r["c0"] == ints(z[f"l{l}_mulpt_c0"]) and r["c1"] == ints(z[f"l{l}_mulpt_c1"])


def test_polymul_known_answers():
    """tests/polymul.c:59-77 inputs; expected leading/trailing coefficients are the ones the
    reference's test prints and tests/polymul.gp:1-23 computes (negacyclic product)."""
    oc = o.Ctx(7, 1 << 61)
    assert oc.primes[:2] == [576460752303434497, 576460752303436801]      # tests/polymul.gp:17-18
    n = 128
    a, b = [i + 2 for i in range(n)], [i + 3 for i in range(n)]
    r = o.poly_mul(oc, a, b, oc.dimub, oc.q)
    exp = [0] * n                     # schoolbook negacyclic product, exact integers
    for i in range(n):
        for j in range(n):
            k = i + j
            if k < n: exp[k] += a[i] * b[j]
            else: exp[k - n] -= a[i] * b[j]
    assert r == exp and r[127] == 382784 and r[0] == -390900
    a = [oc.primes[0] - i - 1 for i in range(n)]
    b = [oc.primes[1] - i - 1 for i in range(n)]
    r = o.poly_mul(oc, a, b, oc.dimub, oc.q)
    assert r[127] == 18559595904 and r[0] == -18266527040


def test_survey_spot_constants():
    """Constants probed from the reference for the KAT rings."""
    L = o.lib()
    p = int(L.orc_next_prime((1 << 59) + 1, 1 << 16))
    assert p == 576460752308273153
    assert int(L.orc_montgomery_inv(p)) == 14490355320300634113
    assert int(L.orc_barrett_inv(p)) == 2305843009194295292
    assert int(L.orc_next_prime(p, 1 << 16)) == 576460752315482113
    assert int(L.orc_next_prime((1 << 59) + 1, 1 << 14)) == 576460752304439297
    oc = o.Ctx(14, 1 << 438, 16, 1 << 50)
    assert (oc.L, oc.dim, oc.dimevk, oc.dimub) == (8, 8, 24, 30)
    assert oc.rns[0].ninv == 1 << 50 and oc.primes[23] == 576460752324427777
    assert oc.P.bit_length() == 473 and oc.PqL.bit_length() == 911
    assert [oc.dim_tensor(l) for l in range(8, -1, -1)] == [16, 14, 12, 11, 9, 7, 5, 4, 2]
    assert [oc.dim_ks(l) for l in range(8, -1, -1)] == [24, 23, 22, 21, 20, 19, 19, 18, 17]


def test_scalar_semantics():
    assert o.mpi_rdiv(6, 4) == 1 and o.mpi_rdiv(7, 4) == 2 and o.mpi_rdiv(-6, 4) == -2   # ties round down
    assert o.mpi_rdiv(-7, 4) == -2 and o.mpi_rdiv(-5, 4) == -1
    assert o.mpi_smod(7, 8, 4) == -1 and o.mpi_smod(4, 8, 4) == -4 and o.mpi_smod(3, 8, 4) == 3
    assert o.mpi_smod(5, 1, 0) == -1          # q = 1: 0 >= qh = 0, minus q


def test_oracle_against_live_reference():
    from oracle import refbind as rb
    if not rb.available():
        pytest.skip("oracle/_ref not built here (the GPU box has no /root/reference)")
    import subprocess, sys, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = subprocess.run([sys.executable, os.path.join(root, "tests", "pin_live_reference.py")],
                          capture_output=True, text=True, timeout=600)
    assert code.returncode == 0, code.stdout + code.stderr
