"""The drop-in libgpqhe.so (dropin/): the reference's own host C + our hot symbols over the C ABI.

CPU tier: the drop-in is linked at run time against the emulated build of the device library
(tests/emul) and compared, in one process, with the reference's own CPU definitions that stay in
the same library as gpqhe_ref_<name> (tests/dropin_diff.c), and the reference's UNMODIFIED test
programs are run against it.  GPU tier: the same programs against the real sm_100a library at the
KAT ring (N=2^14) and at N=2^16.
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "dropin", "_build")
REF_PRESENT = os.path.isdir("/root/reference/src")


def _have_gcrypt():
    import ctypes
    try:
        ctypes.CDLL("libgcrypt.so.20")
        return True
    except OSError:
        return False


def _build_dropin():
    if REF_PRESENT:
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "gpqhe_b200", "csrc"), "-j8"])
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "dropin"), "-j8"])
    if not os.path.exists(os.path.join(BUILD, "libgpqhe.so")):
        pytest.skip("dropin/_build is not built and /root/reference is not mounted here")
    if not _have_gcrypt():
        pytest.skip("libgcrypt.so.20 is not installed on this machine")


@pytest.fixture(scope="module")
def emul_env(emul_lib):
    """LD_LIBRARY_PATH that makes dropin/_build/libgpqhe.so resolve libgpqhe_b200.so to the
    emulated build (LD_LIBRARY_PATH is searched before the library's RUNPATH)."""
    _build_dropin()
    d = os.path.join(ROOT, "tests", "emul", "_as_real")
    os.makedirs(d, exist_ok=True)
    shutil.copyfile(emul_lib, os.path.join(d, "libgpqhe_b200.so"))
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = d + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    return env


def _run(cmd, env=None, timeout=900, cwd=None):
    p = subprocess.run(cmd, env=env, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    return p.returncode, p.stdout


def _diffs(out):
    return re.findall(r"diff = (\S+)", out)


def _verdicts(out):
    """pass/fail markers libpmu prints per case, without the timings"""
    clean = re.sub(r"\x1b\[[0-9;]*m", "", out)
    return re.findall(r"\b(OK|FAILED|PASSED)\]", clean) + re.findall(r"(\d+/\d+) subtask", clean)


@pytest.mark.parametrize("args", [("7", "120", "4", "40", "4"), ("6", "438", "4", "50"), ("8", "200", "8", "40")])
def test_dropin_diff_emulated(emul_env, args):
    rc, out = _run([os.path.join(BUILD, "dropin_diff"), *args], env=emul_env)
    assert rc == 0 and "PASS" in out, out[-2000:]
    assert "direct-limb" in out


def test_dropin_diff_emulated_portable_mpi_path(emul_env):
    env = dict(emul_env, GPQHE_DROPIN_SLOW_MPI="1")
    rc, out = _run([os.path.join(BUILD, "dropin_diff"), "6", "120", "4", "40"], env=env)
    assert rc == 0 and "PASS" in out and "gcry_mpi_print/scan" in out, out[-2000:]


def test_dropin_exports_reference_symbols():
    """the drop-in must export everything the reference library exports."""
    _build_dropin()
    ref = os.path.join(ROOT, "oracle", "_ref", "libgpqhe.so")
    if not os.path.exists(ref):
        pytest.skip("reference build absent")

    def syms(path):
        out = subprocess.check_output(["nm", "-D", "--defined-only", path], text=True)
        return {l.split()[-1] for l in out.splitlines() if l.split()[1] in "TBDR"}
    missing = syms(ref) - syms(os.path.join(BUILD, "libgpqhe.so"))
    assert not missing, sorted(missing)


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")), reason="reference build absent")
@pytest.mark.parametrize("op", ["mul", "rot", "conj", "gemv"])
def test_reference_test_program_same_output_emulated(emul_env, op):
    """The reference's unmodified tests/gpqhe.c prints `diff = ...` per case; against the drop-in
    it must print the very same numbers as against the reference library (bit-exact ciphertexts
    decrypt to identical doubles)."""
    args = [op, "pk", "--logn=7", "--logq=120", "--slots=4", "--logDelta=40"]
    _, ours = _run([os.path.join(BUILD, "test-gpqhe"), *args], env=emul_env)
    _, ref = _run([os.path.join(ROOT, "oracle", "_ref", "test-gpqhe"), *args])
    assert _verdicts(ours) and _verdicts(ours) == _verdicts(ref), (ours[-1500:], ref[-1500:])
    assert _diffs(ours) == _diffs(ref), (ours[-1500:], ref[-1500:])


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")), reason="reference build absent")
@pytest.mark.parametrize("op", ["exp", "rlsin"])
def test_reference_evalmod_program_same_output_emulated(emul_env, op):
    """tests/gpqhe.c `exp` / `rlsin` (logq=438, Delta=2^30, iter=5: he_exp runs 12 he_mul, he_rlsin adds he_conj,
    he_sub, he_mulpt) through the drop-in's single device-resident call: same printed numbers as the reference."""
    args = [op, "pk", "--logn=6"]
    _, ours = _run([os.path.join(BUILD, "test-gpqhe"), *args], env=emul_env)
    _, ref = _run([os.path.join(ROOT, "oracle", "_ref", "test-gpqhe"), *args])
    assert _verdicts(ours) and _verdicts(ours) == _verdicts(ref), (ours[-1500:], ref[-1500:])
    assert _diffs(ours) and _diffs(ours) == _diffs(ref), (ours[-1500:], ref[-1500:])


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")), reason="reference build absent")
def test_reference_coeff2slot_program_same_output_emulated(emul_env):
    """tests/gpqhe.c coeff2slot: the reference's own CoeffToSlot fails its assert; the drop-in must
    fail it with the same printed numbers."""
    args = ["coeff2slot", "pk", "--logn=7", "--logq=109", "--slots=4", "--logDelta=30"]
    _, ours = _run([os.path.join(BUILD, "test-gpqhe"), *args], env=emul_env)
    _, ref = _run([os.path.join(ROOT, "oracle", "_ref", "test-gpqhe"), *args])
    assert _verdicts(ours) and _verdicts(ours) == _verdicts(ref), (ours[-1500:], ref[-1500:])
    assert _diffs(ours) and _diffs(ours) == _diffs(ref), (ours[-1500:], ref[-1500:])


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")), reason="reference build absent")
@pytest.mark.parametrize("op", ["inv", "sqrt", "sigmoid", "nrm2", "sum", "idx", "log", "cmp", "add", "enc"])
def test_reference_algo_program_same_output_emulated(emul_env, op):
    """The compositions of src/he-algo.c that are NOT hand-ported (he_inv :131, he_sqrt :166, he_sigmoid, he_log,
    he_cmp, he_sum, he_idx, he_nrm2) run as the reference wrote them, primitive by primitive, on the device shadows
    of their ciphertexts (he_copy_ct / he_free_ct / he_dec wrapped): same printed numbers as the reference."""
    args = [op, "pk", "--logn=6"]
    _, ours = _run([os.path.join(BUILD, "test-gpqhe"), *args], env=emul_env)
    _, ref = _run([os.path.join(ROOT, "oracle", "_ref", "test-gpqhe"), *args])
    assert _verdicts(ours) and _verdicts(ours) == _verdicts(ref), (ours[-1500:], ref[-1500:])
    assert _diffs(ours) == _diffs(ref), (ours[-1500:], ref[-1500:])


def test_dropin_eager_mode_emulated(emul_env):
    """GPQHE_DROPIN_EAGER=1 writes every result back into the MPIs at once (for callers that read ct->c0.coeffs
    directly without the sync hook): same results"""
    env = dict(emul_env, GPQHE_DROPIN_EAGER="1")
    rc, out = _run([os.path.join(BUILD, "dropin_diff"), "6", "120", "4", "40", "4"], env=env)
    assert rc == 0 and "PASS" in out, out[-2000:]


def test_reference_polymul_program_emulated(emul_env):
    """tests/polymul.c (N=2^7, the polynomial tests/polymul.gp computes): same printout."""
    _, ours = _run([os.path.join(BUILD, "test-polymul")], env=emul_env)
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "test-polymul")
    if not os.path.exists(ref_bin):
        pytest.skip("reference build absent")
    _, ref = _run([ref_bin])
    strip = lambda s: [re.sub(r"\d\d:\d\d:\d\d", "", re.sub(r"\x1b\[[0-9;]*m", "", l)) for l in s.splitlines()
                       if "us " not in l and "ms " not in l and "Kb used" not in l]
    assert strip(ours) == strip(ref)


def _keys(out):
    return [l for l in out.splitlines() if l.startswith("key ")]


def _check_keydump(args, env):
    """tests/keydump.c against the drop-in and against the unmodified reference (same -DSUPERCOP stream): sk, pk and
    every evaluation-key buffer (rlk, ck, rk[0..slots)) must be byte-identical (src/he-kem.c:74-169)"""
    ours_bin, ref_bin = os.path.join(BUILD, "keydump"), os.path.join(ROOT, "oracle", "_ref", "keydump")
    if not (os.path.exists(ours_bin) and os.path.exists(ref_bin)):
        pytest.skip("keydump builds absent")
    rc, ours = _run([ours_bin, *args], env=env, timeout=1800)
    assert rc == 0, ours[-1500:]
    rc, ref = _run([ref_bin, *args], timeout=1800)
    assert rc == 0, ref[-1500:]
    assert len(_keys(ours)) == 4 + int(args[2]) and _keys(ours) == _keys(ref), (ours[-1500:], ref[-1500:])
    return ours, ref


@pytest.mark.parametrize("args", [("7", "120", "4", "40"), ("6", "438", "8", "50")])
def test_keygen_byte_identical_emulated(emul_env, args):
    _check_keydump(args, emul_env)


def _crt_printout(binary, env=None):
    """tests/crt.c (-DTEST_CRT: 9-bit primes) prints the prime table, P, phat, phat_invmp and the residues of
    rns_decompose, then crashes in BOTH builds: it passes `a.coeffs` (MPI *) where rns_reconstruct takes an MPI
    (tests/crt.c:107 vs src/rns.c:60) - a stale test upstream, like tests/misc.c.  Everything printed before that
    point is compared (stdbuf: unbuffered stdout survives the crash)."""
    p = subprocess.run(["stdbuf", "-o0", binary], env=env, cwd="/tmp", stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, errors="replace", timeout=300)
    keep = []
    for l in p.stdout.splitlines():
        l = re.sub(r"\d\d:\d\d:\d\d", "", re.sub(r"\x1b\[[0-9;]*m", "", l))
        if "us " in l or "ms " in l or "Kb used" in l or "mmutable" in l:
            continue
        keep.append(l)
    return keep


def _check_crt_program(env):
    ours_bin = os.path.join(BUILD, "crt", "test-crt")
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "test-crt")
    if not (os.path.exists(ours_bin) and os.path.exists(ref_bin)):
        pytest.skip("test-crt builds absent")
    ours, ref = _crt_printout(ours_bin, env), _crt_printout(ref_bin)
    assert any(l.strip().startswith("[               ] ahat[5] =") for l in ours), ours[-5:]
    assert ours == ref


def test_reference_crt_program_emulated(emul_env):
    """tests/crt.c against the drop-in built with -DTEST_CRT (dropin/_build/crt): same printout as the reference"""
    _check_crt_program(emul_env)


# ------------------------------------------------------------------------------- GPU tier
@pytest.mark.gpu
def test_keygen_byte_identical_gpu():
    """he_keypair / he_genrlk / he_genck / he_genrk at the KAT ring (N=2^14, logq=438): the drop-in's keys (decompose,
    NTT and poly_mul on the GPU, samplers on the host) equal the reference's byte for byte"""
    _build_dropin()
    if not _have_gcrypt():
        pytest.skip("libgcrypt.so.20 is not installed on this machine")
    ours, ref = _check_keydump(("14", "438", "4", "50"), None)
    keep = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(keep):
        with open(os.path.join(keep, "keydump_n14.log"), "w") as f:
            f.write("drop-in:   " + [l for l in ours.splitlines() if l.startswith("time")][-1] + "\n")
            f.write("reference: " + [l for l in ref.splitlines() if l.startswith("time")][-1] + "\n")


@pytest.mark.gpu
def test_reference_crt_program_gpu():
    _build_dropin()
    _check_crt_program(None)



@pytest.mark.gpu
@pytest.mark.parametrize("args", [("14", "438", "16", "50", "16"), ("16", "438", "16", "50", "2"), ("12", "109", "4", "30", "4")])
def test_dropin_diff_gpu(args):
    """nrot >= slots at the KAT ring (N=2^14, slots=16) so that the he_gemv / he_coeff2slot block of
    tests/dropin_diff.c runs against the reference's own host loops at the size configs[3] is quoted on"""
    _build_dropin()
    env = dict(os.environ)
    if args[0] == "16":
        env["DROPIN_DIFF_NO_EXP"] = "1"      # the CPU side of he_exp / he_rlsin takes minutes at N=2^16; N=2^14 covers it
    rc, out = _run([os.path.join(BUILD, "dropin_diff"), *args], env=env, timeout=1800)
    keep = os.path.join(ROOT, "gpurun_out")          # the drop-in vs reference wall times it prints are worth keeping
    if os.path.isdir(keep):
        with open(os.path.join(keep, "dropin_diff_n%s_q%s.log" % (args[0], args[1])), "w") as f:
            f.write(out)
    assert rc == 0 and "PASS" in out, out[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("op", ["mul", "rot", "conj", "gemv", "exp", "rlsin", "inv", "sqrt", "sigmoid", "nrm2"])
def test_reference_kat_program_gpu(op):
    """configs[0]: ./test-gpqhe <op> pk at the default params.h ring (N=2^14, logq=438, Delta=2^50,
    -DSUPERCOP stream), the reference's unmodified test program on the GPU library."""
    _build_dropin()
    rc, out = _run([os.path.join(BUILD, "test-gpqhe"), op, "pk"], timeout=1800)
    clean = re.sub(r"\x1b\[[0-9;]*m", "", out)
    assert "PASSED" in clean and "FAILED" not in clean, clean[-3000:]
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")
    if op in ("mul", "gemv", "inv") and os.path.exists(ref_bin) and _have_gcrypt():
        _, ref = _run([ref_bin, op, "pk"], timeout=1800)
        assert _diffs(out) and _diffs(out) == _diffs(ref), "decoded results differ from the reference run"


@pytest.mark.gpu
def test_reference_coeff2slot_program_gpu():
    """configs[4]: ./test-gpqhe coeff2slot pk with the parameters of tests/gpqhe.c:1096.  The reference's own
    CoeffToSlot is unfinished (he_const_pt(1/n) is an integer division) and FAILS its assert;
    the drop-in must fail it with the very same printed numbers."""
    _build_dropin()
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "test-gpqhe")
    if not (os.path.exists(ref_bin) and _have_gcrypt()):
        pytest.skip("reference build absent")
    args = ["coeff2slot", "pk", "--logn=12", "--logq=109", "--slots=4", "--logDelta=30"]
    _, ours = _run([os.path.join(BUILD, "test-gpqhe"), *args], timeout=1800)
    _, ref = _run([ref_bin, *args], timeout=1800)
    assert _verdicts(ours) and _verdicts(ours) == _verdicts(ref), (ours[-1500:], ref[-1500:])
    assert _diffs(ours) and _diffs(ours) == _diffs(ref), (ours[-1500:], ref[-1500:])
