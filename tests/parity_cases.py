"""Parity checks shared by the CPU tier (library compiled against tests/emul) and the GPU
tier (the real sm_100a library).  Every check goes through the C ABI via gpqhe_b200.he."""
import os
import random

import numpy as np
from pytest import raises as pytest_raises

import gpqhe_b200 as g
from oracle import pyoracle as o

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["pow2_n256_q120", "pow2_n64_q438", "general_n128"]


def words_to_int(w):
    return int.from_bytes(np.asarray(w, dtype="<u8").tobytes(), "little")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return z, int(z["logn"]), words_to_int(z["q"]), int(z["slots"]), int(z["Delta"])


def eq(a, b, what):
    assert np.array_equal(np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64)), what


def check_golden(lib_path, name):
    """Product vs the outputs the unmodified reference produced (tests/golden/*.npz)."""
    z, logn, q, slots, Delta = load_golden(name)
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path)
    try:
        W, n = ctx.W, ctx.n
        assert (ctx.L, ctx.dim, ctx.dimevk, ctx.dimub, W) == tuple(int(z[k]) for k in ("L", "dim", "dimevk", "dimub", "W"))
        eq(ctx.primes, z["primes"], "primes")
        for d in range(ctx.dimub):
            eq(ctx.prime_consts(d), z["prime_consts"][d][:4], f"prime consts {d}")
        for tag in ("a", "b"):
            d = int(z[f"limb_{tag}"])
            x, y = z[f"x_{tag}"], z[f"y_{tag}"]
            eq(ctx.ntt(x, d), z[f"ntt_{tag}"], "ntt")
            eq(ctx.invntt(y, d), z[f"invntt_{tag}"], "invntt")
            eq(ctx.poly_rns_mul(x, y, d), z[f"mul_{tag}"], "poly_rns_mul")
            eq(ctx.poly_rns_add(x, y, d), z[f"add_{tag}"], "poly_rns_add")
            eq(ctx.ntt(z[f"hz_{tag}"], d), z[f"hzntt_{tag}"], "ntt non-canonical q vector")
            assert int((z[f"hzntt_{tag}"] == z["primes"][d]).sum()) > 0
            eq(ctx.invntt(z[f"hzntt_{tag}"], d), z[f"hzinv_{tag}"], "invntt of non-canonical input")
        eq(ctx.rns_decompose(z["big_a"], 0, ctx.dimub), z["decompose_a"], "rns_decompose")
        q2 = words_to_int(z["rns2mpi_q2"])
        eq(ctx.poly_rns2mpi(z["rns2mpi_in"], ctx.dimub, q, W), z["rns2mpi_full"], "poly_rns2mpi full")
        eq(ctx.poly_rns2mpi(z["rns2mpi_in"][:ctx.dim * n], ctx.dim, q2, W), z["rns2mpi_dim"], "poly_rns2mpi dim")
        eq(ctx.poly_mul(z["big_a"], z["big_b"], ctx.dimub, q, W), z["poly_mul"], "poly_mul")
        for rot in (1, 5):
            eq(ctx.poly_rot(z["big_a"], rot), z[f"poly_rot{rot}"], "poly_rot")
        eq(ctx.poly_conj(z["big_a"]), z["poly_conj"], "poly_conj")
        evk = ctx.genevk(z["evk_p0"], z["evk_p1"])
        for l in [int(v) for v in z["levels"]]:
            ct1 = g.HeCt(l, z[f"l{l}_a0"], z[f"l{l}_a1"])
            ct2 = g.HeCt(l, z[f"l{l}_b0"], z[f"l{l}_b1"])
            r = ctx.he_mul(ct1, ct2, evk)
            eq(r.c0, z[f"l{l}_mul_c0"], f"he_mul c0 l={l}"); eq(r.c1, z[f"l{l}_mul_c1"], f"he_mul c1 l={l}")
            r = ctx.he_rs(r)
            assert r.l == l - 1
            eq(r.c0, z[f"l{l}_mulrs_c0"], "he_rs c0"); eq(r.c1, z[f"l{l}_mulrs_c1"], "he_rs c1")
            r = ctx.he_moddown(ct1)
            eq(r.c0, z[f"l{l}_moddown_c0"], "he_moddown"); eq(r.c1, z[f"l{l}_moddown_c1"], "he_moddown")
            for rot in (0, 3):
                r = ctx.he_rot(ct1, rot, {rot: evk})
                eq(r.c0, z[f"l{l}_rot{rot}_c0"], f"he_rot {rot}"); eq(r.c1, z[f"l{l}_rot{rot}_c1"], f"he_rot {rot}")
            r = ctx.he_conj(ct1, evk)
            eq(r.c0, z[f"l{l}_conj_c0"], "he_conj"); eq(r.c1, z[f"l{l}_conj_c1"], "he_conj")
            r = ctx.he_mulpt(ct1, g.HePt(z[f"l{l}_pt_m"], nu=float(1 << 40)))
            eq(r.c0, z[f"l{l}_mulpt_c0"], "he_mulpt"); eq(r.c1, z[f"l{l}_mulpt_c1"], "he_mulpt")
    finally:
        ctx.close()


def rlimbs(rng, primes, n, count):
    return np.concatenate([np.array([rng.randrange(primes[d]) for _ in range(n)], dtype=np.uint64)
                           for d in range(count)])


def check_limb_layer_vs_oracle(lib_path, logn, q, seed=0, limbs=None, logp=59):
    """ntt / invntt / poly_rns_mul / poly_rns_add vs oracle.c on seeded inputs + edge vectors."""
    rng = random.Random(seed)
    ctx = g.HeContext(logn, q, lib_path=lib_path, logp=logp)
    oc = o.Ctx(logn, q, logp=logp)
    try:
        n = ctx.n
        assert ctx.primes == oc.primes
        for d in (limbs or (0, ctx.dimub - 1)):
            nd = oc.rns[d]
            p = nd.p
            assert ctx.prime_consts(d) == [nd.p, nd.pinv_mont, nd.pinv_barr, nd.ninv]
            a = np.array([rng.randrange(p) for _ in range(n)], dtype=np.uint64)
            b = np.array([rng.randrange(p) for _ in range(n)], dtype=np.uint64)
            f = ctx.ntt(a, d)
            eq(f, o.ntt(a, nd), f"ntt logn={logn} limb={d}")
            eq(ctx.invntt(f, d), a, "invntt(ntt(a)) == a")
            eq(ctx.invntt(b, d), o.invntt(b, nd), "invntt")
            eq(ctx.poly_rns_mul(a, b, d), o.poly_rns_mul(a, b, nd), "poly_rns_mul")
            eq(ctx.poly_rns_add(a, b, d), o.poly_rns_add(a, b, nd), "poly_rns_add")
            # edge vectors: all zero, all p-1, unit impulses, and the non-canonical-q case
            for v in (np.zeros(n, dtype=np.uint64), np.full(n, p - 1, dtype=np.uint64)):
                eq(ctx.ntt(v, d), o.ntt(v, nd), "ntt edge")
                eq(ctx.invntt(v, d), o.invntt(v, nd), "invntt edge")
            e = np.zeros(n, dtype=np.uint64); e[n - 1] = 1
            eq(ctx.ntt(e, d), o.ntt(e, nd), "ntt impulse")
            zplain = (int(nd.zetas[1]) * pow(1 << 64, -1, p)) % p
            h = np.zeros(n, dtype=np.uint64); h[n // 2] = 1; h[0] = p - zplain
            hf = o.ntt(h, nd)
            assert int((hf == p).sum()) > 0, "hazard vector must produce the value p"
            eq(ctx.ntt(h, d), hf, "ntt keeps q (not 0) on the + leg")
            eq(ctx.invntt(hf, d), o.invntt(hf, nd), "invntt accepts q")
    finally:
        ctx.close()


def rpoly(rng, qq, n):
    return [rng.randrange(qq) - qq // 2 for _ in range(n)]


def check_he_vs_oracle(lib_path, logn, q, slots, Delta, seed=0, levels=None, rots=(0, 3), logqub=0):
    """MPI layer + he_mul / he_rs / he_moddown / he_rot / he_conj / he_mulpt vs pyoracle."""
    rng = random.Random(seed)
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, slots, Delta, logqub=logqub or None)
    try:
        n, W = ctx.n, ctx.W
        assert (ctx.L, ctx.dim, ctx.dimevk, ctx.dimub) == (oc.L, oc.dim, oc.dimevk, oc.dimub)
        a = rpoly(rng, q, n); aw = g.ints_to_big(a, W)
        b = rpoly(rng, q, n); bw = g.ints_to_big(b, W)
        dec = ctx.rns_decompose(aw, 0, ctx.dimub)
        for d in range(ctx.dimub):
            eq(dec[d], o.rns_decompose(a, oc.rns[d]), f"rns_decompose limb {d}")
        # widest basis the CRT kernels hold: 48 words (3072 bits); the reference's HE path never goes beyond dimevk
        # limbs (47 at logq = 881), only polyctx allocates more primes (dimub)
        dmax = ctx.dimub if ctx.dimub * 59 + 59 <= 3072 else ctx.dimevk
        for dim in sorted({1, ctx.dim, dmax}):
            rh = rlimbs(rng, oc.primes, n, dim)
            for qq in (q, oc.rns[dim - 1].P * 3 + 1, 12345):
                if qq.bit_length() > 2040:        # moduli (q, P*q_L) stay below 2048 bits in every HE path
                    continue
                Wr = max(1, (qq.bit_length() + 63) // 64)
                got = g.big_to_ints(ctx.poly_rns2mpi(rh, dim, qq, Wr))
                assert got == o.poly_rns2mpi(rh, oc.rns[dim - 1], qq, n), ("poly_rns2mpi", dim)
        assert g.big_to_ints(ctx.poly_mul(aw, bw, dmax, q)) == o.poly_mul(oc, a, b, dmax, q)
        for rot in (0, 1, 7):
            assert g.big_to_ints(ctx.poly_rot(aw, rot)) == o.poly_rot(a, rot, n)
        assert g.big_to_ints(ctx.poly_conj(aw)) == o.poly_conj(a, n)
        p0, p1 = rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk)
        evk = ctx.genevk(p0, p1)
        for l in (levels or range(oc.L, 0, -1)):
            ql = oc.ql[l]
            a0, a1, b0, b1 = (rpoly(rng, ql, n) for _ in range(4))
            c1 = g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W))
            c2 = g.HeCt(l, g.ints_to_big(b0, W), g.ints_to_big(b1, W))
            d1, d2 = {"l": l, "c0": a0, "c1": a1}, {"l": l, "c0": b0, "c1": b1}
            r, e = ctx.he_mul(c1, c2, evk), o.he_mul(oc, d1, d2, (p0, p1))
            assert g.big_to_ints(r.c0) == e["c0"] and g.big_to_ints(r.c1) == e["c1"], ("he_mul", l)
            r2, e2 = ctx.he_rs(r), o.he_rs(oc, e)
            assert r2.l == e2["l"] and g.big_to_ints(r2.c0) == e2["c0"] and g.big_to_ints(r2.c1) == e2["c1"], ("he_rs", l)
            # one-crossing he_mul+he_rs and the pipelined batch (3 products: both staging buffers reused)
            rf = ctx.he_mul_rs(c1, c2, evk)
            assert rf.l == e2["l"] and g.big_to_ints(rf.c0) == e2["c0"] and g.big_to_ints(rf.c1) == e2["c1"], ("he_mul_rs", l)
            eb = o.he_rs(oc, o.he_mul(oc, d2, d2, (p0, p1)))
            # 7 products over 3 compute lanes and 6 staging buffers: lanes and buffers are both reused
            for resc, want in ((True, [e2, eb, e2] * 2 + [e2]), (False, [e, None, e] * 2 + [e])):
                rb = ctx.he_mul_batch([(c1, c2), (c2, c2), (c2, c1)] * 2 + [(c1, c2)], evk, rescale=resc)
                assert ctx.lanes == 3
                for got_ct, w in zip(rb, want):
                    if w is not None:
                        assert got_ct.l == w["l"] and g.big_to_ints(got_ct.c0) == w["c0"] and \
                            g.big_to_ints(got_ct.c1) == w["c1"], ("he_mul_batch", l, resc)
            r3, e3 = ctx.he_moddown(c1), o.he_moddown(oc, d1)
            assert g.big_to_ints(r3.c0) == e3["c0"] and g.big_to_ints(r3.c1) == e3["c1"], ("he_moddown", l)
            for rot in rots:
                r4, e4 = ctx.he_rot(c1, rot, {rot: evk}), o.he_rot(oc, d1, rot, (p0, p1))
                assert g.big_to_ints(r4.c0) == e4["c0"] and g.big_to_ints(r4.c1) == e4["c1"], ("he_rot", l, rot)
            # hoisted rotations: decompose + NTT of c1 shared, same bits as separate he_rot calls
            hrots = sorted(set(rots) | {0, 1, 5})
            for rot, r4 in zip(hrots, ctx.he_rot_hoisted(c1, hrots, {r: evk for r in hrots})):
                e4 = o.he_rot(oc, d1, rot, (p0, p1))
                assert g.big_to_ints(r4.c0) == e4["c0"] and g.big_to_ints(r4.c1) == e4["c1"], ("he_rot_hoisted", l, rot)
            r5, e5 = ctx.he_conj(c1, evk), o.he_conj(oc, d1, (p0, p1))
            assert g.big_to_ints(r5.c0) == e5["c0"] and g.big_to_ints(r5.c1) == e5["c1"], ("he_conj", l)
            m = [rng.randrange(1 << 40) - (1 << 39) for _ in range(n)]
            pt = g.HePt(g.ints_to_big(m, W), nu=float(1 << 40))
            r6, e6 = ctx.he_mulpt(c1, pt), o.he_mulpt(oc, d1, m, ctx.mulpt_dim(l, pt.nu))
            assert g.big_to_ints(r6.c0) == e6["c0"] and g.big_to_ints(r6.c1) == e6["c1"], ("he_mulpt", l)
            # coefficient-local ops (src/he-add.c); edge values of the centred range included
            edge = [-(ql // 2) - (ql & 1), ql // 2 - 1, 0, -1, 1, ql // 2 - 1, -(ql // 2) - (ql & 1), ql // 3]
            x0, x1 = list(a0), list(a1)
            y0, y1 = list(b0), list(b1)
            x0[:8], y0[:8] = edge, edge[::-1]
            x1[:8], y1[:8] = edge, edge
            cx, cy = g.HeCt(l, g.ints_to_big(x0, W), g.ints_to_big(x1, W)), g.HeCt(l, g.ints_to_big(y0, W), g.ints_to_big(y1, W))
            dx, dy = {"l": l, "c0": x0, "c1": x1}, {"l": l, "c0": y0, "c1": y1}
            for name, got_ct, w in (("he_add", ctx.he_add(cx, cy), o.he_add(oc, dx, dy)),
                                    ("he_sub", ctx.he_sub(cx, cy), o.he_sub(oc, dx, dy)),
                                    ("he_neg", ctx.he_neg(cx), o.he_neg(oc, dx)),
                                    ("he_addpt", ctx.he_addpt(cx, pt), o.he_addpt(oc, dx, m, 1)),
                                    ("he_subpt", ctx.he_subpt(cx, pt), o.he_addpt(oc, dx, m, -1))):
                assert g.big_to_ints(got_ct.c0) == w["c0"] and g.big_to_ints(got_ct.c1) == w["c1"], (name, l)
            # aliasing contract: he_mul(&x,&x,&x) == he_mul with distinct buffers
            r7 = ctx.he_mul(c1, c1, evk)
            e7 = o.he_mul(oc, d1, d1, (p0, p1))
            assert g.big_to_ints(r7.c0) == e7["c0"], "he_mul(x,x)"
    finally:
        ctx.close()


def check_crt_edges(lib_path, logn=5, q=1 << 438, Delta=1 << 50, logqub=438):
    """Edge values of the centred CRT range (the +P/2 boundary, 0, -1 ...) through poly_rns2mpi and,
    with a crafted switching key, through the key-switch epilogue: these are the inputs on which
    the fast low-word CRT kernels cannot decide the overflow count and must hand the coefficient
    to the exact full-width kernel.  Bit-exact vs the oracle, and the hand-over must have happened."""
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub)
    try:
        n, W = ctx.n, ctx.W
        before = ctx.crt_fixups
        for dim in sorted({2, ctx.dim, ctx.dimevk}):
            P = oc.rns[dim - 1].P
            Ph = P // 2
            edge = [Ph - 1, Ph, Ph + 1, 0, 1, P - 1, Ph - 2, Ph + 2, P - Ph, 12345, P - 12345, Ph - (1 << 64)]
            vals = [edge[i % len(edge)] for i in range(n)]
            rh = np.stack([np.array([v % oc.primes[d] for v in vals], dtype=np.uint64) for d in range(dim)]).reshape(-1)
            for qq in (q, 1 << 64, 1 << 65, 2):
                Wr = max(1, (qq.bit_length() + 63) // 64)
                got = g.big_to_ints(ctx.poly_rns2mpi(rh, dim, qq, Wr))
                assert got == o.poly_rns2mpi(rh, oc.rns[dim - 1], qq, n), ("poly_rns2mpi edge", dim, qq.bit_length())
        assert ctx.crt_fixups > before, "edge values must take the exact list path"
        # key switch: c1 = 1 (constant), evk = NTT(residues of Z): the epilogue then sees exactly Z
        before = ctx.crt_fixups
        l = oc.L
        D2 = ctx.dims(l)[1]
        PB = oc.rns[D2 - 1].P
        Ph = PB // 2
        edge = [Ph - 1, Ph, Ph + 1, 0, PB - 1, 1, Ph - 5, 7]
        # ... and the decision boundaries of T mod P over the special modulus: r = 0, P - 1 (the overflow count of the
        # inner CRT) and r = (P -+ 1) / 2 (the rounding of T / P), for positive and negative T
        Ps = oc.rns[ctx.dim - 1].P
        edge += [3 * Ps + Ps // 2, 3 * Ps + Ps // 2 + 1, 9 * Ps, 9 * Ps - 1, 9 * Ps + 1, PB - 4 * Ps - Ps // 2,
                 PB - 4 * Ps - Ps // 2 - 1, PB - 6 * Ps, PB - 6 * Ps + 1, Ps // 2, Ps // 2 + 1, Ps - 1, Ps]
        edge = [v % PB for v in edge]
        z0 = [edge[i % len(edge)] for i in range(n)]
        z1 = [edge[(i + 3) % len(edge)] for i in range(n)]

        def key(z):
            out = np.zeros((oc.dimevk, n), dtype=np.uint64)
            for d in range(oc.dimevk):
                out[d] = o.ntt(np.array([v % oc.primes[d] for v in z], dtype=np.uint64), oc.rns[d])
            return out.reshape(-1)
        p0, p1 = key(z0), key(z1)
        evk = ctx.genevk(p0, p1)
        one = [1] + [0] * (n - 1)
        c0 = [((i * 977) % 1000) - 500 for i in range(n)]
        ct = g.HeCt(l, g.ints_to_big(c0, W), g.ints_to_big(one, W))
        r = ctx.he_rot(ct, 0, {0: evk})
        e = o.he_rot(oc, {"l": l, "c0": c0, "c1": one}, 0, (p0, p1))
        assert g.big_to_ints(r.c0) == e["c0"] and g.big_to_ints(r.c1) == e["c1"], "key-switch epilogue at the P/2 edge"
        assert ctx.crt_fixups > before, "edge values must take the exact list path (key switch)"
    finally:
        ctx.close()


def check_gemv(lib_path, logn, q, Delta, n1=2, n2=2, logqub=0, seed=3):
    """Device-resident he_gemv (baby-step/giant-step, hoisted baby rotations) vs the oracle's
    restatement of the reference loop, distinct rotation keys, at the top level."""
    rng = random.Random(seed)
    slots = n1 * n2
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, slots, Delta, logqub=logqub or None)
    try:
        n, W, l = ctx.n, ctx.W, ctx.L
        ql = oc.ql[l]
        a0, a1 = rpoly(rng, ql, n), rpoly(rng, ql, n)
        need = sorted(set(range(n1)) | {i * n1 for i in range(n2)})
        keys = {r: (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk)) for r in need}
        dkeys = {r: ctx.genevk(*keys[r]) for r in need}
        nu = float(Delta)
        ms = [[rng.randrange(2 * Delta) - Delta for _ in range(n)] for _ in range(slots)]
        pts = [g.HePt(g.ints_to_big(m, W), nu=nu) for m in ms]
        got = ctx.he_gemv_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts, n1, n2, dkeys)
        want = o.he_gemv_encoded(oc, {"l": l, "c0": a0, "c1": a1}, ms, ctx.mulpt_dim(l, nu), n1, n2, keys)
        assert got.l == want["l"]
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_gemv"
        # with a plaintext bound the n1 products of a giant step are accumulated in the NTT domain (one inverse NTT
        # + CRT per giant step): same bits, fewer launches; a bound too large for any basis falls back
        before = ctx.launch_count
        got = ctx.he_gemv_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts, n1, n2, dkeys)
        per_term = ctx.launch_count - before
        ctx.set_plaintext_bits(ctx.plaintext_bits(pts))
        before = ctx.launch_count
        got = ctx.he_gemv_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts, n1, n2, dkeys)
        accumulated = ctx.launch_count - before
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_gemv, NTT-domain accumulation"
        assert accumulated < per_term, (accumulated, per_term)
        ctx.set_plaintext_bits(64 * W)
        got = ctx.he_gemv_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts, n1, n2, dkeys)
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_gemv, bound too large"
        ctx.set_plaintext_bits(ctx.plaintext_bits(pts))
        # giant steps spread over compute lanes: same bits, twice (scratch of every lane reused)
        ctx.set_lanes(3)
        for _ in range(2):
            got = ctx.he_gemv_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts, n1, n2, dkeys)
            assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_gemv over lanes"
    finally:
        ctx.close()


def check_coeff2slot(lib_path, logn, q, Delta, n1=2, n2=2, logqub=0, seed=5, lanes=1):
    """Device-resident he_coeff2slot (src/he-bootstrap.c:359-409: he_conj, 4 he_gemv, he_add, he_mulpt by
    the constant plaintext, he_rs) vs the oracle's restatement of the same composition."""
    rng = random.Random(seed)
    slots = n1 * n2
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, slots, Delta, logqub=logqub or None)
    try:
        if lanes > 1:
            ctx.set_lanes(lanes)
        n, W, l = ctx.n, ctx.W, ctx.L
        assert l >= 2
        ql = oc.ql[l]
        a0, a1 = rpoly(rng, ql, n), rpoly(rng, ql, n)
        need = sorted(set(range(n1)) | {i * n1 for i in range(n2)})
        keys = {r: (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk)) for r in need}
        ckey = (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk))
        dkeys = {r: ctx.genevk(*keys[r]) for r in need}
        dck = ctx.genevk(*ckey)
        nu = float(Delta)
        ms4 = [[[rng.randrange(2 * Delta) - Delta for _ in range(n)] for _ in range(slots)] for _ in range(4)]
        pts4 = [[g.HePt(g.ints_to_big(m, W), nu=nu) for m in ms] for ms in ms4]
        # he_const_pt (src/he-encode.c:119-125) touches coefficients 0 and n/2 only
        cm = [0] * n
        cm[0], cm[n // 2] = rng.randrange(Delta), -rng.randrange(Delta)
        cpt = g.HePt(g.ints_to_big(cm, W), nu=nu)
        re, im = ctx.he_coeff2slot_encoded(g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W)), pts4, cpt,
                                           n1, n2, dck, dkeys)
        wre, wim = o.he_coeff2slot_encoded(oc, {"l": l, "c0": a0, "c1": a1}, ms4, cm, ctx.mulpt_dim(l, nu),
                                           ctx.mulpt_dim(l - 1, nu), n1, n2, ckey, keys)
        for name, got, want in (("real", re, wre), ("imag", im, wim)):
            assert got.l == want["l"] == l - 2
            assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], ("he_coeff2slot", name)
    finally:
        ctx.close()


def check_rlsin(lib_path, logn, q, Delta, iters=1, logqub=0, seed=7, lanes=1, start_level=None):
    """Device-resident he_exp and he_rlsin (src/he-algo.c:364-458, src/he-bootstrap.c:444-475: the EvalMod
    step of bootstrapping) vs the oracle's restatement of the same compositions, plaintexts given."""
    rng = random.Random(seed)
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub or None)
    try:
        if lanes > 1:
            ctx.set_lanes(lanes)
        n, W = ctx.n, ctx.W
        l = ctx.L if start_level is None else start_level
        assert l >= 5 + iters
        ql = oc.ql[l]
        a0, a1 = rpoly(rng, ql, n), rpoly(rng, ql, n)
        rlk = (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk))
        ckey = (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk))
        drlk, dck = ctx.genevk(*rlk), ctx.genevk(*ckey)
        nu = float(Delta)
        # pts[0], pts[8]: he_ecd output (dense, |m| around Delta); pts[1..7]: he_const_pt (coefficients 0 and n/2)
        dense = lambda: [rng.randrange(2 * Delta) - Delta for _ in range(n)]
        ms = [dense()]
        for k in range(7):
            cm = [0] * n
            cm[0] = rng.randrange(8 * Delta) if k % 2 == 0 else rng.randrange(Delta)
            ms.append(cm)
        ms.append(dense())
        pts = [g.HePt(g.ints_to_big(m, W), nu=nu) for m in ms]
        ct = g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W), nu, 1.0)
        dim_a, dim_t = ctx.mulpt_dim(l, nu), ctx.mulpt_dim(l - 1, nu)
        got = ctx.he_exp_encoded(ct, pts[:8], iters, drlk)
        want = o.he_exp_encoded(oc, {"l": l, "c0": a0, "c1": a1}, ms[:8], dim_a, dim_t, iters, rlk)
        assert got.l == want["l"] == l - 4 - iters
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_exp"
        got = ctx.he_rlsin_encoded(ct, pts, iters, drlk, dck)
        want = o.he_rlsin_encoded(oc, {"l": l, "c0": a0, "c1": a1}, ms, dim_a, dim_t,
                                  ctx.mulpt_dim(l - 4 - iters, nu), iters, rlk, ckey)
        assert got.l == want["l"] == l - 5 - iters
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_rlsin"
        with pytest_raises(g.GpqheError):
            ctx.he_rlsin_encoded(g.HeCt(4 + iters, ct.c0, ct.c1, nu, 1.0), pts, iters, drlk, dck)
    finally:
        ctx.close()


def check_slot2coeff(lib_path, logn, q, Delta, n1=2, n2=2, logqub=0, seed=11, lanes=1):
    """Device-resident he_slot2coeff (src/he-bootstrap.c:414-438) vs the oracle's composition."""
    rng = random.Random(seed)
    slots = n1 * n2
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, slots, Delta, logqub=logqub or None)
    try:
        if lanes > 1:
            ctx.set_lanes(lanes)
        n, W, l = ctx.n, ctx.W, ctx.L
        ql = oc.ql[l]
        cts = [(rpoly(rng, ql, n), rpoly(rng, ql, n)) for _ in range(2)]
        need = sorted(set(range(n1)) | {i * n1 for i in range(n2)})
        keys = {r: (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk)) for r in need}
        dkeys = {r: ctx.genevk(*keys[r]) for r in need}
        nu = float(Delta)
        ms2 = [[[rng.randrange(2 * Delta) - Delta for _ in range(n)] for _ in range(slots)] for _ in range(2)]
        pts2 = [[g.HePt(g.ints_to_big(m, W), nu=nu) for m in ms] for ms in ms2]
        hc = [g.HeCt(l, g.ints_to_big(a, W), g.ints_to_big(b, W), nu, 1.0) for a, b in cts]
        got = ctx.he_slot2coeff_encoded(hc[0], hc[1], pts2, n1, n2, dkeys)
        want = o.he_slot2coeff_encoded(oc, {"l": l, "c0": cts[0][0], "c1": cts[0][1]},
                                       {"l": l, "c0": cts[1][0], "c1": cts[1][1]}, ms2, ctx.mulpt_dim(l, nu),
                                       n1, n2, keys)
        assert got.l == want["l"] == l - 1
        assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], "he_slot2coeff"
    finally:
        ctx.close()


def check_lazy_ntt_vs_exact(lib_path, logn, q, device="cpu", limbs=None, seed=1):
    """The lazy transforms of the internal pipelines (gq_ntt_lazy.cu through gq_stage_ntt_dev: canonical in,
    canonical out, sums left unreduced on the way) against the exact ntt() / invntt() kernels, which are pinned
    to the reference, on vectors that maximise the growth of the unreduced values: all p-1, alternating
    0 / p-1, impulses, random."""
    import contextlib
    import ctypes as C
    import torch
    ctx = g.HeContext(logn, q, lib_path=lib_path, logqub=q.bit_length() - 1)
    try:
        n = ctx.n
        rng = np.random.default_rng(seed)
        cm = contextlib.nullcontext()
        if torch.device(device).type == "cuda":
            cm = torch.cuda.stream(torch.cuda.ExternalStream(ctx._lib.gq_ctx_stream(ctx._h), device=device))
        for d in (limbs or (0, ctx.dimub - 1)):
            p = ctx.primes[d]
            vecs = [np.full(n, p - 1, dtype=np.uint64), np.zeros(n, dtype=np.uint64),
                    rng.integers(0, p, size=n, dtype=np.uint64)]
            alt = np.zeros(n, dtype=np.uint64); alt[::2] = p - 1
            blk = np.zeros(n, dtype=np.uint64); blk[: n // 2] = p - 1
            imp = np.zeros(n, dtype=np.uint64); imp[n - 1] = p - 1
            vecs += [alt, blk, imp]
            for inverse in (0, 1):
                want = [(ctx.invntt(v, d) if inverse else ctx.ntt(v, d)) % np.uint64(p) for v in vecs]
                buf = torch.from_numpy(np.stack(vecs).view(np.int64)).to(device)        # [npolys][1 limb][n]
                with cm:
                    capi_check(ctx, ctx._lib.gq_stage_ntt_dev(ctx._h, d, 1, len(vecs), C.c_void_p(buf.data_ptr()),
                                                              inverse, None))
                    ctx.sync()
                got = buf.cpu().numpy().view(np.uint64)
                for k in range(len(vecs)):
                    eq(got[k], want[k], f"lazy {'inverse' if inverse else 'forward'} NTT logn={logn} limb={d} vector {k}")
    finally:
        ctx.close()


def capi_check(ctx, rc):
    from gpqhe_b200 import capi
    capi.check(ctx._lib, rc)


# ---------------------------------------------------------------------------------------------------
# Compositions at the sizes the configs are quoted on (N = 2^14 .. 2^16), where the Python big-integer
# oracle would take hours: the device-resident composite entry points (he_gemv, hoisted rotations,
# he_coeff2slot, he_exp / he_rlsin, he_slot2coeff) against the SAME compositions of oracle/pyoracle.py
# evaluated over the library's own primitives through the C ABI, one call per primitive.  The primitives
# are pinned to the oracle / the live reference at these sizes elsewhere (test_full_size_n16_he_mul_rs_rot,
# dropin_diff); the composition functions are the oracle's own code, so this closes the chain.
class _GpuPrims:
    """oracle.pyoracle's primitive names over gpqhe_b200.HeContext calls; ciphertexts are dicts holding
    word-major numpy arrays instead of lists of ints (the compositions never look inside)."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.keys = {}            # id(host key tuple) -> HeEvk

    def _ct(self, d):
        return g.HeCt(d["l"], d["c0"], d["c1"], d.get("nu", 1.0), d.get("B", 1.0))

    @staticmethod
    def _d(c):
        return {"l": c.l, "c0": c.c0, "c1": c.c1, "nu": c.nu, "B": c.B}

    def he_mul(self, _, a, b, rlk):
        return self._d(self.ctx.he_mul(self._ct(a), self._ct(b), rlk))

    def he_mulpt(self, _, a, m, dim):
        assert self.ctx.mulpt_dim(a["l"], m.nu) == dim
        return self._d(self.ctx.he_mulpt(self._ct(a), m))

    def he_rs(self, _, a):
        return self._d(self.ctx.he_rs(self._ct(a)))

    def he_moddown(self, _, a):
        return self._d(self.ctx.he_moddown(self._ct(a)))

    def he_rot(self, _, a, rot, rk_rot):
        return self._d(self.ctx.he_rot(self._ct(a), rot, {rot: rk_rot}))

    def he_conj(self, _, a, ck):
        return self._d(self.ctx.he_conj(self._ct(a), ck))

    def he_add(self, _, a, b):
        return self._d(self.ctx.he_add(self._ct(a), self._ct(b)))

    def he_sub(self, _, a, b):
        return self._d(self.ctx.he_sub(self._ct(a), self._ct(b)))

    def he_addpt(self, _, a, m, sign=1):
        return self._d(self.ctx.he_addpt(self._ct(a), m) if sign > 0 else self.ctx.he_subpt(self._ct(a), m))


class _patched_oracle:
    """run the oracle's composition functions with its primitives swapped for `prims`"""
    NAMES = ("he_mul", "he_mulpt", "he_rs", "he_moddown", "he_rot", "he_conj", "he_add", "he_sub", "he_addpt")

    def __init__(self, prims):
        self.prims, self.saved = prims, {}

    def __enter__(self):
        for k in self.NAMES:
            self.saved[k] = getattr(o, k)
            setattr(o, k, getattr(self.prims, k))

    def __exit__(self, *a):
        for k, v in self.saved.items():
            setattr(o, k, v)


def _rand_big(nr, n, W, bits):
    """uniform on the centred range of 2^bits, word-major two's complement"""
    w = nr.integers(0, 1 << 63, size=(W, n), dtype=np.uint64) * np.uint64(2) + nr.integers(0, 2, size=(W, n), dtype=np.uint64)
    sw, sb = (bits - 1) // 64, (bits - 1) % 64
    top = (w[sw].astype(np.int64) << np.int64(63 - sb)) >> np.int64(63 - sb)
    w[sw] = top.astype(np.uint64)
    for k in range(sw + 1, W):
        w[k] = (top >> np.int64(63)).astype(np.uint64)
    return w


def check_compositions_full_size(lib_path, logn=16, q=1 << 438, Delta=1 << 50, n1=4, n2=4, lanes=3, logqub=0,
                                 iters=1, seed=2016):
    nr = np.random.default_rng(seed)
    slots = n1 * n2
    ctx = g.HeContext(logn, q, slots, Delta, lib_path=lib_path, logqub=logqub)
    try:
        if lanes > 1:
            ctx.set_lanes(lanes)
        n, W, L = ctx.n, ctx.W, ctx.L
        ld = Delta.bit_length() - 1
        key = lambda: ctx.genevk(np.stack([nr.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(ctx.dimevk)]),
                                 np.stack([nr.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(ctx.dimevk)]))
        need = sorted(set(range(n1)) | {i * n1 for i in range(n2)})
        rk = {r: key() for r in need}
        rlk, ck = key(), key()
        nu = float(Delta)
        pt = lambda: g.HePt(_rand_big(nr, n, W, ld + 1), nu=nu)
        ct = lambda l: {"l": l, "c0": _rand_big(nr, n, W, ctx.ql[l].bit_length() - 1),
                        "c1": _rand_big(nr, n, W, ctx.ql[l].bit_length() - 1), "nu": nu, "B": 1.0}
        prims = _GpuPrims(ctx)
        same = lambda got, want, what: (eq(got.c0, want["c0"], what + " c0"), eq(got.c1, want["c1"], what + " c1"))
        hc = lambda d: g.HeCt(d["l"], d["c0"], d["c1"], d["nu"], d["B"])

        # hoisted rotations == he_rot, rotation by rotation
        a = ct(L)
        outs = ctx.he_rot_hoisted(hc(a), need, rk)
        for r, got in zip(need, outs):
            want = ctx.he_rot(hc(a), r, rk)
            eq(got.c0, want.c0, f"hoisted rotation {r} c0"); eq(got.c1, want.c1, f"hoisted rotation {r} c1")

        # he_gemv (configs[3]); the plaintext bound switches the device call to NTT-domain accumulation, the
        # composition below still runs one he_mulpt per term
        pts = [pt() for _ in range(slots)]
        ctx.set_plaintext_bits(ctx.plaintext_bits(pts) + 1)
        with _patched_oracle(prims):
            want = o.he_gemv_encoded(None, a, pts, ctx.mulpt_dim(L, nu), n1, n2, rk)
        same(ctx.he_gemv_encoded(hc(a), pts, n1, n2, rk), want, "he_gemv")

        # he_coeff2slot (configs[4], CoeffToSlot)
        pts4 = [[pt() for _ in range(slots)] for _ in range(4)]
        cm = np.zeros((W, n), dtype=np.uint64)
        cm[0, 0], cm[0, n // 2] = 12345, (1 << ld) - 77
        cpt = g.HePt(cm, nu=nu)
        with _patched_oracle(prims):
            wre, wim = o.he_coeff2slot_encoded(None, a, pts4, cpt, ctx.mulpt_dim(L, nu), ctx.mulpt_dim(L - 1, nu),
                                               n1, n2, ck, rk)
        re, im = ctx.he_coeff2slot_encoded(hc(a), pts4, cpt, n1, n2, ck, rk)
        same(re, wre, "he_coeff2slot real"); same(im, wim, "he_coeff2slot imag")

        # he_exp / he_rlsin (configs[4], EvalMod), as deep as the levels allow
        it = min(iters, L - 5) if L >= 5 else -1
        if it >= 0:
            epts = [pt()]
            for k in range(7):
                z = np.zeros((W, n), dtype=np.uint64)
                z[0, 0] = int(nr.integers(1, 1 << (ld + 2)))
                epts.append(g.HePt(z, nu=nu))
            epts.append(pt())
            with _patched_oracle(prims):
                want = o.he_rlsin_encoded(None, a, epts, ctx.mulpt_dim(L, nu), ctx.mulpt_dim(L - 1, nu),
                                          ctx.mulpt_dim(L - 4 - it, nu), it, rlk, ck)
            same(ctx.he_rlsin_encoded(hc(a), epts, it, rlk, ck), want, "he_rlsin")

        # he_slot2coeff (configs[4], SlotToCoeff) at a lower level
        ls = max(1, L - 3)
        b0, b1 = ct(ls), ct(ls)
        with _patched_oracle(prims):
            want = o.he_slot2coeff_encoded(None, b0, b1, pts4[:2], ctx.mulpt_dim(ls, nu), n1, n2, rk)
        same(ctx.he_slot2coeff_encoded(hc(b0), hc(b1), pts4[:2], n1, n2, rk), want, "he_slot2coeff")
    finally:
        ctx.close()


def check_fused_rescale_ties(lib_path, logn=5, q=1 << 438, Delta=1 << 50, logqub=438, seed=9):
    """he_rs fused into the key-switch epilogue (gq_he_mul_rs / gq_he_mul_batch): mpi_rdiv's tie rule
    (src/types.c:115-128: remainder == Delta/2 rounds toward -inf) on products whose remainder modulo Delta is
    exactly Delta/2, Delta/2 +- 1, 0 and Delta - 1, positive and negative.  Multiplying by the trivial ciphertext
    (1, 0) makes he_mul return its other operand unchanged (d2 = 0), so the values entering the division are chosen."""
    rng = random.Random(seed)
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
    oc = o.Ctx(logn, q, 4, Delta, logqub=logqub or None)
    try:
        n, W = ctx.n, ctx.W
        rlk = (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk))
        drlk = ctx.genevk(*rlk)
        for l in sorted({oc.L, max(1, oc.L // 2), 1}, reverse=True):
            ql, half = oc.ql[l], Delta // 2
            rems = [half, half + 1, half - 1, 0, Delta - 1, 1]
            vals = []
            for i in range(n):
                hi = rng.randrange(ql // Delta) - ql // (2 * Delta)
                v = hi * Delta + rems[i % len(rems)]
                vals.append(o.mpi_smod(v % ql, ql, ql // 2))
            a0, a1 = vals, [-v if -v >= -(ql // 2) and -v < ql - ql // 2 else v for v in reversed(vals)]
            a1 = [o.mpi_smod(x % ql, ql, ql // 2) for x in a1]
            one, zero = [1] + [0] * (n - 1), [0] * n
            cta = g.HeCt(l, g.ints_to_big(a0, W), g.ints_to_big(a1, W))
            ctb = g.HeCt(l, g.ints_to_big(one, W), g.ints_to_big(zero, W))
            want = o.he_rs(oc, o.he_mul(oc, {"l": l, "c0": a0, "c1": a1}, {"l": l, "c0": one, "c1": zero}, rlk))
            assert want["c0"] == o.he_rs(oc, {"l": l, "c0": a0, "c1": a1})["c0"]      # the construction holds
            got = ctx.he_mul_rs(cta, ctb, drlk)
            assert got.l == l - 1
            assert g.big_to_ints(got.c0) == want["c0"] and g.big_to_ints(got.c1) == want["c1"], ("fused he_mul+he_rs ties", l)
            got2 = ctx.he_mul_batch([(cta, ctb), (ctb, cta)], drlk, rescale=True)
            for r in got2:
                assert g.big_to_ints(r.c0) == want["c0"] and g.big_to_ints(r.c1) == want["c1"], ("batch ties", l)
            sep = ctx.he_rs(ctx.he_mul(cta, ctb, drlk))                                 # two kernels, same bits
            eq(sep.c0, got.c0, "fused == separate c0"); eq(sep.c1, got.c1, "fused == separate c1")
    finally:
        ctx.close()


def check_fused_vs_separate(lib_path, logn, q=1 << 438, Delta=1 << 50, logqub=438, seed=77, level=None, short=False):
    """The pointwise stage folded into the load of the inverse NTT (gq_ntt_lazy.cu, GQ_INV_*) and the CRT
    pre-multiplication folded into its final scaling, against the separate-kernel pipeline of the same library
    (GPQHE_B200_NO_FUSED_INV / GPQHE_B200_NO_PRESCALE, which the oracle pins at small sizes and the GPU tier at
    full size): he_mul, he_mul+he_rs, he_mulpt, he_rot, he_conj, hoisted rotations, poly_mul - identical bits."""
    nr = np.random.default_rng(seed)
    res = []
    for separate in (False, True, "no_tc"):
        if separate is True:
            os.environ["GPQHE_B200_NO_FUSED_INV"] = "1"
            os.environ["GPQHE_B200_NO_PRESCALE"] = "1"
        elif separate == "no_tc":            # the integer-pipe CRT instead of the tensor-core contraction (gq_crt_tc.cu)
            os.environ["GPQHE_B200_NO_TC_CRT"] = "1"
        try:
            ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=logqub)
            try:
                n, W = ctx.n, ctx.W
                l = ctx.L if level is None else level
                rr = np.random.default_rng(seed)          # the same inputs for both passes
                key = lambda: ctx.genevk(
                    np.stack([rr.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(ctx.dimevk)]),
                    np.stack([rr.integers(0, ctx.primes[d], size=n, dtype=np.uint64) for d in range(ctx.dimevk)]))
                k1, k2 = key(), key()
                bits = ctx.ql[l].bit_length() - 1
                a = g.HeCt(l, _rand_big(rr, n, W, bits), _rand_big(rr, n, W, bits), float(Delta), 1.0)
                b = g.HeCt(l, _rand_big(rr, n, W, bits), _rand_big(rr, n, W, bits), float(Delta), 1.0)
                pt = g.HePt(_rand_big(rr, n, W, Delta.bit_length()), nu=float(Delta))
                out = []
                m = ctx.he_mul_rs(a, b, k1); out += [m.c0, m.c1]
                m = ctx.he_mulpt(a, pt); out += [m.c0, m.c1]
                m = ctx.he_rot(a, 3, {3: k2}); out += [m.c0, m.c1]
                if not short:            # short: one op per loader mode of the two-kernel transforms (CPU tier, n = 2^13)
                    m = ctx.he_mul(a, b, k1); out += [m.c0, m.c1]
                    m = ctx.he_conj(b, k2); out += [m.c0, m.c1]
                    for m in ctx.he_rot_hoisted(a, [0, 1, 5], {0: k1, 1: k2, 5: k1}):
                        out += [m.c0, m.c1]
                    out.append(ctx.poly_mul(a.c0, b.c1, ctx.dim, ctx.ql[l]))
                else:
                    for m in ctx.he_rot_hoisted(a, [1], {1: k2}):
                        out += [m.c0, m.c1]
                res.append(out)
            finally:
                ctx.close()
        finally:
            os.environ.pop("GPQHE_B200_NO_FUSED_INV", None)
            os.environ.pop("GPQHE_B200_NO_PRESCALE", None)
            os.environ.pop("GPQHE_B200_NO_TC_CRT", None)
    assert len(res[0]) == len(res[1]) == len(res[2])
    for k, (x, y, z) in enumerate(zip(res[0], res[1], res[2])):
        eq(x, y, f"fused vs separate pipeline, output {k}")
        eq(x, z, f"tensor-core vs integer-pipe CRT, output {k}")
    del nr
