"""Generate tests/golden/*.npz from the UNMODIFIED reference build (oracle/_ref).

Run in the build container (needs /root/reference compiled by `make -C oracle ref`):
    python tests/golden/make_golden.py
Each fixture holds seeded inputs and the reference's own outputs for the hot path:
ntt / invntt / poly_rns_mul / poly_rns_add (incl. the `ntt()` non-canonical-q vector,
SURVEY 8a N1), rns_decompose, poly_rns2mpi, poly_mul, poly_rot, poly_conj, he_mul,
he_rs, he_moddown, he_rot, he_conj, he_mulpt.  Big integers are stored word-major
two's complement ([W, n] uint64), the interchange format of include/gpqhe_b200.h.
One process can hold one reference context, so each fixture is made in a subprocess.
"""
import ctypes as C
import os
import random
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CASES = {
    # name: (logn, q, slots, Delta, seed)
    "pow2_n256_q120": (8, 1 << 120, 4, 1 << 40, 1),
    "pow2_n64_q438": (6, 1 << 438, 4, 1 << 50, 2),
    "general_n128": (7, (1 << 119) + 12345, 4, (1 << 40) + 7, 3),
}


def words(a, W):
    mask = (1 << (64 * W)) - 1
    buf = b"".join((int(x) & mask).to_bytes(8 * W, "little") for x in a)
    return np.ascontiguousarray(np.frombuffer(buf, dtype="<u8").reshape(len(a), W).T).astype(np.uint64)


def make(name):
    from oracle import refbind as rb
    logn, q, slots, Delta, seed = CASES[name]
    R = rb.Ref()
    R.hectx_init(logn, q, slots, Delta)
    n, h = R.n, R.hectx
    rng = random.Random(seed)
    W = (q.bit_length() + 63) // 64
    L, dim, dimevk, dimub = h.L, h.dim, h.dimevk, R.polyctx.dimub
    primes = R.primes()
    out = dict(logn=logn, q=words([q], W + 1)[:, 0], slots=slots, Delta=np.uint64(Delta), W=W, L=L,
               dim=dim, dimevk=dimevk, dimub=dimub, primes=np.array(primes, dtype=np.uint64))
    consts = []
    for d in range(dimub):
        nd = R.node(d).contents
        consts.append([nd.p, nd.pinv_mont, nd.pinv_barr, nd.ninv, nd.zetas[1], nd.zetas_inv[1]])
    out["prime_consts"] = np.array(consts, dtype=np.uint64)
    ql = [R.to_int(h.q[l]) for l in range(L + 1)]

    def rlimb(d):
        return np.array([rng.randrange(primes[d]) for _ in range(n)], dtype=np.uint64)

    # limb layer on limbs 0 and dimub-1
    for tag, d in (("a", 0), ("b", dimub - 1)):
        nd = R.node(d)
        x, y = rlimb(d), rlimb(d)
        f = x.copy(); R.lib.ntt(rb.ptr(f), nd)
        iv = y.copy(); R.lib.invntt(rb.ptr(iv), nd)
        m = np.empty_like(x); R.lib.poly_rns_mul(rb.ptr(m), rb.ptr(x), rb.ptr(y), nd)
        s = np.empty_like(x); R.lib.poly_rns_add(rb.ptr(s), rb.ptr(x), rb.ptr(y), nd)
        p = primes[d]
        zplain = (int(nd.contents.zetas[1]) * pow(1 << 64, -1, p)) % p
        hz = np.zeros(n, dtype=np.uint64); hz[n // 2] = 1; hz[0] = p - zplain
        hf = hz.copy(); R.lib.ntt(rb.ptr(hf), nd)
        hi = hf.copy(); R.lib.invntt(rb.ptr(hi), nd)
        out.update({f"limb_{tag}": d, f"x_{tag}": x, f"y_{tag}": y, f"ntt_{tag}": f, f"invntt_{tag}": iv,
                    f"mul_{tag}": m, f"add_{tag}": s, f"hz_{tag}": hz, f"hzntt_{tag}": hf, f"hzinv_{tag}": hi})

    def rpoly(qq):
        return [rng.randrange(qq) - qq // 2 for _ in range(n)]

    # rns_decompose / poly_rns2mpi / poly_mul / rot / conj
    a, b = rpoly(q), rpoly(q)
    pa, pb = R.new_poly(a), R.new_poly(b)
    dec = np.empty((dimub, n), dtype=np.uint64)
    for d in range(dimub):
        R.lib.rns_decompose(rb.ptr(dec[d]), pa.coeffs, R.node(d))
    out.update(big_a=words(a, W), big_b=words(b, W), decompose_a=dec)
    rh = np.concatenate([rlimb(d) for d in range(dimub)])
    prh = rb.PolyRns(); prh.coeffs = rb.ptr(rh)
    pr = R.new_poly()
    q2 = q // 3 + 1
    for tag, dd, qq in (("full", dimub, q), ("dim", dim, q2)):
        R.lib.poly_rns2mpi(C.byref(pr), C.byref(prh), R.node(dd - 1), R.mpi(qq))
        out[f"rns2mpi_{tag}"] = words(R.poly_ints(pr), W)
    out["rns2mpi_in"] = rh
    out["rns2mpi_q2"] = words([q2], W + 1)[:, 0]
    R.lib.poly_mul(C.byref(pr), C.byref(pa), C.byref(pb), dimub, R.mpi(q))
    out["poly_mul"] = words(R.poly_ints(pr), W)
    for rot in (1, 5):
        R.lib.poly_rot(C.byref(pr), C.byref(pa), rot)
        out[f"poly_rot{rot}"] = words(R.poly_ints(pr), W)
    R.lib.poly_conj(C.byref(pr), C.byref(pa))
    out["poly_conj"] = words(R.poly_ints(pr), W)

    # HE ops at the top level and at level 1, random evk
    p0 = np.concatenate([rlimb(d) for d in range(dimevk)])
    p1 = np.concatenate([rlimb(d) for d in range(dimevk)])
    evk = R.new_evk(p0, p1)
    out.update(evk_p0=p0, evk_p1=p1)
    for l in sorted({L, 1}):
        a0, a1, b0, b1 = rpoly(ql[l]), rpoly(ql[l]), rpoly(ql[l]), rpoly(ql[l])
        out.update({f"l{l}_a0": words(a0, W), f"l{l}_a1": words(a1, W), f"l{l}_b0": words(b0, W),
                    f"l{l}_b1": words(b1, W)})
        ca, cb, cr = R.new_ct(l, a0, a1), R.new_ct(l, b0, b1), R.new_ct(l, a0, a1)
        R.lib.he_mul(C.byref(cr), C.byref(ca), C.byref(cb), C.byref(evk))
        d = R.ct_dict(cr)
        out.update({f"l{l}_mul_c0": words(d["c0"], W), f"l{l}_mul_c1": words(d["c1"], W)})
        R.lib.he_rs(C.byref(cr)); d = R.ct_dict(cr)
        out.update({f"l{l}_mulrs_c0": words(d["c0"], W), f"l{l}_mulrs_c1": words(d["c1"], W)})
        cm = R.new_ct(l, a0, a1); R.lib.he_moddown(C.byref(cm)); d = R.ct_dict(cm)
        out.update({f"l{l}_moddown_c0": words(d["c0"], W), f"l{l}_moddown_c1": words(d["c1"], W)})
        for rot in (0, 3):
            cx = R.new_ct(l, a0, a1)
            arr = (rb.HeEvk * 4)(); arr[rot] = evk
            R.lib.he_rot(C.byref(cx), rot, arr); d = R.ct_dict(cx)
            out.update({f"l{l}_rot{rot}_c0": words(d["c0"], W), f"l{l}_rot{rot}_c1": words(d["c1"], W)})
        cx = R.new_ct(l, a0, a1); R.lib.he_conj(C.byref(cx), C.byref(evk)); d = R.ct_dict(cx)
        out.update({f"l{l}_conj_c0": words(d["c0"], W), f"l{l}_conj_c1": words(d["c1"], W)})
        m = [rng.randrange(1 << 40) - (1 << 39) for _ in range(n)]
        pt = rb.HePt(); R.lib.he_alloc_pt(C.byref(pt)); pt.nu = float(1 << 40)
        for i, x in enumerate(m):
            R.set_mpi(pt.m.coeffs[i], x)
        cx = R.new_ct(l, a0, a1); cd = R.new_ct(l, a0, a1)
        R.lib.he_mulpt(C.byref(cd), C.byref(cx), C.byref(pt)); d = R.ct_dict(cd)
        out.update({f"l{l}_pt_m": words(m, W), f"l{l}_mulpt_c0": words(d["c0"], W),
                    f"l{l}_mulpt_c1": words(d["c1"], W)})
    out["levels"] = np.array(sorted({L, 1}), dtype=np.int64)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **out)
    print("wrote", name)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        make(sys.argv[1])
    else:
        for k in CASES:
            subprocess.check_call([sys.executable, os.path.abspath(__file__), k])
