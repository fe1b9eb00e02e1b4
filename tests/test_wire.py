"""Wire format (gpqhe_b200/wire.py): round trips, integrity checks, context binding."""
import random

import numpy as np
import pytest

import gpqhe_b200 as g
from gpqhe_b200 import wire
from oracle import pyoracle as o
from parity_cases import rlimbs, rpoly


def _roundtrip(lib_path, logn, q, Delta):
    rng = random.Random(logn)
    ctx = g.HeContext(logn, q, 4, Delta, lib_path=lib_path, logqub=q.bit_length() - 1)
    oc = o.Ctx(logn, q, 4, Delta, logqub=q.bit_length() - 1)
    try:
        n, W, l = ctx.n, ctx.W, ctx.L
        a = [rpoly(rng, oc.ql[l], n) for _ in range(4)]
        ct1 = g.HeCt(l, g.ints_to_big(a[0], W), g.ints_to_big(a[1], W), float(Delta), 3.5)
        ct2 = g.HeCt(l, g.ints_to_big(a[2], W), g.ints_to_big(a[3], W), float(Delta), 1.25)
        key = (rlimbs(rng, oc.primes, n, oc.dimevk), rlimbs(rng, oc.primes, n, oc.dimevk))
        rlk = ctx.genevk(*key)
        # ciphertext / plaintext round trips keep every bit and the metadata
        blob = wire.dump_ct(ctx, ct1)
        assert len(blob) == wire.HEADER_BYTES + 2 * W * n * 8
        back = wire.load_ct(ctx, blob)
        assert (back.l, back.nu, back.B) == (ct1.l, ct1.nu, ct1.B)
        assert np.array_equal(back.c0, ct1.c0) and np.array_equal(back.c1, ct1.c1)
        pt = g.HePt(ct2.c0, 7.0)
        pb = wire.load_pt(ctx, wire.dump_pt(ctx, pt))
        assert pb.nu == 7.0 and np.array_equal(pb.m, pt.m)
        # the key comes back from the device as it went in, and a reloaded key computes the same product
        p0, p1 = wire.export_evk(ctx, rlk)
        assert np.array_equal(p0.reshape(-1), key[0]) and np.array_equal(p1.reshape(-1), key[1])
        rlk2 = wire.load_evk(ctx, wire.dump_evk(ctx, rlk))
        r1 = ctx.he_mul(ct1, ct2, rlk)
        r2 = ctx.he_mul(wire.load_ct(ctx, blob), wire.load_ct(ctx, wire.dump_ct(ctx, ct2)), rlk2)
        assert np.array_equal(r1.c0, r2.c0) and np.array_equal(r1.c1, r2.c1)
        want = o.he_mul(oc, {"l": l, "c0": a[0], "c1": a[1]}, {"l": l, "c0": a[2], "c1": a[3]}, key)
        assert g.big_to_ints(r2.c0) == want["c0"] and g.big_to_ints(r2.c1) == want["c1"]
        # integrity: flipped payload bit, flipped header bit, truncation, wrong kind
        bad = bytearray(blob); bad[wire.HEADER_BYTES + 5] ^= 1
        with pytest.raises(g.GpqheError, match="payload checksum"):
            wire.load_ct(ctx, bad)
        bad = bytearray(blob); bad[24] ^= 1
        with pytest.raises(g.GpqheError, match="header checksum"):
            wire.load_ct(ctx, bad)
        with pytest.raises(g.GpqheError, match="truncated"):
            wire.load_ct(ctx, blob[:-8])
        with pytest.raises(g.GpqheError, match="kind"):
            wire.load_pt(ctx, blob)
        # a residue above its prime is refused before it reaches the device
        k0 = key[0].copy(); k0[3] = np.uint64(int(oc.primes[0]) + 1)
        with pytest.raises(g.GpqheError, match="out of range"):
            wire.load_evk(ctx, wire.dump_evk(ctx, (k0, key[1])))
        return blob
    finally:
        ctx.close()


def test_wire_roundtrip_emulated(emul_lib):
    blob = _roundtrip(emul_lib, 5, 1 << 120, 1 << 40)
    # another modulus chain on the same ring: refused
    other = g.HeContext(5, 1 << 160, 4, 1 << 40, lib_path=emul_lib, logqub=160)
    try:
        with pytest.raises(g.GpqheError, match="another context"):
            wire.load_ct(other, blob)
    finally:
        other.close()


@pytest.mark.gpu
def test_wire_roundtrip_gpu(real_lib):
    _roundtrip(real_lib, 12, 1 << 438, 1 << 50)
