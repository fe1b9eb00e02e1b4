"""Differential check of the oracle restatement against the live reference build
(oracle/_ref); run in a subprocess because the reference holds one global context."""
import sys, random, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import pyoracle as o, refbind as rb
import ctypes as C
R = rb.Ref()
logn, logq, logD = 8, 120, 40
R.hectx_init(logn, 1<<logq, 4, 1<<logD)
ctx = o.Ctx(logn, 1<<logq, 4, 1<<logD)
n = ctx.n
assert R.primes() == ctx.primes, (R.primes()[:3], ctx.primes[:3])
h = R.hectx
print("L dim dimevk", h.L, h.dim, h.dimevk, ctx.L, ctx.dim, ctx.dimevk, "dimub", R.polyctx.dimub, ctx.dimub)
assert (h.L,h.dim,h.dimevk)==(ctx.L,ctx.dim,ctx.dimevk)
assert R.to_int(h.P)==ctx.P and R.to_int(h.PqL)==ctx.PqL
for d in range(ctx.dimub):
    nd = R.node(d).contents; on = ctx.rns[d]
    assert (nd.p, nd.pinv_mont, nd.pinv_barr, nd.ninv)==(on.p,on.pinv_mont,on.pinv_barr,on.ninv)
    assert np.array_equal(np.ctypeslib.as_array(nd.zetas,(n,)), on.zetas)
    assert np.array_equal(np.ctypeslib.as_array(nd.zetas_inv,(n,)), on.zetas_inv)
    assert R.to_int(nd.P)==on.P and R.to_int(nd.P_2)==on.P_2
    assert [R.to_int(nd.phat[k]) for k in range(d+1)]==on.phat
    assert [nd.phat_invmp[k] for k in range(d+1)]==on.phat_invmp
print("tables OK")
rng = random.Random(1)
# ntt / invntt / mul / add
for d in (0, 3, ctx.dimub-1):
    on = ctx.rns[d]; nd = R.node(d)
    a = np.array([rng.randrange(on.p) for _ in range(n)], dtype=np.uint64)
    b = np.array([rng.randrange(on.p) for _ in range(n)], dtype=np.uint64)
    ra = a.copy(); R.lib.ntt(rb.ptr(ra), nd)
    assert np.array_equal(ra, o.ntt(a, on))
    ri = ra.copy(); R.lib.invntt(rb.ptr(ri), nd)
    assert np.array_equal(ri, o.invntt(ra, on)) and np.array_equal(ri, a)
    r = np.empty_like(a); R.lib.poly_rns_mul(rb.ptr(r), rb.ptr(a), rb.ptr(b), nd)
    assert np.array_equal(r, o.poly_rns_mul(a,b,on))
    R.lib.poly_rns_add(rb.ptr(r), rb.ptr(a), rb.ptr(b), nd)
    assert np.array_equal(r, o.poly_rns_add(a,b,on))
# hazard vector: a[N/2]=1, a[0]=p - psi^(N/2)
on = ctx.rns[0]; nd = R.node(0)
zplain = (int(on.zetas[1]) * pow(1<<64, -1, on.p)) % on.p
a = np.zeros(n, dtype=np.uint64); a[n//2]=1; a[0]=on.p - zplain
ra = a.copy(); R.lib.ntt(rb.ptr(ra), nd)
oa = o.ntt(a, on)
print("hazard count of p in output:", int((ra==on.p).sum()))
assert np.array_equal(ra, oa)
print("limb layer OK")
# he ops
L = ctx.L
def rand_ct(l):
    q = ctx.ql[l]
    return [rng.randrange(q)-q//2 for _ in range(n)], [rng.randrange(q)-q//2 for _ in range(n)]
def rand_evk():
    p0 = np.concatenate([np.array([rng.randrange(ctx.primes[d]) for _ in range(n)],dtype=np.uint64) for d in range(ctx.dimevk)])
    p1 = np.concatenate([np.array([rng.randrange(ctx.primes[d]) for _ in range(n)],dtype=np.uint64) for d in range(ctx.dimevk)])
    return p0,p1
evk = rand_evk(); revk = R.new_evk(*evk)
for l in (L, 1):
    a0,a1 = rand_ct(l); b0,b1 = rand_ct(l)
    ca = R.new_ct(l,a0,a1); cb = R.new_ct(l,b0,b1); cr = R.new_ct(l,a0,a1)
    R.lib.he_mul(C.byref(cr), C.byref(ca), C.byref(cb), C.byref(revk))
    ref = R.ct_dict(cr)
    mine = o.he_mul(ctx, {"l":l,"c0":a0,"c1":a1}, {"l":l,"c0":b0,"c1":b1}, evk)
    assert ref["c0"]==mine["c0"] and ref["c1"]==mine["c1"], "he_mul"
    R.lib.he_rs(C.byref(cr)); m2 = o.he_rs(ctx, mine); r2 = R.ct_dict(cr)
    assert r2["c0"]==m2["c0"] and r2["c1"]==m2["c1"] and r2["l"]==m2["l"], "he_rs"
    # rot / conj
    for rot in (0,1,3):
        cr2 = R.new_ct(l,a0,a1)
        evks = (HeEvkArr := (rb.HeEvk*4))()
        evks[rot] = revk
        R.lib.he_rot(C.byref(cr2), rot, evks)
        mm = o.he_rot(ctx, {"l":l,"c0":a0,"c1":a1}, rot, evk); rr = R.ct_dict(cr2)
        assert rr["c0"]==mm["c0"] and rr["c1"]==mm["c1"], "he_rot"
    cr3 = R.new_ct(l,a0,a1); R.lib.he_conj(C.byref(cr3), C.byref(revk))
    mm = o.he_conj(ctx, {"l":l,"c0":a0,"c1":a1}, evk); rr = R.ct_dict(cr3)
    assert rr["c0"]==mm["c0"] and rr["c1"]==mm["c1"], "he_conj"
    if l>0:
        cr4 = R.new_ct(l,a0,a1); R.lib.he_moddown(C.byref(cr4)); rr=R.ct_dict(cr4)
        mm = o.he_moddown(ctx, {"l":l,"c0":a0,"c1":a1}); assert rr["c0"]==mm["c0"] and rr["c1"]==mm["c1"]
print("HE ops OK")
