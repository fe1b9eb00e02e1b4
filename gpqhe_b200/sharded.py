"""Limb-sharded he_mul / he_rs over the GPUs of one node.

One process per GPU (torch.distributed, NCCL over NVLink; gloo on CPU for the tests).  The
path has two domains:

  * limb domain   (NTT, pointwise): independent per RNS limb.  Rank r owns the limbs
                  [r*D/G, (r+1)*D/G) of every polynomial, each over the whole ring degree n.
  * coefficient domain (MPI<->RNS, key-switch epilogue, rescale): independent per coefficient.
                  Rank r owns the coefficients [r*n/G, (r+1)*n/G) of every polynomial.

Ciphertexts live COEFFICIENT-SHARDED between operations: c0, c1 are int64 tensors [W, n/G]
(word-major two's complement, the slice of the big poly of include/gpqhe_b200.h).  The only
data-path collective is the all-to-all that re-lays residues out between the two domains; one
he_mul (src/he-mult.c:88-156) needs four of them:

  decompose(ct1,ct2) --A2A--> NTT, tensor, iNTT --A2A--> CRT(d0,d1,d2), decompose(d2)
                     --A2A--> NTT, x rlk, iNTT  --A2A--> key-switch epilogue (+d0,+d1)

he_rs / he_moddown (src/he-rescale.c:33-70) are coefficient-local: no communication.  Every
stage is the same CUDA kernel the single-GPU path runs, called through the gq_stage_* entry
points of the C ABI on this rank's limb range or coefficient slice, so the result is bit-identical
to the single-GPU result (tests/test_sharded.py).
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import capi
from .capi import GpqheError
from .he import HeContext, HeEvk


class ShardedHe:
    """he_mul / he_rs / he_moddown on coefficient-sharded ciphertexts across `group`."""

    def __init__(self, ctx: HeContext, group=None, device: torch.device | str | None = None):
        self.ctx, self.group = ctx, group
        self._pre_cache = {}
        self.G = dist.get_world_size(group) if dist.is_initialized() else 1
        # above two ranks the per-rank kernels are a few microseconds long: programmatic dependent launch measured slower
        # than plain stream / graph edges there (profiles/r02_scaling.md), so the sharded drivers switch it off
        if self.G > 2:
            ctx._lib.gq_set_pdl(0)
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n, self.W = ctx.n, ctx.W
        if self.n % self.G:
            raise GpqheError("ring degree must be divisible by the number of ranks")
        self.n_loc = self.n // self.G
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self._lib, self._h = ctx._lib, ctx._h
        self.exchanges = 0          # all-to-all calls issued (diagnostics)
        self.exchanged_bytes = 0    # bytes this rank sent through them

    # ------------------------------------------------------------------ helpers
    def _ck(self, rc):
        capi.check(self._lib, rc)

    def _empty(self, *shape):
        return torch.empty(shape, dtype=torch.int64, device=self.device)

    def _limbs(self, D):
        """(first limb, limb count) owned by this rank; ragged when G does not divide D"""
        off = self._offsets(D)
        return off[self.rank], off[self.rank + 1] - off[self.rank]

    def _offsets(self, D):
        base, extra = divmod(D, self.G)
        off = [0]
        for r in range(self.G):
            off.append(off[-1] + base + (1 if r < extra else 0))
        return off

    def shard(self, full: torch.Tensor) -> torch.Tensor:
        """[W, n] -> this rank's coefficient slice [W, n/G]."""
        return full[:, self.rank * self.n_loc:(self.rank + 1) * self.n_loc].contiguous()

    def gather(self, sl: torch.Tensor) -> torch.Tensor:
        """coefficient slices [W, n/G] of all ranks -> [W, n] (for checking / hand-over)."""
        if self.G == 1:
            return sl.clone()
        parts = [torch.empty_like(sl) for _ in range(self.G)]
        dist.all_gather(parts, sl.contiguous(), group=self.group)
        return torch.cat(parts, dim=1)

    def _a2a(self, send: torch.Tensor, in_splits, out_splits) -> torch.Tensor:
        """flat all-to-all: send[in_splits[s]] goes to rank s, recv[out_splits[s]] comes from rank s"""
        recv = torch.empty(sum(out_splits), dtype=send.dtype, device=send.device)
        dist.all_to_all_single(recv, send, output_split_sizes=out_splits, input_split_sizes=in_splits,
                               group=self.group)
        self.exchanges += 1
        self.exchanged_bytes += (send.numel() - in_splits[self.rank]) * 8
        return recv

    def _coef_to_limb(self, x: torch.Tensor, D: int) -> torch.Tensor:
        """x[P, D, n_loc] (all limbs of my coefficients) -> [P, Dl, n] (my limbs, all coefficients)."""
        if self.G == 1:
            return x
        P, G, off = x.shape[0], self.G, self._offsets(D)
        Dl = off[self.rank + 1] - off[self.rank]
        send = torch.cat([x[:, off[s]:off[s + 1], :].reshape(-1) for s in range(G)])
        recv = self._a2a(send, [P * (off[s + 1] - off[s]) * self.n_loc for s in range(G)],
                         [P * Dl * self.n_loc] * G)
        return recv.view(G, P, Dl, self.n_loc).permute(1, 2, 0, 3).contiguous().view(P, Dl, self.n)

    def _limb_to_coef(self, y: torch.Tensor, D: int) -> torch.Tensor:
        """y[P, Dl, n] -> [P, D, n_loc]."""
        if self.G == 1:
            return y
        P, G, off = y.shape[0], self.G, self._offsets(D)
        Dl = off[self.rank + 1] - off[self.rank]
        send = y.view(P, Dl, G, self.n_loc).permute(2, 0, 1, 3).contiguous().view(-1)
        sizes = [P * (off[s + 1] - off[s]) * self.n_loc for s in range(G)]
        recv = self._a2a(send, [P * Dl * self.n_loc] * G, sizes)
        out = torch.empty((P, D, self.n_loc), dtype=y.dtype, device=y.device)
        pos = 0
        for s in range(G):
            out[:, off[s]:off[s + 1], :] = recv[pos:pos + sizes[s]].view(P, off[s + 1] - off[s], self.n_loc)
            pos += sizes[s]
        return out

    def _decompose(self, polys, D):
        out = self._empty(len(polys), D, self.n_loc)
        arr = (C.c_void_p * len(polys))(*[p.data_ptr() for p in polys])
        self._ck(self._lib.gq_stage_decompose_dev(self._h, 0, D, len(polys), arr, self.W, self.n_loc,
                                                  out.data_ptr(), None))
        return out


    # -- stage calls: the pre-scaled forms (inverse NTT ending in the CRT pre-multiplication, CRT stages on the tensor
    #    cores) when the library has them for the basis, else the plain ones.  GPQHE_B200_SHARDED_PLAIN=1 forces plain.
    def _pre(self, D, l, ks):
        key = (D, l, ks)
        if key not in self._pre_cache:
            ok = not os.environ.get("GPQHE_B200_SHARDED_PLAIN") and \
                bool(self._lib.gq_stage_prescale_ok(self._h, D, l, 1 if ks else 0))
            self._pre_cache[key] = ok
        return self._pre_cache[key]

    def _tensor_inv(self, lo, Dl, D, l, y, x):
        if self._pre(D, l, False):
            self._ck(self._lib.gq_stage_tensor_inv_pre_dev(self._h, lo, Dl, D, y.data_ptr(), x.data_ptr(), None))
        else:
            self._ck(self._lib.gq_stage_tensor_inv_dev(self._h, lo, Dl, y.data_ptr(), x.data_ptr(), None))

    def _mul_evk_inv(self, lo, Dl, D, l, y, x, evk):
        if self._pre(D, l, True):
            self._ck(self._lib.gq_stage_mul_evk_inv_pre_dev(self._h, lo, Dl, D, y.data_ptr(), x.data_ptr(), evk._h, None))
        else:
            self._ck(self._lib.gq_stage_mul_evk_inv_dev(self._h, lo, Dl, y.data_ptr(), x.data_ptr(), evk._h, None))

    def _reconstruct(self, D, l, rh_ptr, outs, npolys=3):
        f = self._lib.gq_stage_reconstruct_pre_dev if self._pre(D, l, False) else self._lib.gq_stage_reconstruct_dev
        self._ck(f(self._h, D, l, npolys, rh_ptr, outs, self.n_loc, None))

    def _ks_epilogue(self, D2, l, ch_ptr, add0, add1, c0, c1, rescale):
        a0 = add0.data_ptr() if add0 is not None else None
        a1 = add1.data_ptr() if add1 is not None else None
        if self._pre(D2, l, True):
            self._ck(self._lib.gq_stage_ks_epilogue_pre_dev(self._h, l, ch_ptr, a0, a1, c0.data_ptr(), c1.data_ptr(),
                                                            self.n_loc, 1 if rescale else 0, None))
        else:
            f = self._lib.gq_stage_ks_epilogue_rs_dev if rescale else self._lib.gq_stage_ks_epilogue_dev
            self._ck(f(self._h, l, ch_ptr, a0, a1, c0.data_ptr(), c1.data_ptr(), self.n_loc, None))

    # ------------------------------------------------------------------ operations
    def he_mul(self, l: int, a0, a1, b0, b1, rlk: HeEvk):
        """he_mul + relinearise on coefficient slices; returns (c0, c1) slices at level l."""
        lib, h = self._lib, self._h
        D1, D2 = self.ctx.dims(l)
        lo1, D1l = self._limbs(D1)
        lo2, D2l = self._limbs(D2)
        # hot loop A (he-mult.c:116-138)
        x = self._coef_to_limb(self._decompose([a0, a1, b0, b1], D1), D1)          # [4, D1l, n]
        self._ck(lib.gq_stage_ntt_dev(h, lo1, D1l, 4, x.data_ptr(), 0, None))
        y = self._empty(3, D1l, self.n)
        self._tensor_inv(lo1, D1l, D1, l, y, x)                                   # tensor + inverse NTT
        # hot loop B (he-mult.c:139-141)
        rh = self._limb_to_coef(y, D1)                                            # [3, D1, n_loc]
        d = self._empty(3, self.W, self.n_loc)
        outs = (C.c_void_p * 3)(*[d[k].data_ptr() for k in range(3)])
        self._reconstruct(D1, l, rh.data_ptr(), outs)
        # he_relin (he-mult.c:40-85): key switch of d2 = d[2]
        c0, c1 = self.keyswitch(l, d[2], rlk, d[0], d[1], lo2, D2l, D2)
        return c0, c1

    def keyswitch(self, l, dpoly, evk: HeEvk, add0, add1, lo2=None, D2l=None, D2=None):
        lib, h = self._lib, self._h
        if D2 is None:
            D2 = self.ctx.dims(l)[1]
            lo2, D2l = self._limbs(D2)
        x = self._coef_to_limb(self._decompose([dpoly], D2), D2)                  # [1, D2l, n]
        self._ck(lib.gq_stage_ntt_dev(h, lo2, D2l, 1, x.data_ptr(), 0, None))
        y = self._empty(2, D2l, self.n)
        self._mul_evk_inv(lo2, D2l, D2, l, y, x, evk)                             # x evk + inverse NTT
        ch = self._limb_to_coef(y, D2)                                            # [2, D2, n_loc]
        c0, c1 = self._empty(self.W, self.n_loc), self._empty(self.W, self.n_loc)
        self._ks_epilogue(D2, l, ch.data_ptr(), add0, add1, c0, c1, False)
        return c0, c1

    def he_rs(self, l: int, c0, c1):
        """he_rs in place on the slices; the ciphertext is then at level l-1."""
        self._ck(self._lib.gq_stage_rescale_dev(self._h, l, c0.data_ptr(), c1.data_ptr(), 1, self.n_loc, None))
        return c0, c1

    def he_moddown(self, l: int, c0, c1):
        self._ck(self._lib.gq_stage_rescale_dev(self._h, l, c0.data_ptr(), c1.data_ptr(), 0, self.n_loc, None))
        return c0, c1


class MulPlan:
    """he_mul (+ he_rs) at one level with every buffer allocated once and the work between two
    exchanges recorded as ONE CUDA graph (five graphs, four eager all-to-alls per op).

    The eager path of ShardedHe issues ~15 small launches per exchange from Python and is
    launch-bound at N=2^16; replaying five graphs makes the op GPU- and link-bound instead.  The
    collectives stay outside the graphs on purpose (capturing NCCL work needs cross-rank care and
    buys nothing here: they are four launches)."""

    def __init__(self, sh: ShardedHe, l: int, rlk: HeEvk, rescale: bool = True, graphs: bool | None = None):
        self.sh, self.l, self.rlk, self.rescale = sh, l, rlk, rescale
        G, n, nl, W = sh.G, sh.n, sh.n_loc, sh.W
        self.D1, self.D2 = sh.ctx.dims(l)
        self.off1, self.off2 = sh._offsets(self.D1), sh._offsets(self.D2)
        r = sh.rank
        self.lo1, self.D1l = self.off1[r], self.off1[r + 1] - self.off1[r]
        self.lo2, self.D2l = self.off2[r], self.off2[r + 1] - self.off2[r]
        E = sh._empty
        self.inp = E(4, W, nl)                         # a0, a1, b0, b1 slices (copy inputs here)
        self.xc, self.xl = E(4, self.D1, nl), E(4, self.D1l, n)
        self.yl, self.rh = E(3, self.D1l, n), E(3, self.D1, nl)
        self.d = E(3, W, nl)
        self.x2c, self.x2l = E(1, self.D2, nl), E(1, self.D2l, n)
        self.y2l, self.ch = E(2, self.D2l, n), E(2, self.D2, nl)
        self.out = E(2, W, nl)                         # c0, c1 slices
        # flat exchange buffers: (send, recv, in_splits, out_splits) per exchange
        self.ex = [self._xbuf(4, self.off1, True), self._xbuf(3, self.off1, False),
                   self._xbuf(1, self.off2, True), self._xbuf(2, self.off2, False)]
        self.segs = [self._seg0, self._seg1, self._seg2, self._seg3, self._seg4]
        self.graphs = None
        want = (sh.device.type == "cuda") if graphs is None else graphs
        if want:
            self._capture()

    # -- exchange buffers ------------------------------------------------------------
    def _xbuf(self, P, off, coef_to_limb):
        sh = self.sh
        G, nl, r = sh.G, sh.n_loc, sh.rank
        Dl = off[r + 1] - off[r]
        ragged = [P * (off[s + 1] - off[s]) * nl for s in range(G)]
        even = [P * Dl * nl] * G
        ins, outs = (ragged, even) if coef_to_limb else (even, ragged)
        send = torch.empty(sum(ins), dtype=torch.int64, device=sh.device)
        recv = torch.empty(sum(outs), dtype=torch.int64, device=sh.device)
        return send, recv, ins, outs

    def _pack_c2l(self, k, x, off):           # x[P, D, n_loc] -> send blocks by destination limb range
        send, _, ins, _ = self.ex[k]
        P, pos = x.shape[0], 0
        for s in range(self.sh.G):
            send[pos:pos + ins[s]].view(P, off[s + 1] - off[s], self.sh.n_loc).copy_(x[:, off[s]:off[s + 1], :])
            pos += ins[s]

    def _unpack_c2l(self, k, xl):             # recv[G][P][Dl][n_loc] -> xl[P, Dl, n]
        _, recv, _, _ = self.ex[k]
        P, Dl = xl.shape[0], xl.shape[1]
        xl.view(P, Dl, self.sh.G, self.sh.n_loc).copy_(recv.view(self.sh.G, P, Dl, self.sh.n_loc).permute(1, 2, 0, 3))

    def _pack_l2c(self, k, yl):               # yl[P, Dl, n] -> send[G][P][Dl][n_loc]
        send, _, _, _ = self.ex[k]
        P, Dl = yl.shape[0], yl.shape[1]
        send.view(self.sh.G, P, Dl, self.sh.n_loc).copy_(yl.view(P, Dl, self.sh.G, self.sh.n_loc).permute(2, 0, 1, 3))

    def _unpack_l2c(self, k, out, off):       # recv blocks by source limb range -> out[P, D, n_loc]
        _, recv, _, outs = self.ex[k]
        P, pos = out.shape[0], 0
        for s in range(self.sh.G):
            out[:, off[s]:off[s + 1], :].copy_(recv[pos:pos + outs[s]].view(P, off[s + 1] - off[s], self.sh.n_loc))
            pos += outs[s]

    def _exchange(self, k):
        sh = self.sh
        send, recv, ins, outs = self.ex[k]
        if sh.G == 1:
            recv.copy_(send)
            return
        dist.all_to_all_single(recv, send, output_split_sizes=outs, input_split_sizes=ins, group=sh.group)
        sh.exchanges += 1
        sh.exchanged_bytes += (send.numel() - ins[sh.rank]) * 8

    # -- the five compute segments ----------------------------------------------------
    def _seg0(self):
        sh = self.sh
        arr = (C.c_void_p * 4)(*[self.inp[k].data_ptr() for k in range(4)])
        sh._ck(sh._lib.gq_stage_decompose_dev(sh._h, 0, self.D1, 4, arr, sh.W, sh.n_loc, self.xc.data_ptr(), None))
        self._pack_c2l(0, self.xc, self.off1)

    def _seg1(self):
        sh, lib, h = self.sh, self.sh._lib, self.sh._h
        self._unpack_c2l(0, self.xl)
        sh._ck(lib.gq_stage_ntt_dev(h, self.lo1, self.D1l, 4, self.xl.data_ptr(), 0, None))
        sh._tensor_inv(self.lo1, self.D1l, self.D1, self.l, self.yl, self.xl)
        self._pack_l2c(1, self.yl)

    def _seg2(self):
        sh, lib, h = self.sh, self.sh._lib, self.sh._h
        self._unpack_l2c(1, self.rh, self.off1)
        outs = (C.c_void_p * 3)(*[self.d[k].data_ptr() for k in range(3)])
        sh._reconstruct(self.D1, self.l, self.rh.data_ptr(), outs)
        arr = (C.c_void_p * 1)(self.d[2].data_ptr())
        sh._ck(lib.gq_stage_decompose_dev(h, 0, self.D2, 1, arr, sh.W, sh.n_loc, self.x2c.data_ptr(), None))
        self._pack_c2l(2, self.x2c, self.off2)

    def _seg3(self):
        sh, lib, h = self.sh, self.sh._lib, self.sh._h
        self._unpack_c2l(2, self.x2l)
        sh._ck(lib.gq_stage_ntt_dev(h, self.lo2, self.D2l, 1, self.x2l.data_ptr(), 0, None))
        sh._mul_evk_inv(self.lo2, self.D2l, self.D2, self.l, self.y2l, self.x2l, self.rlk)
        self._pack_l2c(3, self.y2l)

    def _seg4(self):
        sh, lib, h = self.sh, self.sh._lib, self.sh._h
        self._unpack_l2c(3, self.ch, self.off2)
        sh._ks_epilogue(self.D2, self.l, self.ch.data_ptr(), self.d[0], self.d[1], self.out[0], self.out[1],
                        self.rescale)                # he_rs rides in the epilogue

    def _capture(self):
        """record each segment once; must be called with the context's stream current"""
        self.run()                                   # warm-up: allocations, lazy inits
        torch.cuda.current_stream().synchronize()
        stream = torch.cuda.current_stream()
        graphs = []
        for seg in self.segs:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream, capture_error_mode="thread_local"):
                seg()
            graphs.append(g)
        self.graphs = graphs

    def run(self):
        """one he_mul (+ he_rs) on self.inp -> self.out"""
        for k in range(5):
            if self.graphs is not None:
                self.graphs[k].replay()
            else:
                self.segs[k]()
            if k < 4:
                self._exchange(k)
        return self.out[0], self.out[1]


class P2PPlan:
    """he_mul (+ he_rs) sharded over G ranks with the re-layouts written straight into the peers'
    memory (gq_p2p.cu): per exchange one scatter kernel + an epoch signal on the producer and a flag
    wait on the consumer; no NCCL, no pack/unpack copies.  The whole op on one rank is ONE CUDA graph
    (the epoch is read from device memory).

    `slabs[s]` is rank s's receive slab as mapped in this process.  build() wires real ranks through
    CUDA IPC handles exchanged over torch.distributed; loopback() builds G virtual ranks inside one
    process (single GPU or the emulated library), run stage by stage in lockstep - the CPU tier uses
    that to check every index of the scatter kernels."""

    SLOTS = 8

    def __init__(self, ctx: HeContext, G: int, rank: int, l: int, rlk: HeEvk, slabs, rescale=True,
                 device="cpu"):
        self.ctx, self.G, self.rank, self.l, self.rlk, self.rescale = ctx, G, rank, l, rlk, rescale
        self._lib, self._h = ctx._lib, ctx._h
        self.n, self.W, self.nl = ctx.n, ctx.W, ctx.n // G
        self.device = torch.device(device)
        self.D1, self.D2 = ctx.dims(l)
        self.off1, self.off2 = self._offsets(self.D1), self._offsets(self.D2)
        self.lo1, self.D1l = self.off1[rank], self.off1[rank + 1] - self.off1[rank]
        self.lo2, self.D2l = self.off2[rank], self.off2[rank + 1] - self.off2[rank]
        lay = self.layout(ctx, G, l)
        self.lay = lay
        self.slab = slabs[rank]
        self._slabs, self._owned, self._mapped = list(slabs), None, False
        E = lambda *shape: torch.empty(shape, dtype=torch.int64, device=self.device)
        self.inp, self.out = E(4, self.W, self.nl), E(2, self.W, self.nl)
        self.xc, self.yl = E(4, self.D1, self.nl), E(3, max(self.D1l, 1), self.n)
        self.d, self.x2c = E(3, self.W, self.nl), E(1, self.D2, self.nl)
        self.y2l = E(2, max(self.D2l, 1), self.n)
        self.epoch = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
        ptrs = lambda key: (C.c_void_p * G)(*[slabs[s] + 8 * lay[key] for s in range(G)])
        self.p_flags, self.p_xl1, self.p_rh = ptrs("flags"), ptrs("xl1"), ptrs("rh")
        self.p_xl2, self.p_ch = ptrs("xl2"), ptrs("ch")
        self.c_off1 = (C.c_int * (G + 1))(*self.off1)
        self.c_off2 = (C.c_int * (G + 1))(*self.off2)
        self.graph = None
        # fused compute + exchange (gq_p2p.h): the producing kernels store into the peers' slabs and signal from their
        # last block, the consuming kernels poll their flags in their prologue - no scatter / signal / wait launches.
        # Needs the two-kernel transforms (n > 2^12) and the fast CRT / decompose kernels; GPQHE_B200_P2P_UNFUSED=1
        # keeps the round-1 plan for A/B measurements.
        self.fused = bool(self._lib.gq_p2p_fused_ok(self._h, l)) and not os.environ.get("GPQHE_B200_P2P_UNFUSED")
        self.ticket = torch.zeros(8, dtype=torch.int32, device=self.device)
        vp = lambda arr: C.cast(arr, C.c_void_p)

        def xchg(slot, D, dst, off):
            x = capi.GqP2pXchg()
            x.G, x.me, x.slot, x.D = G, rank, slot, D
            x.dst, x.off, x.flags = vp(dst), vp(off), vp(self.p_flags)
            x.flags_local = self._loc("flags")
            x.epoch_dev, x.err_dev = self.epoch.data_ptr(), self.err.data_ptr()
            x.ticket_dev = self.ticket.data_ptr() + 4 * slot
            return x
        self.x = [xchg(0, self.D1, self.p_xl1, self.c_off1), xchg(1, self.D1, self.p_rh, self.c_off1),
                  xchg(2, self.D2, self.p_xl2, self.c_off2), xchg(3, self.D2, self.p_ch, self.c_off2)]
        # limb -> coefficient exchanges carry residues pre-scaled for the CRT (tensor-core CRT stages) when the
        # library has those forms for the two bases; every rank takes the same decision (same context parameters)
        if self.G > 2:                 # see ShardedHe.__init__
            self._lib.gq_set_pdl(0)
        plain = bool(os.environ.get("GPQHE_B200_SHARDED_PLAIN"))
        self.pre1 = not plain and bool(self._lib.gq_stage_prescale_ok(self._h, self.D1, self.l, 0))
        self.pre2 = not plain and bool(self._lib.gq_stage_prescale_ok(self._h, self.D2, self.l, 1))
        self.x[1].pre = 1 if self.pre1 else 0
        self.x[3].pre = 1 if self.pre2 else 0

    def _offsets(self, D):
        base, extra = divmod(D, self.G)
        off = [0]
        for r in range(self.G):
            off.append(off[-1] + base + (1 if r < extra else 0))
        return off

    @staticmethod
    def layout(ctx: HeContext, G: int, l: int):
        """word offsets of the sub-buffers inside a slab (identical on every rank)"""
        n, nl = ctx.n, ctx.n // G
        D1, D2 = ctx.dims(l)
        m1, m2 = -(-D1 // G), -(-D2 // G)
        lay, pos = {}, 0
        for key, words in (("flags", 256), ("xl1", 4 * m1 * n), ("rh", 3 * D1 * nl), ("xl2", m2 * n),
                           ("ch", 2 * D2 * nl)):
            lay[key] = pos
            pos += (words + 31) // 32 * 32
        lay["words"] = pos
        return lay

    def _ck(self, rc):
        capi.check(self._lib, rc)

    def _loc(self, key):
        return self.slab + 8 * self.lay[key]

    # -- exchange primitives ---------------------------------------------------------------
    def _c2l(self, src, P, D, dst, off, slot):
        lib, h = self._lib, self._h
        self._ck(lib.gq_p2p_scatter_c2l_dev(h, src.data_ptr(), P, D, self.nl, self.G, self.rank, dst, off, None))
        self._ck(lib.gq_p2p_signal_dev(h, self.p_flags, slot, self.G, self.rank, self.epoch.data_ptr(), self.err.data_ptr(), None))

    def _l2c(self, src, P, Dl, D, dst, off, slot):
        lib, h = self._lib, self._h
        self._ck(lib.gq_p2p_scatter_l2c_dev(h, src.data_ptr(), P, Dl, D, self.nl, self.G, self.rank, dst, off, None))
        self._ck(lib.gq_p2p_signal_dev(h, self.p_flags, slot, self.G, self.rank, self.epoch.data_ptr(), self.err.data_ptr(), None))

    def _wait(self, slot):
        self._ck(self._lib.gq_p2p_wait_dev(self._h, self._loc("flags"), slot, self.G, self.epoch.data_ptr(),
                                           self.err.data_ptr(), None))

    # -- the five segments (producer side ends with scatter + signal, consumer side starts with wait)
    def seg(self, k):
        if self.fused:
            return self._seg_fused(k)
        lib, h, W, nl = self._lib, self._h, self.W, self.nl
        if k == 0:
            self._ck(lib.gq_p2p_bump_dev(h, self.epoch.data_ptr(), None))
            arr = (C.c_void_p * 4)(*[self.inp[j].data_ptr() for j in range(4)])
            self._ck(lib.gq_stage_decompose_dev(h, 0, self.D1, 4, arr, W, nl, self.xc.data_ptr(), None))
            self._c2l(self.xc, 4, self.D1, self.p_xl1, self.c_off1, 0)
        elif k == 1:
            self._wait(0)
            if self.D1l:
                xl = self._loc("xl1")
                self._ck(lib.gq_stage_ntt_dev(h, self.lo1, self.D1l, 4, xl, 0, None))
                if self.pre1:
                    self._ck(lib.gq_stage_tensor_inv_pre_dev(h, self.lo1, self.D1l, self.D1, self.yl.data_ptr(), xl, None))
                else:
                    self._ck(lib.gq_stage_tensor_inv_dev(h, self.lo1, self.D1l, self.yl.data_ptr(), xl, None))
            self._l2c(self.yl, 3, self.D1l, self.D1, self.p_rh, self.c_off1, 1)
        elif k == 2:
            self._wait(1)
            outs = (C.c_void_p * 3)(*[self.d[j].data_ptr() for j in range(3)])
            rec = lib.gq_stage_reconstruct_pre_dev if self.pre1 else lib.gq_stage_reconstruct_dev
            self._ck(rec(h, self.D1, self.l, 3, self._loc("rh"), outs, nl, None))
            arr = (C.c_void_p * 1)(self.d[2].data_ptr())
            self._ck(lib.gq_stage_decompose_dev(h, 0, self.D2, 1, arr, W, nl, self.x2c.data_ptr(), None))
            self._c2l(self.x2c, 1, self.D2, self.p_xl2, self.c_off2, 2)
        elif k == 3:
            self._wait(2)
            if self.D2l:
                xl = self._loc("xl2")
                self._ck(lib.gq_stage_ntt_dev(h, self.lo2, self.D2l, 1, xl, 0, None))
                if self.pre2:
                    self._ck(lib.gq_stage_mul_evk_inv_pre_dev(h, self.lo2, self.D2l, self.D2, self.y2l.data_ptr(), xl, self.rlk._h, None))
                else:
                    self._ck(lib.gq_stage_mul_evk_inv_dev(h, self.lo2, self.D2l, self.y2l.data_ptr(), xl, self.rlk._h, None))
            self._l2c(self.y2l, 2, self.D2l, self.D2, self.p_ch, self.c_off2, 3)
        else:
            self._wait(3)
            if self.pre2:
                self._ck(lib.gq_stage_ks_epilogue_pre_dev(h, self.l, self._loc("ch"), self.d[0].data_ptr(), self.d[1].data_ptr(),
                                                          self.out[0].data_ptr(), self.out[1].data_ptr(), nl,
                                                          1 if self.rescale else 0, None))
            else:
                ks = lib.gq_stage_ks_epilogue_rs_dev if self.rescale else lib.gq_stage_ks_epilogue_dev
                self._ck(ks(h, self.l, self._loc("ch"), self.d[0].data_ptr(), self.d[1].data_ptr(),
                            self.out[0].data_ptr(), self.out[1].data_ptr(), nl, None))

    def _seg_fused(self, k):
        """the same five segments with the exchanges inside the kernels: 16 launches per op instead of 28"""
        lib, h, W, nl, x = self._lib, self._h, self.W, self.nl, self.x
        ref = C.byref
        if k == 0:
            self._ck(lib.gq_p2p_bump_dev(h, self.epoch.data_ptr(), None))
            arr = (C.c_void_p * 4)(*[self.inp[j].data_ptr() for j in range(4)])
            self._ck(lib.gq_p2p_decompose_dev(h, self.D1, 4, arr, W, nl, ref(x[0]), None))
        elif k == 1:
            if self.D1l:
                xl = self._loc("xl1")
                self._ck(lib.gq_p2p_ntt_fwd_dev(h, self.lo1, self.D1l, 4, xl, ref(x[0]), None))
                self._ck(lib.gq_p2p_tensor_inv_dev(h, self.lo1, self.D1l, self.yl.data_ptr(), xl, ref(x[1]), None))
            else:                      # a rank that owns no limb of this basis still takes part in the exchange
                self._wait(0)
                self._ck(lib.gq_p2p_signal_dev(h, self.p_flags, 1, self.G, self.rank, self.epoch.data_ptr(), self.err.data_ptr(), None))
        elif k == 2:
            outs = (C.c_void_p * 3)(*[self.d[j].data_ptr() for j in range(3)])
            self._ck(lib.gq_p2p_reconstruct_dev(h, self.D1, self.l, 3, self._loc("rh"), outs, nl, ref(x[1]), None))
            arr = (C.c_void_p * 1)(self.d[2].data_ptr())
            self._ck(lib.gq_p2p_decompose_dev(h, self.D2, 1, arr, W, nl, ref(x[2]), None))
        elif k == 3:
            if self.D2l:
                xl = self._loc("xl2")
                self._ck(lib.gq_p2p_ntt_fwd_dev(h, self.lo2, self.D2l, 1, xl, ref(x[2]), None))
                self._ck(lib.gq_p2p_mul_evk_inv_dev(h, self.lo2, self.D2l, self.y2l.data_ptr(), xl, self.rlk._h, ref(x[3]), None))
            else:
                self._wait(2)
                self._ck(lib.gq_p2p_signal_dev(h, self.p_flags, 3, self.G, self.rank, self.epoch.data_ptr(), self.err.data_ptr(), None))
        else:
            self._ck(lib.gq_p2p_ks_epilogue_dev(h, self.l, self._loc("ch"), self.d[0].data_ptr(), self.d[1].data_ptr(),
                                                self.out[0].data_ptr(), self.out[1].data_ptr(), nl,
                                                1 if self.rescale else 0, ref(x[3]), None))

    def run(self, check=True):
        """one he_mul (+ he_rs): self.inp -> self.out (real multi-rank use; every rank calls it).
        check=True waits for the op and raises if this rank, or any peer (poison epoch), timed out on an exchange:
        the outputs are never handed out unchecked.  A caller that pipelines ops passes check=False and must call
        failed() before it uses any output."""
        if self.graph is not None:
            self.graph.replay()
        else:
            for k in range(5):
                self.seg(k)
        if check and self.failed():
            raise GpqheError("P2PPlan: a peer did not reach an exchange in time (GPQHE_B200_P2P_TIMEOUT_MS); "
                             "the plan is poisoned on every rank and must be rebuilt")
        return self.out[0], self.out[1]

    def close(self):
        """unmap the peers' slabs and free this rank's (every rank must have finished using the plan)"""
        lib, h = self.ctx._lib, self.ctx._h
        owned = getattr(self, "_owned", None)
        for s, p in enumerate(getattr(self, "_slabs", [])):
            if p and p != owned and getattr(self, "_mapped", False):
                lib.gq_p2p_close(h, C.c_void_p(p))
        if owned:
            lib.gq_p2p_free(h, C.c_void_p(owned))
        self._owned, self._slabs, self.graph = None, [], None

    def capture(self):
        """record the whole op as one CUDA graph (call with the context's stream current, after a
        warm-up run on every rank)"""
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=torch.cuda.current_stream(), capture_error_mode="thread_local"):
            for k in range(5):
                self.seg(k)
        self.graph = g

    def failed(self) -> bool:
        return bool(self.err.item())

    # -- wiring ----------------------------------------------------------------------------
    @classmethod
    def build(cls, sh: ShardedHe, l: int, rlk: HeEvk, rescale=True):
        """real ranks: allocate my slab, exchange IPC handles over torch.distributed, map the peers"""
        ctx, G, r = sh.ctx, sh.G, sh.rank
        lay = cls.layout(ctx, G, l)
        mine, handle = C.c_void_p(), C.create_string_buffer(64)
        capi.check(ctx._lib, ctx._lib.gq_p2p_alloc(ctx._h, 8 * lay["words"], C.byref(mine), handle))
        handles = [None] * G
        if G > 1:
            dist.all_gather_object(handles, handle.raw, group=sh.group)
        else:
            handles[0] = handle.raw
        slabs = []
        for s in range(G):
            if s == r:
                slabs.append(mine.value)
            else:
                p = C.c_void_p()
                capi.check(ctx._lib, ctx._lib.gq_p2p_open(ctx._h, handles[s], C.byref(p)))
                slabs.append(p.value)
        plan = cls(ctx, G, r, l, rlk, slabs, rescale, sh.device)
        plan._owned, plan._mapped = mine.value, True
        if G > 1:
            dist.barrier(group=sh.group)     # every slab is zeroed and mapped before anyone signals
        return plan

    @classmethod
    def loopback(cls, ctx: HeContext, G: int, l: int, rlk: HeEvk, rescale=True, device="cpu"):
        """G virtual ranks in one process (tests): returns the list of plans"""
        lay = cls.layout(ctx, G, l)
        slabs = []
        for _ in range(G):
            p, hd = C.c_void_p(), C.create_string_buffer(64)
            capi.check(ctx._lib, ctx._lib.gq_p2p_alloc(ctx._h, 8 * lay["words"], C.byref(p), hd))
            slabs.append(p.value)
        plans = [cls(ctx, G, r, l, rlk, slabs, rescale, device) for r in range(G)]
        for r, pl in enumerate(plans):
            pl._owned = slabs[r]          # one address space: every virtual rank frees its own slab in close()
        return plans

    @staticmethod
    def run_lockstep(plans):
        for k in range(5):
            for p in plans:
                p.seg(k)
