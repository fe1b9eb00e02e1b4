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

import torch
import torch.distributed as dist

from . import capi
from .capi import GpqheError
from .he import HeContext, HeEvk


class ShardedHe:
    """he_mul / he_rs / he_moddown on coefficient-sharded ciphertexts across `group`."""

    def __init__(self, ctx: HeContext, group=None, device: torch.device | str | None = None):
        self.ctx, self.group = ctx, group
        self.G = dist.get_world_size(group) if dist.is_initialized() else 1
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
        self._ck(lib.gq_stage_tensor_dev(h, lo1, D1l, y.data_ptr(), x.data_ptr(), None))
        self._ck(lib.gq_stage_ntt_dev(h, lo1, D1l, 3, y.data_ptr(), 1, None))
        # hot loop B (he-mult.c:139-141)
        rh = self._limb_to_coef(y, D1)                                            # [3, D1, n_loc]
        d = self._empty(3, self.W, self.n_loc)
        outs = (C.c_void_p * 3)(*[d[k].data_ptr() for k in range(3)])
        self._ck(lib.gq_stage_reconstruct_dev(h, D1, l, 3, rh.data_ptr(), outs, self.n_loc, None))
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
        self._ck(lib.gq_stage_mul_evk_dev(h, lo2, D2l, y.data_ptr(), x.data_ptr(), evk._h, None))
        self._ck(lib.gq_stage_ntt_dev(h, lo2, D2l, 2, y.data_ptr(), 1, None))
        ch = self._limb_to_coef(y, D2)                                            # [2, D2, n_loc]
        c0, c1 = self._empty(self.W, self.n_loc), self._empty(self.W, self.n_loc)
        self._ck(lib.gq_stage_ks_epilogue_dev(h, l, ch.data_ptr(), add0.data_ptr() if add0 is not None else None,
                                              add1.data_ptr() if add1 is not None else None,
                                              c0.data_ptr(), c1.data_ptr(), self.n_loc, None))
        return c0, c1

    def he_rs(self, l: int, c0, c1):
        """he_rs in place on the slices; the ciphertext is then at level l-1."""
        self._ck(self._lib.gq_stage_rescale_dev(self._h, l, c0.data_ptr(), c1.data_ptr(), 1, self.n_loc, None))
        return c0, c1

    def he_moddown(self, l: int, c0, c1):
        self._ck(self._lib.gq_stage_rescale_dev(self._h, l, c0.data_ptr(), c1.data_ptr(), 0, self.n_loc, None))
        return c0, c1
