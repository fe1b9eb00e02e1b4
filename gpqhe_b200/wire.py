"""Wire format for ciphertexts, plaintexts and evaluation keys.

The reference has none: `kemctx` sizes are computed and never used (src/precomp.c:455-463) and every
object lives in process memory as libgcrypt MPIs.  This module defines the packed dump the device path
already uses internally, so that device state can be checkpointed and fixtures shipped to a GPU box:

  header (64 bytes, little endian)
     0  8s   magic  b"GQHEWIR1"
     8  u32  kind   1 = ciphertext (he_ct_t), 2 = plaintext (he_pt_t), 3 = evaluation key (he_evk_t)
    12  u32  logn
    16  u32  rows   words per coefficient W (kinds 1, 2) / limbs per polynomial dimevk (kind 3)
    20  u32  polys  2 (c0, c1 / p0, p1) or 1 (m)
    24  u32  l      level (kind 1), 0 otherwise
    28  u32  ctx_id CRC-32 of the context's prime list, q and Delta: loading into another ring fails
    32  f64  nu
    40  f64  B
    48  u64  payload bytes
    56  u32  CRC-32 of the payload
    60  u32  CRC-32 of the 60 header bytes before it
  payload
    kinds 1, 2: polys x uint64[W][n], word-major two's complement (the C-ABI layout of a poly_mpi_t)
    kind 3    : 2 x uint64[dimevk][n], limb-major residues in ntt() output order (the layout of he_evk_t,
                src/he-kem.c:103-110)

A C reader needs nothing but this table; `dump_*` / `load_*` below are the reference implementation.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

from .capi import GpqheError
from .he import HeContext, HeCt, HeEvk, HePt, int_to_words

MAGIC = b"GQHEWIR1"
KIND_CT, KIND_PT, KIND_EVK = 1, 2, 3
_HEAD = struct.Struct("<8sIIIIIIddQI")   # 60 bytes, then the header CRC
HEADER_BYTES = 64


def ctx_id(ctx: HeContext) -> int:
    """CRC-32 over the primes, q and Delta of the context (all that fixes the meaning of a payload)."""
    h = zlib.crc32(np.ascontiguousarray(ctx.primes, dtype="<u8").tobytes())
    h = zlib.crc32(int_to_words(int(ctx.q)).astype("<u8").tobytes(), h)
    return zlib.crc32(struct.pack("<Q", int(ctx.Delta or 0)), h)


def _pack(kind, ctx, rows, arrays, l=0, nu=0.0, B=0.0) -> bytes:
    payload = b"".join(np.ascontiguousarray(a, dtype="<u8").tobytes() for a in arrays)
    want = len(arrays) * rows * ctx.n * 8
    if len(payload) != want:
        raise GpqheError(f"wire: payload is {len(payload)} bytes, the ring wants {want}")
    head = _HEAD.pack(MAGIC, kind, ctx.logn, rows, len(arrays), l, ctx_id(ctx), float(nu), float(B), len(payload),
                      zlib.crc32(payload))
    return head + struct.pack("<I", zlib.crc32(head)) + payload


def _unpack(blob, ctx, kind, rows):
    blob = bytes(blob)
    if len(blob) < HEADER_BYTES:
        raise GpqheError("wire: truncated header")
    head = blob[:_HEAD.size]
    magic, k, logn, r, polys, l, cid, nu, B, nbytes, crc = _HEAD.unpack(head)
    (hcrc,) = struct.unpack_from("<I", blob, _HEAD.size)
    if magic != MAGIC:
        raise GpqheError("wire: bad magic")
    if hcrc != zlib.crc32(head):
        raise GpqheError("wire: header checksum mismatch")
    if k != kind:
        raise GpqheError(f"wire: object kind {k}, expected {kind}")
    if logn != ctx.logn or cid != ctx_id(ctx):
        raise GpqheError("wire: object was written for another context (ring, modulus chain or Delta differ)")
    if r != rows or nbytes != polys * rows * ctx.n * 8:
        raise GpqheError("wire: shape does not match the context")
    payload = blob[HEADER_BYTES:HEADER_BYTES + nbytes]
    if len(payload) != nbytes:
        raise GpqheError("wire: truncated payload")
    if zlib.crc32(payload) != crc:
        raise GpqheError("wire: payload checksum mismatch")
    arr = np.frombuffer(payload, dtype="<u8").astype(np.uint64).reshape(polys, rows, ctx.n)
    return arr, l, nu, B


def dump_ct(ctx: HeContext, ct: HeCt) -> bytes:
    if ct.l > ctx.L:
        raise GpqheError("wire: level out of range")
    return _pack(KIND_CT, ctx, ctx.W, [ct.c0, ct.c1], ct.l, ct.nu, ct.B)


def load_ct(ctx: HeContext, blob) -> HeCt:
    a, l, nu, B = _unpack(blob, ctx, KIND_CT, ctx.W)
    if a.shape[0] != 2 or l > ctx.L:
        raise GpqheError("wire: not a ciphertext of this context")
    return HeCt(l, a[0].copy(), a[1].copy(), nu, B)


def dump_pt(ctx: HeContext, pt: HePt) -> bytes:
    return _pack(KIND_PT, ctx, ctx.W, [pt.m], 0, pt.nu, 0.0)


def load_pt(ctx: HeContext, blob) -> HePt:
    a, _, nu, _ = _unpack(blob, ctx, KIND_PT, ctx.W)
    if a.shape[0] != 1:
        raise GpqheError("wire: not a plaintext of this context")
    return HePt(a[0].copy(), nu)


def export_evk(ctx: HeContext, evk: HeEvk):
    """(p0, p1) of a device-resident key, uint64 [dimevk, n] each (gq_evk_export)."""
    from . import capi
    p0 = np.empty((ctx.dimevk, ctx.n), dtype=np.uint64)
    p1 = np.empty_like(p0)
    capi.check(ctx._lib, ctx._lib.gq_evk_export(evk._h, p0.ctypes.data, p1.ctypes.data))
    return p0, p1


def dump_evk(ctx: HeContext, evk) -> bytes:
    """evk: a device-resident HeEvk, or the host pair (p0, p1)."""
    p0, p1 = export_evk(ctx, evk) if isinstance(evk, HeEvk) else evk
    return _pack(KIND_EVK, ctx, ctx.dimevk, [p0, p1])


def load_evk(ctx: HeContext, blob, upload: bool = True):
    """-> HeEvk on the device (upload=True) or the host pair (p0, p1)."""
    a, _, _, _ = _unpack(blob, ctx, KIND_EVK, ctx.dimevk)
    if a.shape[0] != 2:
        raise GpqheError("wire: not an evaluation key of this context")
    for d in range(ctx.dimevk):
        if int(a[:, d].max()) > int(ctx.primes[d]):        # ntt() may store p itself on the sum leg (src/ntt.c:48)
            raise GpqheError(f"wire: residue out of range in limb {d}")
    return ctx.genevk(a[0], a[1]) if upload else (a[0].copy(), a[1].copy())
