// gq_p2p.h - device-side pieces of the limb <-> coefficient exchange of the sharded pipeline that live
// INSIDE the producing and consuming kernels (round 2): the producer stores every residue at its final place in the
// owner's receive slab over NVLink while it computes, the last of its blocks to finish publishes the exchange's epoch
// in every peer's flag array; the consumer's blocks poll their own flags before they touch the slab.  No scatter,
// signal or wait launches (gq_p2p.cu keeps those for parameter sets the fused kernels do not cover).
#pragma once
#include "gq_platform.h"

#define GQ_P2P_MAX_RANKS 8
#define GQ_P2P_POISON (~0ull)

struct GqPeers {
  uint64_t *base[GQ_P2P_MAX_RANKS];   // destination buffers (peer-mapped; base[me] is local)
  int off[GQ_P2P_MAX_RANKS + 1];      // limb ownership: rank s owns limbs [off[s], off[s+1])
};

// producer side of one exchange
struct GqPeerOut {
  GqPeers dst;                        // where the data goes (one buffer per rank, as mapped in this process)
  uint64_t *flags[GQ_P2P_MAX_RANKS];  // flag arrays of the ranks, as mapped in this process
  int G, me, nloc, n, D;              // ranks, my rank, coefficients per rank, ring degree, limbs of the exchange
  int lg_nloc;                        // nloc = 2^lg_nloc (n is a power of two and G divides it)
  int slot;
  const unsigned long long *epoch;    // device memory: the op counter (graph replays read the current value)
  const int *err;                     // my sticky error flag: set -> publish the poison epoch instead
  unsigned *ticket;                   // device counter, zero between launches
};

// consumer side of one exchange (flags == null: nothing to wait for)
struct GqWaitIn {
  const uint64_t *flags;              // MY flag array
  int slot, G;
  const unsigned long long *epoch;
  int *err;
  unsigned long long spin_limit;      // polls of 200 ns before giving up
};

// residue of local coefficient i (of my slice), limb d, polynomial b -> its place at the owner of limb d:
//   dst_s[(b*Dl_s + d - off[s]) * n + me*nloc + i]
GQ_DEV uint64_t *gq_peer_c2l(const GqPeerOut &po, int b, int d, int i) {
  int s = 0;
#pragma unroll
  for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) s += (k < po.G && d >= po.dst.off[k]) ? 1 : 0;
  const int Dl = po.dst.off[s + 1] - po.dst.off[s];
  uint64_t *base = po.dst.base[0];
#pragma unroll
  for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) base = (k == s) ? po.dst.base[k] : base;     // no dynamic indexing of params
  return base + ((size_t)(b * Dl + d - po.dst.off[s]) * po.n + (size_t)po.me * po.nloc + i);
}
// coefficient j (global index) of my limb dl, polynomial b -> its place at the owner of coefficient j:
//   dst_s[(b*D + off[me] + dl) * nloc + j - s*nloc]
GQ_DEV uint64_t *gq_peer_l2c(const GqPeerOut &po, int b, int dl, int j, int my_limb0) {
  const int s = j / po.nloc;
  uint64_t *base = po.dst.base[0];
#pragma unroll
  for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) base = (k == s) ? po.dst.base[k] : base;
  return base + ((size_t)(b * po.D + my_limb0 + dl) * po.nloc + (j - s * po.nloc));
}

// the same through a table of the ranks' base pointers in shared memory (gq_peer_bases), shifts instead of the division
GQ_DEV void gq_peer_bases(const GqPeerOut &po, uint64_t **sb) {
  const int s = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (s < GQ_P2P_MAX_RANKS) {
    uint64_t *base = po.dst.base[0];
#pragma unroll
    for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) base = (k == s) ? po.dst.base[k] : base;
    sb[s] = base;
  }
}
GQ_DEV uint64_t *gq_peer_l2c_sb(const GqPeerOut &po, uint64_t *const *sb, size_t row, int j) {
  return sb[j >> po.lg_nloc] + ((row << po.lg_nloc) + (size_t)(j & (po.nloc - 1)));     // row = b * D + my_limb0 + dl
}

// acquire / release at system scope for the flags
GQ_DEV unsigned long long gq_ld_acquire_sys(const uint64_t *p) {
#if defined(__CUDA_ARCH__)
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
#else
  return *(const volatile unsigned long long *)p;
#endif
}

// End of a producing kernel, called by EVERY thread of EVERY block (no early returns before it): the last block to
// arrive publishes the epoch - or the poison value when this rank's error flag is set - to every peer.
GQ_DEV void gq_peer_signal(const GqPeerOut &po) {
  __shared__ int s_last;
  // Release in two steps (PTX memory model: causality order is transitive across synchronising pairs of different
  // scope).  Per block: the block barrier orders every thread's stores before thread 0, whose DEVICE-scope fence is
  // cumulative over them, then the ticket.  The last block to arrive acquires the ticket at device scope (it now
  // happens-after every other block's stores), and only IT pays the system-scope fence before the flags go out: a
  // system fence costs microseconds, and a kernel of a thousand short blocks paid it a thousand times (profiles/
  // r02_c28: the fused producers ran 15 us longer than their plain forms).
  __syncthreads();
  if (threadIdx.x == 0 && threadIdx.y == 0 && threadIdx.z == 0) {
    const unsigned nb = gridDim.x * gridDim.y * gridDim.z;
#ifdef GQ_EMULATE
    const unsigned t = (*po.ticket)++;
#else
    __threadfence();
    const unsigned t = atomicAdd(po.ticket, 1u);
    __threadfence();
#endif
    s_last = (t == nb - 1);
  }
  __syncthreads();
  if (!s_last) return;
  const int s = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (s < po.G) {
    const unsigned long long epoch = (po.err && *po.err) ? GQ_P2P_POISON : *po.epoch;
#ifndef GQ_EMULATE
    __threadfence_system();
#endif
    uint64_t *f = po.flags[0];
#pragma unroll
    for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) f = (k == s) ? po.flags[k] : f;
    *(volatile unsigned long long *)(f + (size_t)po.slot * GQ_P2P_MAX_RANKS + po.me) = epoch;
  }
  if (s == 0) *po.ticket = 0;         // ready for the next launch (stream order separates the launches)
}

// Start of a consuming kernel, called by every thread of every block before the first read of the slab.
GQ_DEV void gq_peer_wait(const GqWaitIn &w) {
  if (!w.flags) return;
  const int s = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (s < w.G && !*w.err) {           // sticky error: fail fast, never wait again
    const unsigned long long epoch = *w.epoch;
    const uint64_t *f = w.flags + (size_t)w.slot * GQ_P2P_MAX_RANKS + s;
    unsigned long long spins = 0, v;
    bool ok = true;
    // acquire loads at system scope: the slab reads behind the block barrier are ordered after the flag read that saw
    // the epoch (no separate system fence: it cost every block of every consumer several microseconds)
    while ((v = gq_ld_acquire_sys(f)) < epoch) {
#ifdef GQ_EMULATE
      if (++spins > 4) { ok = false; break; }
#else
      __nanosleep(200);
      if (++spins > w.spin_limit) { ok = false; break; }
#endif
    }
    if (!ok) *w.err = 1;
    else if (v == GQ_P2P_POISON) *w.err = 2;
  }
  __syncthreads();
}
