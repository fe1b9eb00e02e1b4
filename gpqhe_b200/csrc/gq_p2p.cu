// gq_p2p.cu - the limb <-> coefficient re-layout of the sharded pipeline written
// straight into the peers' memory over NVLink, instead of pack + NCCL all-to-all + unpack.
//
// Every rank owns one receive slab (cudaMalloc, exported with cudaIpcGetMemHandle and mapped by
// the other ranks of the node with cudaIpcOpenMemHandle).  An exchange is then
//     k_scatter_*   : read my residues once, store each element at its final place in the
//                     owner's slab (coalesced 16-byte stores to peer addresses)
//     k_signal      : system-scope fence, then write the exchange's epoch into flag[slot][me] of
//                     every peer
//     k_wait        : (consumer side) spin until flag[slot][src] >= epoch for every src
// Flow control needs no second buffer set: each exchange is an all-to-all dependency, so a rank
// can only reach exchange k of op t+1 after every peer has consumed exchange k of op t.
// One process per GPU drives one GPU; the kernels that wait run on a different GPU than the
// kernels they wait for.
#include <stdlib.h>
#include <string.h>

#include <string>

#include "../../include/gpqhe_b200.h"
#include "gq_context.h"

#define GQ_P2P_MAX_RANKS 8

struct GqPeers {
  uint64_t *base[GQ_P2P_MAX_RANKS];   // destination buffers (peer-mapped; base[me] is local)
  int off[GQ_P2P_MAX_RANKS + 1];      // limb ownership: rank s owns limbs [off[s], off[s+1])
};

namespace {

GQ_DEV ulonglong2 ld2(const uint64_t *p) { return *reinterpret_cast<const ulonglong2 *>(p); }
GQ_DEV void st2(uint64_t *p, ulonglong2 v) { *reinterpret_cast<ulonglong2 *>(p) = v; }

// src[P][D][nloc] (all limbs of my coefficient slice) -> at the owner s of limb d:
//   dst_s[(b*Dl_s + d - off[s]) * n + me*nloc + i]
__global__ void __launch_bounds__(256)
k_scatter_c2l(GqPeers pr, const uint64_t *__restrict__ src, int P, int D, int nloc, int n, int me, int G) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = nloc >> 1;
  const long long total = (long long)P * D * half;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int i2 = (int)(e % half);
    const int bd = (int)(e / half);
    const int b = bd / D, d = bd % D;
    int s = 0;
#pragma unroll
    for (int k = 1; k < GQ_P2P_MAX_RANKS; k++) s += (k < G && d >= pr.off[k]) ? 1 : 0;
    const int Dl = pr.off[s + 1] - pr.off[s];
    uint64_t *dst = pr.base[s] + ((size_t)(b * Dl + d - pr.off[s]) * n + (size_t)me * nloc + 2 * i2);
    st2(dst, ld2(src + 2 * e));
  }
}

// src[P][Dl][n] (my limbs, every coefficient) -> at the owner s of coefficient j:
//   dst_s[(b*D + off[me] + dl) * nloc + j - s*nloc]
__global__ void __launch_bounds__(256)
k_scatter_l2c(GqPeers pr, const uint64_t *__restrict__ src, int P, int Dl, int D, int nloc, int n, int me) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long total = (long long)P * Dl * half;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int j = 2 * (int)(e % half);
    const int bd = (int)(e / half);
    const int b = bd / Dl, dl = bd % Dl;
    const int s = j / nloc;
    uint64_t *dst = pr.base[s] + ((size_t)(b * D + pr.off[me] + dl) * nloc + (j - s * nloc));
    st2(dst, ld2(src + 2 * e));
  }
}

// the epoch lives in device memory so that the same (graph-captured) launch works for every op
__global__ void k_bump(unsigned long long *epoch) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT(); if (threadIdx.x == 0) *epoch += 1; }

// A rank whose error flag is set (a wait of this or an earlier op timed out: its slab may be half written) does
// not publish the epoch but a POISON value: it passes every peer's `flag >= epoch` test at once, and k_wait turns it
// into the peer's own error flag.  One late rank therefore fails the op on every rank, visibly, instead of letting
// the others consume a half-written slab.  The flag is sticky: every later op of the plan fails fast.
#define GQ_P2P_POISON (~0ull)

__global__ void k_signal(GqPeers flags, int slot, int G, int me, const unsigned long long *epoch_p, const int *err) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int s = threadIdx.x;
  if (s >= G) return;
  const unsigned long long epoch = (err && *err) ? GQ_P2P_POISON : *epoch_p;
#ifndef GQ_EMULATE
  __threadfence_system();
#endif
  volatile unsigned long long *f = (volatile unsigned long long *)(flags.base[s] + (size_t)slot * GQ_P2P_MAX_RANKS + me);
  *f = epoch;
}

// err[0] is set when a peer never showed up within `spin_limit` polls of 200 ns (bounded spin: a lost peer must not
// hang the GPU) or published the poison value.  An acquire fence follows the spin so that the slab reads of the
// consuming kernels are ordered after the flag read on every path, not only across the kernel boundary.
__global__ void k_wait(const uint64_t *flags, int slot, int G, const unsigned long long *epoch_p, int *err,
                       unsigned long long spin_limit) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int s = threadIdx.x;
  if (s >= G) return;
  if (*err) return;                                       // sticky: fail fast, never wait again
  const unsigned long long epoch = *epoch_p;
  const volatile unsigned long long *f = (const volatile unsigned long long *)(flags + (size_t)slot * GQ_P2P_MAX_RANKS + s);
  unsigned long long spins = 0;
  while (*f < epoch) {
#ifdef GQ_EMULATE
    if (++spins > 4) { *err = 1; return; }
#else
    __nanosleep(200);
    if (++spins > spin_limit) { *err = 1; return; }
#endif
  }
  if (*f == GQ_P2P_POISON) *err = 2;
#ifndef GQ_EMULATE
  __threadfence_system();
#endif
}

}  // namespace

#define P2P_FAIL(msg) do { gq_set_error(std::string(__func__) + ": " + (msg)); return 1; } while (0)
#define P2P_CUDA(call)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      gq_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

static int fill_peers(GqPeers *pr, const uint64_t *const *ptrs, const int *off, unsigned G) {
  if (G == 0 || G > GQ_P2P_MAX_RANKS) return 1;
  for (unsigned s = 0; s < GQ_P2P_MAX_RANKS; s++) pr->base[s] = (s < G) ? (uint64_t *)ptrs[s] : nullptr;
  for (unsigned s = 0; s <= GQ_P2P_MAX_RANKS; s++) pr->off[s] = off ? off[s <= G ? s : G] : 0;
  return 0;
}
static cudaStream_t pick(gq_ctx *c, void *s) { return s ? (cudaStream_t)s : c->stream; }
static int grid_for(long long items) {
  long long b = (items + 255) / 256, cap = (long long)GQ_NUM_SMS * 8;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

extern "C" int gq_p2p_alloc(gq_ctx *c, size_t bytes, void **ptr, unsigned char handle[64]) {
  if (!c || !ptr || !handle || bytes == 0) P2P_FAIL("bad argument");
  P2P_CUDA(cudaSetDevice(c->device));
  P2P_CUDA(cudaMalloc(ptr, bytes));
  P2P_CUDA(cudaMemset(*ptr, 0, bytes));
#ifdef GQ_EMULATE
  memset(handle, 0, 64);
  memcpy(handle, ptr, sizeof(void *));      // one address space: the "handle" is the pointer
#else
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  P2P_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle, &h, 64);
#endif
  return 0;
}

extern "C" int gq_p2p_open(gq_ctx *c, const unsigned char handle[64], void **ptr) {
  if (!c || !ptr || !handle) P2P_FAIL("bad argument");
#ifdef GQ_EMULATE
  memcpy(ptr, handle, sizeof(void *));
#else
  P2P_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  P2P_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
#endif
  return 0;
}

extern "C" int gq_p2p_close(gq_ctx *c, void *ptr) {
  if (!c || !ptr) P2P_FAIL("bad argument");
#ifndef GQ_EMULATE
  P2P_CUDA(cudaIpcCloseMemHandle(ptr));
#endif
  return 0;
}

extern "C" int gq_p2p_free(gq_ctx *c, void *ptr) {
  if (!c) P2P_FAIL("bad argument");
  if (ptr) P2P_CUDA(cudaFree(ptr));
  return 0;
}

extern "C" int gq_p2p_scatter_c2l_dev(gq_ctx *c, const uint64_t *src, unsigned P, unsigned D, unsigned nloc,
                                      unsigned G, unsigned me, const uint64_t *const *dst, const int *off,
                                      void *stream) {
  GqPeers pr;
  if (!c || !src || !dst || !off || me >= G || fill_peers(&pr, dst, off, G) || (nloc & 1)) P2P_FAIL("bad argument");
  GQ_LAUNCH(k_scatter_c2l, dim3(grid_for((long long)P * D * (nloc >> 1))), dim3(256), pick(c, stream), pr, src,
            (int)P, (int)D, (int)nloc, (int)c->n, (int)me, (int)G);
  c->launches++;
  P2P_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int gq_p2p_scatter_l2c_dev(gq_ctx *c, const uint64_t *src, unsigned P, unsigned Dl, unsigned D,
                                      unsigned nloc, unsigned G, unsigned me, const uint64_t *const *dst,
                                      const int *off, void *stream) {
  GqPeers pr;
  if (!c || !src || !dst || !off || me >= G || fill_peers(&pr, dst, off, G) || (nloc & 1)) P2P_FAIL("bad argument");
  if (Dl == 0) return 0;
  GQ_LAUNCH(k_scatter_l2c, dim3(grid_for((long long)P * Dl * (c->n >> 1))), dim3(256), pick(c, stream), pr, src,
            (int)P, (int)Dl, (int)D, (int)nloc, (int)c->n, (int)me);
  c->launches++;
  P2P_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int gq_p2p_bump_dev(gq_ctx *c, unsigned long long *epoch_dev, void *stream) {
  if (!c || !epoch_dev) P2P_FAIL("bad argument");
  GQ_LAUNCH(k_bump, dim3(1), dim3(32), pick(c, stream), epoch_dev);
  c->launches++;
  P2P_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int gq_p2p_signal_dev(gq_ctx *c, const uint64_t *const *flags, unsigned slot, unsigned G, unsigned me,
                                 const unsigned long long *epoch, const int *err_dev, void *stream) {
  GqPeers pr;
  if (!c || !flags || me >= G || fill_peers(&pr, flags, nullptr, G)) P2P_FAIL("bad argument");
  GQ_LAUNCH(k_signal, dim3(1), dim3(32), pick(c, stream), pr, (int)slot, (int)G, (int)me, epoch, err_dev);
  c->launches++;
  P2P_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int gq_p2p_wait_dev(gq_ctx *c, const uint64_t *flags_local, unsigned slot, unsigned G,
                               const unsigned long long *epoch, int *err_dev, void *stream) {
  if (!c || !flags_local || !err_dev || G == 0 || G > GQ_P2P_MAX_RANKS) P2P_FAIL("bad argument");
  // GPQHE_B200_P2P_TIMEOUT_MS: how long a rank waits for a peer (default 4000 ms; first use of a plan includes the
  // peers' allocations and graph uploads).  Read when the launch is enqueued / captured.
  unsigned long long limit = 20000000ull;
  if (const char *e = getenv("GPQHE_B200_P2P_TIMEOUT_MS")) {
    const long long ms = atoll(e);
    if (ms > 0) limit = (unsigned long long)ms * 5000ull;      // 200 ns per poll
  }
  GQ_LAUNCH(k_wait, dim3(1), dim3(32), pick(c, stream), flags_local, (int)slot, (int)G, epoch, err_dev, limit);
  c->launches++;
  P2P_CUDA(cudaGetLastError());
  return 0;
}
