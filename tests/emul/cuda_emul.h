// cuda_emul.h - TEST INFRASTRUCTURE: a minimal single-process stand-in for the CUDA
// execution model (blocks run one after another; the threads of a block are
// ucontext fibres scheduled round-robin between barriers) plus a fake subset of the
// CUDA runtime API backed by malloc/memcpy.  It lets the `-m "not gpu"` tests run
// the library's host logic and kernel index arithmetic in a container without a
// GPU.  It is selected only by compiling with -DGQ_EMULATE (tests/emul/Makefile);
// the product build (nvcc, sm_100a) never sees this file.
#pragma once

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__ static
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct ulonglong2 { unsigned long long x, y; } __attribute__((aligned(16)));
static inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) {
  ulonglong2 r; r.x = x; r.y = y; return r;
}

extern dim3 gridDim, blockDim;
extern uint3 blockIdx, threadIdx;
static const int warpSize = 32;

namespace gq_emul {
void launch(dim3 grid, dim3 block, const std::function<void()> &body);
// launch with `bytes` of dynamic shared memory, reachable through dyn_smem() (GQ_DYN_SMEM in gq_platform.h)
void launch_dyn(dim3 grid, dim3 block, size_t bytes, const std::function<void()> &body);
void *dyn_smem();
// warp-wide all-gather of `nwords` 32-bit words per lane: all[lane][0..nwords) for the 32 lanes of the caller's
// warp (every lane of the warp must call it); building block of the tensor-core emulation below
void warp_gather(const uint32_t *mine, int nwords, uint32_t *all);
void yield_barrier();                 // __syncthreads
uint64_t exchange(uint64_t v, unsigned src_lane_in_block);   // shuffle helper
unsigned linear_tid();
extern long long g_launches;          // number of emulated kernel launches
}

static inline void __syncthreads() { gq_emul::yield_barrier(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { gq_emul::yield_barrier(); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}

static inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) {
  return (unsigned long long)(((unsigned __int128)a * b) >> 64);
}
static inline unsigned __umulhi(unsigned a, unsigned b) {
  return (unsigned)(((unsigned long long)a * b) >> 32);
}
static inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline unsigned __brev(unsigned x) {
  x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
  x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
  x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
  x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
  return (x >> 16) | (x << 16);
}
template <typename T> static inline T __ldg(const T *p) { return *p; }

// warp shuffles: every thread of the block must execute the same shuffle sequence
template <typename T> static inline T __shfl_xor_sync(unsigned, T v, int lanemask) {
  static_assert(sizeof(T) <= 8, "shuffle of <= 64-bit values only");
  uint64_t raw = 0; std::memcpy(&raw, &v, sizeof(T));
  unsigned tid = gq_emul::linear_tid();
  raw = gq_emul::exchange(raw, (tid & ~31u) | ((tid ^ (unsigned)lanemask) & 31u));
  T r; std::memcpy(&r, &raw, sizeof(T)); return r;
}
template <typename T> static inline T __shfl_sync(unsigned, T v, int srclane) {
  uint64_t raw = 0; std::memcpy(&raw, &v, sizeof(T));
  unsigned tid = gq_emul::linear_tid();
  raw = gq_emul::exchange(raw, (tid & ~31u) | ((unsigned)srclane & 31u));
  T r; std::memcpy(&r, &raw, sizeof(T)); return r;
}

// mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 (PTX ISA fragment layouts): g = lane / 4, t = lane % 4
//   A: a0..a3 = rows (g, g+8, g, g+8) x k-bytes (4t.., 4t.., 16+4t.., 16+4t..);  B: b0, b1 = k-bytes (4t.., 16+4t..) x col g
//   C: c0, c1 = row g, cols 2t, 2t+1;  c2, c3 = row g+8
static inline void gq_emul_mma_u8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  uint32_t mine[6] = {a[0], a[1], a[2], a[3], b0, b1}, all[32][6];
  gq_emul::warp_gather(mine, 6, &all[0][0]);
  const unsigned lane = gq_emul::linear_tid() & 31u, g = lane >> 2, t = lane & 3u;
  for (int e = 0; e < 4; e++) {
    const unsigned row = g + ((e & 2) ? 8 : 0), col = 2 * t + (e & 1);
    long long acc = c[e];
    for (unsigned k = 0; k < 32; k++) {
      const unsigned tk = (k & 15) >> 2, byte = k & 3;
      const uint32_t ar = all[(row & 7) * 4 + tk][(row >= 8 ? 1 : 0) + (k >= 16 ? 2 : 0)];
      const uint32_t br = all[col * 4 + tk][k >= 16 ? 5 : 4];
      acc += (long long)((ar >> (8 * byte)) & 0xff) * (long long)((br >> (8 * byte)) & 0xff);
    }
    c[e] = (int)acc;
  }
}

// ---------------------------------------------------------------- fake runtime
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1 };
typedef struct gq_emul_stream *cudaStream_t;
typedef struct gq_emul_event *cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1,
                      cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3,
                      cudaMemcpyDefault = 4 };
enum { cudaStreamNonBlocking = 1, cudaEventDefault = 0, cudaEventDisableTiming = 2, cudaHostAllocDefault = 0 };

cudaError_t cudaMalloc(void **p, size_t n);
cudaError_t cudaFree(void *p);
cudaError_t cudaMallocHost(void **p, size_t n);
cudaError_t cudaFreeHost(void *p);
cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind k);
cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind k, cudaStream_t st);
cudaError_t cudaMemset(void *d, int v, size_t n);
cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t st);
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned flags);
cudaError_t cudaStreamDestroy(cudaStream_t s);
cudaError_t cudaStreamSynchronize(cudaStream_t s);
cudaError_t cudaDeviceSynchronize();
cudaError_t cudaGetLastError();
cudaError_t cudaPeekAtLastError();
const char *cudaGetErrorString(cudaError_t e);
cudaError_t cudaSetDevice(int d);
cudaError_t cudaGetDevice(int *d);
cudaError_t cudaGetDeviceCount(int *n);
cudaError_t cudaEventCreate(cudaEvent_t *e);
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { return cudaEventCreate(e); }
cudaError_t cudaEventDestroy(cudaEvent_t e);
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s);
cudaError_t cudaEventSynchronize(cudaEvent_t e);
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b);
cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned flags);
// a __device__ variable is a plain host object here: its "device address" is its address
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &symbol) {
  *p = (void *)&symbol; return cudaSuccess;
}
template <typename T> static inline cudaError_t cudaMalloc(T **p, size_t n) {
  return cudaMalloc((void **)p, n);
}
template <typename T> static inline cudaError_t cudaMallocHost(T **p, size_t n) {
  return cudaMallocHost((void **)p, n);
}
