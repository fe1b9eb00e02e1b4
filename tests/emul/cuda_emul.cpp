// cuda_emul.cpp - TEST INFRASTRUCTURE, see cuda_emul.h.
#include "cuda_emul.h"

#include <time.h>
#include <ucontext.h>

#include <cstdio>
#include <vector>

dim3 gridDim, blockDim;
uint3 blockIdx, threadIdx;

namespace gq_emul {

long long g_launches = 0;

namespace {
constexpr size_t kStack = 256 * 1024;

struct Fibre {
  ucontext_t ctx;
  uint3 tid;
  bool done;
};

ucontext_t g_sched;
std::vector<Fibre> g_fibres;
std::vector<char *> g_stacks;
std::vector<uint64_t> g_slots;
const std::function<void()> *g_body = nullptr;
int g_cur = -1;

void trampoline() {
  (*g_body)();
  g_fibres[g_cur].done = true;
  swapcontext(&g_fibres[g_cur].ctx, &g_sched);
}
}  // namespace

unsigned linear_tid() {
  return threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
}

void yield_barrier() { swapcontext(&g_fibres[g_cur].ctx, &g_sched); }

uint64_t exchange(uint64_t v, unsigned src) {
  g_slots[linear_tid()] = v;
  yield_barrier();
  uint64_t r = g_slots[src];
  yield_barrier();
  return r;
}

static std::vector<uint32_t> g_wide;
static std::vector<char> g_dyn;
void *dyn_smem() { return g_dyn.data(); }
void launch_dyn(dim3 grid, dim3 block, size_t bytes, const std::function<void()> &body) {
  if (g_dyn.size() < bytes + 16) g_dyn.resize(bytes + 16);
  launch(grid, block, body);
}
void warp_gather(const uint32_t *mine, int nwords, uint32_t *all) {
  const unsigned tid = linear_tid(), nt = blockDim.x * blockDim.y * blockDim.z;
  if (g_wide.size() < (size_t)nt * 8) g_wide.resize((size_t)nt * 8 + 256);
  for (int k = 0; k < nwords; k++) g_wide[(size_t)tid * 8 + k] = mine[k];
  yield_barrier();
  const unsigned w0 = tid & ~31u;
  for (unsigned l = 0; l < 32; l++)
    for (int k = 0; k < nwords; k++) all[l * nwords + k] = (w0 + l < nt) ? g_wide[(size_t)(w0 + l) * 8 + k] : 0;
  yield_barrier();
}

void launch(dim3 grid, dim3 block, const std::function<void()> &body) {
  g_launches++;
  const unsigned nt = block.x * block.y * block.z;
  gridDim = grid;
  blockDim = block;
  g_body = &body;
  if (g_fibres.size() < nt) g_fibres.resize(nt);
  if (g_slots.size() < nt) g_slots.resize(nt);
  while (g_stacks.size() < nt) g_stacks.push_back((char *)malloc(kStack));
  for (unsigned bz = 0; bz < grid.z; bz++)
    for (unsigned by = 0; by < grid.y; by++)
      for (unsigned bx = 0; bx < grid.x; bx++) {
        blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
        unsigned t = 0;
        for (unsigned tz = 0; tz < block.z; tz++)
          for (unsigned ty = 0; ty < block.y; ty++)
            for (unsigned tx = 0; tx < block.x; tx++, t++) {
              Fibre &f = g_fibres[t];
              f.tid.x = tx; f.tid.y = ty; f.tid.z = tz;
              f.done = false;
              getcontext(&f.ctx);
              f.ctx.uc_stack.ss_sp = g_stacks[t];
              f.ctx.uc_stack.ss_size = kStack;
              f.ctx.uc_link = &g_sched;
              makecontext(&f.ctx, trampoline, 0);
            }
        unsigned live = nt;
        while (live) {
          live = 0;
          for (unsigned i = 0; i < nt; i++) {
            Fibre &f = g_fibres[i];
            if (f.done) continue;
            g_cur = (int)i;
            threadIdx = f.tid;
            swapcontext(&g_sched, &f.ctx);
            if (!f.done) live++;
          }
        }
      }
  g_body = nullptr;
  g_cur = -1;
}

}  // namespace gq_emul

// ---------------------------------------------------------------- fake runtime
struct gq_emul_stream { int dummy; };
struct gq_emul_event { struct timespec ts; };

cudaError_t cudaMalloc(void **p, size_t n) {
  *p = aligned_alloc(256, (n + 255) / 256 * 256 + 256);
  return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
cudaError_t cudaFree(void *p) { free(p); return cudaSuccess; }
cudaError_t cudaMallocHost(void **p, size_t n) { return cudaMalloc(p, n); }
cudaError_t cudaFreeHost(void *p) { free(p); return cudaSuccess; }
cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) {
  memmove(d, s, n); return cudaSuccess;
}
cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t) {
  memmove(d, s, n); return cudaSuccess;
}
cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t) {
  memset(d, v, n); return cudaSuccess;
}
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) {
  *s = new gq_emul_stream(); return cudaSuccess;
}
cudaError_t cudaStreamDestroy(cudaStream_t s) { delete s; return cudaSuccess; }
cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
cudaError_t cudaGetLastError() { return cudaSuccess; }
cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
const char *cudaGetErrorString(cudaError_t e) { return e ? "emulated error" : "no error"; }
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return cudaSuccess; }
cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new gq_emul_event(); return cudaSuccess; }
cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) {
  clock_gettime(CLOCK_MONOTONIC, &e->ts); return cudaSuccess;
}
cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = (float)((b->ts.tv_sec - a->ts.tv_sec) * 1e3 + (b->ts.tv_nsec - a->ts.tv_nsec) * 1e-6);
  return cudaSuccess;
}
cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
