// gq_noalloc.cu - gq_stage_reconstruct_noalloc_dev: the CRT reconstruction stage (K5, poly_rns2mpi of
// src/poly.c:109-120) for callers that must not allocate: code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
//
// The fast CRT kernels (gq_crt.cu: k_reconstruct_fast + k_reconstruct in list mode) need scratch memory: a counter
// and a list of the coefficient indices whose overflow count the fixed-point estimate could not decide (values on
// the +P/2 edge).  Every other entry point takes that scratch from the context, which obtained it with cudaMalloc
// (gq_ctx_create) - and may grow other scratch slots on the way.  This variant takes it from the file-scope
// __device__ array below instead and calls no cudaMalloc / cudaFree / allocating routine:
//   * GQ_NOALLOC_MAX_ENTRIES = 262 144 list entries (4 polynomials x 65 536 coefficients) + a 4-int header =
//     1 MiB + 16 B of device memory that every GPU which loads the library holds until the process exits;
//     npolys * ncoef beyond that is rejected with an error;
//   * the array is ONE per device and shared by every call and every context on that device; it keeps the last
//     call's contents.  The header (pending count, lifetime total) is cleared on the caller's stream at the start of
//     every call; the list needs no clearing (only entries below the count are read).  Two calls that overlap on
//     the GPU (different streams, no ordering between them) RACE on it: order such calls, or use the context-owned
//     scratch of gq_stage_reconstruct_dev on separate lanes;
//   * in host code `g_noalloc_fix` is a host-side stand-in for the symbol, not a device address: the device address
//     is taken once per device, with that device current, by cudaGetSymbolAddress below and passed to the kernels.
// Inputs and outputs are the caller's, exactly as for gq_stage_reconstruct_dev; the kernels are the same, so the
// results are bit-identical.
#include <string>

#include "../../include/gpqhe_b200.h"
#include "gq_context.h"

#define GQ_NOALLOC_MAX_ENTRIES (4 * 65536)
#define GQ_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) int g_noalloc_fix[4 + GQ_NOALLOC_MAX_ENTRIES];

// device address of g_noalloc_fix per device ordinal (null = not looked up yet)
static int *g_noalloc_ptr[GQ_NOALLOC_MAX_DEVICES];

#define NA_FAIL(msg) do { gq_set_error(std::string(__func__) + ": " + (msg)); return 1; } while (0)
#define NA_CUDA(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      gq_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

static int noalloc_scratch(int device, int **out) {
  if (device < 0 || device >= GQ_NOALLOC_MAX_DEVICES) NA_FAIL("device ordinal beyond GQ_NOALLOC_MAX_DEVICES");
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    NA_CUDA(cudaSetDevice(device));                                   // the symbol has one instance per device
    NA_CUDA(cudaGetSymbolAddress(&p, g_noalloc_fix));
    if (!p) NA_FAIL("cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[device] = (int *)p;
  }
  *out = g_noalloc_ptr[device];
  return 0;
}

extern "C" int gq_stage_reconstruct_noalloc_dev(gq_ctx *c, unsigned D, unsigned l, unsigned npolys,
                                                const uint64_t *rhat, uint64_t *const *out, unsigned ncoef,
                                                void *stream) {
  if (!c) NA_FAIL("null context");
  if (!c->has_he) NA_FAIL("context was created without Delta (polyctx only)");
  if (l > c->L) NA_FAIL("level out of range");
  if (!rhat || !out || npolys == 0 || npolys > 4) NA_FAIL("bad argument");
  for (unsigned k = 0; k < npolys; k++) if (!out[k]) NA_FAIL("null output polynomial");
  if (ncoef == 0 || ncoef > c->n) NA_FAIL("coefficient count out of range");
  if ((unsigned long long)npolys * ncoef > GQ_NOALLOC_MAX_ENTRIES)
    NA_FAIL("npolys * ncoef exceeds the static scratch (GQ_NOALLOC_MAX_ENTRIES = 262144 coefficients per call)");
  if (D == 0 || D > c->dimub) NA_FAIL("basis size out of range");
  if (c->basis[D - 1].dev.WT == 0) NA_FAIL("RNS basis wider than 3072 bits is not supported by the CRT kernels");
  cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
  int *fix = nullptr;
  if (c->fast_crt) {
    if (noalloc_scratch(c->device, &fix)) return 1;
    // pending count and lifetime total start at zero for this call (the array keeps the last call's contents)
    NA_CUDA(cudaMemsetAsync(fix, 0, 4 * sizeof(int), st));
  }
  GqPtrs4 o;
  for (unsigned k = 0; k < 4; k++) o.p[k] = k < npolys ? out[k] : nullptr;
  const int r = gq_launch_reconstruct(c->basis[D - 1].dev, c->d_primes, (int)ncoef, rhat, c->qmod[l].dev, o, (int)c->W,
                                      (int)npolys, st, fix, 0, false);
  if (r < 0) NA_FAIL("no reconstruct kernel for this width");
  c->launches += (unsigned)r;
  NA_CUDA(cudaGetLastError());
  return 0;
}
