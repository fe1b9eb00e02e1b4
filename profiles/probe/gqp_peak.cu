// gqp_peak.cu - MEASUREMENT ONLY (profiles/probe/libgqprobe.so, loaded by bench.py; not part of the product library).
// Measured integer-pipe ceilings for the roofline of the NTT / CRT kernels (SURVEY 8d:
// "measure the IMAD peak on the box with a dependent-free IMAD microbenchmark").
//   k_peak_butterfly : the lazy Harvey butterfly of gq_ntt_lazy.cu (one Shoup multiplication by a
//                      fixed twiddle + add / sub), operands in registers, 4 independent chains per
//                      thread, no memory traffic: the speed of light of a butterfly on the SM.
//   k_peak_madwide   : dependent-free 32x32+64 multiply-adds (IMAD.WIDE.U32), the raw multiplier rate.
// Both report operations per second over the whole GPU; bench.py divides what the NTT kernels achieve
// by the first one.
#include "gqp_common.h"

namespace {

__global__ void __launch_bounds__(256)
k_peak_butterfly(uint64_t *sink, uint64_t p, uint64_t w, uint64_t wsh, int iters) {
  const uint64_t p2 = 2 * p;
  uint64_t x[4], y[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    x[k] = (uint64_t)(threadIdx.x + 1) * 0x9E3779B97F4A7C15ull % p;
    y[k] = (uint64_t)(blockIdx.x + k + 1) * 0xD1B54A32D192ED03ull % p;
  }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      // forward lazy butterfly: t = y*w in [0,2p); x' = x + t; y' = x - t + 2p; then the conditional
      // subtraction the radix-8 round amortises over three stages is done every time here
      const uint64_t q = gq_mulhi(y[k], wsh);
      const uint64_t t = y[k] * w - q * p;
      uint64_t a = x[k] + t, b = x[k] - t + p2;
      x[k] = (a >= p2) ? a - p2 : a;
      y[k] = (b >= p2) ? b - p2 : b;
    }
  }
  uint64_t acc = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) acc ^= x[k] ^ y[k];
  if (acc == 0x1234567) sink[0] = acc;   // keeps the loop alive, practically never taken
}

// the same butterfly with an APPROXIMATE quotient: hi64(y * wsh) without the lo x lo partial product and without the
// carry of the two cross terms (3 IMAD.WIDE instead of 4 and no carry chain): q' in {q-2, q-1, q}, so t = y*w - q'*p lies
// in [0,4p) instead of [0,2p) - still y*w modulo p.  Measures what that buys on the multiply pipe.
__device__ __forceinline__ uint64_t mulhi_approx(uint64_t a, uint64_t b) {
  const uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
  const uint64_t m1 = (uint64_t)a1 * b0, m2 = (uint64_t)a0 * b1;
  return gq_mad_wide(a1, b1, (m1 >> 32) + (m2 >> 32));
}
__global__ void __launch_bounds__(256)
k_peak_butterfly_approx(uint64_t *sink, uint64_t p, uint64_t w, uint64_t wsh, int iters) {
  const uint64_t p4 = 4 * p, p16 = 16 * p;
  uint64_t x[4], y[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    x[k] = (uint64_t)(threadIdx.x + 1) * 0x9E3779B97F4A7C15ull % p;
    y[k] = (uint64_t)(blockIdx.x + k + 1) * 0xD1B54A32D192ED03ull % p;
  }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint64_t q = mulhi_approx(y[k], wsh);
      const uint64_t t = y[k] * w - q * p;                 // [0,4p)
      uint64_t a = x[k] + t, b = x[k] - t + p4;            // grow by 4p per stage
      // one conditional subtraction of 16p per value every third stage in the real kernels; here every time, as in
      // k_peak_butterfly (same ALU work, so the two rates compare the multiply pipe)
      x[k] = (a >= p16) ? a - p16 : a;
      y[k] = (b >= p16) ? b - p16 : b;
    }
  }
  uint64_t acc = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) acc ^= x[k] ^ y[k];
  if (acc == 0x1234567) sink[0] = acc;
}

__global__ void __launch_bounds__(256)
k_peak_madwide(uint64_t *sink, uint32_t m, int iters) {
  uint64_t a[8];
  uint32_t b[8];
#pragma unroll
  for (int k = 0; k < 8; k++) { a[k] = threadIdx.x + k; b[k] = (uint32_t)(blockIdx.x * 8 + k) | 1u; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = (uint64_t)((uint32_t)a[k] ^ b[k]) * m + a[k];   // IMAD.WIDE.U32 (+ one LOP3)
  }
  uint64_t acc = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) acc ^= a[k];
  if (acc == 0x1234567) sink[0] = acc;
}

}  // namespace

static double g_last_approx_rate = 0;
// butterflies per second of the approximate-quotient variant measured by the last gqp_bench_intpipe call
extern "C" GQP_EXPORT double gqp_last_butterfly_approx_rate(void) { return g_last_approx_rate; }

#define PK_CUDA(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      gqp_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

// p: a prime below 2^60, w: any residue, wsh = floor(w * 2^64 / p)
extern "C" GQP_EXPORT int gqp_bench_intpipe(int device, uint64_t p, uint64_t w, uint64_t wsh, double *butterflies_per_s,
                                            double *madwide_per_s) {
  if (!butterflies_per_s || !madwide_per_s || p < 3) { gqp_set_error("gqp_bench_intpipe: bad argument"); return 1; }
  PK_CUDA(cudaSetDevice(device));
  uint64_t *sink = nullptr;
  PK_CUDA(cudaMalloc(&sink, 8));
  cudaStream_t stream = 0;
  const int blocks = GQ_NUM_SMS * 8, threads = 256;
  cudaEvent_t e0, e1;
  PK_CUDA(cudaEventCreate(&e0));
  PK_CUDA(cudaEventCreate(&e1));
  float ms = 0.f;
  int iters = 2000;
#ifdef GQ_EMULATE
  iters = 2;
#endif
  for (int rep = 0; rep < 2; rep++) {   // first pass warms up
    PK_CUDA(cudaEventRecord(e0, stream));
    GQ_LAUNCH(k_peak_butterfly, dim3(blocks), dim3(threads), stream, sink, p, w, wsh, iters);
    PK_CUDA(cudaEventRecord(e1, stream));
    PK_CUDA(cudaEventSynchronize(e1));
    PK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  }
  *butterflies_per_s = (double)blocks * threads * 4.0 * iters / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  for (int rep = 0; rep < 2; rep++) {
    PK_CUDA(cudaEventRecord(e0, stream));
    GQ_LAUNCH(k_peak_madwide, dim3(blocks), dim3(threads), stream, sink, 0x9E3779B1u, iters * 4);
    PK_CUDA(cudaEventRecord(e1, stream));
    PK_CUDA(cudaEventSynchronize(e1));
    PK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  }
  *madwide_per_s = (double)blocks * threads * 8.0 * iters * 4 / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  for (int rep = 0; rep < 2; rep++) {
    PK_CUDA(cudaEventRecord(e0, stream));
    GQ_LAUNCH(k_peak_butterfly_approx, dim3(blocks), dim3(threads), stream, sink, p, w, wsh, iters);
    PK_CUDA(cudaEventRecord(e1, stream));
    PK_CUDA(cudaEventSynchronize(e1));
    PK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  }
  g_last_approx_rate = (double)blocks * threads * 4.0 * iters / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(sink);
  return 0;
}
