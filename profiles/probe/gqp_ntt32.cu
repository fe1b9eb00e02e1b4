// gqp_ntt32.cu - MEASUREMENT ONLY (profiles/probe/libgqprobe.so, loaded by bench.py; not part of the product library): the forward negacyclic NTT over 30-bit primes on 32-bit words.
//
// Not on any product path.  DESIGN.md section 4 ("what ncu says is left") plans a 31-bit-class internal RNS
// basis for the polynomial products: only exact integers cross the stage boundaries of he_mul / he_relin, so
// the residues in between may live in any basis whose product is large enough, and a 32-bit Shoup butterfly
// is 3 multiply-pipe instructions against 10.6 for the 64-bit one.  This file measures what that buys on this
// GPU before the pipelines are rebuilt around it: gq_bench_ntt32 runs the same two-kernel structure as
// gq_ntt_lazy.cu (4 strided register stages + 4096-point shared-memory tiles, radix-8 rounds) on uint32 data,
// checks one transform against a plain host implementation, and reports limb-NTTs per second.
//
// Primes: the largest p < 2^30 with p == 1 (mod 2n) (4p < 2^32: Harvey butterflies with values in [0,4p)).
// Twiddles: psi^brv(k) in the reference's order (src/precomp.c:255-263), plain form + Shoup companion
// floor(w * 2^32 / p).  Butterfly: x <- x mod 2p; t = w*y - hi32(w'*y)*p in [0,2p); (x, y) <- (x + t, x - t + 2p).
#include <vector>

#include "gqp_common.h"

namespace {

GQ_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

GQ_HD void bf32(uint32_t &x, uint32_t &y, uint32_t w, uint32_t wsh, uint32_t p, uint32_t p2) {
  const uint32_t xr = (x >= p2) ? x - p2 : x;
  const uint32_t t = y * w - mulhi32(y, wsh) * p;
  y = xr - t + p2;
  x = xr + t;
}

// 4-byte elements, 32 banks: low 3 bits take bits 5-7, bits 3-4 take bits 6-7 (XOR-linear; every radix-8 round
// of a 4096-point tile and the row accesses touch 32 distinct banks per warp)
GQ_DEV int swz32(int i) { return i ^ ((i >> 5) & 7) ^ (((i >> 6) & 3) << 3); }
GQ_DEV int idx8(int t, int lo) { return ((t >> lo) << (lo + 3)) | (t & ((1 << lo) - 1)); }

struct alignas(8) Tw32 { uint32_t x, y; };      // {w, floor(w*2^32/p)}, one 8-byte load

struct Tab32 {
  const Tw32 *tw;        // [nlimbs][n] {w, floor(w*2^32/p)}
  const uint32_t *p;      // [nlimbs]
  int n, logn;
};

GQ_DEV void round8(uint32_t *v, int lo, int kq, const Tw32 *__restrict__ tw, uint32_t p, uint32_t p2) {
  {
    const Tw32 w = tw[kq >> (lo + 3)];
#pragma unroll
    for (int e = 0; e < 4; e++) bf32(v[e], v[e + 4], w.x, w.y, p, p2);
  }
  {
    const int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const Tw32 w = tw[kb + g];
      bf32(v[4 * g + 0], v[4 * g + 2], w.x, w.y, p, p2);
      bf32(v[4 * g + 1], v[4 * g + 3], w.x, w.y, p, p2);
    }
  }
  {
    const int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      const Tw32 w = tw[kb + g];
      bf32(v[2 * g], v[2 * g + 1], w.x, w.y, p, p2);
    }
  }
}

// last 12 stages on 4096 contiguous points; 512 threads x 8 values
__global__ void __launch_bounds__(512, 3)
k_ntt32_fwd_tile(Tab32 tb, uint32_t *__restrict__ data, int nlimbs) {
  constexpr int LOGT = 12, T = 1 << LOGT, NT = T / 8;
  __shared__ uint32_t sm[T];
  const int t = threadIdx.x, st = swz32(t);
  const int m = blockIdx.y, d = m % nlimbs;
  const uint32_t p = tb.p[d], p2 = 2 * p;
  const Tw32 *tw = tb.tw + (size_t)d * tb.n;
  uint32_t *a = data + (size_t)m * tb.n + (size_t)blockIdx.x * T;
  const int jbase = tb.n + blockIdx.x * T;
  uint32_t v[8];
#pragma unroll
  for (int e = 0; e < 8; e++) v[e] = a[e * NT + t];
  round8(v, LOGT - 3, jbase + t, tw, p, p2);
#pragma unroll
  for (int e = 0; e < 8; e++) sm[st ^ swz32(e * NT)] = v[e];
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int lo = LOGT - 3 - 3 * (r + 1);
    const int base = idx8(t, lo), sb = swz32(base);
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = sm[sb ^ swz32(e << lo)];
    round8(v, lo, jbase + base, tw, p, p2);
    if (r == 2) {
#pragma unroll
      for (int e = 0; e < 8; e++) {
        uint32_t x = v[e];
        x = (x >= p2) ? x - p2 : x;
        v[e] = (x >= p) ? x - p : x;
      }
    }
#pragma unroll
    for (int e = 0; e < 8; e++) sm[sb ^ swz32(e << lo)] = v[e];
    __syncthreads();
  }
#pragma unroll
  for (int e = 0; e < 8; e++) a[e * NT + t] = sm[st ^ swz32(e * NT)];
}

// first SA stages: 2^SA values per thread at stride n >> SA
template <int SA>
__global__ void __launch_bounds__(256)
k_ntt32_fwd_strided(Tab32 tb, uint32_t *__restrict__ data, int nlimbs) {
  constexpr int R = 1 << SA;
  __shared__ Tw32 stw[R];
  const int cols = tb.n >> SA;
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y, d = m % nlimbs;
  if (threadIdx.x < R) stw[threadIdx.x] = tb.tw[(size_t)d * tb.n + threadIdx.x];
  __syncthreads();
  if (col >= cols) return;
  const uint32_t p = tb.p[d], p2 = 2 * p;
  uint32_t *a = data + (size_t)m * tb.n + col;
  uint32_t v[R];
#pragma unroll
  for (int r = 0; r < R; r++) v[r] = a[(size_t)r * cols];
#pragma unroll
  for (int s = 0; s < SA; s++) {
    const int mb = SA - 1 - s;
#pragma unroll
    for (int g = 0; g < (1 << s); g++) {
      const Tw32 w = stw[(1 << s) + g];
#pragma unroll
      for (int i = 0; i < (1 << mb); i++) {
        const int r = (g << (mb + 1)) | i;
        bf32(v[r], v[r | (1 << mb)], w.x, w.y, p, p2);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; r++) a[(size_t)r * cols] = v[r];
}

}  // namespace

// ------------------------------------------------------------------------------- host side
#define N32_CUDA(call)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      gqp_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

static uint32_t powm32(uint32_t b, uint64_t e, uint32_t p) {
  uint64_t r = 1, x = b % p;
  while (e) { if (e & 1) r = r * x % p; x = x * x % p; e >>= 1; }
  return (uint32_t)r;
}
static bool is_prime32(uint32_t n) {
  if (n < 2) return false;
  for (uint32_t q : {2u, 3u, 5u, 7u, 11u, 13u, 17u, 19u, 23u, 29u, 31u, 37u}) {
    if (n % q == 0) return n == q;
  }
  uint32_t d = n - 1; int s = 0;
  while (!(d & 1)) { d >>= 1; s++; }
  for (uint32_t a : {2u, 7u, 61u}) {               // deterministic Miller-Rabin below 4 759 123 141
    uint64_t x = powm32(a % n, d, n);
    if (x == 0 || x == 1 || x == n - 1) continue;
    bool comp = true;
    for (int r = 1; r < s && comp; r++) { x = x * x % n; if (x == n - 1) comp = false; }
    if (comp) return false;
  }
  return true;
}
static unsigned brv(unsigned x, int bits) {
  unsigned r = 0;
  for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
  return r;
}

// limb-NTTs per second of the forward transform on 30-bit primes, nlimbs x npolys limb-polys per launch;
// *mismatches = differences of one device transform against the host implementation below plus failed
// evaluation checks (the output at index j must be a(psi^(2*brv(j)+1)))
extern "C" GQP_EXPORT int gqp_bench_ntt32(int device, unsigned logn, unsigned nlimbs, unsigned npolys, unsigned reps,
                                          double *limb_ntt_per_s, int *mismatches) {
  cudaStream_t stream = 0;
  if (!limb_ntt_per_s || !mismatches || nlimbs == 0 || npolys == 0 || reps == 0) {
    gqp_set_error("gqp_bench_ntt32: bad argument"); return 1;
  }
  if (logn < 13 || logn > 16) { gqp_set_error("gqp_bench_ntt32: logn must be 13..16 (two-kernel transform)"); return 1; }
  N32_CUDA(cudaSetDevice(device));
  const unsigned n = 1u << logn, two_n = 2 * n;
  std::vector<uint32_t> primes;
  for (uint32_t cand = ((1u << 30) - 1) / two_n * two_n + 1; primes.size() < nlimbs && cand > two_n; cand -= two_n)
    if (cand < (1u << 30) && is_prime32(cand)) primes.push_back(cand);
  if (primes.size() < nlimbs) { gqp_set_error("gqp_bench_ntt32: not enough 30-bit primes for this ring"); return 1; }
  std::vector<Tw32> tw((size_t)nlimbs * n);
  std::vector<uint32_t> psis(nlimbs);
  for (unsigned d = 0; d < nlimbs; d++) {
    const uint32_t p = primes[d];
    uint32_t psi = 0;
    for (uint32_t x = 2; x < 1000 && !psi; x++) {
      const uint32_t y = powm32(x, (p - 1) / two_n, p);
      if (powm32(y, n, p) == p - 1) psi = y;           // order exactly 2n
    }
    if (!psi) { gqp_set_error("gqp_bench_ntt32: no primitive root found"); return 1; }
    psis[d] = psi;
    std::vector<uint32_t> pw(n);
    uint64_t v = 1;
    for (unsigned i = 0; i < n; i++) { pw[i] = (uint32_t)v; v = v * psi % p; }
    for (unsigned k = 0; k < n; k++) {
      const uint32_t w = pw[brv(k, (int)logn)];
      tw[(size_t)d * n + k] = Tw32{w, (uint32_t)(((uint64_t)w << 32) / p)};
    }
  }
  const size_t count = (size_t)nlimbs * npolys * n;
  std::vector<uint32_t> host(count);
  uint64_t s = 0x9E3779B97F4A7C15ull;
  for (size_t i = 0; i < count; i++) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    host[i] = (uint32_t)((s >> 33) % primes[(i / n) % nlimbs]);
  }
  uint32_t *d_data = nullptr, *d_p = nullptr;
  Tw32 *d_tw = nullptr;
  N32_CUDA(cudaMalloc((void **)&d_data, count * 4));
  N32_CUDA(cudaMalloc((void **)&d_p, nlimbs * 4));
  N32_CUDA(cudaMalloc((void **)&d_tw, tw.size() * sizeof(Tw32)));
  N32_CUDA(cudaMemcpy(d_data, host.data(), count * 4, cudaMemcpyHostToDevice));
  N32_CUDA(cudaMemcpy(d_p, primes.data(), nlimbs * 4, cudaMemcpyHostToDevice));
  N32_CUDA(cudaMemcpy(d_tw, tw.data(), tw.size() * sizeof(Tw32), cudaMemcpyHostToDevice));
  Tab32 tb; tb.tw = d_tw; tb.p = d_p; tb.n = (int)n; tb.logn = (int)logn;
  const int sa = (int)logn - 12, M = (int)(nlimbs * npolys);
  const dim3 gt(n >> 12, M), gs(((n >> sa) + 255) / 256, M);
  auto run = [&]() {
    switch (sa) {
      case 1: GQ_LAUNCH(k_ntt32_fwd_strided<1>, gs, dim3(256), stream, tb, d_data, (int)nlimbs); break;
      case 2: GQ_LAUNCH(k_ntt32_fwd_strided<2>, gs, dim3(256), stream, tb, d_data, (int)nlimbs); break;
      case 3: GQ_LAUNCH(k_ntt32_fwd_strided<3>, gs, dim3(256), stream, tb, d_data, (int)nlimbs); break;
      default: GQ_LAUNCH(k_ntt32_fwd_strided<4>, gs, dim3(256), stream, tb, d_data, (int)nlimbs); break;
    }
    GQ_LAUNCH(k_ntt32_fwd_tile, gt, dim3(512), stream, tb, d_data, (int)nlimbs);
  };
  run();
  N32_CUDA(cudaStreamSynchronize(stream));
  // check limb-poly 0 of the last limb against a plain host transform and against direct evaluation
  int bad = 0;
  {
    const unsigned d = nlimbs - 1;
    const uint32_t p = primes[d];
    std::vector<uint32_t> ref(host.begin() + (size_t)d * n, host.begin() + (size_t)(d + 1) * n), got(n);
    unsigned k = 1;
    for (unsigned len = n / 2; len >= 1; len /= 2)
      for (unsigned start = 0; start < n; start += 2 * len) {
        const uint64_t z = tw[(size_t)d * n + k++].x;
        for (unsigned j = start; j < start + len; j++) {
          const uint32_t t = (uint32_t)(z * ref[j + len] % p);
          ref[j + len] = (ref[j] + p - t) % p;
          ref[j] = (ref[j] + t) % p;
        }
      }
    N32_CUDA(cudaMemcpy(got.data(), d_data + (size_t)d * n, n * 4, cudaMemcpyDeviceToHost));
    for (unsigned j = 0; j < n; j++) bad += got[j] != ref[j];
    for (unsigned j : {0u, 1u, n / 3, n - 1}) {          // output j = a(psi^(2*brv(j)+1))
      const uint64_t x = powm32(psis[d], 2ull * brv(j, (int)logn) + 1, p);
      uint64_t acc = 0;
      for (unsigned i = n; i-- > 0;) acc = (acc * x + host[(size_t)d * n + i]) % p;
      bad += (uint32_t)acc != got[j];
    }
  }
  *mismatches = bad;
  cudaEvent_t e0, e1;
  N32_CUDA(cudaEventCreate(&e0));
  N32_CUDA(cudaEventCreate(&e1));
  N32_CUDA(cudaEventRecord(e0, stream));
  for (unsigned r = 0; r < reps; r++) run();      // transforms of transforms: values stay below p, timing is data independent
  N32_CUDA(cudaEventRecord(e1, stream));
  N32_CUDA(cudaEventSynchronize(e1));
  float ms = 0.f;
  N32_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *limb_ntt_per_s = (double)M * reps / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d_data); cudaFree(d_p); cudaFree(d_tw);
  return 0;
}
