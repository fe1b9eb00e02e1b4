// gqp_imma.cu - MEASUREMENT ONLY (profiles/probe/libgqprobe.so): can the dense integer contractions of the CRT stages
// run on the tensor cores?  (north_star: "Tensor cores are used only if the ncu evidence shows this step is a dense
// contraction worth the limb-splitting cost")
//
// rns_decompose (src/rns.c:37-48) is residues[i][d] = sum_j x[i][j] * (2^(8j) mod p_d) mod p_d: with the coefficient
// split into bytes x[i][j] and every table entry into 8 bytes it is an (N x 56) * (56 x 8D) u8 GEMM with exact int32
// accumulation (56 * 255 * 255 < 2^22), followed by a per-residue recombination of 8 byte columns and one modular
// fold.  This file measures
//   gqp_bench_imma_raw        the raw rate of mma.sync.m16n8k32 u8*u8+s32 on this GPU (register operands, no memory)
//   gqp_bench_imma_decompose  a complete prototype of that decomposition (W = 7 words, D limbs, every prime
//                             2^59 + c): A fragments straight from the word-major coefficients, B from shared memory,
//                             recombination + pseudo-Mersenne fold in the epilogue, checked against the plain
//                             evaluation of the same residues, timed in the shape he_mul uses (4 polys x 65536
//                             coefficients x 16 limbs).
// Fragment layouts (PTX ISA, mma.m16n8k32 with .u8): A: a0..a3 = rows (g, g+8) x k-bytes (4t..4t+3, 16+4t..16+4t+3);
// B: b0, b1 = k-bytes (4t..4t+3, 16+4t..) x column g; C: c0, c1 = row g, columns 2t, 2t+1; c2, c3 = row g+8
// (g = lane / 4, t = lane % 4).  The K and N orders are ours to choose (the table is a constant), so
//   * k-slot (4t..4t+3, 16+4t..16+4t+3) of k-chunk c carries the 8 bytes of WORD 4c + t of the coefficient: an A
//     fragment is one 64-bit load per row, coalesced over the 8 rows of a quad column;
//   * n-tile v of limb group G carries, in columns 2t and 2t+1, bytes 2v and 2v+1 of the table entries of limb
//     4G + t: after the 4 n-tiles of a group thread (g, t) holds all 8 byte columns of (row g, limb 4G + t) and of
//     (row g + 8, limb 4G + t) - the recombination is thread-local, no shuffles.
#include <vector>

#include "gqp_common.h"

namespace {

GQ_DEV void mma_u8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#if defined(__CUDA_ARCH__)
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
#elif defined(GQ_EMULATE)
  gq_emul_mma_u8(c, a, b0, b1);
#endif
}

__global__ void __launch_bounds__(256)
k_imma_raw(int *sink, int iters) {
  int acc[8][4];
  uint32_t a[4], b0 = threadIdx.x * 0x01010101u + 0x01020304u, b1 = blockIdx.x * 0x00010001u + 0x05060708u;
#pragma unroll
  for (int k = 0; k < 4; k++) a[k] = (threadIdx.x + k) * 0x9E3779B1u;
#pragma unroll
  for (int j = 0; j < 8; j++)
#pragma unroll
    for (int k = 0; k < 4; k++) acc[j][k] = j + k;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < 8; j++) mma_u8(acc[j], a, b0 + j, b1);     // 8 independent accumulators
  }
  int s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++)
#pragma unroll
    for (int k = 0; k < 4; k++) s ^= acc[j][k];
  if (s == 0x12345) sink[0] = s;
}

// ---- decomposition prototype ---------------------------------------------------------------------------
constexpr int PW = 7;            // words per coefficient
constexpr int KC = 2;            // k-chunks of 32 bytes: words 0-3, words 4-6 + one zero word
struct ImmaDec {
  const uint64_t *x;             // [4 polys][PW][n] word-major coefficients (non-negative here)
  uint64_t *out;                 // [4][D][n]
  const uint32_t *btab;          // [D/4 groups][KC][4 n-tiles][32 lanes][2] B fragments
  const uint64_t *primes;        // [D]
  int n, D;
};

// V = sum_b acc[b] * 2^(8b), acc[b] < 2^22: 86 bits in (lo, hi)
GQ_DEV void recombine(const int *acc, uint64_t &lo, uint64_t &hi) {
  lo = 0; hi = 0;
#pragma unroll
  for (int b = 0; b < 8; b++) {
    const uint64_t v = (uint64_t)(uint32_t)acc[b];
    const uint64_t add = v << (8 * b);
    lo += add;
    hi += (b >= 6 ? (v >> (64 - 8 * b)) : 0) + (lo < add ? 1u : 0u);
  }
}
// (hi, lo) < 2^86 -> canonical residue mod p = 2^59 + c, c < 2^28 (the fold of k_decompose28, gq_crt.cu)
GQ_DEV uint64_t fold59(uint64_t lo, uint64_t hi, uint64_t p) {
  const uint64_t mask = (((uint64_t)1) << 59) - 1;
  const uint32_t cc = (uint32_t)(p - (mask + 1));
  const uint64_t Vh = (lo >> 59) | (hi << 5);                 // < 2^27
  const uint64_t t = Vh * cc;                                  // < 2^55
  uint64_t r = (lo & mask) + p - t;                            // in (0, 2p)
  r = (r >= p) ? r - p : r;
  return r;
}

__global__ void __launch_bounds__(128)
k_imma_decompose(ImmaDec a) {
  GQ_DYN_SMEM(uint32_t, sB);                                   // the whole B table of this launch
  const int nb = (a.D / 4) * KC * 4 * 64;
  for (int j = threadIdx.x; j < nb; j += blockDim.x) sB[j] = a.btab[j];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int poly = blockIdx.y;
  const uint64_t *x = a.x + (size_t)poly * PW * a.n;
  uint64_t *out = a.out + (size_t)poly * a.D * a.n;
  // one warp = 16 coefficients per step
  for (int i0 = (blockIdx.x * 4 + warp) * 16; i0 < a.n; i0 += gridDim.x * 64) {
    uint32_t A[KC][4];
#pragma unroll
    for (int c = 0; c < KC; c++) {
      const int w = 4 * c + t;
      const uint64_t v0 = (w < PW) ? x[(size_t)w * a.n + i0 + g] : 0;
      const uint64_t v1 = (w < PW) ? x[(size_t)w * a.n + i0 + g + 8] : 0;
      A[c][0] = (uint32_t)v0; A[c][2] = (uint32_t)(v0 >> 32);
      A[c][1] = (uint32_t)v1; A[c][3] = (uint32_t)(v1 >> 32);
    }
    for (int G = 0; G < a.D / 4; G++) {
      int acc[4][4];
#pragma unroll
      for (int v = 0; v < 4; v++)
#pragma unroll
        for (int k = 0; k < 4; k++) acc[v][k] = 0;
#pragma unroll
      for (int c = 0; c < KC; c++)
#pragma unroll
        for (int v = 0; v < 4; v++) {
          const uint64_t b = *reinterpret_cast<const uint64_t *>(&sB[(((G * KC + c) * 4 + v) * 32 + lane) * 2]);
          mma_u8(acc[v], A[c], (uint32_t)b, (uint32_t)(b >> 32));
        }
      const uint64_t p = a.primes[4 * G + t];
      int r0[8], r1[8];
#pragma unroll
      for (int v = 0; v < 4; v++) { r0[2 * v] = acc[v][0]; r0[2 * v + 1] = acc[v][1]; r1[2 * v] = acc[v][2]; r1[2 * v + 1] = acc[v][3]; }
      uint64_t lo, hi;
      recombine(r0, lo, hi);
      out[(size_t)(4 * G + t) * a.n + i0 + g] = fold59(lo, hi, p);
      recombine(r1, lo, hi);
      out[(size_t)(4 * G + t) * a.n + i0 + g + 8] = fold59(lo, hi, p);
    }
  }
}

}  // namespace

#define IM_CUDA(call)                                                                    \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      gqp_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

// u8 multiply-accumulates per second (one m16n8k32 = 4096 of them) with operands in registers
extern "C" GQP_EXPORT int gqp_bench_imma_raw(int device, double *mac_per_s) {
  if (!mac_per_s) { gqp_set_error("gqp_bench_imma_raw: null argument"); return 1; }
  IM_CUDA(cudaSetDevice(device));
  int *sink = nullptr;
  IM_CUDA(cudaMalloc((void **)&sink, 8));
  cudaStream_t stream = 0;
  cudaEvent_t e0, e1;
  IM_CUDA(cudaEventCreate(&e0));
  IM_CUDA(cudaEventCreate(&e1));
  int iters = 4000;
#ifdef GQ_EMULATE
  iters = 1;
#endif
  const int blocks = GQ_NUM_SMS * 8, threads = 256;
  float ms = 0.f;
  for (int rep = 0; rep < 2; rep++) {
    IM_CUDA(cudaEventRecord(e0, stream));
    GQ_LAUNCH(k_imma_raw, dim3(blocks), dim3(threads), stream, sink, iters);
    IM_CUDA(cudaEventRecord(e1, stream));
    IM_CUDA(cudaEventSynchronize(e1));
    IM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  }
  *mac_per_s = (double)blocks * (threads / 32) * 8.0 * iters * 4096.0 / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  return 0;
}

// Decomposition of 4 polys x n coefficients (7 words, non-negative) into D residues (D a multiple of 4, primes
// 2^59 + 1 + k*2n as the reference finds them) on the tensor cores: residues per second and the number of
// mismatches against the plain host evaluation of a sample of coefficients.
extern "C" GQP_EXPORT int gqp_bench_imma_decompose(int device, unsigned logn, unsigned D, const uint64_t *primes,
                                                   unsigned reps, double *residues_per_s, double *ms_per_launch,
                                                   int *mismatches) {
  if (!primes || !residues_per_s || !mismatches || !ms_per_launch || D == 0 || D % 4 || D > 32 || logn < 4 || logn > 17 || reps == 0) {
    gqp_set_error("gqp_bench_imma_decompose: bad argument"); return 1;
  }
  IM_CUDA(cudaSetDevice(device));
  const unsigned n = 1u << logn;
  for (unsigned d = 0; d < D; d++)
    if ((primes[d] >> 59) != 1 || (primes[d] & ((1ull << 59) - 1)) >= (1ull << 28)) {
      gqp_set_error("gqp_bench_imma_decompose: primes must be 2^59 + c, c < 2^28"); return 1;
    }
  // B fragments: entry (limb, k-byte) = byte `e` of U = 2^(8*(8w+a)) mod p with k-byte = (word w, byte a)
  std::vector<uint32_t> bt((size_t)(D / 4) * KC * 4 * 64, 0);
  for (unsigned G = 0; G < D / 4; G++)
    for (int c = 0; c < KC; c++)
      for (int v = 0; v < 4; v++)
        for (int lane = 0; lane < 32; lane++) {
          const int g = lane >> 2, t = lane & 3;            // B fragment: column g, k-bytes 4t.. and 16+4t..
          // column g of n-tile v: thread t' = g / 2 owns columns 2t', 2t'+1 -> limb 4G + g/2, byte 2v + g%2
          const unsigned limb = 4 * G + g / 2, ebyte = 2 * v + g % 2;
          const uint64_t p = primes[limb];
          uint32_t reg[2] = {0, 0};
          for (int half = 0; half < 2; half++)
            for (int kb = 0; kb < 4; kb++) {
              // k-slot 4t+kb (half 0) / 16+4t+kb (half 1) = byte kb / 4+kb of word 4c + t
              const int w = 4 * c + t, a = 4 * half + kb;
              uint64_t U = 0;
              if (w < PW) {
                unsigned __int128 pw = 1;                   // 2^(8*(8w+a)) mod p
                for (int s = 0; s < 8 * w + a; s++) pw = (pw << 8) % p;
                U = (uint64_t)pw;
              }
              reg[half] |= (uint32_t)((U >> (8 * ebyte)) & 0xff) << (8 * kb);
            }
          bt[((((size_t)G * KC + c) * 4 + v) * 32 + lane) * 2 + 0] = reg[0];
          bt[((((size_t)G * KC + c) * 4 + v) * 32 + lane) * 2 + 1] = reg[1];
        }
  std::vector<uint64_t> hx((size_t)4 * PW * n);
  uint64_t s = 0x9E3779B97F4A7C15ull;
  for (auto &v : hx) { s = s * 6364136223846793005ull + 1442695040888963407ull; v = s ^ (s >> 29); }
  for (unsigned poly = 0; poly < 4; poly++)                 // 439-bit magnitudes: top word below 2^55
    for (unsigned i = 0; i < n; i++) hx[((size_t)poly * PW + PW - 1) * n + i] &= (1ull << 55) - 1;
  uint64_t *dx = nullptr, *dout = nullptr, *dp = nullptr;
  uint32_t *db = nullptr;
  IM_CUDA(cudaMalloc((void **)&dx, hx.size() * 8));
  IM_CUDA(cudaMalloc((void **)&dout, (size_t)4 * D * n * 8));
  IM_CUDA(cudaMalloc((void **)&dp, D * 8));
  IM_CUDA(cudaMalloc((void **)&db, bt.size() * 4));
  IM_CUDA(cudaMemcpy(dx, hx.data(), hx.size() * 8, cudaMemcpyHostToDevice));
  IM_CUDA(cudaMemcpy(dp, primes, D * 8, cudaMemcpyHostToDevice));
  IM_CUDA(cudaMemcpy(db, bt.data(), bt.size() * 4, cudaMemcpyHostToDevice));
  ImmaDec a; a.x = dx; a.out = dout; a.btab = db; a.primes = dp; a.n = (int)n; a.D = (int)D;
  const size_t smem = bt.size() * 4;
  cudaStream_t stream = 0;
  int gx = (int)((n + 63) / 64);
  if (gx > GQ_NUM_SMS * 4) gx = GQ_NUM_SMS * 4;
  const dim3 grid(gx, 4), block(128);
#ifndef GQ_EMULATE
  IM_CUDA(cudaFuncSetAttribute(k_imma_decompose, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_imma_decompose<<<grid, block, smem, stream>>>(a);
#else
  gq_emul::launch_dyn(grid, block, smem, [&]() { k_imma_decompose(a); });
#endif
  IM_CUDA(cudaStreamSynchronize(stream));
  IM_CUDA(cudaGetLastError());
  std::vector<uint64_t> ho((size_t)4 * D * n);
  IM_CUDA(cudaMemcpy(ho.data(), dout, ho.size() * 8, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (unsigned poly = 0; poly < 4; poly++)
    for (unsigned i = 0; i < n; i += (n > 4096 ? 97 : 1))
      for (unsigned d = 0; d < D; d++) {
        unsigned __int128 r = 0;
        for (int w = PW - 1; w >= 0; w--) r = ((r << 64) + hx[((size_t)poly * PW + w) * n + i]) % primes[d];
        bad += (uint64_t)r != ho[((size_t)poly * D + d) * n + i];
      }
  *mismatches = bad;
  cudaEvent_t e0, e1;
  IM_CUDA(cudaEventCreate(&e0));
  IM_CUDA(cudaEventCreate(&e1));
  float ms = 0.f;
  IM_CUDA(cudaEventRecord(e0, stream));
  for (unsigned r = 0; r < reps; r++) {
#ifndef GQ_EMULATE
    k_imma_decompose<<<grid, block, smem, stream>>>(a);
#else
    gq_emul::launch_dyn(grid, block, smem, [&]() { k_imma_decompose(a); });
#endif
  }
  IM_CUDA(cudaEventRecord(e1, stream));
  IM_CUDA(cudaEventSynchronize(e1));
  IM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *ms_per_launch = ms / reps;
  *residues_per_s = 4.0 * n * D * reps / ((ms > 0.f ? ms : 1e-3) * 1e-3);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(dx); cudaFree(dout); cudaFree(dp); cudaFree(db);
  return 0;
}
