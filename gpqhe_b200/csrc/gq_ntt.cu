// gq_ntt.cu - K1/K2: batched negacyclic NTT / inverse NTT, bit-exact with
// ntt()/invntt() of src/ntt.c:37-73 for inputs in [0,p].
//
// Structure (N = 2^logn, forward = Cooley-Tukey, natural in -> bit-reversed out):
//   logn <= 12 : one "tile" kernel, the whole limb-poly lives in one CTA's shared memory.
//   logn  > 12 : a register-only "strided" kernel does the first SA = logn-12 stages
//                (each thread owns the 2^SA elements of one column, loads coalesced across
//                columns, twiddles warp-uniform), then the tile kernel finishes the
//                remaining 12 stages on contiguous 4096-element tiles.
//   The inverse runs the same two kernels in the opposite order (Gentleman-Sande) and
//   folds the final N^-1 scaling into its last stage.
// Tile kernel: 2^LOGT elements, 2^LOGT/8 threads, 8 elements per thread, radix-8
// rounds in registers, one shared-memory exchange per round with an XOR swizzle
// (i ^ ((i>>4)&7)) that keeps both the strided and the contiguous-8 access patterns
// conflict-free for 64-bit words.
// Twiddle index of the butterfly that touches global index j at the stage whose
// partner distance is 2^b:   k = (N + j) >> (b+1)   (same for zetas and zetas_inv).
#include "gq_kernels.h"

namespace {

// shared-memory index swizzle (8-byte elements, 16 per 128-byte wavefront): the low 3 bits take bits 4-6 and
// bit 3 takes bit 7, so that the warp access patterns of all three radix-8 rounds (element strides 1, 8
// and 64) touch every 8-byte bank exactly twice; without the bit-3 term the stride-8 round was 4-way
// (ncu: 0.54 M conflicts on 2.1 M wavefronts)
GQ_DEV int swz(int i) { return i ^ ((i >> 4) & 7) ^ (((i >> 7) & 1) << 3); }
GQ_DEV int idx8(int t, int lo) { return ((t >> lo) << (lo + 3)) | (t & ((1 << lo) - 1)); }

// ---- radix-8 register rounds -------------------------------------------------------
// The thread's 8 values sit at local indices base | (e << lo), e = 0..7.  kq = jbase+base
// where jbase = N + (global index of local index 0).
// forward: stages on bits lo+2 .. lo+3-NB (descending)
template <int NB>
GQ_DEV void fwd_round(uint64_t *v, int lo, int kq, const ulonglong2 *__restrict__ tw, uint64_t p) {
  {
    ulonglong2 w = __ldg(&tw[kq >> (lo + 3)]);
#pragma unroll
    for (int e = 0; e < 4; e++) gq_bfly_fwd(v[e], v[e + 4], w.x, w.y, p);
  }
  if (NB >= 2) {
    int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      ulonglong2 w = __ldg(&tw[kb + g]);
      gq_bfly_fwd(v[4 * g + 0], v[4 * g + 2], w.x, w.y, p);
      gq_bfly_fwd(v[4 * g + 1], v[4 * g + 3], w.x, w.y, p);
    }
  }
  if (NB >= 3) {
    int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      ulonglong2 w = __ldg(&tw[kb + g]);
      gq_bfly_fwd(v[2 * g], v[2 * g + 1], w.x, w.y, p);
    }
  }
}

// inverse: stages on bits lo+3-NB .. lo+2 (ascending).  If LAST, the stage on bit lo+2 is
// the transform's final stage (len = N/2) and carries the folded N^-1 scaling.
template <int NB, bool LAST>
GQ_DEV void inv_round(uint64_t *v, int lo, int kq, const ulonglong2 *__restrict__ tw, uint64_t p,
                      ulonglong2 wlast, uint64_t ninv, uint64_t ninv_sh) {
  if (NB >= 3) {
    int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      ulonglong2 w = __ldg(&tw[kb + g]);
      gq_bfly_inv(v[2 * g], v[2 * g + 1], w.x, w.y, p);
    }
  }
  if (NB >= 2) {
    int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      ulonglong2 w = __ldg(&tw[kb + g]);
      gq_bfly_inv(v[4 * g + 0], v[4 * g + 2], w.x, w.y, p);
      gq_bfly_inv(v[4 * g + 1], v[4 * g + 3], w.x, w.y, p);
    }
  }
  if (LAST) {
#pragma unroll
    for (int e = 0; e < 4; e++) {
      gq_bfly_inv(v[e], v[e + 4], wlast.x, wlast.y, p);
      v[e] = gq_mul_shoup(v[e], ninv, ninv_sh, p);
    }
  } else {
    ulonglong2 w = __ldg(&tw[kq >> (lo + 3)]);
#pragma unroll
    for (int e = 0; e < 4; e++) gq_bfly_inv(v[e], v[e + 4], w.x, w.y, p);
  }
}

template <int LOGT>
struct TileCfg {
  static constexpr int T = 1 << LOGT;
  static constexpr int NT = T / 8;
  static constexpr int NB0 = (LOGT % 3 == 0) ? 3 : (LOGT % 3);  // stages of the partial (top) round
  static constexpr int NFULL = (LOGT - NB0) / 3;                 // full rounds below it
};

// ---- forward tile kernel -----------------------------------------------------------
template <int LOGT>
__global__ void __launch_bounds__(TileCfg<LOGT>::NT)
k_ntt_fwd_tile(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  typedef TileCfg<LOGT> C;
  __shared__ uint64_t sm[C::T];
  const int t = threadIdx.x;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  const uint64_t p = tb.primes[d].p;
  const ulonglong2 *tw = tb.tw_fwd + (size_t)d * tb.n;
  uint64_t *a = data + (size_t)m * tb.n + (size_t)blockIdx.x * C::T;
  const int jbase = tb.n + blockIdx.x * C::T;
  uint64_t v[8];

  // top (possibly partial) round straight from global memory: idx = e*NT + t
#pragma unroll
  for (int e = 0; e < 8; e++) v[e] = a[e * C::NT + t];
  fwd_round<C::NB0>(v, LOGT - 3, jbase + t, tw, p);
  if (C::NFULL == 0) {
#pragma unroll
    for (int e = 0; e < 8; e++) a[e * C::NT + t] = v[e];
    return;
  }
#pragma unroll
  for (int e = 0; e < 8; e++) sm[swz(t) ^ swz(e * C::NT)] = v[e];
  __syncthreads();
#pragma unroll
  for (int r = 0; r < C::NFULL; r++) {
    const int lo = LOGT - C::NB0 - 3 * (r + 1);
    const int base = idx8(t, lo);
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = sm[swz(base) ^ swz(e << lo)];
    fwd_round<3>(v, lo, jbase + base, tw, p);
#pragma unroll
    for (int e = 0; e < 8; e++) sm[swz(base) ^ swz(e << lo)] = v[e];
    __syncthreads();
  }
#pragma unroll
  for (int e = 0; e < 8; e++) a[e * C::NT + t] = sm[swz(t) ^ swz(e * C::NT)];
}

// ---- inverse tile kernel -----------------------------------------------------------
// WHOLE: the tile is the whole transform (logn == LOGT), so the top stage is the final one.
template <int LOGT, bool WHOLE>
__global__ void __launch_bounds__(TileCfg<LOGT>::NT)
k_ntt_inv_tile(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  typedef TileCfg<LOGT> C;
  __shared__ uint64_t sm[C::T];
  const int t = threadIdx.x;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  const GqPrime pr = tb.primes[d];
  const uint64_t p = pr.p;
  const ulonglong2 *tw = tb.tw_inv + (size_t)d * tb.n;
  const ulonglong2 wlast = tb.tw_inv_last[d];
  uint64_t *a = data + (size_t)m * tb.n + (size_t)blockIdx.x * C::T;
  const int jbase = tb.n + blockIdx.x * C::T;
  uint64_t v[8];

  if (C::NFULL > 0) {
#pragma unroll
    for (int e = 0; e < 8; e++) sm[swz(t) ^ swz(e * C::NT)] = a[e * C::NT + t];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < C::NFULL; r++) {
      const int lo = 3 * r;
      const int base = idx8(t, lo);
#pragma unroll
      for (int e = 0; e < 8; e++) v[e] = sm[swz(base) ^ swz(e << lo)];
      inv_round<3, (WHOLE && C::NB0 == 0)>(v, lo, jbase + base, tw, p, wlast, pr.ninv, pr.ninv_sh);
#pragma unroll
      for (int e = 0; e < 8; e++) sm[swz(base) ^ swz(e << lo)] = v[e];
      __syncthreads();
    }
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = sm[swz(t) ^ swz(e * C::NT)];
  } else {
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = a[e * C::NT + t];
  }
  inv_round<C::NB0, WHOLE>(v, LOGT - 3, jbase + t, tw, p, wlast, pr.ninv, pr.ninv_sh);
#pragma unroll
  for (int e = 0; e < 8; e++) a[e * C::NT + t] = v[e];
}

// ---- strided register kernels (first SA forward stages / last SA inverse stages) ----
template <int SA>
__global__ void __launch_bounds__(256)
k_ntt_fwd_strided(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  constexpr int R = 1 << SA;
  const int cols = tb.n >> SA;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  const uint64_t p = tb.primes[d].p;
  const ulonglong2 *tw = tb.tw_fwd + (size_t)d * tb.n;
  uint64_t *a = data + (size_t)m * tb.n + c;
  uint64_t v[R];
#pragma unroll
  for (int r = 0; r < R; r++) v[r] = a[(size_t)r * cols];
#pragma unroll
  for (int s = 0; s < SA; s++) {
    const int mb = SA - 1 - s;
#pragma unroll
    for (int g = 0; g < (1 << s); g++) {
      ulonglong2 w = __ldg(&tw[(1 << s) + g]);
#pragma unroll
      for (int i = 0; i < (1 << mb); i++) {
        const int r = (g << (mb + 1)) | i;
        gq_bfly_fwd(v[r], v[r | (1 << mb)], w.x, w.y, p);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; r++) a[(size_t)r * cols] = v[r];
}

template <int SA>
__global__ void __launch_bounds__(256)
k_ntt_inv_strided(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  constexpr int R = 1 << SA;
  const int cols = tb.n >> SA;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  const GqPrime pr = tb.primes[d];
  const uint64_t p = pr.p;
  const ulonglong2 *tw = tb.tw_inv + (size_t)d * tb.n;
  const ulonglong2 wlast = tb.tw_inv_last[d];
  uint64_t *a = data + (size_t)m * tb.n + c;
  uint64_t v[R];
#pragma unroll
  for (int r = 0; r < R; r++) v[r] = a[(size_t)r * cols];
#pragma unroll
  for (int s = SA - 1; s >= 1; s--) {
    const int mb = SA - 1 - s;
#pragma unroll
    for (int g = 0; g < (1 << s); g++) {
      ulonglong2 w = __ldg(&tw[(1 << s) + g]);
#pragma unroll
      for (int i = 0; i < (1 << mb); i++) {
        const int r = (g << (mb + 1)) | i;
        gq_bfly_inv(v[r], v[r | (1 << mb)], w.x, w.y, p);
      }
    }
  }
  // final stage (len = N/2) with the N^-1 scaling folded in
#pragma unroll
  for (int r = 0; r < R / 2; r++) {
    gq_bfly_inv(v[r], v[r + R / 2], wlast.x, wlast.y, p);
    v[r] = gq_mul_shoup(v[r], pr.ninv, pr.ninv_sh, p);
  }
#pragma unroll
  for (int r = 0; r < R; r++) a[(size_t)r * cols] = v[r];
}

template <int LOGT>
void launch_fwd_tile(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                     cudaStream_t st) {
  dim3 grid(tb.n >> LOGT, nlimbs * npolys);
  GQ_LAUNCH(k_ntt_fwd_tile<LOGT>, grid, dim3(TileCfg<LOGT>::NT), st, tb, data, limb0, nlimbs);
}
template <int LOGT, bool WHOLE>
void launch_inv_tile(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                     cudaStream_t st) {
  dim3 grid(tb.n >> LOGT, nlimbs * npolys);
  GQ_LAUNCH((k_ntt_inv_tile<LOGT, WHOLE>), grid, dim3(TileCfg<LOGT>::NT), st, tb, data, limb0, nlimbs);
}
template <int SA>
void launch_fwd_strided(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                        cudaStream_t st) {
  int cols = tb.n >> SA;
  dim3 grid((cols + 255) / 256, nlimbs * npolys);
  GQ_LAUNCH(k_ntt_fwd_strided<SA>, grid, dim3(256), st, tb, data, limb0, nlimbs);
}
template <int SA>
void launch_inv_strided(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                        cudaStream_t st) {
  int cols = tb.n >> SA;
  dim3 grid((cols + 255) / 256, nlimbs * npolys);
  GQ_LAUNCH(k_ntt_inv_strided<SA>, grid, dim3(256), st, tb, data, limb0, nlimbs);
}

}  // namespace

int gq_launch_ntt(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                  bool inverse, cudaStream_t st, int only) {
  // only: 0 = whole transform; 1 = tile kernel only, 2 = strided kernel only (profiling)
  const int logn = tb.logn;
  if (logn < 3 || logn > 17) return -1;
  int launches = 0;
  if (only && logn > 12) {
    if (only == 1) {
      if (inverse) launch_inv_tile<12, false>(tb, data, limb0, nlimbs, npolys, st);
      else launch_fwd_tile<12>(tb, data, limb0, nlimbs, npolys, st);
    } else if (logn == 16) {
      if (inverse) launch_inv_strided<4>(tb, data, limb0, nlimbs, npolys, st);
      else launch_fwd_strided<4>(tb, data, limb0, nlimbs, npolys, st);
    } else if (logn == 14) {
      if (inverse) launch_inv_strided<2>(tb, data, limb0, nlimbs, npolys, st);
      else launch_fwd_strided<2>(tb, data, limb0, nlimbs, npolys, st);
    } else {
      return -1;
    }
    return 1;
  }
  if (!inverse) {
    if (logn > 12) {
      switch (logn - 12) {
        case 1: launch_fwd_strided<1>(tb, data, limb0, nlimbs, npolys, st); break;
        case 2: launch_fwd_strided<2>(tb, data, limb0, nlimbs, npolys, st); break;
        case 3: launch_fwd_strided<3>(tb, data, limb0, nlimbs, npolys, st); break;
        case 4: launch_fwd_strided<4>(tb, data, limb0, nlimbs, npolys, st); break;
        case 5: launch_fwd_strided<5>(tb, data, limb0, nlimbs, npolys, st); break;
      }
      launches++;
    }
    switch (logn > 12 ? 12 : logn) {
#define GQ_CASE(L) case L: launch_fwd_tile<L>(tb, data, limb0, nlimbs, npolys, st); break;
      GQ_CASE(3) GQ_CASE(4) GQ_CASE(5) GQ_CASE(6) GQ_CASE(7) GQ_CASE(8) GQ_CASE(9)
      GQ_CASE(10) GQ_CASE(11) GQ_CASE(12)
#undef GQ_CASE
    }
    launches++;
  } else {
    if (logn > 12) {
      launch_inv_tile<12, false>(tb, data, limb0, nlimbs, npolys, st);
      switch (logn - 12) {
        case 1: launch_inv_strided<1>(tb, data, limb0, nlimbs, npolys, st); break;
        case 2: launch_inv_strided<2>(tb, data, limb0, nlimbs, npolys, st); break;
        case 3: launch_inv_strided<3>(tb, data, limb0, nlimbs, npolys, st); break;
        case 4: launch_inv_strided<4>(tb, data, limb0, nlimbs, npolys, st); break;
        case 5: launch_inv_strided<5>(tb, data, limb0, nlimbs, npolys, st); break;
      }
      launches += 2;
    } else {
      switch (logn) {
#define GQ_CASE(L) case L: launch_inv_tile<L, true>(tb, data, limb0, nlimbs, npolys, st); break;
        GQ_CASE(3) GQ_CASE(4) GQ_CASE(5) GQ_CASE(6) GQ_CASE(7) GQ_CASE(8) GQ_CASE(9)
        GQ_CASE(10) GQ_CASE(11) GQ_CASE(12)
#undef GQ_CASE
      }
      launches++;
    }
  }
  return launches;
}
