// gq_ntt_lazy.cu - K1/K2 for the INTERNAL pipelines (poly_mul, he_mul, he_mulpt, key switch):
// the same transforms as gq_ntt.cu, with lazy (Harvey) butterflies.
//
// Why two sets.  ntt() of the reference may store p instead of 0 on the sum leg
// (src/ntt.c:48); the public gq_ntt / gq_ntt_batch entry points reproduce that bit for bit
// with the exact butterflies of gq_ntt.cu.  Inside a polynomial product the forward NTT only
// feeds poly_rns_mul, whose Barrett step maps 0 and p to the same canonical residue
// (src/poly.c:77-82), and invntt() always ends canonical (src/ntt.c:71-72).  So the internal
// transforms only have to be right modulo p at their boundaries:
//   forward : canonical in  -> canonical out   (intermediate values < 14p)
//   inverse : canonical in  -> canonical out   (intermediate values < 2p, N^-1 folded in)
// which removes two of the three compare/select pairs of every butterfly (ncu of the exact
// kernels: ALU pipe 59 %, math-pipe throttle the top stall - instruction count is the bound).
//
// Ranges (p = 2^sh + c with 32c < 2^sh and p < 2^60 is checked on the host, else the exact kernels are
// used; so 30p < 2^64):
//   forward CT butterfly: t = y*w - hi(y*wsh)*p in [0,2p) for ANY y < 2^64;
//     x' = x + t, y' = x - t + 2p: every value grows by 2p per stage and nothing is reduced on the way:
//     canonical input, logn stages -> (1 + 2 logn) p, at most 25p up to n = 2^12.  The two-kernel
//     transforms (n >= 2^13) take ONE conditional subtraction of 16p per value when the last radix-8
//     round loads its inputs (< 29p -> < 16p, + 6p = 22p at the end); the last round ends with an exact
//     reduction to [0,p) (full_reduce: v < 27p).
//   inverse GS butterfly: s = x + y, d = x - y + B*p (B*p = bound of y), y' = lazy d*w in [0,2p):
//     differences come back below 2p, sums double their bound per stage (2p, 4p, 8p, 16p < 2^64), so a
//     radix-8 round with inputs < 2p ends with bounds {16,8,4,4,2,2,2,2} p and only positions 0..3 are
//     reduced (7 conditional subtractions per round instead of one per butterfly = 12).
#include "gq_kernels.h"
#include "gq_p2p.h"

namespace {

// shared-memory index swizzle (8-byte elements, 16 per 128-byte wavefront): the low 3 bits take bits 4-6 and
// bit 3 takes bit 7, so that the warp access patterns of all three radix-8 rounds (element strides 1, 8
// and 64) touch every 8-byte bank exactly twice; without the bit-3 term the stride-8 round was 4-way
// (ncu: 0.54 M conflicts on 2.1 M wavefronts)
GQ_DEV int swz(int i) { return i ^ ((i >> 4) & 7) ^ (((i >> 7) & 1) << 3); }
GQ_DEV int idx8(int t, int lo) { return ((t >> lo) << (lo + 3)) | (t & ((1 << lo) - 1)); }

GQ_DEV uint64_t csub(uint64_t x, uint64_t m) { return (x >= m) ? x - m : x; }

// x*w mod p in [0,2p), any x
GQ_DEV uint64_t mul_lazy(uint64_t x, uint64_t w, uint64_t wsh, uint64_t p) {
  return x * w - gq_mulhi(x, wsh) * p;
}

GQ_DEV void bf_fwd(uint64_t &x, uint64_t &y, uint64_t w, uint64_t wsh, uint64_t p, uint64_t p2) {
  const uint64_t t = mul_lazy(y, w, wsh, p);
  y = x - t + p2;
  x = x + t;
}
GQ_DEV void bf_inv(uint64_t &x, uint64_t &y, uint64_t w, uint64_t wsh, uint64_t p, uint64_t p2) {
  const uint64_t s = csub(x + y, p2);
  const uint64_t d = x - y + p2;
  x = s;
  y = mul_lazy(d, w, wsh, p);
}
// v < 27p  ->  v mod p.  p = 2^sh + c with 32c < 2^sh, so k = v >> sh (<= 27) is floor(v/p) or one more:
// v - k*p = (v mod 2^sh) - k*c > -p.
GQ_DEV uint64_t full_reduce(uint64_t v, uint64_t p, int sh) {
  const uint64_t k = v >> sh;
  const uint64_t r = v - k * p;
  return ((int64_t)r < 0) ? r + p : r;
}

// resident CTAs the strided kernels are compiled for: 3 x 256 threads (80 registers, no spills) up to 16
// points per thread; the 32-point form (n = 2^17) needs its 200+ registers
#ifndef GQ_TILE_CTAS
#define GQ_TILE_CTAS 2
#endif
#ifndef GQ_STRIDED_CTAS
#define GQ_STRIDED_CTAS 3
#endif
#define GQ_STRIDED_MIN_CTAS (SA <= 4 ? GQ_STRIDED_CTAS : 1)
struct PC { uint64_t p, p2, p4, p8, p16; int sh; };
GQ_DEV PC load_pc(const GqPrime &P) {
  PC c; c.p = P.p; c.p2 = 2 * P.p; c.p4 = 4 * P.p; c.p8 = 8 * P.p; c.p16 = 16 * P.p; c.sh = (int)P.bits - 1; return c;
}

// lazy GS butterfly on values with the SAME known bound b*p (b = 2, 4, 8; compile-time after unrolling):
// sum < 2b*p unreduced, difference + b*p multiplied lazily back into [0,2p)
GQ_DEV uint64_t pmul(const PC &c, int b) { return b == 2 ? c.p2 : b == 4 ? c.p4 : c.p8; }
GQ_DEV void bf_inv_lz(uint64_t &x, uint64_t &y, int b, uint64_t w, uint64_t wsh, const PC &c) {
  const uint64_t d = x - y + pmul(c, b);
  x = x + y;
  y = mul_lazy(d, w, wsh, c.p);
}
// v < b*p -> v < 2p
GQ_DEV uint64_t norm2p(uint64_t v, int b, const PC &c) {
  if (b >= 16) v = csub(v, c.p8);
  if (b >= 8) v = csub(v, c.p4);
  if (b >= 4) v = csub(v, c.p2);
  return v;
}

template <int NB>
GQ_DEV void fwd_round(uint64_t *v, int lo, int kq, const ulonglong2 *__restrict__ tw, const PC &c) {
  {
    const ulonglong2 w = __ldg(&tw[kq >> (lo + 3)]);
#pragma unroll
    for (int e = 0; e < 4; e++) bf_fwd(v[e], v[e + 4], w.x, w.y, c.p, c.p2);
  }
  if (NB >= 2) {
    const int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_fwd(v[4 * g + 0], v[4 * g + 2], w.x, w.y, c.p, c.p2);
      bf_fwd(v[4 * g + 1], v[4 * g + 3], w.x, w.y, c.p, c.p2);
    }
  }
  if (NB >= 3) {
    const int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_fwd(v[2 * g], v[2 * g + 1], w.x, w.y, c.p, c.p2);
    }
  }
}

// radix-8 inverse round, inputs < 2p, outputs < 2p, sums left unreduced inside the round
GQ_DEV void inv_round8_lz(uint64_t *v, int lo, int kq, const ulonglong2 *__restrict__ tw, const PC &c) {
  {
    const int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_inv_lz(v[2 * g], v[2 * g + 1], 2, w.x, w.y, c);            // bounds -> {4,2,4,2,4,2,4,2}
    }
  }
  {
    const int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_inv_lz(v[4 * g + 0], v[4 * g + 2], 4, w.x, w.y, c);        // -> {8,4,2,2,8,4,2,2}
      bf_inv_lz(v[4 * g + 1], v[4 * g + 3], 2, w.x, w.y, c);
    }
  }
  {
    const ulonglong2 w = __ldg(&tw[kq >> (lo + 3)]);
    bf_inv_lz(v[0], v[4], 8, w.x, w.y, c);                          // -> {16,8,4,4,2,2,2,2}
    bf_inv_lz(v[1], v[5], 4, w.x, w.y, c);
    bf_inv_lz(v[2], v[6], 2, w.x, w.y, c);
    bf_inv_lz(v[3], v[7], 2, w.x, w.y, c);
  }
  v[0] = norm2p(v[0], 16, c);
  v[1] = norm2p(v[1], 8, c);
  v[2] = norm2p(v[2], 4, c);
  v[3] = norm2p(v[3], 4, c);
}

template <int NB, bool LAST>
GQ_DEV void inv_round(uint64_t *v, int lo, int kq, const ulonglong2 *__restrict__ tw, const PC &c,
                      ulonglong2 wlast, ulonglong2 fs) {
  const uint64_t ninv = fs.x, ninv_sh = fs.y;
  if (NB == 3 && !LAST) { inv_round8_lz(v, lo, kq, tw, c); return; }
  if (NB >= 3) {
    const int kb = kq >> (lo + 1);
#pragma unroll
    for (int g = 0; g < 4; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_inv(v[2 * g], v[2 * g + 1], w.x, w.y, c.p, c.p2);
    }
  }
  if (NB >= 2) {
    const int kb = kq >> (lo + 2);
#pragma unroll
    for (int g = 0; g < 2; g++) {
      const ulonglong2 w = __ldg(&tw[kb + g]);
      bf_inv(v[4 * g + 0], v[4 * g + 2], w.x, w.y, c.p, c.p2);
      bf_inv(v[4 * g + 1], v[4 * g + 3], w.x, w.y, c.p, c.p2);
    }
  }
  if (LAST) {   // final stage: N^-1 folded into both legs, canonical out
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const uint64_t s = v[e] + v[e + 4], d = v[e] - v[e + 4] + c.p2;
      v[e] = csub(mul_lazy(s, ninv, ninv_sh, c.p), c.p);
      v[e + 4] = csub(mul_lazy(d, wlast.x, wlast.y, c.p), c.p);
    }
  } else {
    const ulonglong2 w = __ldg(&tw[kq >> (lo + 3)]);
#pragma unroll
    for (int e = 0; e < 4; e++) bf_inv(v[e], v[e + 4], w.x, w.y, c.p, c.p2);
  }
}

template <int LOGT>
struct TileCfg {
  static constexpr int T = 1 << LOGT;
  static constexpr int NT = T / 8;
  static constexpr int NB0 = (LOGT % 3 == 0) ? 3 : (LOGT % 3);
  static constexpr int NFULL = (LOGT - NB0) / 3;
};

// ---- forward tile kernel: 2^LOGT contiguous elements, the last LOGT stages ----------
// PRE: inputs come from the strided kernel (up to 11p); see the range notes at the top.
template <int LOGT, bool PRE>
__global__ void __launch_bounds__(TileCfg<LOGT>::NT, (TileCfg<LOGT>::NT >= 512 ? GQ_TILE_CTAS : 1))
k_nttl_fwd_tile(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  typedef TileCfg<LOGT> C;
  __shared__ uint64_t sm[C::T];
  const int t = threadIdx.x;
  const int st = swz(t);      // swz is XOR-linear: swz(x | y) = swz(x) ^ swz(y) for disjoint bits, and the second
                              // term is a compile-time constant in every access below
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  const PC c = load_pc(tb.primes[d]);
  const ulonglong2 *tw = tb.tw_fwd + (size_t)d * tb.n;
  uint64_t *a = data + (size_t)m * tb.n + (size_t)blockIdx.x * C::T;
  const int jbase = tb.n + blockIdx.x * C::T;
  uint64_t v[8];
#pragma unroll
  for (int e = 0; e < 8; e++) v[e] = a[e * C::NT + t];
  fwd_round<C::NB0>(v, LOGT - 3, jbase + t, tw, c);
  if (C::NFULL > 0) {
#pragma unroll
    for (int e = 0; e < 8; e++) sm[st ^ swz(e * C::NT)] = v[e];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < C::NFULL; r++) {
      const int lo = LOGT - C::NB0 - 3 * (r + 1);
      const int base = idx8(t, lo), sb = swz(base);
#pragma unroll
      for (int e = 0; e < 8; e++) v[e] = sm[sb ^ swz(e << lo)];
      if (PRE && r == C::NFULL - 1) {       // two-kernel transform: < 29p here, the only reduction on the way
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = csub(v[e], c.p16);
      }
      fwd_round<3>(v, lo, jbase + base, tw, c);
      if (r == C::NFULL - 1) {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = full_reduce(v[e], c.p, c.sh);
      }
#pragma unroll
      for (int e = 0; e < 8; e++) sm[sb ^ swz(e << lo)] = v[e];
      __syncthreads();
    }
#pragma unroll
    for (int e = 0; e < 8; e++) a[e * C::NT + t] = sm[st ^ swz(e * C::NT)];
  } else {
#pragma unroll
    for (int e = 0; e < 8; e++) a[e * C::NT + t] = full_reduce(v[e], c.p, c.sh);
  }
}

// ---- inverse tile kernel: the first LOGT stages; WHOLE: also the final one ------------
// MODE selects what the kernel transforms (the pointwise stage of every pipeline sits exactly between a forward
// transform that ends on these tiles and this kernel, so it is done while the tile is loaded; K3 as separate
// launches is gone from he_mul / he_rot / he_mulpt / poly_mul):
//   GQ_INV_PLAIN   y[m] in place (x == y)
//   GQ_INV_TENSOR  x[4][D][n] = NTT(a0, a1, b0, b1) -> y[3][D][n] = a0*b0, a0*b1 + a1*b0, a1*b1  (he-mult.c:125-136)
//   GQ_INV_EVK     x[D][n], key halves e0, e1 [D][n] -> y[2][D][n] = x*e0, x*e1                   (he-mult.c:61-64)
//   GQ_INV_EVKPERM the same with x read through the slot permutation of the Galois map g (hoisted rotations)
//   GQ_INV_PT      x[3][D][n] = NTT(c0, c1, m) -> y[2][D][n] = c0*m, c1*m                          (he-mult.c:186-189)
//   GQ_INV_MUL     x[2][D][n] = NTT(a, b) -> y[1][D][n] = a*b                                       (poly.c:96-101)
//   GQ_INV_GEMV    x[T][2][D][n] = NTT(c0, c1) of T rotated ciphertexts, e0[T][D][n] = NTT of T plaintexts ->
//                  y[2][D][n] = sum_t c0_t*m_t, sum_t c1_t*m_t: one giant step of he_gemv (he-algo.c:65-79)
//                  accumulated in the NTT domain, one inverse transform for T he_mulpt calls (T = S.g)
// Products are the pseudo-Mersenne ones of gq_arith.h (canonical, the residue poly_rns_mul stores); the host
// selects these modes only when every prime of the range has that shape (pm_ok), else the separate kernels run.
enum { GQ_INV_PLAIN = 0, GQ_INV_TENSOR, GQ_INV_EVK, GQ_INV_EVKPERM, GQ_INV_PT, GQ_INV_MUL, GQ_INV_GEMV };

struct InvSrc {
  const uint64_t *x;        // inputs (limb-major batches as above)
  const uint64_t *e0, *e1;  // key halves, already offset to the first limb of the range
  unsigned g;               // Galois element (GQ_INV_EVKPERM)
};

template <int MODE>
GQ_DEV uint64_t inv_load(const InvSrc &S, const GqPM &pm, int n, int logn, int nlimbs, int b, int dd, int j) {
  const size_t lp = (size_t)dd * n + j, stride = (size_t)nlimbs * n;
  if (MODE == GQ_INV_TENSOR) {
    if (b == 0) return gq_mulm_pm(S.x[lp], S.x[2 * stride + lp], pm);
    if (b == 2) return gq_mulm_pm(S.x[stride + lp], S.x[3 * stride + lp], pm);
    return gq_mul2add_pm(S.x[lp], S.x[3 * stride + lp], S.x[stride + lp], S.x[2 * stride + lp], pm);
  }
  if (MODE == GQ_INV_EVK) return gq_mulm_pm(S.x[lp], (b ? S.e1 : S.e0)[lp], pm);
  if (MODE == GQ_INV_EVKPERM) {
    // slot j of ntt()'s output holds the value at psi^(2*brv(j)+1); sigma_g moves it to the slot of exponent e*g
    const unsigned e = ((2u * (__brev((unsigned)j) >> (32 - logn)) + 1u) * S.g) & (2u * (unsigned)n - 1u);
    const unsigned src = __brev((e - 1u) >> 1) >> (32 - logn);
    return gq_mulm_pm(S.x[(size_t)dd * n + src], (b ? S.e1 : S.e0)[lp], pm);
  }
  if (MODE == GQ_INV_PT) return gq_mulm_pm(S.x[(size_t)b * stride + lp], S.x[2 * stride + lp], pm);
  if (MODE == GQ_INV_MUL) return gq_mulm_pm(S.x[lp], S.x[stride + lp], pm);
  if (MODE == GQ_INV_GEMV) {
    // products below p^2 < 2^118 (1 + 2^-30): four of them stay below the 2^(2s+3) gq_reduce_pm accepts, so the
    // 128-bit sum is folded every four terms and the canonical partial result rides along into the next group
    GqWide acc; acc.lo = 0; acc.hi = 0;
    const int T = (int)S.g;
    for (int t = 0; t < T; t++) {
      acc = gq_add_wide(acc, gq_mul_wide(S.x[(size_t)(2 * t + b) * stride + lp], S.e0[(size_t)t * stride + lp]));
      if ((t & 3) == 3 || t == T - 1) { acc.lo = gq_reduce_pm(acc, pm.p, pm.c, pm.s); acc.hi = 0; }
    }
    return acc.lo;
  }
  return S.x[(size_t)b * stride + lp];
}

template <int LOGT, bool WHOLE, int MODE>
__global__ void __launch_bounds__(TileCfg<LOGT>::NT, (TileCfg<LOGT>::NT >= 512 ? GQ_TILE_CTAS : 1))
k_nttl_inv_tile(GqNttTables tb, InvSrc src, uint64_t *__restrict__ y, int limb0, int nlimbs) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  typedef TileCfg<LOGT> C;
  __shared__ uint64_t sm[C::T];
  const int t = threadIdx.x;
  const int st = swz(t);
  const int m = blockIdx.y;
  const int dd = m % nlimbs, b = m / nlimbs;
  const int d = limb0 + dd;
  const GqPrime pr = tb.primes[d];
  const PC c = load_pc(pr);
  const ulonglong2 *tw = tb.tw_inv + (size_t)d * tb.n;
  const int tile0 = blockIdx.x * C::T;
  uint64_t *a = y + (size_t)m * tb.n + tile0;
  const int jbase = tb.n + tile0;
  uint64_t v[8];
  {
    const GqPM pm = gq_load_pm(pr);
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = inv_load<MODE>(src, pm, tb.n, tb.logn, nlimbs, b, dd, tile0 + e * C::NT + t);
  }
  if (C::NFULL > 0) {
#pragma unroll
    for (int e = 0; e < 8; e++) sm[st ^ swz(e * C::NT)] = v[e];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < C::NFULL; r++) {
      const int lo = 3 * r;
      const int base = idx8(t, lo), sb = swz(base);
#pragma unroll
      for (int e = 0; e < 8; e++) v[e] = sm[sb ^ swz(e << lo)];
      inv_round<3, false>(v, lo, jbase + base, tw, c, make_ulonglong2(0, 0), make_ulonglong2(0, 0));
#pragma unroll
      for (int e = 0; e < 8; e++) sm[sb ^ swz(e << lo)] = v[e];
      __syncthreads();
    }
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = sm[st ^ swz(e * C::NT)];
  }
  inv_round<C::NB0, WHOLE>(v, LOGT - 3, jbase + t, tw, c, tb.fin_w[d], tb.fin_s[d]);
#pragma unroll
  for (int e = 0; e < 8; e++) a[e * C::NT + t] = v[e];
}

// ---- strided register kernels ------------------------------------------------------
// The 2^SA - 1 twiddles of these stages are the same for every thread of the grid row: they are staged in
// shared memory once per CTA (broadcast reads at the point of use) instead of being held in 4 registers
// each from the top of the kernel - ncu had these kernels at 112-128 registers, 2 CTAs per SM, 22 % of
// the warp slots and 36-40 % issue utilisation.
// wt.flags != null: `data` is a receive slab of the sharded pipeline - wait for every peer's epoch first (gq_p2p.h)
template <int SA>
__global__ void __launch_bounds__(256, GQ_STRIDED_MIN_CTAS)
k_nttl_fwd_strided(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs, GqWaitIn wt) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  constexpr int R = 1 << SA;
  __shared__ ulonglong2 stw[R];
  const int cols = tb.n >> SA;
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  if (threadIdx.x < R) stw[threadIdx.x] = tb.tw_fwd[(size_t)d * tb.n + threadIdx.x];
  gq_peer_wait(wt);
  __syncthreads();
  if (col >= cols) return;
  const PC c = load_pc(tb.primes[d]);
  const ulonglong2 *tw = stw;
  uint64_t *a = data + (size_t)m * tb.n + col;
  uint64_t v[R];
#pragma unroll
  for (int r = 0; r < R; r++) v[r] = a[(size_t)r * cols];
#pragma unroll
  for (int s = 0; s < SA; s++) {       // canonical in: bound 1 + 2*SA <= 11 p, no reduction
    const int mb = SA - 1 - s;
#pragma unroll
    for (int g = 0; g < (1 << s); g++) {
      const ulonglong2 w = tw[(1 << s) + g];
#pragma unroll
      for (int i = 0; i < (1 << mb); i++) {
        const int r = (g << (mb + 1)) | i;
        bf_fwd(v[r], v[r | (1 << mb)], w.x, w.y, c.p, c.p2);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; r++) a[(size_t)r * cols] = v[r];
}

// PEER: the finished coefficients are not written back to `data` but into the receive slabs of the ranks that own
// them (sharded pipeline, gq_p2p.h); the last block to finish publishes the exchange's epoch.
template <int SA, bool PEER>
__global__ void __launch_bounds__(256, GQ_STRIDED_MIN_CTAS)
k_nttl_inv_strided(GqNttTables tb, uint64_t *__restrict__ data, int limb0, int nlimbs, GqPeerOut po) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  constexpr int R = 1 << SA;
  __shared__ ulonglong2 stw[R];
  __shared__ uint64_t *sbase[GQ_P2P_MAX_RANKS];
  const int cols = tb.n >> SA;
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y;
  const int d = limb0 + m % nlimbs;
  if (threadIdx.x < R) stw[threadIdx.x] = tb.tw_inv[(size_t)d * tb.n + threadIdx.x];
  if (PEER) gq_peer_bases(po, sbase);
  __syncthreads();
  const bool live = col < cols;
  if (!PEER && !live) return;
  const GqPrime pr = tb.primes[d];
  const PC c = load_pc(pr);
  const ulonglong2 *tw = stw;
  const ulonglong2 wlast = tb.fin_w[d], fs = tb.fin_s[d];
  uint64_t *a = data + (size_t)m * tb.n + (live ? col : 0);
  uint64_t v[R];
#pragma unroll
  for (int r = 0; r < R; r++) v[r] = a[(size_t)r * cols];
  // inputs < 2p (the tile kernel's outputs).  Lazy stages: both legs of a butterfly carry the same bound
  // bnd[r]*p (it depends only on the index bits already processed); sums double it, differences return to
  // 2p; a pair at 16p is brought back to 8p first so that no sum reaches 32p.  All indices and bounds are
  // compile-time constants after unrolling.
  int bnd[R];
#pragma unroll
  for (int r = 0; r < R; r++) bnd[r] = 2;
#pragma unroll
  for (int s = SA - 1; s >= 1; s--) {
    const int mb = SA - 1 - s;
#pragma unroll
    for (int g = 0; g < (1 << s); g++) {
      const ulonglong2 w = tw[(1 << s) + g];
#pragma unroll
      for (int i = 0; i < (1 << mb); i++) {
        const int r = (g << (mb + 1)) | i, q = r | (1 << mb);
        if (bnd[r] >= 16) { v[r] = csub(v[r], c.p8); v[q] = csub(v[q], c.p8); bnd[r] = 8; }
        bf_inv_lz(v[r], v[q], bnd[r], w.x, w.y, c);
        bnd[r] *= 2; bnd[q] = 2;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R / 2; r++) {
    const int q = r + R / 2;
    if (bnd[r] >= 16) { v[r] = csub(v[r], c.p8); v[q] = csub(v[q], c.p8); bnd[r] = 8; }
    const uint64_t s = v[r] + v[q], dd = v[r] - v[q] + pmul(c, bnd[r]);
    v[r] = csub(mul_lazy(s, fs.x, fs.y, c.p), c.p);
    v[q] = csub(mul_lazy(dd, wlast.x, wlast.y, c.p), c.p);
  }
  if (PEER) {
    if (live) {
      const size_t row = (size_t)(m / nlimbs) * po.D + limb0 + m % nlimbs;
#pragma unroll
      for (int r = 0; r < R; r++) *gq_peer_l2c_sb(po, sbase, row, col + r * cols) = v[r];
    }
    gq_peer_signal(po);
  } else {
#pragma unroll
    for (int r = 0; r < R; r++) a[(size_t)r * cols] = v[r];
  }
}

}  // namespace

template <int MODE>
static int launch_inv(const GqNttTables &tb, const InvSrc &src, uint64_t *y, int limb0, int nlimbs, int npolys,
                      cudaStream_t st, int only, const GqPeerOut *po = nullptr) {
  const int logn = tb.logn;
  const int lt = logn > 12 ? 12 : logn, sa = logn - lt;
  const dim3 gt(tb.n >> lt, nlimbs * npolys), gs(((tb.n >> (sa ? sa : 1)) + 255) / 256, nlimbs * npolys);
  if (sa) {
    if (only != 2) GQ_LAUNCH((k_nttl_inv_tile<12, false, MODE>), gt, dim3(TileCfg<12>::NT), st, tb, src, y, limb0, nlimbs);
    if (only != 1) {
      const GqPeerOut none = GqPeerOut();
#define GQ_INVS(SAV) do { if (po) GQ_LAUNCH((k_nttl_inv_strided<SAV, true>), gs, dim3(256), st, tb, y, limb0, nlimbs, *po); \
                          else GQ_LAUNCH((k_nttl_inv_strided<SAV, false>), gs, dim3(256), st, tb, y, limb0, nlimbs, none); } while (0)
      switch (sa) {
        case 1: GQ_INVS(1); break;
        case 2: GQ_INVS(2); break;
        case 3: GQ_INVS(3); break;
        case 4: GQ_INVS(4); break;
        case 5: GQ_INVS(5); break;
      }
#undef GQ_INVS
    }
    return only ? 1 : 2;
  }
  if (po) return -2;          // one-kernel transforms (n <= 2^12) have no peer-output form
  switch (lt) {
#define GQ_CASE(L) case L: GQ_LAUNCH((k_nttl_inv_tile<L, true, MODE>), gt, dim3(TileCfg<L>::NT), st, tb, src, y, limb0, nlimbs); break;
    GQ_CASE(3) GQ_CASE(4) GQ_CASE(5) GQ_CASE(6) GQ_CASE(7) GQ_CASE(8) GQ_CASE(9) GQ_CASE(10) GQ_CASE(11) GQ_CASE(12)
#undef GQ_CASE
  }
  return 1;
}

// Same contract as gq_launch_ntt, canonical in / canonical out only.
int gq_launch_ntt_lazy(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                       bool inverse, cudaStream_t st, int only, const GqWaitIn *wt) {
  // only: 0 = whole transform; 1 = tile kernel only, 2 = strided kernel only (profiling: the
  // output is then not a transform)
  const int logn = tb.logn;
  if (logn < 3 || logn > 17) return -1;
  if (only && logn <= 12) return -1;
  if (inverse) {
    if (wt) return -2;
    InvSrc src; src.x = data; src.e0 = src.e1 = nullptr; src.g = 1;
    return launch_inv<GQ_INV_PLAIN>(tb, src, data, limb0, nlimbs, npolys, st, only);
  }
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  const int lt = logn > 12 ? 12 : logn, sa = logn - lt;
  const dim3 gt(tb.n >> lt, nlimbs * npolys), gs(((tb.n >> (sa ? sa : 1)) + 255) / 256, nlimbs * npolys);
  if (sa) {
    if (only != 1) {
      switch (sa) {
        case 1: GQ_LAUNCH(k_nttl_fwd_strided<1>, gs, dim3(256), st, tb, data, limb0, nlimbs, w0); break;
        case 2: GQ_LAUNCH(k_nttl_fwd_strided<2>, gs, dim3(256), st, tb, data, limb0, nlimbs, w0); break;
        case 3: GQ_LAUNCH(k_nttl_fwd_strided<3>, gs, dim3(256), st, tb, data, limb0, nlimbs, w0); break;
        case 4: GQ_LAUNCH(k_nttl_fwd_strided<4>, gs, dim3(256), st, tb, data, limb0, nlimbs, w0); break;
        case 5: GQ_LAUNCH(k_nttl_fwd_strided<5>, gs, dim3(256), st, tb, data, limb0, nlimbs, w0); break;
      }
    }
    if (only != 2) GQ_LAUNCH((k_nttl_fwd_tile<12, true>), gt, dim3(TileCfg<12>::NT), st, tb, data, limb0, nlimbs);
    return only ? 1 : 2;
  }
  if (wt) return -2;          // one-kernel transforms (n <= 2^12) have no flag-wait form
  switch (lt) {
#define GQ_CASE(L) case L: GQ_LAUNCH((k_nttl_fwd_tile<L, false>), gt, dim3(TileCfg<L>::NT), st, tb, data, limb0, nlimbs); break;
    GQ_CASE(3) GQ_CASE(4) GQ_CASE(5) GQ_CASE(6) GQ_CASE(7) GQ_CASE(8) GQ_CASE(9) GQ_CASE(10) GQ_CASE(11) GQ_CASE(12)
#undef GQ_CASE
  }
  return 1;
}

// The pointwise stage + inverse transform of the internal pipelines in one pass: y[nout][nlimbs][n] from x (and
// the key halves e0, e1 at the same limb range; g = Galois element for the permuted form).  mode = GQ_NTTF_*.
// Every prime of the range must be pseudo-Mersenne (the caller checks pm_ok).  Returns kernels launched.
int gq_launch_ntt_lazy_fused(const GqNttTables &tb, int mode, const uint64_t *x, uint64_t *y, const uint64_t *e0,
                             const uint64_t *e1, unsigned g, int limb0, int nlimbs, cudaStream_t st,
                             const GqPeerOut *po) {
  const int logn = tb.logn;
  if (logn < 3 || logn > 17) return -1;
  InvSrc src; src.x = x; src.e0 = e0; src.e1 = e1; src.g = g;
  switch (mode) {
    case GQ_NTTF_TENSOR: return launch_inv<GQ_INV_TENSOR>(tb, src, y, limb0, nlimbs, 3, st, 0, po);
    case GQ_NTTF_EVK: return launch_inv<GQ_INV_EVK>(tb, src, y, limb0, nlimbs, 2, st, 0, po);
    case GQ_NTTF_EVKPERM: return launch_inv<GQ_INV_EVKPERM>(tb, src, y, limb0, nlimbs, 2, st, 0);
    case GQ_NTTF_PT: return launch_inv<GQ_INV_PT>(tb, src, y, limb0, nlimbs, 2, st, 0);
    case GQ_NTTF_MUL: return launch_inv<GQ_INV_MUL>(tb, src, y, limb0, nlimbs, 1, st, 0);
    case GQ_NTTF_GEMV: return g ? launch_inv<GQ_INV_GEMV>(tb, src, y, limb0, nlimbs, 2, st, 0) : -1;
  }
  return -1;
}
