// gq_arith.h - 64-bit modular arithmetic for the RNS limb layer.
//
// Replaces src/reduce.c (montgomery_reduce :59, barrett_reduce :88).  The reference
// keeps twiddles in Montgomery form and reduces every product with a 128-bit
// Montgomery step; the outputs that matter are the *canonical residues* it stores.
// We reach the same canonical residues with Shoup multiplication by precomputed
// constants (w, floor(w*2^64/p)): one mulhi + two mullo on the integer pipe and a
// shorter dependency chain than Montgomery.  Products of two variables use the
// reference's own Barrett step so that poly_rns_mul/add agree for *any* input the
// reference accepts (a < 2^(2*bits(p))).
#pragma once
#include "gq_platform.h"

GQ_HD uint64_t gq_mulhi(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
  return __umul64hi(a, b);
#else
  return (uint64_t)(((gq_u128)a * b) >> 64);
#endif
}

// a*b + c with 32-bit factors: one IMAD.WIDE.U32.  Written as PTX because the C form
// (uint64_t)a * (uint64_t)b + c leaves an add of the (zero) high cross term behind every multiply in the SASS.
GQ_HD uint64_t gq_mad_wide(uint32_t a, uint32_t b, uint64_t c) {
#if defined(__CUDA_ARCH__)
  uint64_t r;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
  return r;
#else
  return (uint64_t)a * (uint64_t)b + c;
#endif
}

// Per-prime constants (device image of one struct rns_ctx node, src/poly.h:28-41).
struct GqPrime {
  uint64_t p;          // the prime
  uint64_t pinv_mont;  // p^-1 mod 2^64                   (montgomery_inv, src/reduce.c:36)
  uint64_t pinv_barr;  // floor(2^(2*bits(p)) / p)        (barrett_inv,    src/reduce.c:75)
  uint64_t ninv_mont;  // N^-1 * 2^64 mod p               (ntt_init,       src/precomp.c:248)
  uint64_t ninv;       // N^-1 mod p (plain)
  uint64_t ninv_sh;    // floor(ninv * 2^64 / p)
  uint32_t barr_shift; // 2*bits(p) - 64
  uint32_t bits;       // bits(p)
};

// x*w mod p, canonical, for any x < 2^64, w < p < 2^63, wsh = floor(w*2^64/p).
GQ_HD uint64_t gq_mul_shoup(uint64_t x, uint64_t w, uint64_t wsh, uint64_t p) {
  uint64_t q = gq_mulhi(x, wsh);
  uint64_t r = x * w - q * p;   // in [0, 2p)
  return (r >= p) ? r - p : r;
}

// Same value as barrett_reduce(a, p, pinv_barr) of src/reduce.c:88-106 for a = hi:lo.
GQ_HD uint64_t gq_barrett128(uint64_t lo, uint64_t hi, const GqPrime &P) {
  // t = ((lo*mu)>>64) + hi*mu   (fits 128 bits for a < 2^(2k))
  uint64_t t_lo = gq_mulhi(lo, P.pinv_barr);
  uint64_t h_lo = hi * P.pinv_barr;
  uint64_t h_hi = gq_mulhi(hi, P.pinv_barr);
  uint64_t s_lo = t_lo + h_lo;
  uint64_t s_hi = h_hi + (s_lo < t_lo ? 1u : 0u);
  // t >>= shift  (0 <= shift < 64); only the low 64 bits of t*p are needed below
  uint64_t t = (P.barr_shift == 0) ? s_lo : ((s_lo >> P.barr_shift) | (s_hi << (64 - P.barr_shift)));
  uint64_t r = lo - t * P.p;
  return (r < P.p) ? r : r - P.p;
}

GQ_HD uint64_t gq_mulmod_barrett(uint64_t a, uint64_t b, const GqPrime &P) {
  return gq_barrett128(a * b, gq_mulhi(a, b), P);
}

GQ_HD uint64_t gq_addmod_barrett(uint64_t a, uint64_t b, const GqPrime &P) {
  uint64_t lo = a + b;
  uint64_t hi = (lo < a) ? 1u : 0u;
  return gq_barrett128(lo, hi, P);
}

// montgomery_reduce of src/reduce.c:59-66 (used on the host to convert the
// reference's Montgomery-form tables to plain residues).
GQ_HD uint64_t gq_mont_reduce(uint64_t lo, uint64_t hi, uint64_t p, uint64_t pinv) {
  uint64_t u = lo * pinv;
  uint64_t t = gq_mulhi(u, p);
  return (hi < t) ? hi - t + p : hi - t;
}

// Forward (Cooley-Tukey) butterfly with the reference's exact range rules
// (src/ntt.c:46-48): the difference leg is canonical, the sum leg may hold p.
GQ_HD void gq_bfly_fwd(uint64_t &x, uint64_t &y, uint64_t w, uint64_t wsh, uint64_t p) {
  uint64_t t = gq_mul_shoup(y, w, wsh, p);
  uint64_t d = (x >= t) ? x - t : x - t + p;
  uint64_t s = (x <= p - t) ? x + t : x + t - p;
  x = s;
  y = d;
}

// Inverse (Gentleman-Sande) butterfly, src/ntt.c:63-67.
GQ_HD void gq_bfly_inv(uint64_t &x, uint64_t &y, uint64_t w, uint64_t wsh, uint64_t p) {
  uint64_t t = x, u = y;
  uint64_t s = (u <= p - t) ? t + u : t + u - p;
  uint64_t d = (u <= t) ? t - u : t - u + p;
  x = s;
  y = gq_mul_shoup(d, w, wsh, p);
}

// ---- products of two variables for primes p = 2^s + c with c < 2^28 (every GPQHE prime is
// 2^59 + 1 + k*2N).  2^s == -c (mod p), so a 2s+3-bit value z = zh*2^s + zl folds as
//   z == zl - c*zh,   c*zh = uh*2^s + ul   ==>   z == zl - ul + c*uh   in (-2^s, 2^(s+1))
// with three 32x32 products instead of Barrett's ~14; the result is then made canonical.
// Same canonical residue as barrett_reduce (src/reduce.c:88-106); used by the internal pointwise
// kernels only (the public poly_rns_mul keeps the reference's Barrett domain).
struct GqWide { uint64_t lo, hi; };

GQ_HD GqWide gq_mul_wide(uint64_t a, uint64_t b) {
  const uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
  const uint64_t p00 = (uint64_t)a0 * b0, p01 = (uint64_t)a0 * b1, p10 = (uint64_t)a1 * b0,
                 p11 = (uint64_t)a1 * b1;
  const uint64_t mid = (p00 >> 32) + (uint32_t)p01 + (uint32_t)p10;      // < 3 * 2^32
  GqWide r;
  r.lo = (mid << 32) | (uint32_t)p00;
  r.hi = p11 + (p01 >> 32) + (p10 >> 32) + (mid >> 32);
  return r;
}
GQ_HD GqWide gq_add_wide(GqWide x, GqWide y) {
  GqWide r;
  r.lo = x.lo + y.lo;
  r.hi = x.hi + y.hi + (r.lo < x.lo ? 1u : 0u);
  return r;
}
// z < 2^(2s+3), s = bits(p)-1 in [33,60], c = p - 2^s < 2^28  ->  z mod p, canonical
GQ_HD uint64_t gq_reduce_pm(GqWide z, uint64_t p, uint32_t c, int s) {
  const uint64_t mask = (((uint64_t)1) << s) - 1;
  const uint64_t zl = z.lo & mask;
  const uint64_t zh = (z.lo >> s) | (z.hi << (64 - s));                  // < 2^(s+3) <= 2^63
  const uint64_t u0 = (uint64_t)c * (uint32_t)zh, u1 = (uint64_t)c * (uint32_t)(zh >> 32);
  // u = u0 + u1*2^32 < 2^91
  const uint64_t ulo = u0 + (u1 << 32);
  const uint64_t uhi = (u1 >> 32) + (ulo < u0 ? 1u : 0u);
  const uint64_t ul = ulo & mask;
  const uint64_t uh = (ulo >> s) | (uhi << (64 - s));                    // < 2^31
  uint64_t r = zl - ul + (uint64_t)c * (uint32_t)uh;                     // in (-2^s, 2^(s+1)) as signed
  r = ((int64_t)r < 0) ? r + p : r;
  return (r >= p) ? r - p : r;
}

// per-prime constants of the pseudo-Mersenne products (p = 2^s + c) and the products of two variables the
// internal pointwise stages use (gq_pointwise.cu, and the loaders of the fused inverse NTT in gq_ntt_lazy.cu)
struct GqPM { uint64_t p; uint32_t c; int s; };
GQ_HD GqPM gq_load_pm(const GqPrime &P) {
  GqPM m; m.p = P.p; m.s = (int)P.bits - 1; m.c = (uint32_t)(P.p - (((uint64_t)1) << m.s)); return m;
}
GQ_HD uint64_t gq_mulm_pm(uint64_t a, uint64_t b, const GqPM &m) { return gq_reduce_pm(gq_mul_wide(a, b), m.p, m.c, m.s); }
// a*b + c2*d mod p (the cross term of the tensor product, one reduction for both products)
GQ_HD uint64_t gq_mul2add_pm(uint64_t a, uint64_t b, uint64_t c2, uint64_t d, const GqPM &m) {
  return gq_reduce_pm(gq_add_wide(gq_mul_wide(a, b), gq_mul_wide(c2, d)), m.p, m.c, m.s);
}
