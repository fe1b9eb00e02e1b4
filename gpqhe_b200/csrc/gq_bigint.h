// gq_bigint.h - fixed-width multiword integers kept in registers (compile-time W,
// fully unrolled), little-endian 64-bit words, two's complement where signed.
//
// These replace the libgcrypt calls on the reference's hot path:
//   gcry_mpi_mulm/addm in rns_reconstruct        (src/rns.c:63-74)
//   mpi_smod                                      (src/types.c:108-113)
//   mpi_rdiv                                      (src/types.c:115-128)
// All results are the exact integers libgcrypt produces; only the representation
// (fixed-width two's complement instead of sign+magnitude) differs.
#pragma once
#include "gq_platform.h"

// r += y * c   (r has W+1 words; returns nothing, top carry lands in r[W])
template <int W>
GQ_HD void mw_mac(uint64_t *r, uint64_t y, const uint64_t *c) {
  uint64_t carry = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    gq_u128 t = (gq_u128)y * c[w] + r[w] + carry;
    r[w] = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
  }
  r[W] += carry;
}

// r += y * c over the low W words only (mod 2^(64W)); the last column needs no high product
template <int W>
GQ_HD void mw_mac_low(uint64_t *r, uint64_t y, const uint64_t *c) {
  uint64_t carry = 0;
#pragma unroll
  for (int w = 0; w < W - 1; w++) {
    gq_u128 t = (gq_u128)y * c[w] + r[w] + carry;
    r[w] = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
  }
  r[W - 1] += y * c[W - 1] + carry;
}

// r -= k * c  over W words (mod 2^(64W)); returns nothing
template <int W>
GQ_HD void mw_mulsub(uint64_t *r, uint64_t k, const uint64_t *c) {
  uint64_t carry = 0, borrow = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    gq_u128 t = (gq_u128)k * c[w] + carry;
    uint64_t lo = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
    uint64_t a = r[w];
    uint64_t d = a - lo;
    uint64_t b1 = a < lo;
    uint64_t d2 = d - borrow;
    uint64_t b2 = d < borrow;
    r[w] = d2;
    borrow = b1 | b2;
  }
}

template <int W>
GQ_HD uint64_t mw_add(uint64_t *r, const uint64_t *a, const uint64_t *b) {
  uint64_t carry = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    gq_u128 t = (gq_u128)a[w] + b[w] + carry;
    r[w] = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
  }
  return carry;
}

template <int W>
GQ_HD uint64_t mw_sub(uint64_t *r, const uint64_t *a, const uint64_t *b) {
  uint64_t borrow = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t x = a[w], y = b[w];
    uint64_t d = x - y;
    uint64_t b1 = x < y;
    uint64_t d2 = d - borrow;
    uint64_t b2 = d < borrow;
    r[w] = d2;
    borrow = b1 | b2;
  }
  return borrow;
}

// unsigned a >= b
template <int W>
GQ_HD bool mw_geq(const uint64_t *a, const uint64_t *b) {
  // borrow of a - b is clear  <=>  a >= b
  uint64_t borrow = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t x = a[w], y = b[w];
    uint64_t d = x - y;
    uint64_t b1 = x < y;
    uint64_t b2 = d < borrow;
    borrow = b1 | b2;
  }
  return borrow == 0;
}

// two's-complement negate in place
template <int W>
GQ_HD void mw_neg(uint64_t *r) {
  uint64_t carry = 1;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t x = ~r[w] + carry;
    carry = (carry && x == 0) ? 1u : 0u;
    r[w] = x;
  }
}

// Sign-extend r (W words) from bit position k-1: the result is the representative of
// r mod 2^k in [-2^(k-1), 2^(k-1)) == mpi_smod(r, 2^k, 2^(k-1)) for any integer r.
// 0 <= k <= 64*W.  k == 0 (q == 1, qh == 0) gives -1, as the reference does: r mod 1 = 0,
// 0 >= qh, so it subtracts q.
template <int W>
GQ_HD void mw_smod_pow2(uint64_t *r, unsigned k) {
  if (k == 0) {
#pragma unroll
    for (int w = 0; w < W; w++) r[w] = ~(uint64_t)0;
    return;
  }
  const unsigned sw = (k - 1) >> 6, sb = (k - 1) & 63;
  // the usual case (k is the same for every thread of a launch: a uniform branch): the sign bit lies in the top
  // word, only that word changes
  if (sw == (unsigned)(W - 1)) {
    r[W - 1] = (uint64_t)(((int64_t)(r[W - 1] << (63 - sb))) >> (63 - sb));
    return;
  }
  if (W >= 2 && sw == (unsigned)(W - 2)) {
    const uint64_t x = (uint64_t)(((int64_t)(r[W >= 2 ? W - 2 : 0] << (63 - sb))) >> (63 - sb));
    r[W >= 2 ? W - 2 : 0] = x;
    r[W - 1] = (uint64_t)((int64_t)x >> 63);
    return;
  }
  // Word selection is done with masks, never `if (w == sw) x = r[w]`: the compiler turns
  // such chains into a dynamic index, which forces the whole array into local memory.
  uint64_t sel = 0;
#pragma unroll
  for (int w = 0; w < W; w++) sel |= r[w] & (0 - (uint64_t)((unsigned)w == sw));
  const uint64_t fill = 0 - ((sel >> sb) & 1u);
  const uint64_t keep = (sb == 63) ? ~(uint64_t)0 : ((((uint64_t)1) << (sb + 1)) - 1);
#pragma unroll
  for (int w = 0; w < W; w++) {
    const uint64_t is_sw = 0 - (uint64_t)((unsigned)w == sw);
    const uint64_t above = 0 - (uint64_t)((unsigned)w > sw);
    const uint64_t at = (r[w] & keep) | (fill & ~keep);
    r[w] = (above & fill) | (is_sw & at) | (~(above | is_sw) & r[w]);
  }
}

// Rounded division by 2^j, j in [1, 64*W): floor(r / 2^j) + (r mod 2^j > 2^(j-1)),
// == mpi_rdiv(r, 2^j) (an exact tie rounds toward -inf).  In place, signed.
template <int W>
GQ_HD void mw_rdiv_pow2(uint64_t *r, unsigned j) {
  if (j < 64) {     // the usual case (Delta below 2^64; j is uniform over the launch): word-local funnel shifts
    const uint64_t half = (r[0] >> (j - 1)) & 1u;
    const uint64_t lower = (j > 1) ? (r[0] & ((((uint64_t)1) << (j - 1)) - 1)) : 0;
    const uint64_t sign = 0 - (r[W - 1] >> 63);
    uint64_t carry = (half && lower) ? 1u : 0u;        // remainder > 2^(j-1): an exact tie rounds toward -inf
#pragma unroll
    for (int w = 0; w < W; w++) {
      const uint64_t hi = (w + 1 < W) ? r[w + 1] : sign;
      const uint64_t x = ((r[w] >> j) | (hi << (64 - j))) + carry;
      carry = (x < carry) ? 1u : 0u;
      r[w] = x;
    }
    return;
  }
  const unsigned jw = j >> 6, jb = j & 63;
  // remainder > 2^(j-1)  <=>  bit j-1 set and some lower bit set
  const unsigned hb = j - 1, hw = hb >> 6, hbb = hb & 63;
  uint64_t hsel = 0, lower = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    lower |= r[w] & (0 - (uint64_t)((unsigned)w < hw));
    hsel |= r[w] & (0 - (uint64_t)((unsigned)w == hw));
  }
  const uint64_t half = (hsel >> hbb) & 1u;
  lower |= hbb ? (hsel & ((((uint64_t)1) << hbb) - 1)) : 0;
  const uint64_t sign = 0 - (r[W - 1] >> 63);
  uint64_t t[W];
#pragma unroll
  for (int w = 0; w < W; w++) {
    // word w of (r >> j) is built from words w+jw and w+jw+1 of r (sign beyond the top)
    uint64_t lo = 0, hi = 0;
#pragma unroll
    for (int v = 0; v < W; v++) {
      lo |= r[v] & (0 - (uint64_t)((unsigned)v == (unsigned)w + jw));
      hi |= r[v] & (0 - (uint64_t)((unsigned)v == (unsigned)w + jw + 1));
    }
    if ((unsigned)w + jw >= (unsigned)W) lo = sign;
    if ((unsigned)w + jw + 1 >= (unsigned)W) hi = sign;
    t[w] = (jb == 0) ? lo : ((lo >> jb) | (hi << (64 - jb)));
  }
  uint64_t carry = (half && lower) ? 1u : 0u;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t x = t[w] + carry;
    carry = (x < carry) ? 1u : 0u;
    r[w] = x;
  }
}

// ------------------------------------------------------------------------------
// Runtime-width helpers for the general-modulus slow path (arbitrary q, Delta).
// Arrays live in local memory; this path is exact, not fast.
// ------------------------------------------------------------------------------

// 128-by-64 division step: (hi:lo) / d with hi < d.  Returns quotient, writes remainder.
GQ_HD uint64_t gq_div128(uint64_t hi, uint64_t lo, uint64_t d, uint64_t *rem) {
  gq_u128 n = ((gq_u128)hi << 64) | lo;
  uint64_t q = (uint64_t)(n / d);
  *rem = (uint64_t)(n - (gq_u128)q * d);
  return q;
}

// Knuth algorithm D remainder: u (nu words, unsigned, modified in place) mod v (nv
// words, top word non-zero).  On return u[0..nv) holds the remainder and u[nv..nu)
// is zero.  `quot` (nu-nv+1 words) receives the quotient when not null.
GQ_HD void gq_divrem(uint64_t *u, int nu, const uint64_t *v, int nv, uint64_t *quot) {
  if (quot) for (int i = 0; i < nu - nv + 1; i++) quot[i] = 0;
  if (nu < nv) return;
  if (nv == 1) {
    uint64_t rem = 0;
    for (int i = nu - 1; i >= 0; i--) {
      uint64_t q = gq_div128(rem, u[i], v[0], &rem);
      if (quot) quot[i] = q;
      u[i] = 0;
    }
    u[0] = rem;
    return;
  }
  // normalise so the divisor's top bit is set
#if defined(__CUDA_ARCH__)
  int s = __clzll((long long)v[nv - 1]);
#else
  int s = __builtin_clzll(v[nv - 1]);
#endif
  uint64_t vt = v[nv - 1], vs = v[nv - 2], vss = (nv >= 3) ? v[nv - 3] : 0;
  uint64_t vn1 = s ? ((vt << s) | (vs >> (64 - s))) : vt;
  uint64_t vn2 = s ? ((vs << s) | (vss >> (64 - s))) : vs;
  for (int j = nu - nv; j >= 0; j--) {
    // top three words of the current (normalised) partial remainder
    uint64_t a2 = (j + nv < nu) ? u[j + nv] : 0;
    uint64_t a1 = u[j + nv - 1], a0 = u[j + nv - 2];
    uint64_t am = (j + nv - 3 >= 0) ? u[j + nv - 3] : 0;
    uint64_t n2 = s ? ((a2 << s) | (a1 >> (64 - s))) : a2;
    uint64_t n1 = s ? ((a1 << s) | (a0 >> (64 - s))) : a1;
    uint64_t n0 = s ? ((a0 << s) | (am >> (64 - s))) : a0;
    uint64_t qhat, rhat;
    bool rh_over = false;
    if (n2 >= vn1) {
      qhat = ~(uint64_t)0;
      rhat = n1 + vn1;               // (n2:n1) - qhat*vn1 with n2 == vn1
      rh_over = rhat < n1;
    } else {
      qhat = gq_div128(n2, n1, vn1, &rhat);
    }
    while (!rh_over) {
      gq_u128 lhs = (gq_u128)qhat * vn2;
      gq_u128 rhs = ((gq_u128)rhat << 64) | n0;
      if (lhs <= rhs) break;
      qhat--;
      uint64_t nr = rhat + vn1;
      rh_over = nr < rhat;
      rhat = nr;
    }
    // u[j..j+nv] -= qhat * v
    uint64_t carry = 0, borrow = 0;
    for (int i = 0; i < nv; i++) {
      gq_u128 t = (gq_u128)qhat * v[i] + carry;
      uint64_t lo = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
      uint64_t x = u[j + i];
      uint64_t d = x - lo;
      uint64_t b1 = x < lo;
      uint64_t d2 = d - borrow;
      uint64_t b2 = d < borrow;
      u[j + i] = d2;
      borrow = b1 | b2;
    }
    uint64_t top = (j + nv < nu) ? u[j + nv] : 0;
    uint64_t need = carry + borrow;          // cannot overflow: carry <= 2^64-2
    bool negative = top < need;
    if (j + nv < nu) u[j + nv] = top - need;
    if (negative) {                           // qhat was one too large: add back
      qhat--;
      uint64_t c = 0;
      for (int i = 0; i < nv; i++) {
        gq_u128 t = (gq_u128)u[j + i] + v[i] + c;
        u[j + i] = (uint64_t)t;
        c = (uint64_t)(t >> 64);
      }
      if (j + nv < nu) u[j + nv] += c;
    }
    if (quot) quot[j] = qhat;
  }
}
