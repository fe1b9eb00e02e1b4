// gq_hostbig.h - small unsigned big-integer class for HOST-side context setup only
// (products of primes, P/p_d, q_l tables, P^-1 mod 2^k).  It replaces the libgcrypt
// calls of src/precomp.c:266-293,386-409; nothing here runs per coefficient.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <vector>

#include "gq_bigint.h"

struct BigU {
  std::vector<uint64_t> w;  // little-endian, no leading zero words (empty == 0)

  BigU() {}
  explicit BigU(uint64_t x) { if (x) w.push_back(x); }
  BigU(const uint64_t *p, size_t n) : w(p, p + n) { trim(); }
  void trim() { while (!w.empty() && w.back() == 0) w.pop_back(); }
  bool is_zero() const { return w.empty(); }
  unsigned bits() const {
    if (w.empty()) return 0;
    return (unsigned)(64 * (w.size() - 1) + 64 - __builtin_clzll(w.back()));
  }
  bool is_pow2() const {
    if (w.empty()) return false;
    for (size_t i = 0; i + 1 < w.size(); i++) if (w[i]) return false;
    return (w.back() & (w.back() - 1)) == 0;
  }
  uint64_t word(size_t i) const { return i < w.size() ? w[i] : 0; }
  // copy into a fixed-width buffer (zero padded)
  void store(uint64_t *dst, size_t n) const { for (size_t i = 0; i < n; i++) dst[i] = word(i); }

  static int cmp(const BigU &a, const BigU &b) {
    if (a.w.size() != b.w.size()) return a.w.size() < b.w.size() ? -1 : 1;
    for (size_t i = a.w.size(); i-- > 0;)
      if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    return 0;
  }
  BigU mul_u64(uint64_t m) const {
    BigU r; r.w.resize(w.size() + 1);
    uint64_t c = 0;
    for (size_t i = 0; i < w.size(); i++) {
      gq_u128 t = (gq_u128)w[i] * m + c;
      r.w[i] = (uint64_t)t; c = (uint64_t)(t >> 64);
    }
    r.w[w.size()] = c; r.trim(); return r;
  }
  BigU mul(const BigU &b) const {
    BigU r; r.w.assign(w.size() + b.w.size() + 1, 0);
    for (size_t i = 0; i < w.size(); i++) {
      uint64_t c = 0;
      for (size_t j = 0; j < b.w.size(); j++) {
        gq_u128 t = (gq_u128)w[i] * b.w[j] + r.w[i + j] + c;
        r.w[i + j] = (uint64_t)t; c = (uint64_t)(t >> 64);
      }
      r.w[i + b.w.size()] += c;
    }
    r.trim(); return r;
  }
  // floor(this / d), remainder through *rem
  BigU div_u64(uint64_t d, uint64_t *rem = nullptr) const {
    BigU r; r.w.resize(w.size());
    uint64_t c = 0;
    for (size_t i = w.size(); i-- > 0;) {
      gq_u128 t = ((gq_u128)c << 64) | w[i];
      r.w[i] = (uint64_t)(t / d); c = (uint64_t)(t % d);
    }
    if (rem) *rem = c;
    r.trim(); return r;
  }
  uint64_t mod_u64(uint64_t d) const { uint64_t r; div_u64(d, &r); return r; }
  BigU shr(unsigned s) const {
    BigU r; size_t ws = s / 64; unsigned bs = s % 64;
    if (ws >= w.size()) return r;
    r.w.resize(w.size() - ws);
    for (size_t i = 0; i < r.w.size(); i++) {
      uint64_t lo = w[i + ws], hi = (i + ws + 1 < w.size()) ? w[i + ws + 1] : 0;
      r.w[i] = bs ? ((lo >> bs) | (hi << (64 - bs))) : lo;
    }
    r.trim(); return r;
  }
  // floor(this / b) for general b (Knuth D through gq_divrem)
  BigU div(const BigU &b, BigU *rem = nullptr) const {
    BigU q;
    if (cmp(*this, b) < 0) { if (rem) *rem = *this; return q; }
    std::vector<uint64_t> u(w);
    q.w.assign(u.size() - b.w.size() + 1, 0);
    gq_divrem(u.data(), (int)u.size(), b.w.data(), (int)b.w.size(), q.w.data());
    if (rem) { rem->w.assign(u.begin(), u.begin() + b.w.size()); rem->trim(); }
    q.trim(); return q;
  }
};
