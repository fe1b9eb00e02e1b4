/*
 * oracle.c - CPU restatement of GPQHE's 64-bit RNS layer.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (gpqhe_b200/, include/,
 * dropin/) may include, link or call this file; only tests/, the smoke check in
 * __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs do, and
 * only as the checker / CPU baseline.
 *
 * What it restates (reference file:line, commit 5e70953c):
 *   orc_montgomery_inv     src/reduce.c:36-49
 *   orc_montgomery_reduce  src/reduce.c:59-66
 *   orc_barrett_inv        src/reduce.c:75-78
 *   orc_barrett_reduce     src/reduce.c:88-106
 *   orc_ntt / orc_invntt   src/ntt.c:37-52 / 54-73   (incl. the `<= q-t` rule that
 *                          can store q instead of 0 on the + leg)
 *   orc_poly_rns_add/mul   src/poly.c:71-82
 *   orc_next_prime         src/precomp.c:153-191,370-376 (p += 2N until prime)
 *   orc_root_of_unity      src/precomp.c:193-242 (smallest generator g, g^((p-1)/m))
 *   orc_ntt_tables         src/precomp.c:244-264 (Montgomery-form, bit-reversed index)
 *
 * Parity status: PINNED - tests/test_oracle.py checks every function here against
 * outputs of the unmodified reference (oracle/_ref/libgpqhe.so, built by
 * oracle/Makefile) in this container and against the fixtures committed under
 * tests/golden/ (generated from the reference by tests/golden/make_golden.py).
 *
 * The reference's primality test draws Miller-Rabin witnesses from libc rand();
 * a prime passes for every witness, and a composite 60-bit number surviving 50
 * random witnesses has probability < 4^-50, so the deterministic 12-base test
 * below selects the same primes (checked against the reference's actual primes).
 */
#include <stdint.h>
#include <stddef.h>

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ reduce.c */

uint64_t orc_montgomery_inv(uint64_t q)
{
  /* q^(2^64-1) mod 2^64 by 64 square-and-multiply rounds == q^-1 mod 2^64 */
  uint64_t qinv = 1;
  for (unsigned k = 64; k > 0; k--) {
    qinv *= q;
    q *= q;
  }
  return qinv;
}

uint64_t orc_montgomery_reduce(u128 a, uint64_t q, uint64_t qinv)
{
  uint64_t lo = (uint64_t)a;
  uint64_t hi = (uint64_t)(a >> 64);
  uint64_t u = lo * qinv;
  uint64_t t = (uint64_t)(((u128)u * q) >> 64);
  return (hi < t) ? hi - t + q : hi - t;
}

static unsigned nbits64(uint64_t q) { return 64u - (unsigned)__builtin_clzll(q); }

uint64_t orc_barrett_inv(uint64_t q)
{
  return (uint64_t)(((u128)1 << (2 * nbits64(q))) / (u128)q);
}

/* returns 0 and sets *err when the reference would abort (2*bits(q) < 64) */
uint64_t orc_barrett_reduce128(uint64_t alo, uint64_t ahi, uint64_t q, uint64_t qinv, int *err)
{
  u128 a = ((u128)ahi << 64) | alo;
  u128 t = (((u128)alo * qinv) >> 64) + (u128)ahi * qinv;
  int shift = 2 * (int)nbits64(q) - 64;
  if (shift < 0) { if (err) *err = 1; return 0; }
  t >>= shift;
  a -= t * q;
  uint64_t r = (uint64_t)a;
  return (r < q) ? r : r - q;
}

static inline uint64_t barrett(u128 a, uint64_t q, uint64_t qinv)
{
  return orc_barrett_reduce128((uint64_t)a, (uint64_t)(a >> 64), q, qinv, NULL);
}

/* --------------------------------------------------------------------- ntt.c */

static inline uint64_t fqmul(uint64_t a, uint64_t b, uint64_t q, uint64_t qinv)
{
  return orc_montgomery_reduce((u128)a * b, q, qinv);
}

void orc_ntt(uint64_t *a, unsigned n, uint64_t q, uint64_t qinv, const uint64_t *zetas)
{
  unsigned len, start, j = 0, k = 1;
  for (len = n / 2; len >= 1; len >>= 1) {
    for (start = 0; start < n; start = j + len) {
      uint64_t zeta = zetas[k++];
      for (j = start; j < start + len; j++) {
        uint64_t t = fqmul(a[j + len], zeta, q, qinv);
        a[j + len] = (a[j] >= t) ? a[j] - t : a[j] - t + q;
        a[j]       = (a[j] <= q - t) ? a[j] + t : a[j] + t - q;
      }
    }
  }
}

void orc_invntt(uint64_t *a, unsigned n, uint64_t q, uint64_t qinv, uint64_t ninv,
                const uint64_t *zetas_inv)
{
  unsigned len, start, j = 0, k;
  for (len = 1; len <= n / 2; len <<= 1) {
    k = n / (2 * len);
    for (start = 0; start < n; start = j + len) {
      uint64_t zi = zetas_inv[k++];
      for (j = start; j < start + len; j++) {
        uint64_t t = a[j];
        uint64_t u = a[j + len];
        a[j] = (u <= q - t) ? t + u : t + u - q;
        uint64_t d = (u <= t) ? t - u : t - u + q;
        a[j + len] = fqmul(d, zi, q, qinv);
      }
    }
  }
  for (unsigned i = 0; i < n; i++)
    a[i] = fqmul(a[i], ninv, q, qinv);
}

/* -------------------------------------------------------------------- poly.c */

void orc_poly_rns_add(uint64_t *r, const uint64_t *a, const uint64_t *b, unsigned n,
                      uint64_t q, uint64_t qinv_barr)
{
  for (unsigned i = 0; i < n; i++)
    r[i] = barrett((u128)a[i] + (u128)b[i], q, qinv_barr);
}

void orc_poly_rns_mul(uint64_t *r, const uint64_t *a, const uint64_t *b, unsigned n,
                      uint64_t q, uint64_t qinv_barr)
{
  for (unsigned i = 0; i < n; i++)
    r[i] = barrett((u128)a[i] * (u128)b[i], q, qinv_barr);
}

/* ----------------------------------------------------------------- precomp.c */

static uint64_t mulm(uint64_t a, uint64_t b, uint64_t m) { return (uint64_t)(((u128)a * b) % m); }

uint64_t orc_powm(uint64_t a, uint64_t b, uint64_t m)
{
  uint64_t r = 1;
  while (b > 0) {
    if (b & 1) r = mulm(r, a, m);
    b >>= 1;
    a = mulm(a, a, m);
  }
  return r;
}

int orc_isprime(uint64_t p)
{
  static const uint64_t bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
  if (p < 2) return 0;
  for (int i = 0; i < 12; i++) {
    if (p == bases[i]) return 1;
    if (p % bases[i] == 0) return 0;
  }
  uint64_t u = p - 1; unsigned t = 0;
  while (!(u & 1)) { u >>= 1; t++; }
  for (int i = 0; i < 12; i++) {
    uint64_t x = orc_powm(bases[i], u, p);
    if (x == 1 || x == p - 1) continue;
    int comp = 1;
    for (unsigned r = 1; r < t; r++) {
      x = mulm(x, x, p);
      if (x == p - 1) { comp = 0; break; }
    }
    if (comp) return 0;
  }
  return 1;
}

/* precomp.c:370-376: candidates start from the previous prime (or 2^logp+1) and step 2N */
uint64_t orc_next_prime(uint64_t p, unsigned n)
{
  for (;;) {
    p += 2ull * n;
    if (orc_isprime(p)) return p;
  }
}

static uint64_t generator(uint64_t p)
{
  uint64_t phi = p - 1, fac[64]; size_t nf = 0;
  uint64_t m = phi;
  for (uint64_t f = 2; f * f <= m; f++) {
    if (m % f == 0) { fac[nf++] = f; while (m % f == 0) m /= f; }
  }
  if (m > 1) fac[nf++] = m;
  for (uint64_t g = 2; g <= phi; g++) {
    int ok = 1;
    for (size_t j = 0; j < nf; j++)
      if (orc_powm(g, phi / fac[j], p) == 1) { ok = 0; break; }
    if (ok) return g;
  }
  return 0;
}

uint64_t orc_root_of_unity(uint64_t m, uint64_t p)
{
  return orc_powm(generator(p), (p - 1) / m, p);
}

static uint32_t bitrev32(uint32_t a)
{
  a = ((a & 0xaaaaaaaau) >> 1) | ((a & 0x55555555u) << 1);
  a = ((a & 0xccccccccu) >> 2) | ((a & 0x33333333u) << 2);
  a = ((a & 0xf0f0f0f0u) >> 4) | ((a & 0x0f0f0f0fu) << 4);
  a = ((a & 0xff00ff00u) >> 8) | ((a & 0x00ff00ffu) << 8);
  return (a >> 16) | (a << 16);
}

/* fills zetas[n], zetas_inv[n]; returns ninv = n^-1 * 2^64 mod p through *ninv */
void orc_ntt_tables(uint64_t p, unsigned logn, uint64_t *zetas, uint64_t *zetas_inv,
                    uint64_t *ninv)
{
  unsigned n = 1u << logn;
  u128 R = (u128)1 << 64;
  *ninv = (uint64_t)(((u128)orc_powm(n, p - 2, p) * R) % p);
  uint64_t root = orc_root_of_unity(2ull * n, p);
  uint64_t rootinv = orc_powm(root, p - 2, p);
  uint64_t pw = 1, pwi = 1;
  for (unsigned i = 0; i < n; i++) {
    uint32_t j = bitrev32(i) >> (32 - logn);
    zetas[j]     = (uint64_t)(((u128)pw  * R) % p);
    zetas_inv[j] = (uint64_t)(((u128)pwi * R) % p);
    pw  = mulm(pw, root, p);
    pwi = mulm(pwi, rootinv, p);
  }
}
