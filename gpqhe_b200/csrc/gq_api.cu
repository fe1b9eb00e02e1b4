// gq_api.cu - the C ABI (include/gpqhe_b200.h): context construction (the host part of
// src/precomp.c: prime search, roots, CRT constants, level table), table upload, and
// the orchestration of the per-stage kernels for each reference entry point.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/gpqhe_b200.h"
#include "gq_context.h"
#include "gq_p2p.h"

// ------------------------------------------------------------------------------ errors
static thread_local std::string g_err;
void gq_set_error(const std::string &msg) { g_err = msg; }
extern "C" const char *gq_last_error(void) { return g_err.c_str(); }
extern "C" int gq_abi_version(void) { return 1; }

#define GQ_FAIL(msg) do { gq_set_error(std::string(__func__) + ": " + (msg)); return 1; } while (0)
#define GQ_CUDA(call)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      gq_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_)); \
      return 1;                                                                        \
    }                                                                                  \
  } while (0)
#define GQ_CHECK_LAUNCH(ctx)                                                           \
  do {                                                                                 \
    cudaError_t e_ = cudaGetLastError();                                               \
    if (e_ != cudaSuccess) {                                                           \
      gq_set_error(std::string(__func__) + ": kernel launch -> " + cudaGetErrorString(e_)); \
      return 1;                                                                        \
    }                                                                                  \
  } while (0)

// --------------------------------------------------- number theory (src/precomp.c:119-242)
static uint64_t mulm(uint64_t a, uint64_t b, uint64_t m) { return (uint64_t)(((gq_u128)a * b) % m); }
static uint64_t powm(uint64_t a, uint64_t e, uint64_t m) {
  uint64_t r = 1 % m;
  a %= m;
  while (e) { if (e & 1) r = mulm(r, a, m); a = mulm(a, a, m); e >>= 1; }
  return r;
}
// The reference runs 50 Miller-Rabin rounds with rand() witnesses (src/precomp.c:153-191);
// the first 12 prime bases are a proof for every n < 3.3e24, so the same numbers pass.
static bool is_prime(uint64_t p) {
  static const uint64_t B[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
  if (p < 2) return false;
  for (uint64_t b : B) { if (p == b) return true; if (p % b == 0) return false; }
  uint64_t u = p - 1; int t = 0;
  while (!(u & 1)) { u >>= 1; t++; }
  for (uint64_t b : B) {
    uint64_t x = powm(b, u, p);
    if (x == 1 || x == p - 1) continue;
    bool witness = true;
    for (int r = 1; r < t && witness; r++) { x = mulm(x, x, p); if (x == p - 1) witness = false; }
    if (witness) return false;
  }
  return true;
}
// smallest primitive root (src/precomp.c:205-225), then g^((p-1)/m) (:234-242)
static uint64_t root_of_unity(uint64_t m, uint64_t p) {
  uint64_t phi = p - 1, rest = phi;
  std::vector<uint64_t> fac;
  for (uint64_t f = 2; f * f <= rest; f++)
    if (rest % f == 0) { fac.push_back(f); while (rest % f == 0) rest /= f; }
  if (rest > 1) fac.push_back(rest);
  for (uint64_t g = 2; g <= phi; g++) {
    bool ok = true;
    for (uint64_t f : fac) if (powm(g, phi / f, p) == 1) { ok = false; break; }
    if (ok) return powm(g, phi / m, p);
  }
  return 0;
}
static uint32_t bitrev(uint32_t a, unsigned bits) {
  uint32_t r = 0;
  for (unsigned i = 0; i < bits; i++) r |= ((a >> i) & 1u) << (bits - 1 - i);
  return r;
}
static unsigned std_logqub(unsigned logn) {   // src/precomp.c:52-61 (classical, 128 bit)
  switch (logn) { case 10: return 27; case 11: return 54; case 12: return 109;
                  case 13: return 218; case 14: return 438; case 15: return 881; }
  return 0;
}
static unsigned round_wt(unsigned words) { return (words + 7) / 8 * 8; }

// ------------------------------------------------------------------------ device helpers
template <typename T>
static int upload(T **dst, const T *src, size_t count) {
  GQ_CUDA(cudaMalloc((void **)dst, count * sizeof(T) + 16));
  GQ_CUDA(cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

static int ensure_slot(gq_ctx *c, int s, size_t bytes) {
  if (c->slot_bytes[s] >= bytes) return 0;
  if (c->slot[s]) { GQ_CUDA(cudaStreamSynchronize(c->stream)); GQ_CUDA(cudaFree(c->slot[s])); c->slot[s] = nullptr; }
  size_t want = bytes + bytes / 4 + 256;
  GQ_CUDA(cudaMalloc(&c->slot[s], want));
  c->slot_bytes[s] = want;
  return 0;
}

static int build_modulus(GqHostModulus &m, const BigU &q) {
  m.q = q;
  m.dev.is_pow2 = q.is_pow2() ? 1 : 0;
  m.dev.k = q.bits() ? q.bits() - 1 : 0;
  m.dev.nwords = (int)q.w.size();
  std::vector<uint64_t> buf(2 * GQ_MAX_W, 0);
  q.store(buf.data(), GQ_MAX_W);
  q.shr(1).store(buf.data() + GQ_MAX_W, GQ_MAX_W);
  uint64_t *d = nullptr;
  if (upload(&d, buf.data(), buf.size())) return 1;
  m.blob = d;
  m.dev.q = d;
  m.dev.qh = d + GQ_MAX_W;
  return 0;
}

// CRT constants of the basis {p_0..p_{D-1}} (rns_init, src/precomp.c:266-293)
static int build_basis(gq_ctx *c, unsigned D, const BigU &P) {
  GqHostBasis &hb = c->basis[D - 1];
  hb.P = P;
  const unsigned bitsP = P.bits();
  const unsigned WT = round_wt((bitsP + 1 + 63) / 64);
  if (WT > GQ_MAX_WT || D > WT + WT / 8 + 2) {
    // wider than the register-resident CRT kernels support (3072 bits): the context is still
    // usable for the limb layer and for narrower bases; CRT over this basis fails loudly.
    hb.dev.D = (int)D; hb.dev.WT = 0;
    return 0;
  }
  std::vector<uint64_t> buf((size_t)(2 + D) * WT + 4 * D + 2, 0);
  uint64_t *hP = buf.data(), *hPh = hP + WT, *hphat = hPh + WT, *hinv = hphat + (size_t)D * WT,
           *hinvsh = hinv + D, *hu = hinvsh + D, *hhc = hu + 2 * D;
  P.store(hP, WT);
  P.shr(1).store(hPh, WT);
  for (unsigned d = 0; d < D; d++) {
    const uint64_t pd = c->primes[d].p;
    BigU phat = P.div_u64(pd);
    phat.store(hphat + (size_t)d * WT, WT);
    uint64_t inv = powm(phat.mod_u64(pd), pd - 2, pd);   // precomp.c:289
    hinv[d] = inv;
    hinvsh[d] = (uint64_t)(((gq_u128)inv << 64) / pd);
    hb.inv.push_back(inv);
  }
  hb.phat.assign(hphat, hphat + (size_t)D * WT);
  // fixed point for the fast CRT kernels: F = 64 + bits(p) - 1 when every prime has the same length
  int F = 64 + (int)c->primes[0].bits - 1;
  for (unsigned d = 0; d < D; d++) if (c->primes[d].bits != c->primes[0].bits) F = 0;
  if (F > 126 || c->primes[0].bits < 33) F = 0;
  if (F) {
    std::vector<uint64_t> pw(3, 0);
    pw[F / 64] = 1ull << (F % 64);
    for (unsigned d = 0; d < D; d++) BigU(pw.data(), 3).div_u64(c->primes[d].p).store(hu + 2 * d, 2);
    // floor(2^F * ceil(P/2) / P); P is odd, so ceil(P/2) = floor(P/2) + 1
    BigU half = P.shr(1);
    std::vector<uint64_t> hv(half.w.size() + 1, 0);
    uint64_t cy = 1;
    for (size_t i = 0; i < hv.size(); i++) { uint64_t t = half.word(i) + cy; cy = (cy && t == 0); hv[i] = t; }
    BigU(hv.data(), hv.size()).mul(BigU(pw.data(), 3)).div(P).store(hhc, 2);
    for (unsigned d = 0; d < D; d++) hb.u.push_back(hu[2 * d]);
  }
  uint64_t *dv = nullptr;
  if (upload(&dv, buf.data(), buf.size())) return 1;
  hb.blob = dv;
  GqCrtBasis &B = hb.dev;
  B.D = (int)D; B.WT = (int)WT; B.bitsP = bitsP;
  B.P = dv; B.Ph = dv + WT; B.phat = dv + 2 * WT; B.inv = B.phat + (size_t)D * WT; B.inv_sh = B.inv + D;
  B.u = B.inv_sh + D; B.hc = B.u + 2 * D; B.F = F;
  if (bitsP >= 58) {
    B.sh = (int)bitsP - 58;
    BigU top = P.shr((unsigned)B.sh);
    uint64_t t1 = top.word(0) + 1;
    gq_u128 num = (((gq_u128)1) << 120) - 1;
    B.recip = (uint64_t)(num / t1);
  } else {
    B.sh = -1; B.recip = 0;
  }
  return 0;
}

// Final-scale tables of the lazy inverse NTT with the CRT pre-multiplication folded in, one pair per prefix basis
// (K x K Shoup pairs, a few KB), and for the key switch the factors that turn the pre-scaled residues of a basis D
// into those of the special-modulus basis {p_0..p_(dim-1)}.  Built once per context (lanes share them).
static int build_prescale(gq_ctx *c) {
  const unsigned K = c->dimub;
  if (getenv("GPQHE_B200_NO_PRESCALE")) return 0;
  for (unsigned D = 1; D <= K; D++) {
    GqHostBasis &hb = c->basis[D - 1];
    if (hb.dev.WT == 0 || hb.dev.F == 0 || hb.inv.size() != D) continue;
    std::vector<ulonglong2> fs(K, make_ulonglong2(0, 0)), fw(K, make_ulonglong2(0, 0));
    for (unsigned d = 0; d < D; d++) {
      const GqPrime &P = c->primes[d];
      const uint64_t sv = mulm(P.ninv, hb.inv[d], P.p), wv = mulm(c->h_inv1[d], sv, P.p);
      fs[d] = make_ulonglong2(sv, (uint64_t)(((gq_u128)sv << 64) / P.p));
      fw[d] = make_ulonglong2(wv, (uint64_t)(((gq_u128)wv << 64) / P.p));
    }
    if (upload(&hb.fin_s, fs.data(), fs.size()) || upload(&hb.fin_w, fw.data(), fw.size())) return 1;
    if (c->has_he && D >= c->dim && c->basis[c->dim - 1].inv.size() == c->dim) {
      std::vector<ulonglong2> rt(c->dim);
      for (unsigned d = 0; d < c->dim; d++) {
        const GqPrime &P = c->primes[d];
        const uint64_t r = mulm(c->basis[c->dim - 1].inv[d], powm(hb.inv[d], P.p - 2, P.p), P.p);
        rt[d] = make_ulonglong2(r, (uint64_t)(((gq_u128)r << 64) / P.p));
      }
      if (upload(&hb.ratio, rt.data(), rt.size())) return 1;
    }
  }
  return 0;
}

// ------------------------------------------------------------------------------- create
#ifndef GQ_EMULATE
int gq_pdl_enabled = getenv("GPQHE_B200_NO_PDL") ? 0 : 1;      // programmatic dependent launch (gq_platform.h)
#endif
// process-wide switch of the programmatic dependent launch; returns the previous setting.  On by default: it hides
// the launch gaps of an eagerly issued op (one he_mul + he_rs: 296 -> 270 us).  The sharded plans switch it off above
// two ranks: their per-rank kernels are a few microseconds long and already replayed from graphs, and there the early
// launch + in-kernel wait measured SLOWER than plain graph edges (8 GPUs: 0.136 against 0.113 ms per op).
extern "C" int gq_set_pdl(int on) {
#ifndef GQ_EMULATE
  const int prev = gq_pdl_enabled;
  if (!getenv("GPQHE_B200_NO_PDL")) gq_pdl_enabled = on ? 1 : 0;
  return prev;
#else
  (void)on;
  return 0;
#endif
}
static void tc_prebuild(gq_ctx *c);
extern "C" int gq_ctx_create(gq_ctx **out, unsigned logn, const uint64_t *qw, unsigned q_words,
                             unsigned slots, uint64_t Delta, unsigned logp, unsigned logqub,
                             int device, const uint64_t *primes_in, const uint64_t *const *zetas_in,
                             const uint64_t *const *zetas_inv_in) {
  if (!out || !qw || q_words == 0) GQ_FAIL("null argument");
  if (logn < 3 || logn > 17) GQ_FAIL("logn must be in [3,17]");
  if (logp == 0) logp = 59;
  if (logp < 8 || logp > 61) GQ_FAIL("logp must be in [8,61]");
  int ndev = 0;
  GQ_CUDA(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) GQ_FAIL("no CUDA device: gpqhe_b200 has no CPU path");
  if (device < 0 || device >= ndev) GQ_FAIL("bad device index");
  GQ_CUDA(cudaSetDevice(device));

  gq_ctx *c = new gq_ctx();
  c->device = device;
  c->logn = logn; c->n = 1u << logn; c->logp = logp;
  c->q = BigU(qw, q_words);
  if (c->q.bits() < 2) { delete c; GQ_FAIL("modulus q too small"); }
  const unsigned logq = c->q.bits() - 1;
  if (logqub == 0) {                                   // precomp.c:338-340
    logqub = std_logqub(logn);
    if (logn < 10 || logn > 15) logqub = logq;
  }
  if (logq > logqub) { delete c; GQ_FAIL("The input modulus q is too large. Must guarantee log(q)<=log(qub)."); }
  c->logqub = logqub;
  c->dimub = (1 + logn + 4 * logqub) / logp + 1;       // precomp.c:357
  if (c->dimub > GQ_MAX_PRIMES) { delete c; GQ_FAIL("dimub exceeds GQ_MAX_PRIMES"); }
  // from here on every failure goes through gq_ctx_destroy(c): nothing allocated so far may leak
#define GQ_CUDA_CTX(call)                                                                \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      gq_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_));  \
      gq_ctx_destroy(c);                                                                 \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)
  GQ_CUDA_CTX(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->lazy_ntt = getenv("GPQHE_B200_EXACT_NTT") == nullptr;   // A/B switch for tests and profiling

  const unsigned n = c->n, K = c->dimub;
  // primes: p = 2^logp + 1 + k*2N ascending (precomp.c:358-376), or the caller's
  c->primes.resize(K);
  uint64_t p = (1ull << logp) + 1;
  for (unsigned d = 0; d < K; d++) {
    if (primes_in) p = primes_in[d];
    else do { p += 2ull * n; } while (!is_prime(p));
    if ((p - 1) % (2ull * n) != 0 || !is_prime(p)) { delete c; GQ_FAIL("supplied modulus is not an NTT prime"); }
    GqPrime &P = c->primes[d];
    P.p = p;
    P.bits = 64 - (unsigned)__builtin_clzll(p);
    uint64_t inv = 1, sq = p;                              // montgomery_inv, reduce.c:36-49
    for (int k = 0; k < 64; k++) { inv *= sq; sq *= sq; }
    P.pinv_mont = inv;
    P.pinv_barr = (2 * P.bits >= 128) ? 0 : (uint64_t)((((gq_u128)1) << (2 * P.bits)) / p);
    P.barr_shift = (2 * P.bits >= 64) ? 2 * P.bits - 64 : 0;
    P.ninv = powm(n, p - 2, p);
    P.ninv_mont = (uint64_t)((((gq_u128)P.ninv) << 64) % p);
    P.ninv_sh = (uint64_t)((((gq_u128)P.ninv) << 64) / p);
  }
  // twiddles: {w, floor(w*2^64/p)} indexed like zetas[] / zetas_inv[] (precomp.c:244-264)
  {
    std::vector<ulonglong2> fwd((size_t)K * n), inv((size_t)K * n), last(K);
    for (unsigned d = 0; d < K; d++) {
      const GqPrime &P = c->primes[d];
      ulonglong2 *f = &fwd[(size_t)d * n], *iv = &inv[(size_t)d * n];
      if (zetas_in && zetas_inv_in) {
        for (unsigned k = 0; k < n; k++) {      // Montgomery form -> plain residue
          uint64_t w = gq_mont_reduce(zetas_in[d][k], 0, P.p, P.pinv_mont);
          uint64_t wi = gq_mont_reduce(zetas_inv_in[d][k], 0, P.p, P.pinv_mont);
          f[k] = make_ulonglong2(w, (uint64_t)(((gq_u128)w << 64) / P.p));
          iv[k] = make_ulonglong2(wi, (uint64_t)(((gq_u128)wi << 64) / P.p));
        }
      } else {
        uint64_t root = root_of_unity(2ull * n, P.p), rinv = powm(root, P.p - 2, P.p);
        uint64_t pw = 1, pwi = 1;
        for (unsigned i = 0; i < n; i++) {
          uint32_t j = bitrev(i, logn);
          f[j] = make_ulonglong2(pw, (uint64_t)(((gq_u128)pw << 64) / P.p));
          iv[j] = make_ulonglong2(pwi, (uint64_t)(((gq_u128)pwi << 64) / P.p));
          pw = mulm(pw, root, P.p);
          pwi = mulm(pwi, rinv, P.p);
        }
      }
      uint64_t wl = mulm(iv[1].x, P.ninv, P.p);
      last[d] = make_ulonglong2(wl, (uint64_t)(((gq_u128)wl << 64) / P.p));
      c->h_inv1.push_back(iv[1].x);
    }
    std::vector<ulonglong2> fs(K);
    for (unsigned d = 0; d < K; d++) fs[d] = make_ulonglong2(c->primes[d].ninv, c->primes[d].ninv_sh);
    if (upload(&c->d_tw_fwd, fwd.data(), fwd.size()) || upload(&c->d_tw_inv, inv.data(), inv.size()) ||
        upload(&c->d_tw_inv_last, last.data(), last.size()) || upload(&c->d_fin_s, fs.data(), fs.size())) {
      gq_ctx_destroy(c); return 1;
    }
  }
  if (upload(&c->d_primes, c->primes.data(), K)) { gq_ctx_destroy(c); return 1; }
  {   // decompose table: 2^(64w) mod p
    std::vector<uint64_t> pw((size_t)K * GQ_DEC_W), t1(K);
    for (unsigned d = 0; d < K; d++) {
      const uint64_t pd = c->primes[d].p;
      const uint64_t b = (uint64_t)((((gq_u128)1) << 64) % pd);
      uint64_t v = 1 % pd;
      for (unsigned w = 0; w < GQ_DEC_W; w++) { pw[(size_t)d * GQ_DEC_W + w] = v; v = mulm(v, b, pd); }
      t1[d] = (uint64_t)(((gq_u128)b << 64) / pd);
    }
    std::vector<uint64_t> p32((size_t)K * GQ_DEC32), osh(K);
    for (unsigned d = 0; d < K; d++) {
      const uint64_t pd = c->primes[d].p;
      uint64_t v = 1 % pd;
      const uint64_t b32 = (1ull << 32) % pd;
      for (unsigned j = 0; j < GQ_DEC32; j++) { p32[(size_t)d * GQ_DEC32 + j] = v; v = mulm(v, b32, pd); }
      osh[d] = (uint64_t)((((gq_u128)1) << 64) / pd);
    }
    std::vector<uint64_t> p28((size_t)K * GQ_DEC28);
    for (unsigned d = 0; d < K; d++) {
      const uint64_t pd = c->primes[d].p;
      uint64_t v = 1 % pd;
      const uint64_t b28 = (1ull << 28) % pd;
      for (unsigned j = 0; j < GQ_DEC28; j++) { p28[(size_t)d * GQ_DEC28 + j] = v; v = mulm(v, b28, pd); }
    }
    if (upload(&c->d_pow64, pw.data(), pw.size()) || upload(&c->d_t1sh, t1.data(), t1.size()) ||
        upload(&c->d_pow28, p28.data(), p28.size()) ||
        upload(&c->d_pow32, p32.data(), p32.size()) || upload(&c->d_onesh, osh.data(), osh.size())) {
      gq_ctx_destroy(c); return 1;
    }
  }
  // CRT constants for every prefix basis
  c->basis.resize(K);
  {
    BigU P(1);
    for (unsigned D = 1; D <= K; D++) {
      P = P.mul_u64(c->primes[D - 1].p);
      if (build_basis(c, D, P)) { gq_ctx_destroy(c); return 1; }
    }
  }
  GQ_CUDA_CTX(cudaMalloc((void **)&c->d_adhoc, 2 * GQ_MAX_W * sizeof(uint64_t)));
  GQ_CUDA_CTX(cudaMalloc((void **)&c->d_fix, (4 + 4 * (size_t)n) * sizeof(int)));
  GQ_CUDA_CTX(cudaMemset(c->d_fix, 0, 4 * sizeof(int)));
#undef GQ_CUDA_CTX
  c->fast_crt = getenv("GPQHE_B200_EXACT_CRT") == nullptr;    // A/B switch for tests and profiling

  // hectx_init / qtable_init (precomp.c:386-452)
  if (Delta) {
    if (slots == 0 || (slots & (slots - 1))) { gq_ctx_destroy(c); GQ_FAIL("The slots must be the power of 2."); }
    if (slots > n / 2) { gq_ctx_destroy(c); GQ_FAIL("Must guarantee slots<=(n/2)."); }
    c->has_he = 1; c->slots = slots; c->Delta = Delta;
    const unsigned bitsD = 64 - (unsigned)__builtin_clzll(Delta);
    const unsigned logDelta = bitsD - 1;
    if (logDelta == 0) { gq_ctx_destroy(c); GQ_FAIL("Delta must be at least 2"); }
    c->L = logq / logDelta;                               // precomp.c:391 (integer division)
    c->ql.resize(c->L + 1); c->qh.resize(c->L + 1); c->qmod.resize(c->L + 1);
    BigU q = c->q;
    bool pow2 = (Delta & (Delta - 1)) == 0;
    for (int l = (int)c->L; l >= 0; l--) {
      c->ql[l] = q; c->qh[l] = q.shr(1);
      if (q.is_zero()) { gq_ctx_destroy(c); GQ_FAIL("level modulus reached zero"); }
      pow2 = pow2 && q.is_pow2();
      if (build_modulus(c->qmod[l], q)) { gq_ctx_destroy(c); return 1; }
      q = q.div_u64(Delta);
    }
    c->q_is_pow2 = pow2 ? 1 : 0;
    c->logDelta = logDelta;
    c->dim = (c->ql[c->L].bits() + logn) / logp + 1;      // precomp.c:401
    if (c->dim > K) { gq_ctx_destroy(c); GQ_FAIL("hectx.dim exceeds dimub"); }
    c->P = c->basis[c->dim - 1].P;
    c->PqL = c->P.mul(c->q);
    c->dimevk = (c->ql[c->L].bits() + c->PqL.bits() + logn) / logp + 1;   // precomp.c:407
    // dimevk > dimub is not rejected here: the reference's hectx_init accepts it and only
    // the key-switch entry points (which would walk off its prime list) fail.
    c->W = (c->q.bits() + 63) / 64;
    // P^-1 mod 2^(64*GQ_MAX_W) by word-serial exact division of 1 by P
    {
      std::vector<uint64_t> Pw(GQ_MAX_W), x(GQ_MAX_W, 0), pinv(GQ_MAX_W);
      c->P.store(Pw.data(), GQ_MAX_W);
      uint64_t inv0 = 1, sq = Pw[0];
      for (int k = 0; k < 64; k++) { inv0 *= sq; sq *= sq; }
      x[0] = 1;
      for (int i = 0; i < GQ_MAX_W; i++) {
        uint64_t qi = x[i] * inv0;
        pinv[i] = qi;
        uint64_t carry = 0, borrow = 0;          // x[i..] -= qi * P
        for (int j = 0; i + j < GQ_MAX_W; j++) {
          gq_u128 t = (gq_u128)qi * Pw[j] + carry;
          uint64_t lo = (uint64_t)t; carry = (uint64_t)(t >> 64);
          uint64_t a = x[i + j], dd = a - lo, b1 = a < lo, d2 = dd - borrow, b2 = dd < borrow;
          x[i + j] = d2; borrow = b1 | b2;
        }
      }
      std::vector<uint64_t> pp(2 * GQ_MAX_W);
      c->P.store(pp.data(), GQ_MAX_W);
      c->P.shr(1).store(pp.data() + GQ_MAX_W, GQ_MAX_W);
      c->h_Pinv = pinv;
      if (upload(&c->d_Pinv, pinv.data(), pinv.size()) || upload(&c->d_Pwords, pp.data(), pp.size())) {
        gq_ctx_destroy(c); return 1;
      }
    }
  } else {
    c->W = (c->q.bits() + 63) / 64;
  }
  if (build_prescale(c)) { gq_ctx_destroy(c); return 1; }
  tc_prebuild(c);
  *out = c;
  return 0;
}

extern "C" void gq_ctx_destroy(gq_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (gq_ctx *l : c->lanes) gq_ctx_destroy(l);
  c->lanes.clear();
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (!c->parent) {     // a lane shares the tables of its parent
    cudaFree(c->d_primes); cudaFree(c->d_tw_fwd); cudaFree(c->d_tw_inv); cudaFree(c->d_tw_inv_last); cudaFree(c->d_fin_s);
    for (auto &b : c->basis) { cudaFree(b.fin_s); cudaFree(b.fin_w); cudaFree(b.ratio); }
    for (auto &kv : c->tc_tabs) cudaFree(kv.second);
    cudaFree(c->d_pow64); cudaFree(c->d_t1sh); cudaFree(c->d_pow32); cudaFree(c->d_pow28); cudaFree(c->d_onesh); cudaFree(c->d_Pinv); cudaFree(c->d_Pwords);
    for (auto &b : c->basis) cudaFree(b.blob);
    for (auto &m : c->qmod) cudaFree(m.blob);
  }
  cudaFree(c->d_adhoc); cudaFree(c->d_fix);
  for (int s = 0; s < GQ_SLOT_COUNT; s++) cudaFree(c->slot[s]);
  for (auto &e : c->ev_pipe) if (e) cudaEventDestroy(e);
  for (auto &e : c->ev_batch) if (e) cudaEventDestroy(e);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->s_in) cudaStreamDestroy(c->s_in);
  if (c->s_out) cudaStreamDestroy(c->s_out);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

// ------------------------------------------------------------------------------- lanes
static int lane_create(gq_ctx *c, gq_ctx **out) {
  const size_t n = c->n;
  gq_ctx *l = new gq_ctx(*c);        // tables, level moduli, CRT bases: shared device pointers
  l->parent = c;
  l->lanes.clear();
  l->stream = nullptr; l->s_in = nullptr; l->s_out = nullptr;
  for (auto &e : l->ev_pipe) e = nullptr;
  l->ev_batch.clear();
  l->ev_fork = nullptr; l->ev_join = nullptr;
  for (int s = 0; s < GQ_SLOT_COUNT; s++) { l->slot[s] = nullptr; l->slot_bytes[s] = 0; }
  l->d_adhoc = nullptr; l->d_fix = nullptr;
  l->launches = 0;
  if (cudaStreamCreateWithFlags(&l->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&l->ev_join, cudaEventDisableTiming) != cudaSuccess ||
      cudaMalloc((void **)&l->d_adhoc, 2 * GQ_MAX_W * sizeof(uint64_t)) != cudaSuccess ||
      cudaMalloc((void **)&l->d_fix, (4 + 4 * n) * sizeof(int)) != cudaSuccess ||
      cudaMemset(l->d_fix, 0, 4 * sizeof(int)) != cudaSuccess) {
    gq_ctx_destroy(l);
    GQ_FAIL("cannot allocate a compute lane");
  }
  *out = l;
  return 0;
}
extern "C" int gq_ctx_set_lanes(gq_ctx *c, unsigned lanes) {
  if (!c || c->parent) GQ_FAIL("not a root context");
  if (lanes < 1 || lanes > 8) GQ_FAIL("lanes must be in [1,8]");
  GQ_CUDA(cudaSetDevice(c->device));
  if (!c->ev_fork) {
    GQ_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    GQ_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  }
  while (c->lanes.size() + 1 < lanes) {
    gq_ctx *l = nullptr;
    if (lane_create(c, &l)) return 1;
    c->lanes.push_back(l);
  }
  return 0;
}
extern "C" unsigned gq_ctx_lane_count(const gq_ctx *c) { return c ? (unsigned)c->lanes.size() + 1 : 0; }
extern "C" gq_ctx *gq_ctx_lane(gq_ctx *c, unsigned i) {
  if (!c) return nullptr;
  if (i == 0) return c;
  return i <= c->lanes.size() ? c->lanes[i - 1] : nullptr;
}
extern "C" int gq_ctx_lanes_fork(gq_ctx *c) {
  if (!c || c->parent) GQ_FAIL("not a root context");
  if (c->lanes.empty()) return 0;
  GQ_CUDA(cudaEventRecord(c->ev_fork, c->stream));
  for (gq_ctx *l : c->lanes) GQ_CUDA(cudaStreamWaitEvent(l->stream, c->ev_fork, 0));
  return 0;
}
extern "C" int gq_ctx_lanes_join(gq_ctx *c) {
  if (!c || c->parent) GQ_FAIL("not a root context");
  for (gq_ctx *l : c->lanes) {
    GQ_CUDA(cudaEventRecord(l->ev_join, l->stream));
    GQ_CUDA(cudaStreamWaitEvent(c->stream, l->ev_join, 0));
  }
  return 0;
}

// ------------------------------------------------------------------------------ queries
extern "C" int gq_ctx_info(const gq_ctx *c, gq_info *o) {
  if (!c || !o) GQ_FAIL("null argument");
  o->logn = c->logn; o->n = c->n; o->logp = c->logp; o->dimub = c->dimub; o->has_he = c->has_he;
  o->L = c->L; o->dim = c->dim; o->dimevk = c->dimevk; o->W = c->W; o->q_is_pow2 = c->q_is_pow2;
  o->device = c->device;
  return 0;
}
extern "C" int gq_ctx_primes(const gq_ctx *c, uint64_t *out) {
  if (!c || !out) GQ_FAIL("null argument");
  for (unsigned d = 0; d < c->dimub; d++) out[d] = c->primes[d].p;
  return 0;
}
extern "C" int gq_ctx_prime_consts(const gq_ctx *c, unsigned limb, uint64_t out[4]) {
  if (!c || !out || limb >= c->dimub) GQ_FAIL("bad argument");
  const GqPrime &P = c->primes[limb];
  out[0] = P.p; out[1] = P.pinv_mont; out[2] = P.pinv_barr; out[3] = P.ninv_mont;
  return 0;
}
static unsigned dim_tensor(const gq_ctx *c, unsigned l) {      // he-mult.c:99
  return (c->ql[l].bits() * 2 + c->logn) / c->logp + 1;
}
static unsigned dim_ks(const gq_ctx *c, unsigned l) {          // he-mult.c:51
  return (c->ql[l].bits() + c->PqL.bits() + c->logn) / c->logp + 1;
}
extern "C" int gq_ctx_dims(const gq_ctx *c, unsigned l, unsigned *dt, unsigned *dk) {
  if (!c || !c->has_he || l > c->L) GQ_FAIL("bad argument");
  if (dt) *dt = dim_tensor(c, l);
  if (dk) *dk = dim_ks(c, l);
  return 0;
}
// Bound on the plaintext coefficients handed to gq_he_gemv / gq_he_coeff2slot / gq_he_slot2coeff from now on:
// |m| < 2^bits for every coefficient of every encoded diagonal; 0 = unknown (the per-term path of he_gemv runs).
extern "C" int gq_ctx_set_plaintext_bits(gq_ctx *c, unsigned bits) {
  if (!c) GQ_FAIL("null context");
  if (bits > 64u * GQ_MAX_W) GQ_FAIL("bound out of range");
  c->pt_bits = bits;
  return 0;
}
extern "C" int gq_ctx_sync(gq_ctx *c) {
  if (!c) GQ_FAIL("null context");
  for (gq_ctx *l : c->lanes) GQ_CUDA(cudaStreamSynchronize(l->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
extern "C" unsigned long long gq_ctx_launch_count(const gq_ctx *c) {
  if (!c) return 0;
  unsigned long long t = c->launches;
  for (const gq_ctx *l : c->lanes) t += l->launches;
  return t;
}
extern "C" long long gq_ctx_crt_fixups(gq_ctx *c) {
  if (!c || !c->d_fix) return -1;
  int v[2] = {0, 0};
  if (cudaStreamSynchronize(c->stream) != cudaSuccess) return -1;
  if (cudaMemcpy(v, c->d_fix, sizeof v, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  long long t = v[1];
  for (gq_ctx *l : c->lanes) {
    const long long u = gq_ctx_crt_fixups(l);
    if (u < 0) return -1;
    t += u;
  }
  return t;
}
extern "C" void *gq_ctx_stream(gq_ctx *c) { return c ? (void *)c->stream : nullptr; }
extern "C" int gq_host_alloc(void **p, size_t bytes) {
  if (!p) GQ_FAIL("null argument");
  GQ_CUDA(cudaMallocHost(p, bytes));
  return 0;
}
extern "C" void gq_host_free(void *p) { if (p) cudaFreeHost(p); }

static cudaStream_t pick(gq_ctx *c, void *s) { return s ? (cudaStream_t)s : c->stream; }

// ---- device memory for hosts that do not link the CUDA runtime (dropin/: plain C) -------------------
extern "C" int gq_dev_alloc(gq_ctx *c, size_t bytes, void **p) {
  if (!c || !p || bytes == 0) GQ_FAIL("bad argument");
  GQ_CUDA(cudaSetDevice(c->device));
  GQ_CUDA(cudaMalloc(p, bytes));
  return 0;
}
extern "C" void gq_dev_free(gq_ctx *c, void *p) {
  if (!c || !p) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  cudaFree(p);
}
extern "C" int gq_dev_upload(gq_ctx *c, void *dst_dev, const void *src_host, size_t bytes, void *stream) {
  if (!c || !dst_dev || !src_host) GQ_FAIL("null argument");
  GQ_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, pick(c, stream)));
  return 0;
}
extern "C" int gq_dev_download(gq_ctx *c, void *dst_host, const void *src_dev, size_t bytes, void *stream) {
  if (!c || !dst_host || !src_dev) GQ_FAIL("null argument");
  GQ_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, pick(c, stream)));
  return 0;
}
extern "C" int gq_dev_copy(gq_ctx *c, void *dst_dev, const void *src_dev, size_t bytes, void *stream) {
  if (!c || !dst_dev || !src_dev) GQ_FAIL("null argument");
  GQ_CUDA(cudaMemcpyAsync(dst_dev, src_dev, bytes, cudaMemcpyDeviceToDevice, pick(c, stream)));
  return 0;
}
static bool ntt_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs) {
  if (limb0 + nlimbs > c->dimub || nlimbs == 0) return false;
  for (unsigned d = limb0; d < limb0 + nlimbs; d++)
    if (c->primes[d].bits < 32) return false;   // the reference aborts too (reduce.c:94-100)
  return true;
}

// --------------------------------------------------------------------------- limb layer
extern "C" int gq_ntt_batch_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys,
                                uint64_t *data, int inverse, void *stream) {
  if (!c || !data || npolys == 0) GQ_FAIL("bad argument");
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  // inverse: bit 0 = inverse transform; bits 1,2 = profiling only (run just the tile /
  // just the strided kernel of a two-kernel transform; the output is then not an NTT)
  int r = gq_launch_ntt(c->ntt_tables(), data, (int)limb0, (int)nlimbs, (int)npolys, (inverse & 1) != 0,
                        pick(c, stream), (inverse >> 1) & 3);
  if (r < 0) GQ_FAIL("unsupported ring degree");
  c->launches += (unsigned)r;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// every prime of the range is 2^s + c, c < 2^28, 33 <= s <= 59: the pseudo-Mersenne pointwise kernels apply
static bool pm_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs) {
  if (!c->lazy_ntt) return false;
  for (unsigned d = limb0; d < limb0 + nlimbs; d++) {
    const GqPrime &P = c->primes[d];
    if (P.bits < 34 || P.bits > 60) return false;
    const unsigned sh = P.bits - 1, cb = (sh - 31 < 28) ? sh - 31 : 28;   // c*uh (uh < 2^31) must stay below 2^s
    if (P.p - (1ull << sh) >= (1ull << cb)) return false;
  }
  return true;
}

// NTT of the internal pipelines: canonical in / canonical out is all they need, so the lazy
// kernels are used whenever every prime of the range has the shape they assume.
static bool lazy_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs) {
  bool lazy = c->lazy_ntt && limb0 + nlimbs <= c->dimub;
  for (unsigned d = limb0; lazy && d < limb0 + nlimbs; d++) {
    const GqPrime &P = c->primes[d];
    if (P.bits < 33 || P.bits > 60) { lazy = false; break; }
    const uint64_t top = 1ull << (P.bits - 1), cc = P.p - top;
    if (cc >= (top >> 5)) lazy = false;      // gq_ntt_lazy.cu: 32c < 2^sh
  }
  return lazy;
}
// scaleD > 0 (lazy kernels only; the caller checked crt_pre_ok): the inverse transform ends with the CRT
// pre-multiplication of the basis {p_0..p_(scaleD-1)} folded into N^-1
static int ntt_internal(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys, uint64_t *data,
                        bool inverse, cudaStream_t st, int only = 0, unsigned scaleD = 0,
                        const GqWaitIn *wt = nullptr) {
  const bool lazy = lazy_ok(c, limb0, nlimbs);
  if (!lazy) {
    if (scaleD || wt) GQ_FAIL("internal: pre-scaled / flag-waiting NTT needs the lazy kernels");
    return gq_ntt_batch_dev(c, limb0, nlimbs, npolys, data, (inverse ? 1 : 0) | (only << 1), (void *)st);
  }
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  int r = gq_launch_ntt_lazy(c->ntt_tables(inverse ? scaleD : 0), data, (int)limb0, (int)nlimbs, (int)npolys, inverse, st, only, wt);
  if (r < 0) GQ_FAIL(r == -2 ? "no flag-waiting transform for this ring degree" : "unsupported ring degree");
  c->launches += (unsigned)r;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// The residues entering the CRT over {p_0..p_(D-1)} may arrive pre-multiplied by (P/p_d)^-1 (folded into the final
// scaling of the lazy inverse NTT) when the fast CRT kernels will consume them: power-of-two modulus, fixed-point
// constants available, tables built.
static bool crt_pre_ok(const gq_ctx *c, unsigned D, const GqModulus &q) {
  if (D == 0 || D > c->dimub || !c->fast_crt || !lazy_ok(c, 0, D)) return false;
  const GqHostBasis &hb = c->basis[D - 1];
  return hb.fin_s && hb.dev.F && hb.dev.WT && q.is_pow2 && q.k >= 1 && q.k <= 64u * GQ_MAX_W;
}

// pointwise stage + inverse NTT in one pass when the lazy kernels and the pseudo-Mersenne products apply to the
// whole range (gq_launch_ntt_lazy_fused); false = the caller runs the separate pointwise kernel and ntt_internal
static bool fused_inv_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs);
static int ntt_inv_fused(gq_ctx *c, int mode, unsigned limb0, unsigned nlimbs, const uint64_t *x, uint64_t *y,
                         const uint64_t *e0, const uint64_t *e1, uint64_t g, cudaStream_t st, unsigned scaleD,
                         const GqPeerOut *po = nullptr) {
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  int r = gq_launch_ntt_lazy_fused(c->ntt_tables(scaleD), mode, x, y, e0, e1, (unsigned)g, (int)limb0, (int)nlimbs, st, po);
  if (r < 0) GQ_FAIL(r == -2 ? "no peer-output transform for this ring degree" : "unsupported ring degree");
  c->launches += (unsigned)r;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

extern "C" int gq_poly_rns_mul_batch_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys,
                                         uint64_t *r, const uint64_t *a, const uint64_t *b,
                                         void *stream) {
  if (!c || !r || !a || !b || npolys == 0) GQ_FAIL("bad argument");
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  gq_launch_binary(c->d_primes, (int)c->n, (int)limb0, (int)nlimbs, (int)npolys, true, r, a, b,
                   pick(c, stream));
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

extern "C" int gq_ntt_batch(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys,
                            uint64_t *data, int inverse) {
  if (!c || !data) GQ_FAIL("bad argument");
  const size_t bytes = (size_t)nlimbs * npolys * c->n * 8;
  if (ensure_slot(c, GQ_SLOT_LIMB_A, bytes)) return 1;
  uint64_t *d = (uint64_t *)c->slot[GQ_SLOT_LIMB_A];
  GQ_CUDA(cudaMemcpyAsync(d, data, bytes, cudaMemcpyHostToDevice, c->stream));
  if (gq_ntt_batch_dev(c, limb0, nlimbs, npolys, d, inverse, nullptr)) return 1;
  GQ_CUDA(cudaMemcpyAsync(data, d, bytes, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
extern "C" int gq_ntt(gq_ctx *c, unsigned limb, uint64_t *a) { return gq_ntt_batch(c, limb, 1, 1, a, 0); }
extern "C" int gq_invntt(gq_ctx *c, unsigned limb, uint64_t *a) { return gq_ntt_batch(c, limb, 1, 1, a, 1); }

static int binary_host(gq_ctx *c, unsigned limb, uint64_t *r, const uint64_t *a, const uint64_t *b,
                       bool mul) {
  if (!c || !r || !a || !b) GQ_FAIL("bad argument");
  if (!ntt_ok(c, limb, 1)) GQ_FAIL("limb invalid or modulus below 32 bits");
  const size_t bytes = (size_t)c->n * 8;
  if (ensure_slot(c, GQ_SLOT_LIMB_A, 3 * bytes)) return 1;
  uint64_t *d = (uint64_t *)c->slot[GQ_SLOT_LIMB_A];
  GQ_CUDA(cudaMemcpyAsync(d, a, bytes, cudaMemcpyHostToDevice, c->stream));
  GQ_CUDA(cudaMemcpyAsync(d + c->n, b, bytes, cudaMemcpyHostToDevice, c->stream));
  gq_launch_binary(c->d_primes, (int)c->n, (int)limb, 1, 1, mul, d + 2 * c->n, d, d + c->n, c->stream);
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  GQ_CUDA(cudaMemcpyAsync(r, d + 2 * c->n, bytes, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
extern "C" int gq_poly_rns_add(gq_ctx *c, unsigned limb, uint64_t *r, const uint64_t *a, const uint64_t *b) {
  return binary_host(c, limb, r, a, b, false);
}
extern "C" int gq_poly_rns_mul(gq_ctx *c, unsigned limb, uint64_t *r, const uint64_t *a, const uint64_t *b) {
  return binary_host(c, limb, r, a, b, true);
}

static bool pm_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs);
static bool fused_inv_ok(const gq_ctx *c, unsigned limb0, unsigned nlimbs) {
  return !getenv("GPQHE_B200_NO_FUSED_INV") && lazy_ok(c, limb0, nlimbs) && pm_ok(c, limb0, nlimbs);
}

// ------------------------------------------------------------------- MPI <-> RNS (device)
static GqPtrs4 ptrs(const uint64_t *a, const uint64_t *b = nullptr, const uint64_t *c2 = nullptr,
                    const uint64_t *d = nullptr) {
  GqPtrs4 p;
  p.p[0] = (uint64_t *)a; p.p[1] = (uint64_t *)b; p.p[2] = (uint64_t *)c2; p.p[3] = (uint64_t *)d;
  return p;
}

// B-fragment table of the tensor-core decomposition for coefficients of W words (every prime of the context), built
// on first use and kept by the root context
static const unsigned long long *dec_tc_table(gq_ctx *c, unsigned W) {
  if (getenv("GPQHE_B200_NO_TC_CRT") || W == 0 || W > 7) return nullptr;
  gq_ctx *root = c->parent ? c->parent : c;
  const unsigned key = 0x20000u + W;
  auto it = root->tc_tabs.find(key);
  if (it != root->tc_tabs.end()) return (const unsigned long long *)it->second;
  std::vector<uint64_t> p(root->dimub);
  for (unsigned d = 0; d < root->dimub; d++) p[d] = root->primes[d].p;
  std::vector<unsigned long long> tab(gq_dec_tc_table_words((int)root->dimub));
  gq_dec_tc_table(tab.data(), (int)root->dimub, p.data(), (int)W);
  uint32_t *dv = nullptr;
  if (cudaMalloc((void **)&dv, tab.size() * 8) != cudaSuccess) return nullptr;
  if (cudaMemcpy(dv, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(dv); return nullptr; }
  root->tc_tabs[key] = dv;
  return (const unsigned long long *)dv;
}

// ncoef: coefficients per polynomial in this call (0 = the whole ring); the coefficient-local
// stages work on any slice because one thread owns one coefficient
static int decompose_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, const GqPtrs4 &in, unsigned W,
                         unsigned npolys, uint64_t *out, cudaStream_t st, unsigned ncoef = 0,
                         const GqPeerOut *po = nullptr) {
  if (ncoef == 0) ncoef = c->n;
  if (limb0 + nlimbs > c->dimub) GQ_FAIL("limb range exceeds dimub");
  if (W == 0 || W > GQ_MAX_W) GQ_FAIL("word count out of range");
  bool dig = c->fast_crt && W <= 8;         // A1 of k_decompose32 stays below 2^64 for <= 15 digits
  for (unsigned d = limb0; dig && d < limb0 + nlimbs; d++)
    dig = c->primes[d].bits >= 33 && c->primes[d].bits <= 60;
  // 0 = word form, 1 = 32-bit digits, 2 = 28-bit digits (at most 16 of them: carry-free column sums),
  // 3 = 28-bit digits with the pseudo-Mersenne fold
  const int digits = !dig ? 0 : (W <= 7 && !getenv("GPQHE_B200_DEC32")) ? (pm_ok(c, limb0, nlimbs) ? 3 : 2) : 1;
  if (digits == 3) {          // pseudo-Mersenne primes, at most 7 words: the u8 tensor-core contraction (gq_crt_tc.cu)
    const unsigned long long *tt = dec_tc_table(c, W);
    if (tt && gq_launch_decompose_tc(tt, c->d_primes, (int)ncoef, (int)limb0, (int)nlimbs, in, (int)W, out, (int)npolys,
                                     st, po) == 1) {
      c->launches++;
      GQ_CHECK_LAUNCH(c);
      return 0;
    }
  }
  if (gq_launch_decompose(c->dec_tables(), (int)ncoef, (int)limb0, (int)nlimbs, in, (int)W, out,
                          (int)npolys, st, digits, po) < 0)
    GQ_FAIL(po ? "no peer-output decompose kernel for these parameters" : "no decompose kernel for this width");
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// r = a * b mod 2^512 (8 words each)
static void mul_lo8(uint64_t *r, const uint64_t *a, const uint64_t *b) {
  uint64_t t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < 8; i++) {
    uint64_t cy = 0;
    for (int j = 0; i + j < 8; j++) {
      gq_u128 x = (gq_u128)a[i] * b[j] + t[i + j] + cy;
      t[i + j] = (uint64_t)x; cy = (uint64_t)(x >> 64);
    }
  }
  for (int i = 0; i < 8; i++) r[i] = t[i];
}

static const uint32_t *tc_upload(gq_ctx *root, unsigned key, int slots, const std::vector<uint64_t> &K,
                                 const std::vector<uint64_t> &u, const uint64_t *cw) {
  std::vector<uint32_t> tab(gq_crt_tc_table_words(slots));
  gq_crt_tc_table(tab.data(), slots, K.data(), u.data(), cw);
  uint32_t *dv = nullptr;
  if (cudaMalloc((void **)&dv, tab.size() * 4) != cudaSuccess) return nullptr;
  // synchronous copy: every stream that will read the table is ordered after this host call
  if (cudaMemcpy(dv, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(dv); return nullptr; }
  root->tc_tabs[key] = dv;
  return dv;
}

static bool tc_basis_ok(const GqHostBasis &hb, unsigned D) {
  return hb.dev.WT && hb.dev.F >= 96 && hb.dev.F <= 126 && hb.u.size() == D && hb.phat.size() == (size_t)D * hb.dev.WT;
}

// B-fragment table of the tensor-core CRT over the basis D (results of at most 8 words), built once and kept by the
// root context
static const uint32_t *tc_table(gq_ctx *c, unsigned D, unsigned W) {
  const bool off = getenv("GPQHE_B200_NO_TC_CRT") != nullptr;      // A/B switch for tests and profiling
  gq_ctx *root = c->parent ? c->parent : c;
  if (off || W == 0 || W > 8 || D == 0 || D > root->dimub) return nullptr;
  const GqHostBasis &hb = root->basis[D - 1];
  if (!tc_basis_ok(hb, D)) return nullptr;
  auto it = root->tc_tabs.find(D);
  if (it != root->tc_tabs.end()) return it->second;
  const int slots = 4 * (int)((D + 3) / 4);
  std::vector<uint64_t> K((size_t)slots * 8, 0), u(slots, 0);
  for (unsigned d = 0; d < D; d++) {
    for (unsigned w = 0; w < 8 && w < (unsigned)hb.dev.WT; w++) K[(size_t)d * 8 + w] = hb.phat[(size_t)d * hb.dev.WT + w];
    u[d] = hb.u[d];
  }
  uint64_t cw[8];
  hb.P.store(cw, 8);
  return tc_upload(root, D, slots, K, u, cw);
}

// ... of the key-switch epilogue over the basis D2 (y part) and the special basis {p_0..p_(dim-1)} (z part):
// every constant carries the factor P^-1 mod 2^512 of the exact quotient (gq_crt_tc.cu)
static const uint32_t *tc_table_ks(gq_ctx *c, unsigned D2, unsigned k) {
  const bool off = getenv("GPQHE_B200_NO_TC_CRT") != nullptr;
  gq_ctx *root = c->parent ? c->parent : c;
  if (off || k == 0 || k > 512 || D2 == 0 || D2 > root->dimub || !root->has_he || root->dim == 0 || root->dim > D2) return nullptr;
  const GqHostBasis &hb = root->basis[D2 - 1], &hp = root->basis[root->dim - 1];
  if (!tc_basis_ok(hb, D2) || !tc_basis_ok(hp, root->dim) || root->h_Pinv.size() < 8) return nullptr;
  const unsigned key = 0x10000u + D2;
  auto it = root->tc_tabs.find(key);
  if (it != root->tc_tabs.end()) return it->second;
  const unsigned dim = root->dim;
  const int ys = 4 * (int)((D2 + 3) / 4), zs = 4 * (int)((dim + 3) / 4);
  std::vector<uint64_t> K((size_t)(ys + zs) * 8, 0), u(ys + zs, 0);
  const uint64_t *pinv = root->h_Pinv.data();
  uint64_t lo[8];
  for (unsigned d = 0; d < D2; d++) {              // (PQ / p_d) * P^-1
    for (unsigned w = 0; w < 8; w++) lo[w] = w < (unsigned)hb.dev.WT ? hb.phat[(size_t)d * hb.dev.WT + w] : 0;
    mul_lo8(&K[(size_t)d * 8], lo, pinv);
    u[d] = hb.u[d];
  }
  for (unsigned i = 0; i < dim; i++) {             // -(P / p_i) * P^-1
    for (unsigned w = 0; w < 8; w++) lo[w] = w < (unsigned)hp.dev.WT ? hp.phat[(size_t)i * hp.dev.WT + w] : 0;
    uint64_t *Kz = &K[(size_t)(ys + i) * 8];
    mul_lo8(Kz, lo, pinv);
    uint64_t cy = 1;
    for (unsigned w = 0; w < 8; w++) { const uint64_t x = ~Kz[w] + cy; cy = (cy && x == 0) ? 1 : 0; Kz[w] = x; }
    u[ys + i] = hp.u[i];
  }
  uint64_t cw[8];
  hb.P.store(lo, 8);
  mul_lo8(cw, lo, pinv);                           // PQ * P^-1 = Q
  return tc_upload(root, key, ys + zs, K, u, cw);
}

// The tables of the tensor-core kernels for every basis / level the HE entry points use, built with the context so
// that no later call allocates or synchronises for them (a first use inside a stream capture would fail otherwise).
// Best effort: a table that cannot be built leaves that shape on the integer kernels.
static void tc_prebuild(gq_ctx *c) {
  if (c->parent || !c->fast_crt || getenv("GPQHE_B200_NO_TC_CRT")) return;
  for (unsigned D = 1; D <= c->dimub; D++)
    if (c->basis[D - 1].fin_s) tc_table(c, D, 8);
  if (!c->has_he) return;
  for (unsigned l = 0; l <= c->L; l++) {
    const unsigned D2 = dim_ks(c, l);
    if (D2 <= c->dimub && c->basis[D2 - 1].ratio && c->qmod[l].dev.is_pow2 && c->qmod[l].dev.k >= 1 && c->qmod[l].dev.k <= 512)
      tc_table_ks(c, D2, c->qmod[l].dev.k);
  }
  if (c->W <= 7) dec_tc_table(c, c->W);
}

// rs_j > 0: he_rs fused behind the CRT (power-of-two q and Delta; the caller checks with rs_fusable)
static int reconstruct_dev(gq_ctx *c, unsigned D, const uint64_t *rhat, const GqModulus &q,
                           const GqPtrs4 &out, unsigned Wout, unsigned npolys, cudaStream_t st,
                           unsigned ncoef = 0, unsigned rs_j = 0, bool pre = false, const GqWaitIn *wt = nullptr) {
  if (ncoef == 0) ncoef = c->n;
  if (D == 0 || D > c->dimub) GQ_FAIL("basis size out of range");
  if (c->basis[D - 1].dev.WT == 0) GQ_FAIL("RNS basis wider than 3072 bits is not supported by the CRT kernels");
  if (npolys > 4) GQ_FAIL("at most 4 polynomials per reconstruct launch");
  const uint32_t *tct = (pre && c->fast_crt && q.is_pow2 && q.k >= 1) ? tc_table(c, D, (q.k + 63) / 64) : nullptr;
  const int r = gq_launch_reconstruct(c->basis[D - 1].dev, c->d_primes, (int)ncoef, rhat, q, out, (int)Wout,
                                      (int)npolys, st, c->fast_crt ? c->d_fix : nullptr, rs_j, pre, wt, tct);
  if (r < 0) GQ_FAIL(r == -2 ? "no flag-waiting reconstruct kernel for these parameters" : "no reconstruct kernel for this width");
  c->launches += (unsigned)r;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// ad-hoc modulus for poly_mul / poly_rns2mpi (the caller's q, any value)
static int adhoc_modulus(gq_ctx *c, const uint64_t *qw, unsigned q_words, GqModulus *m) {
  BigU q(qw, q_words);
  if (q.is_zero()) GQ_FAIL("modulus is zero");
  if (q.w.size() > GQ_MAX_W) GQ_FAIL("modulus wider than GQ_MAX_W words");
  m->is_pow2 = q.is_pow2() ? 1 : 0;
  m->k = q.bits() - 1;
  m->nwords = (int)q.w.size();
  m->q = c->d_adhoc; m->qh = c->d_adhoc + GQ_MAX_W;
  if (!m->is_pow2) {
    uint64_t buf[2 * GQ_MAX_W];
    q.store(buf, GQ_MAX_W);
    q.shr(1).store(buf + GQ_MAX_W, GQ_MAX_W);
    GQ_CUDA(cudaStreamSynchronize(c->stream));
    GQ_CUDA(cudaMemcpy(c->d_adhoc, buf, sizeof(buf), cudaMemcpyHostToDevice));
  }
  return 0;
}

extern "C" int gq_rns_decompose(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *ahat,
                                const uint64_t *a, unsigned W) {
  if (!c || !ahat || !a || nlimbs == 0) GQ_FAIL("bad argument");
  const size_t nb = (size_t)W * c->n * 8, ob = (size_t)nlimbs * c->n * 8;
  if (ensure_slot(c, GQ_SLOT_IO, nb) || ensure_slot(c, GQ_SLOT_LIMB_A, ob)) return 1;
  uint64_t *din = (uint64_t *)c->slot[GQ_SLOT_IO], *dout = (uint64_t *)c->slot[GQ_SLOT_LIMB_A];
  GQ_CUDA(cudaMemcpyAsync(din, a, nb, cudaMemcpyHostToDevice, c->stream));
  if (decompose_dev(c, limb0, nlimbs, ptrs(din), W, 1, dout, c->stream)) return 1;
  GQ_CUDA(cudaMemcpyAsync(ahat, dout, ob, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int gq_poly_rns2mpi(gq_ctx *c, unsigned dim, uint64_t *r, unsigned Wr, const uint64_t *rhat,
                               const uint64_t *qw, unsigned q_words) {
  if (!c || !r || !rhat || !qw || Wr == 0) GQ_FAIL("bad argument");
  GqModulus m;
  if (adhoc_modulus(c, qw, q_words, &m)) return 1;
  const size_t ib = (size_t)dim * c->n * 8, ob = (size_t)Wr * c->n * 8;
  if (ensure_slot(c, GQ_SLOT_LIMB_A, ib) || ensure_slot(c, GQ_SLOT_IO, ob)) return 1;
  uint64_t *din = (uint64_t *)c->slot[GQ_SLOT_LIMB_A], *dout = (uint64_t *)c->slot[GQ_SLOT_IO];
  GQ_CUDA(cudaMemcpyAsync(din, rhat, ib, cudaMemcpyHostToDevice, c->stream));
  if (reconstruct_dev(c, dim, din, m, ptrs(dout), Wr, 1, c->stream)) return 1;
  GQ_CUDA(cudaMemcpyAsync(r, dout, ob, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int gq_poly_mul(gq_ctx *c, uint64_t *r, unsigned Wr, const uint64_t *a, const uint64_t *b,
                           unsigned W, unsigned dim, const uint64_t *qw, unsigned q_words) {
  if (!c || !r || !a || !b || !qw) GQ_FAIL("bad argument");
  if (!ntt_ok(c, 0, dim)) GQ_FAIL("dim invalid or modulus below 32 bits");
  GqModulus m;
  if (adhoc_modulus(c, qw, q_words, &m)) return 1;
  const size_t n = c->n, ib = (size_t)W * n * 8, ob = (size_t)Wr * n * 8;
  if (ensure_slot(c, GQ_SLOT_IO, 2 * ib + ob) || ensure_slot(c, GQ_SLOT_LIMB_A, 2 * dim * n * 8) ||
      ensure_slot(c, GQ_SLOT_LIMB_B, dim * n * 8)) return 1;
  uint64_t *io = (uint64_t *)c->slot[GQ_SLOT_IO], *A = (uint64_t *)c->slot[GQ_SLOT_LIMB_A],
           *B = (uint64_t *)c->slot[GQ_SLOT_LIMB_B];
  uint64_t *da = io, *db = io + (size_t)W * n, *dr = io + 2 * (size_t)W * n;
  cudaStream_t st = c->stream;
  GQ_CUDA(cudaMemcpyAsync(da, a, ib, cudaMemcpyHostToDevice, st));
  GQ_CUDA(cudaMemcpyAsync(db, b, ib, cudaMemcpyHostToDevice, st));
  const bool pre = crt_pre_ok(c, dim, m);
  if (decompose_dev(c, 0, dim, ptrs(da, db), W, 2, A, st)) return 1;
  if (ntt_internal(c, 0, dim, 2, A, false, st)) return 1;
  if (fused_inv_ok(c, 0, dim)) {
    if (ntt_inv_fused(c, GQ_NTTF_MUL, 0, dim, A, B, nullptr, nullptr, 1, st, pre ? dim : 0)) return 1;
  } else {
    gq_launch_binary(c->d_primes, (int)n, 0, (int)dim, 1, true, B, A, A + (size_t)dim * n, st);   // poly_mul: Barrett as is
    c->launches++;
    if (ntt_internal(c, 0, dim, 1, B, true, st, 0, pre ? dim : 0)) return 1;
  }
  if (reconstruct_dev(c, dim, B, m, ptrs(dr), Wr, 1, st, 0, 0, pre)) return 1;
  GQ_CUDA(cudaMemcpyAsync(r, dr, ob, cudaMemcpyDeviceToHost, st));
  GQ_CUDA(cudaStreamSynchronize(st));
  return 0;
}

static uint64_t galois_rot(const gq_ctx *c, uint64_t rot) {   // 5^rot mod 2n (poly.c:265-268)
  uint64_t g = 1, m = 2ull * c->n;
  uint64_t b = 5, e = rot;
  while (e) { if (e & 1) g = (g * b) % m; b = (b * b) % m; e >>= 1; }
  return g;
}

static int automorph_host(gq_ctx *c, uint64_t *r, const uint64_t *a, unsigned W, uint64_t g) {
  if (!c || !r || !a || W == 0 || W > GQ_MAX_W) GQ_FAIL("bad argument");
  const size_t nb = (size_t)W * c->n * 8;
  if (ensure_slot(c, GQ_SLOT_IO, 2 * nb)) return 1;
  uint64_t *din = (uint64_t *)c->slot[GQ_SLOT_IO], *dout = din + (size_t)W * c->n;
  GQ_CUDA(cudaMemcpyAsync(din, a, nb, cudaMemcpyHostToDevice, c->stream));
  gq_launch_automorph((int)c->n, (int)W, ptrs(dout), ptrs(din), 1, g, c->stream);
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  GQ_CUDA(cudaMemcpyAsync(r, dout, nb, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
extern "C" int gq_poly_rot(gq_ctx *c, uint64_t *r, const uint64_t *a, unsigned W, uint64_t rot) {
  if (!c) GQ_FAIL("null context");
  return automorph_host(c, r, a, W, galois_rot(c, rot));
}
extern "C" int gq_poly_conj(gq_ctx *c, uint64_t *r, const uint64_t *a, unsigned W) {
  if (!c) GQ_FAIL("null context");
  return automorph_host(c, r, a, W, 2ull * c->n - 1);
}

// ---------------------------------------------------------------------- evaluation keys
extern "C" int gq_evk_create(gq_ctx *c, gq_evk **evk, const uint64_t *p0, const uint64_t *p1) {
  if (!c || !evk || !p0 || !p1 || !c->has_he) GQ_FAIL("bad argument");
  if (c->dimevk > c->dimub) GQ_FAIL("hectx.dimevk exceeds polyctx.dimub: no evaluation key exists for these parameters");
  const size_t bytes = (size_t)c->dimevk * c->n * 8;
  gq_evk *e = new gq_evk();
  e->ctx = c;
  GQ_CUDA(cudaMalloc((void **)&e->p0, bytes));
  GQ_CUDA(cudaMalloc((void **)&e->p1, bytes));
  GQ_CUDA(cudaMemcpy(e->p0, p0, bytes, cudaMemcpyHostToDevice));
  GQ_CUDA(cudaMemcpy(e->p1, p1, bytes, cudaMemcpyHostToDevice));
  *evk = e;
  return 0;
}
extern "C" int gq_evk_create_dev(gq_ctx *c, gq_evk **evk, const uint64_t *p0, const uint64_t *p1) {
  if (!c || !evk || !p0 || !p1 || !c->has_he) GQ_FAIL("bad argument");
  gq_evk *e = new gq_evk();
  e->ctx = c; e->p0 = (uint64_t *)p0; e->p1 = (uint64_t *)p1; e->owned = false;
  *evk = e;
  return 0;
}
// wire format support: the device copy of a key back to the host, he_evk_t layout
extern "C" int gq_evk_export(const gq_evk *e, uint64_t *p0, uint64_t *p1) {
  if (!e || !e->ctx || !p0 || !p1) GQ_FAIL("null argument");
  const gq_ctx *c = e->ctx;
  const size_t bytes = (size_t)c->dimevk * c->n * 8;
  GQ_CUDA(cudaSetDevice(c->device));
  GQ_CUDA(cudaMemcpy(p0, e->p0, bytes, cudaMemcpyDeviceToHost));
  GQ_CUDA(cudaMemcpy(p1, e->p1, bytes, cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" void gq_evk_destroy(gq_evk *e) {
  if (!e) return;
  if (e->owned) { cudaFree(e->p0); cudaFree(e->p1); }
  delete e;
}

// ------------------------------------------------------------------ HE primitives (device)
// K6 on a coefficient slice: chat[2][D2][ncoef] -> c0, c1 (W words each), + add0/add1 (may be null)
static int ks_epilogue_dev(gq_ctx *c, unsigned l, unsigned D2, const uint64_t *B, const uint64_t *add0,
                           const uint64_t *add1, uint64_t *c0, uint64_t *c1, unsigned ncoef,
                           cudaStream_t st, unsigned rs_j = 0, bool pre = false, const GqWaitIn *wt = nullptr) {
  if (ncoef == 0) ncoef = c->n;
  const unsigned n = ncoef, W = c->W;
  if (D2 == 0 || D2 > c->dimub || c->basis[D2 - 1].dev.WT == 0) GQ_FAIL("key-switch basis not supported");
  const GqHostModulus &qm = c->qmod[l];
  int r;
  if (qm.dev.is_pow2) {
    r = gq_launch_ks_epilogue(c->basis[D2 - 1].dev, c->basis[c->dim - 1].dev, c->d_Pinv, c->d_primes,
                              (int)n, B, qm.dev.k, ptrs(add0, add1), (int)W, ptrs(c0, c1), (int)W, 2, st,
                              c->fast_crt ? c->d_fix : nullptr, rs_j, pre ? c->basis[D2 - 1].ratio : nullptr, wt,
                              (pre && c->fast_crt) ? tc_table_ks(c, D2, qm.dev.k) : nullptr);
  } else {
    if (rs_j || pre || wt) GQ_FAIL("fused rescale / pre-scaled residues / flag wait need power-of-two moduli");
    r = gq_launch_ks_epilogue_general(c->basis[D2 - 1].dev, c->d_Pwords, (int)c->P.w.size(),
                                      c->d_Pwords + GQ_MAX_W, c->d_primes, (int)n, B, qm.dev,
                                      ptrs(add0, add1), (int)W, ptrs(c0, c1), (int)W, 2, st);
  }
  if (r < 0) GQ_FAIL("no key-switch epilogue kernel for these widths");
  c->launches += (unsigned)r;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// Key switch shared by he_relin (he-mult.c:40-85) and he_swk (he-automorphism.c:40-85):
//   out_k = smod_ql( rdiv(d * evk.p_k, P) + add_k ),  k = 0,1 ; add_k may be null.
static int keyswitch_dev(gq_ctx *c, unsigned l, const uint64_t *d, const gq_evk *evk,
                         const uint64_t *add0, const uint64_t *add1, uint64_t *c0, uint64_t *c1,
                         cudaStream_t st, unsigned rs_j = 0) {
  const unsigned D2 = dim_ks(c, l), n = c->n, W = c->W;
  if (D2 > c->dimevk) GQ_FAIL("key-switch basis exceeds dimevk");
  if (!ntt_ok(c, 0, D2)) GQ_FAIL("modulus below 32 bits");
  if (c->basis[D2 - 1].dev.WT == 0) GQ_FAIL("RNS basis wider than 3072 bits is not supported by the CRT kernels");
  if (ensure_slot(c, GQ_SLOT_LIMB_A, (size_t)4 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_LIMB_B, (size_t)3 * c->dimub * n * 8)) return 1;
  uint64_t *A = (uint64_t *)c->slot[GQ_SLOT_LIMB_A], *B = (uint64_t *)c->slot[GQ_SLOT_LIMB_B];
  const bool pre = crt_pre_ok(c, D2, c->qmod[l].dev) && c->basis[D2 - 1].ratio && D2 >= c->dim;
  if (decompose_dev(c, 0, D2, ptrs(d), W, 1, A, st)) return 1;
  if (ntt_internal(c, 0, D2, 1, A, false, st)) return 1;
  if (fused_inv_ok(c, 0, D2)) {
    if (ntt_inv_fused(c, GQ_NTTF_EVK, 0, D2, A, B, evk->p0, evk->p1, 1, st, pre ? D2 : 0)) return 1;
  } else {
    gq_launch_mul_evk(c->d_primes, (int)n, (int)D2, B, A, evk->p0, evk->p1, st, pm_ok(c, 0, D2));
    c->launches++;
    if (ntt_internal(c, 0, D2, 2, B, true, st, 0, pre ? D2 : 0)) return 1;
  }
  return ks_epilogue_dev(c, l, D2, B, add0, add1, c0, c1, 0, st, rs_j, pre);
}

static int he_check(const gq_ctx *c, unsigned l) {
  if (!c) GQ_FAIL("null context");
  if (!c->has_he) GQ_FAIL("context was created without Delta (polyctx only)");
  if (l > c->L) GQ_FAIL("level out of range");
  return 0;
}

// he_rs can ride on the CRT kernel that produces the ciphertext when q_l, q_(l-1) and Delta are powers of two
// (every shipped parameter set): log2(Delta) then, 0 otherwise (the caller launches the rescale kernel)
static unsigned rs_fusable(const gq_ctx *c, unsigned l) {
  if (l == 0 || !c->fast_crt) return 0;
  if ((c->Delta & (c->Delta - 1)) != 0) return 0;
  const GqModulus &a = c->qmod[l].dev, &b = c->qmod[l - 1].dev;
  if (!a.is_pow2 || !b.is_pow2 || a.k != b.k + c->logDelta) return 0;
  return c->logDelta;
}
static int rescale_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, bool divide, void *stream, unsigned ncoef);

// he_mul (src/he-mult.c:88-156), optionally with he_rs (src/he-rescale.c:33-54) on the product
static int he_mul_impl(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                       const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                       const gq_evk *rlk, void *stream, bool rescale) {
  if (he_check(c, l)) return 1;
  if (rescale && l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1 || !rlk) GQ_FAIL("null argument");
  cudaStream_t st = pick(c, stream);
  const unsigned rs_j = rescale ? rs_fusable(c, l) : 0;
  const unsigned D1 = dim_tensor(c, l), n = c->n, W = c->W;
  if (D1 > c->dimub) GQ_FAIL("tensor basis exceeds dimub");
  if (!ntt_ok(c, 0, D1)) GQ_FAIL("modulus below 32 bits");
  if (ensure_slot(c, GQ_SLOT_LIMB_A, (size_t)4 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_LIMB_B, (size_t)3 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_BIG, (size_t)5 * W * n * 8)) return 1;
  uint64_t *A = (uint64_t *)c->slot[GQ_SLOT_LIMB_A], *B = (uint64_t *)c->slot[GQ_SLOT_LIMB_B];
  uint64_t *big = (uint64_t *)c->slot[GQ_SLOT_BIG];
  uint64_t *d0 = big, *d1 = big + (size_t)W * n, *d2 = big + 2 * (size_t)W * n;
  // hot loop A (he-mult.c:116-138): decompose, forward NTT, tensor, inverse NTT
  const bool pre = crt_pre_ok(c, D1, c->qmod[l].dev);
  if (decompose_dev(c, 0, D1, ptrs(a0, a1, b0, b1), W, 4, A, st)) return 1;
  if (ntt_internal(c, 0, D1, 4, A, false, st)) return 1;
  if (fused_inv_ok(c, 0, D1)) {      // tensor product formed while the inverse transform loads its tiles
    if (ntt_inv_fused(c, GQ_NTTF_TENSOR, 0, D1, A, B, nullptr, nullptr, 1, st, pre ? D1 : 0)) return 1;
  } else {
    gq_launch_tensor(c->d_primes, (int)n, (int)D1, B, A, st, pm_ok(c, 0, D1));
    c->launches++;
    if (ntt_internal(c, 0, D1, 3, B, true, st, 0, pre ? D1 : 0)) return 1;
  }
  // hot loop B (he-mult.c:139-141): d0, d1, d2 mod q_l
  if (reconstruct_dev(c, D1, B, c->qmod[l].dev, ptrs(d0, d1, d2), W, 3, st, 0, 0, pre)) return 1;
  // he_relin (he-mult.c:143); he_rs in the same epilogue kernel when it can be fused
  if (keyswitch_dev(c, l, d2, rlk, d0, d1, c0, c1, st, rs_j)) return 1;
  if (rescale && !rs_j) return rescale_dev(c, l, c0, c1, true, stream, 0);
  return 0;
}
extern "C" int gq_he_mul_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                             const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                             const gq_evk *rlk, void *stream) {
  return he_mul_impl(c, l, c0, c1, a0, a1, b0, b1, rlk, stream, false);
}
extern "C" int gq_he_mul_rs_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                                const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                                const gq_evk *rlk, void *stream) {
  return he_mul_impl(c, l, c0, c1, a0, a1, b0, b1, rlk, stream, true);
}

static int he_mulpt_impl(gq_ctx *c, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                         const uint64_t *a0, const uint64_t *a1, const uint64_t *m, void *stream, bool rescale) {
  if (he_check(c, l)) return 1;
  if (rescale && l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !m) GQ_FAIL("null argument");
  const unsigned rs_j = rescale ? rs_fusable(c, l) : 0;
  if (dim == 0 || dim > c->dimub) GQ_FAIL("dim out of range");
  if (!ntt_ok(c, 0, dim)) GQ_FAIL("modulus below 32 bits");
  cudaStream_t st = pick(c, stream);
  const unsigned n = c->n, W = c->W;
  if (ensure_slot(c, GQ_SLOT_LIMB_A, (size_t)4 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_LIMB_B, (size_t)3 * c->dimub * n * 8)) return 1;
  uint64_t *A = (uint64_t *)c->slot[GQ_SLOT_LIMB_A], *B = (uint64_t *)c->slot[GQ_SLOT_LIMB_B];
  const bool pre = crt_pre_ok(c, dim, c->qmod[l].dev);
  if (decompose_dev(c, 0, dim, ptrs(a0, a1, m), W, 3, A, st)) return 1;
  if (ntt_internal(c, 0, dim, 3, A, false, st)) return 1;
  if (fused_inv_ok(c, 0, dim)) {
    if (ntt_inv_fused(c, GQ_NTTF_PT, 0, dim, A, B, nullptr, nullptr, 1, st, pre ? dim : 0)) return 1;
  } else {
    gq_launch_mul_pt(c->d_primes, (int)n, (int)dim, B, A, st, pm_ok(c, 0, dim));
    c->launches++;
    if (ntt_internal(c, 0, dim, 2, B, true, st, 0, pre ? dim : 0)) return 1;
  }
  if (reconstruct_dev(c, dim, B, c->qmod[l].dev, ptrs(c0, c1), W, 2, st, 0, rs_j, pre)) return 1;
  if (rescale && !rs_j) return rescale_dev(c, l, c0, c1, true, stream, 0);
  return 0;
}
extern "C" int gq_he_mulpt_dev(gq_ctx *c, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                               const uint64_t *a0, const uint64_t *a1, const uint64_t *m, void *stream) {
  return he_mulpt_impl(c, l, dim, c0, c1, a0, a1, m, stream, false);
}
extern "C" int gq_he_mulpt_rs_dev(gq_ctx *c, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                                  const uint64_t *a0, const uint64_t *a1, const uint64_t *m, void *stream) {
  return he_mulpt_impl(c, l, dim, c0, c1, a0, a1, m, stream, true);
}

static int rescale_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, bool divide, void *stream,
                       unsigned ncoef) {
  if (c && ncoef == 0) ncoef = c->n;
  if (he_check(c, l)) return 1;
  if (l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1) GQ_FAIL("null argument");
  cudaStream_t st = pick(c, stream);
  const GqHostModulus &qm = c->qmod[l - 1];
  const bool dpow2 = (c->Delta & (c->Delta - 1)) == 0;
  if (qm.dev.is_pow2 && (dpow2 || !divide))
    gq_launch_rescale_pow2((int)ncoef, (int)c->W, ptrs(c0, c1), 2, divide ? c->logDelta : 0, qm.dev.k, st);
  else
    gq_launch_rescale_general((int)ncoef, (int)c->W, ptrs(c0, c1), 2, divide ? c->Delta : 1, qm.dev, st);
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}
extern "C" int gq_he_rs_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, void *s) {
  return rescale_dev(c, l, c0, c1, true, s, 0);
}
extern "C" int gq_he_moddown_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, void *s) {
  return rescale_dev(c, l, c0, c1, false, s, 0);
}

// he_add / he_sub / he_addpt / he_subpt / he_neg (src/he-add.c:32-142): coefficient-local, one launch
static int linear_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                      const uint64_t *b0, const uint64_t *b1, int op0, int op1, void *stream) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || (op0 < 2 && !b0) || (op1 < 2 && !b1)) GQ_FAIL("null argument");
  gq_launch_linear((int)c->n, (int)c->W, ptrs(c0, c1), ptrs(a0, a1), ptrs(b0 ? b0 : a0, b1 ? b1 : a1), 2,
                   op0, op1, c->qmod[l].dev, pick(c, stream));
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}
extern "C" int gq_he_add_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                             const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, void *s) {
  return linear_dev(c, l, c0, c1, a0, a1, b0, b1, 0, 0, s);
}
extern "C" int gq_he_sub_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                             const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, void *s) {
  return linear_dev(c, l, c0, c1, a0, a1, b0, b1, 1, 1, s);
}
extern "C" int gq_he_addpt_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                               const uint64_t *a1, const uint64_t *m, void *s) {
  return linear_dev(c, l, c0, c1, a0, a1, m, nullptr, 0, 3, s);
}
extern "C" int gq_he_subpt_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                               const uint64_t *a1, const uint64_t *m, void *s) {
  return linear_dev(c, l, c0, c1, a0, a1, m, nullptr, 1, 3, s);
}
extern "C" int gq_he_neg_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, void *s) {
  return linear_dev(c, l, c0, c1, c0, c1, nullptr, nullptr, 2, 2, s);
}

static int automorph_ks_dev(gq_ctx *c, unsigned l, uint64_t g, uint64_t *c0, uint64_t *c1,
                            const gq_evk *swk, void *stream) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !swk) GQ_FAIL("null argument");
  cudaStream_t st = pick(c, stream);
  const unsigned n = c->n, W = c->W;
  if (ensure_slot(c, GQ_SLOT_BIG, (size_t)5 * W * n * 8)) return 1;
  uint64_t *big = (uint64_t *)c->slot[GQ_SLOT_BIG];
  uint64_t *d0 = big + 3 * (size_t)W * n, *d1 = big + 4 * (size_t)W * n;
  gq_launch_automorph((int)n, (int)W, ptrs(d0, d1), ptrs(c0, c1), 2, g, st);
  c->launches++;
  // he_swk: only c0 receives the additive term (he-automorphism.c:73)
  return keyswitch_dev(c, l, d1, swk, d0, nullptr, c0, c1, st);
}
extern "C" int gq_he_rot_dev(gq_ctx *c, unsigned l, unsigned rot, uint64_t *c0, uint64_t *c1,
                             const gq_evk *rk, void *s) {
  if (!c) GQ_FAIL("null context");
  return automorph_ks_dev(c, l, galois_rot(c, rot), c0, c1, rk, s);
}
extern "C" int gq_he_conj_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const gq_evk *ck, void *s) {
  if (!c) GQ_FAIL("null context");
  return automorph_ks_dev(c, l, 2ull * c->n - 1, c0, c1, ck, s);
}

// Several rotations of ONE ciphertext (the baby steps of he_gemv, src/he-algo.c:62-66): identical bits
// to `count` separate he_rot calls, with decompose + forward NTT of c1 done once (gq_pointwise.cu,
// k_mul_evk_perm).  Outputs must not overlap the inputs.
extern "C" int gq_he_rot_hoisted_dev(gq_ctx *c, unsigned l, unsigned count, const unsigned *rots,
                                     uint64_t *const *c0_out, uint64_t *const *c1_out, const uint64_t *c0,
                                     const uint64_t *c1, const gq_evk *const *rk, void *stream) {
  if (he_check(c, l)) return 1;
  if (!rots || !c0_out || !c1_out || !c0 || !c1 || !rk) GQ_FAIL("null argument");
  for (unsigned k = 0; k < count; k++)
    if (!c0_out[k] || !c1_out[k] || !rk[k] || c0_out[k] == c0 || c1_out[k] == c1 || c0_out[k] == c1 || c1_out[k] == c0)
      GQ_FAIL("null or aliased output");
  if (count == 0) return 0;
  cudaStream_t st = pick(c, stream);
  const unsigned D2 = dim_ks(c, l), n = c->n, W = c->W;
  if (D2 > c->dimevk) GQ_FAIL("key-switch basis exceeds dimevk");
  if (!ntt_ok(c, 0, D2)) GQ_FAIL("modulus below 32 bits");
  if (c->basis[D2 - 1].dev.WT == 0) GQ_FAIL("RNS basis wider than 3072 bits is not supported by the CRT kernels");
  if (ensure_slot(c, GQ_SLOT_LIMB_A, (size_t)4 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_LIMB_B, (size_t)3 * c->dimub * n * 8) ||
      ensure_slot(c, GQ_SLOT_BIG, (size_t)5 * W * n * 8)) return 1;
  uint64_t *A = (uint64_t *)c->slot[GQ_SLOT_LIMB_A], *B = (uint64_t *)c->slot[GQ_SLOT_LIMB_B];
  uint64_t *d0 = (uint64_t *)c->slot[GQ_SLOT_BIG] + 3 * (size_t)W * n;
  if (decompose_dev(c, 0, D2, ptrs(c1), W, 1, A, st)) return 1;
  if (ntt_internal(c, 0, D2, 1, A, false, st)) return 1;
  const bool pre = crt_pre_ok(c, D2, c->qmod[l].dev) && c->basis[D2 - 1].ratio && D2 >= c->dim;
  const bool fused = fused_inv_ok(c, 0, D2);
  for (unsigned k = 0; k < count; k++) {
    const uint64_t g = galois_rot(c, rots[k]);
    gq_launch_automorph((int)n, (int)W, ptrs(d0), ptrs(c0), 1, g, st);
    c->launches++;
    if (fused) {
      if (ntt_inv_fused(c, GQ_NTTF_EVKPERM, 0, D2, A, B, rk[k]->p0, rk[k]->p1, g, st, pre ? D2 : 0)) return 1;
    } else {
      gq_launch_mul_evk_perm(c->d_primes, (int)n, (int)c->logn, (int)D2, B, A, rk[k]->p0, rk[k]->p1, g, st,
                             pm_ok(c, 0, D2));
      c->launches++;
      if (ntt_internal(c, 0, D2, 2, B, true, st, 0, pre ? D2 : 0)) return 1;
    }
    if (ks_epilogue_dev(c, l, D2, B, d0, nullptr, c0_out[k], c1_out[k], 0, st, 0, pre)) return 1;
  }
  return 0;
}

// he_gemv's ciphertext arithmetic (src/he-algo.c:47-90), device resident: baby-step / giant-step
// diagonal method with the ENCODED diagonals supplied by the caller (he_ecd is host floating point):
//   for i < n2: inner = sum_j he_mulpt(he_rot(ct, j), pts[i*n1+j]);  he_rot(inner, i*n1);  outer += inner
//   he_rs(outer)
// Same operations in the same order as the reference, each reduced exactly as there, so the result is
// bit-identical; the n1 baby rotations are computed once (the reference recomputes them for every i)
// and hoisted.  rk is indexed by rotation amount (entries 0..n1-1 and i*n1 are used).
extern "C" int gq_he_gemv_dev(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                              uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                              const gq_evk *const *rk, void *stream) {
  if (he_check(c, l)) return 1;
  if (l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rk || n1 == 0 || n2 == 0) GQ_FAIL("bad argument");
  for (unsigned j = 0; j < n1; j++) if (!rk[j]) GQ_FAIL("missing baby-step rotation key");
  for (unsigned i = 0; i < n2; i++) if (!rk[i * n1]) GQ_FAIL("missing giant-step rotation key");
  for (unsigned k = 0; k < n1 * n2; k++) if (!pts[k]) GQ_FAIL("missing diagonal");
  const size_t poly = (size_t)c->W * c->n;
  const size_t n = c->n;
  // giant steps are independent until they are summed: with compute lanes (gq_ctx_set_lanes) and the
  // context's own stream, giant step i runs on lane i % K; the sums are taken afterwards in the reference's
  // order on the context's stream
  const unsigned K = (c->parent || stream) ? 1u : (unsigned)c->lanes.size() + 1;
  // NTT-domain accumulation.  The reference computes inner_i = smod(sum_j smod(he_mulpt(rot_j(ct), pt_ij))) with one
  // decompose / NTT / inverse NTT / CRT per term (src/he-algo.c:65-79, src/he-mult.c:159-196).  Each he_mulpt is
  // smod_q(X_j), X_j the exact integer product, as long as |X_j| < P_dim / 2; smod_q of a sum of representatives is
  // smod_q of the exact sum, so inner_i = smod_q(sum_j X_j), and that sum is what ONE inverse NTT + CRT of
  // sum_j NTT(c_j) * NTT(m_ij) returns provided |sum_j X_j| < P_dimX / 2.  Both bounds follow from |c| <= q_l / 2 and
  // the caller's plaintext bound |m| < 2^pt_bits (gq_ctx_set_plaintext_bits; the drop-in measures it while it
  // marshals the diagonals): |X_j| <= n * 2^(bits(q_l) - 2) * 2^pt_bits.  Without a bound, or when a bound fails (the
  // reference's own he_mulpt would wrap modulo P_dim: reproduced only by the per-term path), the per-term path runs.
  unsigned dimX = 0;
  if (c->pt_bits && !getenv("GPQHE_B200_GEMV_PER_TERM") && dim >= 1 && dim <= c->dimub) {
    unsigned lg = 0;
    while ((1u << lg) < n1) lg++;
    const unsigned term = c->logn + c->ql[l].bits() - 1 + c->pt_bits;      // |X_j| < 2^term (one spare bit)
    const bool ref_exact = c->basis[dim - 1].dev.WT && term + 3 <= c->basis[dim - 1].dev.bitsP;
    for (unsigned D = dim; ref_exact && D <= c->dimub && !dimX; D++)
      if (c->basis[D - 1].dev.WT && term + lg + 3 <= c->basis[D - 1].dev.bitsP) dimX = D;
    if (dimX && (!fused_inv_ok(c, 0, dimX) || (2 * (size_t)n1 + (size_t)n1 * n2) * dimX * n * 8 > ((size_t)6 << 30))) dimX = 0;
  }
  const size_t hat_words = dimX ? (2 * (size_t)n1 + (size_t)n1 * n2) * dimX * n : 0;
  if (ensure_slot(c, GQ_SLOT_GEMV, ((2 * (size_t)n1 + 2 * (size_t)n2 + 2 * (size_t)K) * poly + hat_words) * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_GEMV];
  uint64_t *accs = base + 2 * (size_t)n1 * poly, *Ts = accs + 2 * (size_t)n2 * poly;
  uint64_t *Xhat = Ts + 2 * (size_t)K * poly, *Mhat = Xhat + 2 * (size_t)n1 * dimX * n;
  std::vector<unsigned> rots(n1);
  std::vector<uint64_t *> r0(n1), r1(n1);
  for (unsigned j = 0; j < n1; j++) { rots[j] = j; r0[j] = base + 2 * (size_t)j * poly; r1[j] = r0[j] + poly; }
  if (gq_he_rot_hoisted_dev(c, l, n1, rots.data(), r0.data(), r1.data(), a0, a1, rk, stream)) return 1;
  if (dimX) {
    // NTT(decompose(.)) of the 2*n1 rotated polynomials and of the n1*n2 encoded diagonals, over dimX limbs, once
    cudaStream_t st0 = pick(c, stream);
    std::vector<const uint64_t *> src;
    for (unsigned j = 0; j < n1; j++) { src.push_back(r0[j]); src.push_back(r1[j]); }
    for (unsigned k = 0; k < n1 * n2; k++) src.push_back(pts[k]);
    for (size_t g0 = 0; g0 < src.size(); g0 += 4) {
      const unsigned cnt = (unsigned)std::min<size_t>(4, src.size() - g0);
      if (decompose_dev(c, 0, dimX, ptrs(src[g0], cnt > 1 ? src[g0 + 1] : nullptr, cnt > 2 ? src[g0 + 2] : nullptr,
                                        cnt > 3 ? src[g0 + 3] : nullptr),
                        c->W, cnt, Xhat + g0 * dimX * n, st0)) return 1;
    }
    if (ntt_internal(c, 0, dimX, (unsigned)src.size(), Xhat, false, st0)) return 1;
  }
  if (K > 1 && gq_ctx_lanes_fork(c)) return 1;
  for (unsigned i = 0; i < n2; i++) {
    gq_ctx *ln = (K > 1) ? gq_ctx_lane(c, i % K) : c;
    void *ls = (K > 1) ? nullptr : stream;
    uint64_t *acc = accs + 2 * (size_t)i * poly, *T = Ts + 2 * (size_t)(i % K) * poly;
    if (dimX) {
      // one giant step: sum_j NTT(rot_j) * NTT(pt_ij) formed in the load of ONE inverse transform, one CRT
      cudaStream_t lst = pick(ln, ls);
      if (ensure_slot(ln, GQ_SLOT_LIMB_B, (size_t)3 * c->dimub * n * 8)) return 1;
      uint64_t *Y = (uint64_t *)ln->slot[GQ_SLOT_LIMB_B];
      const bool pre = crt_pre_ok(c, dimX, c->qmod[l].dev);
      if (ntt_inv_fused(ln, GQ_NTTF_GEMV, 0, dimX, Xhat, Y, Mhat + (size_t)i * n1 * dimX * n, nullptr, n1, lst,
                        pre ? dimX : 0)) return 1;
      if (reconstruct_dev(ln, dimX, Y, c->qmod[l].dev, ptrs(acc, acc + poly), c->W, 2, lst, 0, 0, pre)) return 1;
    } else {
      for (unsigned j = 0; j < n1; j++) {
        uint64_t *dst = (j == 0) ? acc : T;
        if (gq_he_mulpt_dev(ln, l, dim, dst, dst + poly, r0[j], r1[j], pts[i * n1 + j], ls)) return 1;
        if (j && linear_dev(ln, l, acc, acc + poly, acc, acc + poly, T, T + poly, 0, 0, ls)) return 1;
      }
    }
    if (gq_he_rot_dev(ln, l, i * n1, acc, acc + poly, rk[i * n1], ls)) return 1;
  }
  if (K > 1 && gq_ctx_lanes_join(c)) return 1;
  uint64_t *outer = accs;
  for (unsigned i = 1; i < n2; i++) {
    uint64_t *inner = accs + 2 * (size_t)i * poly;
    if (linear_dev(c, l, outer, outer + poly, outer, outer + poly, inner, inner + poly, 0, 0, stream)) return 1;
  }
  if (rescale_dev(c, l, outer, outer + poly, true, stream, 0)) return 1;
  cudaStream_t st = pick(c, stream);
  GQ_CUDA(cudaMemcpyAsync(c0, outer, poly * 8, cudaMemcpyDeviceToDevice, st));
  GQ_CUDA(cudaMemcpyAsync(c1, outer + poly, poly * 8, cudaMemcpyDeviceToDevice, st));
  return 0;
}

extern "C" int gq_he_gemv(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                          uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                          const gq_evk *const *rk) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rk || n1 == 0 || n2 == 0) GQ_FAIL("bad argument");
  const size_t poly = (size_t)c->W * c->n, np = (size_t)n1 * n2;
  if (ensure_slot(c, GQ_SLOT_PIPE, (4 + np) * poly * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_PIPE];
  GQ_CUDA(cudaMemcpyAsync(base, a0, poly * 8, cudaMemcpyHostToDevice, c->stream));
  GQ_CUDA(cudaMemcpyAsync(base + poly, a1, poly * 8, cudaMemcpyHostToDevice, c->stream));
  std::vector<const uint64_t *> dp(np);
  for (size_t k = 0; k < np; k++) {
    if (!pts[k]) GQ_FAIL("missing diagonal");
    uint64_t *d = base + (4 + k) * poly;
    GQ_CUDA(cudaMemcpyAsync(d, pts[k], poly * 8, cudaMemcpyHostToDevice, c->stream));
    dp[k] = d;
  }
  if (gq_he_gemv_dev(c, l, n1, n2, dim, base + 2 * poly, base + 3 * poly, base, base + poly, dp.data(), rk, nullptr))
    return 1;
  GQ_CUDA(cudaMemcpyAsync(c0, base + 2 * poly, poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaMemcpyAsync(c1, base + 3 * poly, poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

// he_coeff2slot, src/he-bootstrap.c:359-409 (the linear-transform half of bootstrapping as the reference
// composes it):  ct_conj = he_conj(ct);  for part in {real, imag}:
//   he_gemv(U_conjT, ct) + he_gemv(U_T, ct_conj) -> he_add -> he_mulpt(const pt) -> he_rs
// The same operations in the same order, every intermediate ciphertext in HBM.  pts holds the encoded
// diagonals of U0_conjT, U0_T, U1_conjT, U1_T (4 x n1*n2, each set laid out as gq_he_gemv wants), cpt the
// encoded constant plaintext; dim_gemv / dim_pt are the he_mulpt limb counts (he-mult.c:168) at levels l and
// l-1.  Outputs are at level l-2.
extern "C" int gq_he_coeff2slot_dev(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim_gemv,
                                    unsigned dim_pt, uint64_t *re0, uint64_t *re1, uint64_t *im0, uint64_t *im1,
                                    const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                                    const uint64_t *cpt, const gq_evk *ck, const gq_evk *const *rk, void *stream) {
  if (he_check(c, l)) return 1;
  if (l < 2) GQ_FAIL("needs two levels");
  if (!re0 || !re1 || !im0 || !im1 || !a0 || !a1 || !pts || !cpt || !ck || !rk) GQ_FAIL("null argument");
  const size_t poly = (size_t)c->W * c->n, m = (size_t)n1 * n2;
  if (ensure_slot(c, GQ_SLOT_C2S, 6 * poly * 8)) return 1;
  uint64_t *cj = (uint64_t *)c->slot[GQ_SLOT_C2S], *t0 = cj + 2 * poly, *t1 = t0 + 2 * poly;
  cudaStream_t st = pick(c, stream);
  GQ_CUDA(cudaMemcpyAsync(cj, a0, poly * 8, cudaMemcpyDeviceToDevice, st));
  GQ_CUDA(cudaMemcpyAsync(cj + poly, a1, poly * 8, cudaMemcpyDeviceToDevice, st));
  if (gq_he_conj_dev(c, l, cj, cj + poly, ck, stream)) return 1;
  for (int part = 0; part < 2; part++) {
    uint64_t *o0 = part ? im0 : re0, *o1 = part ? im1 : re1;
    if (gq_he_gemv_dev(c, l, n1, n2, dim_gemv, t0, t0 + poly, a0, a1, pts + (2 * part) * m, rk, stream)) return 1;
    if (gq_he_gemv_dev(c, l, n1, n2, dim_gemv, t1, t1 + poly, cj, cj + poly, pts + (2 * part + 1) * m, rk, stream)) return 1;
    if (linear_dev(c, l - 1, o0, o1, t0, t0 + poly, t1, t1 + poly, 0, 0, stream)) return 1;
    if (he_mulpt_impl(c, l - 1, dim_pt, o0, o1, o0, o1, cpt, stream, true)) return 1;
  }
  return 0;
}

extern "C" int gq_he_coeff2slot(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim_gemv,
                                unsigned dim_pt, uint64_t *re0, uint64_t *re1, uint64_t *im0, uint64_t *im1,
                                const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                                const uint64_t *cpt, const gq_evk *ck, const gq_evk *const *rk) {
  if (he_check(c, l)) return 1;
  if (!re0 || !re1 || !im0 || !im1 || !a0 || !a1 || !pts || !cpt || !ck || !rk || n1 == 0 || n2 == 0)
    GQ_FAIL("bad argument");
  const size_t poly = (size_t)c->W * c->n, np = 4 * (size_t)n1 * n2;
  if (ensure_slot(c, GQ_SLOT_PIPE, (7 + np) * poly * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_PIPE];
  GQ_CUDA(cudaMemcpyAsync(base, a0, poly * 8, cudaMemcpyHostToDevice, c->stream));
  GQ_CUDA(cudaMemcpyAsync(base + poly, a1, poly * 8, cudaMemcpyHostToDevice, c->stream));
  GQ_CUDA(cudaMemcpyAsync(base + 2 * poly, cpt, poly * 8, cudaMemcpyHostToDevice, c->stream));
  std::vector<const uint64_t *> dp(np);
  for (size_t k = 0; k < np; k++) {
    if (!pts[k]) GQ_FAIL("missing diagonal");
    uint64_t *d = base + (7 + k) * poly;
    GQ_CUDA(cudaMemcpyAsync(d, pts[k], poly * 8, cudaMemcpyHostToDevice, c->stream));
    dp[k] = d;
  }
  uint64_t *o = base + 3 * poly;
  if (gq_he_coeff2slot_dev(c, l, n1, n2, dim_gemv, dim_pt, o, o + poly, o + 2 * poly, o + 3 * poly, base, base + poly,
                           dp.data(), base + 2 * poly, ck, rk, nullptr)) return 1;
  uint64_t *dst[4] = {re0, re1, im0, im1};
  for (int k = 0; k < 4; k++)
    GQ_CUDA(cudaMemcpyAsync(dst[k], o + (size_t)k * poly, poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

// he_exp (src/he-algo.c:363-458), device resident: act = he_rs(he_mulpt(ct, pts[0])); the degree-7 Taylor
// polynomial of he_exp_taylor (:364-433) in the reference's own evaluation order (ct2, ct4, ct01, ct23, ct0123,
// ct45, ct67, ct4567 - every he_mul / he_mulpt / he_rs / he_moddown / he_add / he_addpt at the level it has
// there); then `iter` squarings.  pts (encoded by the caller, host floating point): 0 = he_ecd of a/2^iter in
// every slot, 1..7 = he_const_pt of 1, 3, 1/6, 5, 1/120, 7, 1/5040.  dim_a / dim_t = he_mulpt limb counts
// (he-mult.c:168) at levels l and l-1.  Result at level l-4-iter.  The output may alias the input.
extern "C" int gq_he_exp_dev(gq_ctx *c, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, uint64_t *c0,
                             uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                             const gq_evk *rlk, void *stream) {
  if (he_check(c, l)) return 1;
  if (l < 4 + iter) GQ_FAIL("he_exp needs 4 + iter levels");
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rlk) GQ_FAIL("null argument");
  for (int k = 0; k < 8; k++) if (!pts[k]) GQ_FAIL("missing plaintext");
  const size_t poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_EXP, 18 * poly * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_EXP];
  struct Ct { uint64_t *p0, *p1; };
  auto ct = [&](int k) { Ct x{base + 2 * (size_t)k * poly, base + (2 * (size_t)k + 1) * poly}; return x; };
  const Ct act = ct(0), ct2 = ct(1), ct4 = ct(2), ct01 = ct(3), ct23 = ct(4), ct0123 = ct(5), ct45 = ct(6),
           ct67 = ct(7), ct4567 = ct(8);
  void *s = stream;
  // every he_mul of he_exp_taylor is followed by he_rs on its result: one fused epilogue
  auto mulrs = [&](unsigned lv, Ct d, Ct x, Ct y) { return he_mul_impl(c, lv, d.p0, d.p1, x.p0, x.p1, y.p0, y.p1, rlk, s, true); };
  auto md = [&](unsigned lv, Ct x) { return rescale_dev(c, lv, x.p0, x.p1, false, s, 0); };
  auto add = [&](unsigned lv, Ct d, Ct x, Ct y) { return linear_dev(c, lv, d.p0, d.p1, x.p0, x.p1, y.p0, y.p1, 0, 0, s); };
  // dest = he_rs(he_mulpt(he_addpt(act, pts[ka]), pts[km]))  (he-algo.c:385-388 and its three siblings)
  auto lin = [&](unsigned lv, Ct d, int ka, int km) {
    if (linear_dev(c, lv, d.p0, d.p1, act.p0, act.p1, pts[ka], nullptr, 0, 3, s)) return 1;
    return he_mulpt_impl(c, lv, dim_t, d.p0, d.p1, d.p0, d.p1, pts[km], s, true);
  };
  if (he_mulpt_impl(c, l, dim_a, act.p0, act.p1, a0, a1, pts[0], s, true)) return 1;   // he-algo.c:448-449
  const unsigned k = l - 1;
  if (mulrs(k, ct2, act, act)) return 1;                               // ct^2, level k-1
  if (mulrs(k - 1, ct4, ct2, ct2)) return 1;                           // ct^4, level k-2
  if (lin(k, ct01, 1, 1) || md(k - 1, ct01)) return 1;                 // (ct+1)*1, level k-2
  if (lin(k, ct23, 2, 3) || mulrs(k - 1, ct23, ct2, ct23)) return 1;
  if (add(k - 2, ct0123, ct01, ct23) || md(k - 2, ct0123)) return 1;   // level k-3
  if (lin(k, ct45, 4, 5) || md(k - 1, ct45)) return 1;
  if (lin(k, ct67, 6, 7) || mulrs(k - 1, ct67, ct2, ct67)) return 1;
  if (add(k - 2, ct4567, ct45, ct67) || mulrs(k - 2, ct4567, ct4, ct4567)) return 1;
  const Ct out{c0, c1};
  if (add(k - 3, out, ct0123, ct4567)) return 1;
  for (unsigned t = 0; t < iter; t++)                                  // he-algo.c:451-454
    if (mulrs(k - 3 - t, out, out, out)) return 1;
  return 0;
}

// he_rlsin (src/he-bootstrap.c:444-475; the EvalMod step of he_bootstrap, :518-520):
//   e = he_exp(i*a*ct);  dest = he_rs(he_mulpt(he_sub(e, he_conj(e)), pts[8]))
// pts[0..7] as gq_he_exp_dev (pts[0] encodes i*a/2^iter), pts[8] = he_ecd of 1/(2*i*a) in every slot; dim_f =
// he_mulpt limb count at level l-4-iter.  Result at level l-5-iter.  The output may alias the input.
extern "C" int gq_he_rlsin_dev(gq_ctx *c, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, unsigned dim_f,
                               uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                               const uint64_t *const *pts, const gq_evk *rlk, const gq_evk *ck, void *stream) {
  if (he_check(c, l)) return 1;
  if (l < 5 + iter) GQ_FAIL("he_rlsin needs 5 + iter levels");
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rlk || !ck) GQ_FAIL("null argument");
  if (!pts[8]) GQ_FAIL("missing plaintext");
  const size_t poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_RLSIN, 4 * poly * 8)) return 1;
  uint64_t *e = (uint64_t *)c->slot[GQ_SLOT_RLSIN], *en = e + 2 * poly;
  const unsigned m = l - 4 - iter;
  cudaStream_t st = pick(c, stream);
  if (gq_he_exp_dev(c, l, iter, dim_a, dim_t, e, e + poly, a0, a1, pts, rlk, stream)) return 1;
  GQ_CUDA(cudaMemcpyAsync(en, e, 2 * poly * 8, cudaMemcpyDeviceToDevice, st));
  if (gq_he_conj_dev(c, m, en, en + poly, ck, stream)) return 1;
  if (linear_dev(c, m, c0, c1, e, e + poly, en, en + poly, 1, 1, stream)) return 1;
  return he_mulpt_impl(c, m, dim_f, c0, c1, c0, c1, pts[8], stream, true);
}

// he_slot2coeff (src/he-bootstrap.c:414-438): he_gemv(U0, ct0) + he_gemv(U1, ct1); pts = the encoded diagonals
// of U0 then U1 (2 x n1*n2).  Result at level l-1.
extern "C" int gq_he_slot2coeff_dev(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                                    uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0,
                                    const uint64_t *b1, const uint64_t *const *pts, const gq_evk *const *rk,
                                    void *stream) {
  if (he_check(c, l)) return 1;
  if (l < 1) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1 || !pts || !rk) GQ_FAIL("null argument");
  const size_t poly = (size_t)c->W * c->n, m = (size_t)n1 * n2;
  if (ensure_slot(c, GQ_SLOT_C2S, 6 * poly * 8)) return 1;
  uint64_t *t0 = (uint64_t *)c->slot[GQ_SLOT_C2S] + 2 * poly, *t1 = t0 + 2 * poly;
  if (gq_he_gemv_dev(c, l, n1, n2, dim, t0, t0 + poly, a0, a1, pts, rk, stream)) return 1;
  if (gq_he_gemv_dev(c, l, n1, n2, dim, t1, t1 + poly, b0, b1, pts + m, rk, stream)) return 1;
  return linear_dev(c, l - 1, c0, c1, t0, t0 + poly, t1, t1 + poly, 0, 0, stream);
}

// host-pointer forms: npts plaintexts and nin input polys up, 2 result polys down, one crossing
static int staged_call(gq_ctx *c, unsigned nin, const uint64_t *const *in, unsigned npts, const uint64_t *const *pts,
                       uint64_t *c0, uint64_t *c1, int (*body)(gq_ctx *, uint64_t *const *, const uint64_t *const *, void *),
                       void *arg) {
  const size_t poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_PIPE, (2 + (size_t)nin + npts) * poly * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_PIPE];
  std::vector<uint64_t *> io(2 + nin);
  std::vector<const uint64_t *> dp(npts);
  for (unsigned k = 0; k < 2 + nin; k++) io[k] = base + (size_t)k * poly;
  for (unsigned k = 0; k < nin; k++) {
    if (!in[k]) GQ_FAIL("null argument");
    GQ_CUDA(cudaMemcpyAsync(io[2 + k], in[k], poly * 8, cudaMemcpyHostToDevice, c->stream));
  }
  for (unsigned k = 0; k < npts; k++) {
    if (!pts[k]) GQ_FAIL("missing plaintext");
    uint64_t *d = base + (2 + (size_t)nin + k) * poly;
    GQ_CUDA(cudaMemcpyAsync(d, pts[k], poly * 8, cudaMemcpyHostToDevice, c->stream));
    dp[k] = d;
  }
  if (body(c, io.data(), dp.data(), arg)) return 1;
  GQ_CUDA(cudaMemcpyAsync(c0, io[0], poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaMemcpyAsync(c1, io[1], poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
struct ExpArgs { unsigned l, iter, dim_a, dim_t, dim_f, n1, n2; const gq_evk *rlk, *ck; const gq_evk *const *rk; };
extern "C" int gq_he_exp(gq_ctx *c, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, uint64_t *c0,
                         uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                         const gq_evk *rlk) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rlk) GQ_FAIL("null argument");
  ExpArgs a{l, iter, dim_a, dim_t, 0, 0, 0, rlk, nullptr, nullptr};
  const uint64_t *in[2] = {a0, a1};
  return staged_call(c, 2, in, 8, pts, c0, c1, [](gq_ctx *c, uint64_t *const *io, const uint64_t *const *dp, void *p) {
    const ExpArgs &a = *(const ExpArgs *)p;
    return gq_he_exp_dev(c, a.l, a.iter, a.dim_a, a.dim_t, io[0], io[1], io[2], io[3], dp, a.rlk, nullptr);
  }, &a);
}
extern "C" int gq_he_rlsin(gq_ctx *c, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, unsigned dim_f,
                           uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                           const uint64_t *const *pts, const gq_evk *rlk, const gq_evk *ck) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !pts || !rlk || !ck) GQ_FAIL("null argument");
  ExpArgs a{l, iter, dim_a, dim_t, dim_f, 0, 0, rlk, ck, nullptr};
  const uint64_t *in[2] = {a0, a1};
  return staged_call(c, 2, in, 9, pts, c0, c1, [](gq_ctx *c, uint64_t *const *io, const uint64_t *const *dp, void *p) {
    const ExpArgs &a = *(const ExpArgs *)p;
    return gq_he_rlsin_dev(c, a.l, a.iter, a.dim_a, a.dim_t, a.dim_f, io[0], io[1], io[2], io[3], dp, a.rlk, a.ck, nullptr);
  }, &a);
}
extern "C" int gq_he_slot2coeff(gq_ctx *c, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                                uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0,
                                const uint64_t *b1, const uint64_t *const *pts, const gq_evk *const *rk) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1 || !pts || !rk || n1 == 0 || n2 == 0) GQ_FAIL("bad argument");
  ExpArgs a{l, 0, dim, 0, 0, n1, n2, nullptr, nullptr, rk};
  const uint64_t *in[4] = {a0, a1, b0, b1};
  return staged_call(c, 4, in, 2 * n1 * n2, pts, c0, c1, [](gq_ctx *c, uint64_t *const *io, const uint64_t *const *dp, void *p) {
    const ExpArgs &a = *(const ExpArgs *)p;
    return gq_he_slot2coeff_dev(c, a.l, a.n1, a.n2, a.dim_a, io[0], io[1], io[2], io[3], io[4], io[5], dp, a.rk, nullptr);
  }, &a);
}

// -------------------------------------------------------------- HE primitives (host buffers)
// stage up to 5 big polys in, run, copy up to 2 out
struct HostIO {
  gq_ctx *c; uint64_t *base; size_t poly;
  uint64_t *in(int k) const { return base + (size_t)k * poly; }
};
static int io_begin(gq_ctx *c, HostIO *io, const uint64_t *const *src, int nsrc) {
  io->c = c; io->poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_IO, 7 * io->poly * 8)) return 1;
  io->base = (uint64_t *)c->slot[GQ_SLOT_IO];
  for (int k = 0; k < nsrc; k++)
    GQ_CUDA(cudaMemcpyAsync(io->in(k), src[k], io->poly * 8, cudaMemcpyHostToDevice, c->stream));
  return 0;
}
static int io_end(HostIO *io, uint64_t *c0, uint64_t *c1, int k0, int k1) {
  gq_ctx *c = io->c;
  GQ_CUDA(cudaMemcpyAsync(c0, io->in(k0), io->poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaMemcpyAsync(c1, io->in(k1), io->poly * 8, cudaMemcpyDeviceToHost, c->stream));
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int gq_he_mul(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                         const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1) GQ_FAIL("null argument");
  HostIO io; const uint64_t *src[4] = {a0, a1, b0, b1};
  if (io_begin(c, &io, src, 4)) return 1;
  if (gq_he_mul_dev(c, l, io.in(5), io.in(6), io.in(0), io.in(1), io.in(2), io.in(3), rlk, nullptr)) return 1;
  return io_end(&io, c0, c1, 5, 6);
}
// he_mul followed by he_rs on its result in ONE crossing (the pair every CKKS product makes:
// src/he-mult.c:100-150 then src/he-rescale.c:30-60)
extern "C" int gq_he_mul_rs(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                            const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk) {
  if (he_check(c, l)) return 1;
  if (l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1) GQ_FAIL("null argument");
  HostIO io; const uint64_t *src[4] = {a0, a1, b0, b1};
  if (io_begin(c, &io, src, 4)) return 1;
  if (he_mul_impl(c, l, io.in(5), io.in(6), io.in(0), io.in(1), io.in(2), io.in(3), rlk, nullptr, true)) return 1;
  return io_end(&io, c0, c1, 5, 6);
}

// `count` independent products (the row loop of he_gemv / a batch of ciphertext pairs), software
// pipelined over the two copy engines: the upload of product i+1 and the download of product i-1
// overlap the kernels of product i.  Two staging buffers per direction; host buffers should be
// page-locked (gq_host_alloc) for the copies to be asynchronous.  Synchronous for the caller.
static int pipe_init(gq_ctx *c) {
  if (c->s_in) return 0;
  GQ_CUDA(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
  GQ_CUDA(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
  for (auto &e : c->ev_pipe) GQ_CUDA(cudaEventCreate(&e));
  return 0;
}
extern "C" int gq_he_mul_batch(gq_ctx *c, unsigned l, unsigned count, uint64_t *const *c0, uint64_t *const *c1,
                               const uint64_t *const *a0, const uint64_t *const *a1,
                               const uint64_t *const *b0, const uint64_t *const *b1, const gq_evk *rlk,
                               int rescale) {
  if (he_check(c, l)) return 1;
  if (rescale && l == 0) GQ_FAIL("cannot go below level 0");
  if (!c0 || !c1 || !a0 || !a1 || !b0 || !b1 || !rlk) GQ_FAIL("null argument");
  for (unsigned i = 0; i < count; i++)
    if (!c0[i] || !c1[i] || !a0[i] || !a1[i] || !b0[i] || !b1[i]) GQ_FAIL("null polynomial pointer");
  if (count == 0) return 0;
  GQ_CUDA(cudaSetDevice(c->device));
  if (c->parent) GQ_FAIL("not a root context");
  if (pipe_init(c)) return 1;
  // products go round-robin over the compute lanes (3 unless the caller chose with gq_ctx_set_lanes or
  // GPQHE_B200_LANES); 2*K staging buffers, so buffer b always belongs to lane b % K and a lane never
  // waits for a buffer that another lane still computes on
  if (c->lanes.empty()) {
    const char *e = getenv("GPQHE_B200_LANES");
    unsigned want = e ? (unsigned)atoi(e) : 3u;
    if (want < 1 || want > 8) want = 3;          // not clamped to `count`: a first call with few products must not
    if (want > 1 && gq_ctx_set_lanes(c, want)) return 1;   // freeze the lane count of every later batch
  }
  const unsigned K = (unsigned)c->lanes.size() + 1, NB = 2 * K;
  const size_t poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_PIPE, (size_t)NB * 6 * poly * 8)) return 1;
  while (c->ev_batch.size() < 4 * (size_t)NB) {
    cudaEvent_t e = nullptr;
    GQ_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->ev_batch.push_back(e);
  }
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_PIPE];
  cudaEvent_t *ev = c->ev_batch.data();   // [4*b + {0: in_ready, 1: in_free, 2: out_ready, 3: out_free}]
  int rc = 0;
  // every exit drains the copy streams and the lanes first: copies and kernels in flight reference the caller's
  // host buffers and the shared staging slot, and the next call starts without waits
  auto drain = [&]() {
    cudaStreamSynchronize(c->s_in);
    for (unsigned k = 0; k < K; k++) cudaStreamSynchronize(gq_ctx_lane(c, k)->stream);
    return cudaStreamSynchronize(c->s_out);
  };
#define GQ_CUDA_B(call)                                                                  \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      gq_set_error(std::string(__func__) + ": " #call " -> " + cudaGetErrorString(e_));  \
      drain();                                                                           \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)
  for (unsigned i = 0; i < count && !rc; i++) {
    const unsigned bf = i % NB;
    gq_ctx *ln = gq_ctx_lane(c, i % K);
    uint64_t *in = base + (size_t)bf * 6 * poly, *out = in + 4 * poly;
    const uint64_t *src[4] = {a0[i], a1[i], b0[i], b1[i]};
    if (i >= NB) GQ_CUDA_B(cudaStreamWaitEvent(c->s_in, ev[4 * bf + 1], 0));
    for (int j = 0; j < 4;) {          // polys that are adjacent in host memory travel as one copy
      int e = j + 1;
      while (e < 4 && src[e] == src[e - 1] + poly) e++;
      GQ_CUDA_B(cudaMemcpyAsync(in + j * poly, src[j], (size_t)(e - j) * poly * 8, cudaMemcpyHostToDevice, c->s_in));
      j = e;
    }
    GQ_CUDA_B(cudaEventRecord(ev[4 * bf + 0], c->s_in));
    GQ_CUDA_B(cudaStreamWaitEvent(ln->stream, ev[4 * bf + 0], 0));
    if (i >= NB) GQ_CUDA_B(cudaStreamWaitEvent(ln->stream, ev[4 * bf + 3], 0));
    rc = he_mul_impl(ln, l, out, out + poly, in, in + poly, in + 2 * poly, in + 3 * poly, rlk, nullptr, rescale != 0);
    if (rc) break;
    GQ_CUDA_B(cudaEventRecord(ev[4 * bf + 1], ln->stream));
    GQ_CUDA_B(cudaEventRecord(ev[4 * bf + 2], ln->stream));
    GQ_CUDA_B(cudaStreamWaitEvent(c->s_out, ev[4 * bf + 2], 0));
    if (c1[i] == c0[i] + poly) {
      GQ_CUDA_B(cudaMemcpyAsync(c0[i], out, 2 * poly * 8, cudaMemcpyDeviceToHost, c->s_out));
    } else {
      GQ_CUDA_B(cudaMemcpyAsync(c0[i], out, poly * 8, cudaMemcpyDeviceToHost, c->s_out));
      GQ_CUDA_B(cudaMemcpyAsync(c1[i], out + poly, poly * 8, cudaMemcpyDeviceToHost, c->s_out));
    }
    GQ_CUDA_B(cudaEventRecord(ev[4 * bf + 3], c->s_out));
  }
  const cudaError_t de = drain();
#undef GQ_CUDA_B
  if (!rc && de != cudaSuccess) { gq_set_error(std::string(__func__) + ": " + cudaGetErrorString(de)); return 1; }
  return rc;
}

extern "C" int gq_he_mulpt(gq_ctx *c, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                           const uint64_t *a0, const uint64_t *a1, const uint64_t *m) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || !m) GQ_FAIL("null argument");
  HostIO io; const uint64_t *src[3] = {a0, a1, m};
  if (io_begin(c, &io, src, 3)) return 1;
  if (gq_he_mulpt_dev(c, l, dim, io.in(5), io.in(6), io.in(0), io.in(1), io.in(2), nullptr)) return 1;
  return io_end(&io, c0, c1, 5, 6);
}
extern "C" int gq_he_rot_hoisted(gq_ctx *c, unsigned l, unsigned count, const unsigned *rots,
                                 uint64_t *const *c0_out, uint64_t *const *c1_out, const uint64_t *c0,
                                 const uint64_t *c1, const gq_evk *const *rk) {
  if (he_check(c, l)) return 1;
  if (!rots || !c0_out || !c1_out || !c0 || !c1 || !rk) GQ_FAIL("null argument");
  if (count == 0) return 0;
  const size_t poly = (size_t)c->W * c->n;
  if (ensure_slot(c, GQ_SLOT_PIPE, (2 + 2 * (size_t)count) * poly * 8)) return 1;
  uint64_t *base = (uint64_t *)c->slot[GQ_SLOT_PIPE];
  GQ_CUDA(cudaMemcpyAsync(base, c0, poly * 8, cudaMemcpyHostToDevice, c->stream));
  GQ_CUDA(cudaMemcpyAsync(base + poly, c1, poly * 8, cudaMemcpyHostToDevice, c->stream));
  std::vector<uint64_t *> o0(count), o1(count);
  for (unsigned k = 0; k < count; k++) { o0[k] = base + (2 + 2 * (size_t)k) * poly; o1[k] = o0[k] + poly; }
  if (gq_he_rot_hoisted_dev(c, l, count, rots, o0.data(), o1.data(), base, base + poly, rk, nullptr)) return 1;
  for (unsigned k = 0; k < count; k++) {
    if (!c0_out[k] || !c1_out[k]) GQ_FAIL("null output");
    GQ_CUDA(cudaMemcpyAsync(c0_out[k], o0[k], poly * 8, cudaMemcpyDeviceToHost, c->stream));
    GQ_CUDA(cudaMemcpyAsync(c1_out[k], o1[k], poly * 8, cudaMemcpyDeviceToHost, c->stream));
  }
  GQ_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

static int linear_host(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                       const uint64_t *b0, const uint64_t *b1, int op0, int op1) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1 || !a0 || !a1 || (op0 < 2 && !b0) || (op1 < 2 && !b1)) GQ_FAIL("null argument");
  HostIO io; const uint64_t *src[4] = {a0, a1, b0, b1};
  int nsrc = 2;
  if (b0) { src[2] = b0; nsrc = 3; }
  if (b1) { src[3] = b1; nsrc = 4; }
  if (io_begin(c, &io, src, nsrc)) return 1;
  if (linear_dev(c, l, io.in(5), io.in(6), io.in(0), io.in(1), b0 ? io.in(2) : nullptr, b1 ? io.in(3) : nullptr,
                 op0, op1, nullptr)) return 1;
  return io_end(&io, c0, c1, 5, 6);
}
extern "C" int gq_he_add(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                         const uint64_t *b0, const uint64_t *b1) { return linear_host(c, l, c0, c1, a0, a1, b0, b1, 0, 0); }
extern "C" int gq_he_sub(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                         const uint64_t *b0, const uint64_t *b1) { return linear_host(c, l, c0, c1, a0, a1, b0, b1, 1, 1); }
extern "C" int gq_he_addpt(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                           const uint64_t *m) { return linear_host(c, l, c0, c1, a0, a1, m, nullptr, 0, 3); }
extern "C" int gq_he_subpt(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                           const uint64_t *m) { return linear_host(c, l, c0, c1, a0, a1, m, nullptr, 1, 3); }
extern "C" int gq_he_neg(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1) {
  return linear_host(c, l, c0, c1, c0, c1, nullptr, nullptr, 2, 2);
}

static int inplace_host(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, int op, unsigned rot,
                        const gq_evk *k) {
  if (he_check(c, l)) return 1;
  if (!c0 || !c1) GQ_FAIL("null argument");
  HostIO io; const uint64_t *src[2] = {c0, c1};
  if (io_begin(c, &io, src, 2)) return 1;
  int r = 1;
  switch (op) {
    case 0: r = gq_he_rs_dev(c, l, io.in(0), io.in(1), nullptr); break;
    case 1: r = gq_he_moddown_dev(c, l, io.in(0), io.in(1), nullptr); break;
    case 2: r = gq_he_rot_dev(c, l, rot, io.in(0), io.in(1), k, nullptr); break;
    case 3: r = gq_he_conj_dev(c, l, io.in(0), io.in(1), k, nullptr); break;
  }
  if (r) return 1;
  return io_end(&io, c0, c1, 0, 1);
}
extern "C" int gq_he_rs(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1) { return inplace_host(c, l, c0, c1, 0, 0, nullptr); }
extern "C" int gq_he_moddown(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1) { return inplace_host(c, l, c0, c1, 1, 0, nullptr); }
extern "C" int gq_he_rot(gq_ctx *c, unsigned l, unsigned rot, uint64_t *c0, uint64_t *c1, const gq_evk *rk) {
  return inplace_host(c, l, c0, c1, 2, rot, rk);
}
extern "C" int gq_he_conj(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, const gq_evk *ck) {
  return inplace_host(c, l, c0, c1, 3, 0, ck);
}

// -------------------------------------------------- stage-level entry points (device pointers)
// One call = one pipeline stage on a limb range (NTT domain) or on a coefficient slice (big-integer
// domain).  They exist for the limb-sharded multi-GPU pipeline (gpqhe_b200/sharded.py): every GPU
// runs the limb-local stages on the limbs it owns and the coefficient-local stages on its
// coefficient slice, with an all-to-all between the two domains.
extern "C" int gq_stage_decompose_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys,
                                      const uint64_t *const *in, unsigned W, unsigned ncoef,
                                      uint64_t *out, void *stream) {
  if (!c || !in || !out || npolys == 0 || npolys > 4 || nlimbs == 0) GQ_FAIL("bad argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  GqPtrs4 p = ptrs(in[0], npolys > 1 ? in[1] : nullptr, npolys > 2 ? in[2] : nullptr, npolys > 3 ? in[3] : nullptr);
  return decompose_dev(c, limb0, nlimbs, p, W, npolys, out, pick(c, stream), ncoef);
}

extern "C" int gq_stage_ntt_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys,
                                uint64_t *data, int inverse, void *stream) {
  if (!c || !data || npolys == 0) GQ_FAIL("bad argument");
  // inverse: bit 0 = inverse transform; +2 / +4 = profiling only (just the tile / just the strided
  // kernel of a two-kernel transform)
  return ntt_internal(c, limb0, nlimbs, npolys, data, (inverse & 1) != 0, pick(c, stream), (inverse >> 1) & 3);
}

extern "C" int gq_stage_tensor_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y,
                                   const uint64_t *x, void *stream) {
  if (!c || !y || !x) GQ_FAIL("bad argument");
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  gq_launch_tensor(c->d_primes + limb0, (int)c->n, (int)nlimbs, y, x, pick(c, stream), pm_ok(c, limb0, nlimbs));
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

extern "C" int gq_stage_mul_evk_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y,
                                    const uint64_t *x, const gq_evk *evk, void *stream) {
  if (!c || !y || !x || !evk) GQ_FAIL("bad argument");
  if (limb0 + nlimbs > c->dimevk || !ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid");
  const size_t off = (size_t)limb0 * c->n;
  gq_launch_mul_evk(c->d_primes + limb0, (int)c->n, (int)nlimbs, y, x, evk->p0 + off, evk->p1 + off,
                    pick(c, stream), pm_ok(c, limb0, nlimbs));
  c->launches++;
  GQ_CHECK_LAUNCH(c);
  return 0;
}

// tensor product + inverse NTT of a limb range in one pass (the pointwise stage rides on the load of the inverse
// transform's tiles): x[4][nlimbs][n] (forward-transformed) -> y[3][nlimbs][n] (coefficient domain, canonical)
extern "C" int gq_stage_tensor_inv_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y,
                                       const uint64_t *x, void *stream) {
  if (!c || !y || !x) GQ_FAIL("bad argument");
  if (!ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or modulus below 32 bits");
  cudaStream_t st = pick(c, stream);
  if (fused_inv_ok(c, limb0, nlimbs))
    return ntt_inv_fused(c, GQ_NTTF_TENSOR, limb0, nlimbs, x, y, nullptr, nullptr, 1, st, 0);
  if (gq_stage_tensor_dev(c, limb0, nlimbs, y, x, stream)) return 1;
  return ntt_internal(c, limb0, nlimbs, 3, y, true, st);
}
// x * evk.p0, x * evk.p1 + inverse NTT of a limb range in one pass: x[nlimbs][n] -> y[2][nlimbs][n]
extern "C" int gq_stage_mul_evk_inv_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y,
                                        const uint64_t *x, const gq_evk *evk, void *stream) {
  if (!c || !y || !x || !evk) GQ_FAIL("bad argument");
  if (limb0 + nlimbs > c->dimevk || !ntt_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid");
  cudaStream_t st = pick(c, stream);
  const size_t off = (size_t)limb0 * c->n;
  if (fused_inv_ok(c, limb0, nlimbs))
    return ntt_inv_fused(c, GQ_NTTF_EVK, limb0, nlimbs, x, y, evk->p0 + off, evk->p1 + off, 1, st, 0);
  if (gq_stage_mul_evk_dev(c, limb0, nlimbs, y, x, evk, stream)) return 1;
  return ntt_internal(c, limb0, nlimbs, 2, y, true, st);
}

extern "C" int gq_stage_reconstruct_dev(gq_ctx *c, unsigned D, unsigned l, unsigned npolys,
                                        const uint64_t *rhat, uint64_t *const *out, unsigned ncoef,
                                        void *stream) {
  if (he_check(c, l)) return 1;
  if (!rhat || !out || npolys == 0 || npolys > 4) GQ_FAIL("bad argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  GqPtrs4 p = ptrs(out[0], npolys > 1 ? out[1] : nullptr, npolys > 2 ? out[2] : nullptr, npolys > 3 ? out[3] : nullptr);
  return reconstruct_dev(c, D, rhat, c->qmod[l].dev, p, c->W, npolys, pick(c, stream), ncoef);
}

extern "C" int gq_stage_ks_epilogue_dev(gq_ctx *c, unsigned l, const uint64_t *chat,
                                        const uint64_t *add0, const uint64_t *add1, uint64_t *c0,
                                        uint64_t *c1, unsigned ncoef, void *stream) {
  if (he_check(c, l)) return 1;
  if (!chat || !c0 || !c1) GQ_FAIL("null argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  return ks_epilogue_dev(c, l, dim_ks(c, l), chat, add0, add1, c0, c1, ncoef, pick(c, stream));
}

// key-switch epilogue followed by he_rs on the slice, one kernel when the moduli are powers of two
extern "C" int gq_stage_ks_epilogue_rs_dev(gq_ctx *c, unsigned l, const uint64_t *chat,
                                           const uint64_t *add0, const uint64_t *add1, uint64_t *c0,
                                           uint64_t *c1, unsigned ncoef, void *stream) {
  if (he_check(c, l)) return 1;
  if (l == 0) GQ_FAIL("cannot go below level 0");
  if (!chat || !c0 || !c1) GQ_FAIL("null argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  const unsigned rs_j = rs_fusable(c, l);
  if (ks_epilogue_dev(c, l, dim_ks(c, l), chat, add0, add1, c0, c1, ncoef, pick(c, stream), rs_j)) return 1;
  if (!rs_j) return rescale_dev(c, l, c0, c1, true, stream, ncoef);
  return 0;
}

// ---- the same stages on PRE-SCALED residues: the inverse transforms end with the CRT pre-multiplication of the basis
// {p_0..p_(D-1)} folded into N^-1 (y_d = a_d (P/p_d)^-1 mod p_d), the CRT stages take such residues and run on the
// tensor cores where gq_crt_tc.cu applies.  gq_stage_prescale_ok says whether the forms exist for (D, l).
extern "C" int gq_stage_prescale_ok(gq_ctx *c, unsigned D, unsigned l, int keyswitch) {
  if (!c || !c->has_he || l > c->L || D == 0 || D > c->dimub) return 0;
  if (!crt_pre_ok(c, D, c->qmod[l].dev) || !fused_inv_ok(c, 0, D)) return 0;
  if (keyswitch && (!c->basis[D - 1].ratio || D < c->dim || D != dim_ks(c, l))) return 0;
  return 1;
}
extern "C" int gq_stage_tensor_inv_pre_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned D, uint64_t *y,
                                           const uint64_t *x, void *stream) {
  if (!c || !y || !x) GQ_FAIL("bad argument");
  if (limb0 + nlimbs > D || !fused_inv_ok(c, limb0, nlimbs) || D > c->dimub || !c->basis[D - 1].fin_s)
    GQ_FAIL("no pre-scaled inverse transform for this limb range / basis (gq_stage_prescale_ok)");
  return ntt_inv_fused(c, GQ_NTTF_TENSOR, limb0, nlimbs, x, y, nullptr, nullptr, 1, pick(c, stream), D);
}
extern "C" int gq_stage_mul_evk_inv_pre_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned D, uint64_t *y,
                                            const uint64_t *x, const gq_evk *evk, void *stream) {
  if (!c || !y || !x || !evk) GQ_FAIL("bad argument");
  if (limb0 + nlimbs > D || limb0 + nlimbs > c->dimevk || !fused_inv_ok(c, limb0, nlimbs) || D > c->dimub ||
      !c->basis[D - 1].fin_s)
    GQ_FAIL("no pre-scaled inverse transform for this limb range / basis (gq_stage_prescale_ok)");
  const size_t off = (size_t)limb0 * c->n;
  return ntt_inv_fused(c, GQ_NTTF_EVK, limb0, nlimbs, x, y, evk->p0 + off, evk->p1 + off, 1, pick(c, stream), D);
}
extern "C" int gq_stage_reconstruct_pre_dev(gq_ctx *c, unsigned D, unsigned l, unsigned npolys,
                                            const uint64_t *rhat, uint64_t *const *out, unsigned ncoef,
                                            void *stream) {
  if (he_check(c, l)) return 1;
  if (!rhat || !out || npolys == 0 || npolys > 4) GQ_FAIL("bad argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  if (!gq_stage_prescale_ok(c, D, l, 0)) GQ_FAIL("pre-scaled residues are not supported for this basis / level");
  GqPtrs4 p = ptrs(out[0], npolys > 1 ? out[1] : nullptr, npolys > 2 ? out[2] : nullptr, npolys > 3 ? out[3] : nullptr);
  return reconstruct_dev(c, D, rhat, c->qmod[l].dev, p, c->W, npolys, pick(c, stream), ncoef, 0, true);
}
extern "C" int gq_stage_ks_epilogue_pre_dev(gq_ctx *c, unsigned l, const uint64_t *chat, const uint64_t *add0,
                                            const uint64_t *add1, uint64_t *c0, uint64_t *c1, unsigned ncoef,
                                            int rescale, void *stream) {
  if (he_check(c, l)) return 1;
  if (rescale && l == 0) GQ_FAIL("cannot go below level 0");
  if (!chat || !c0 || !c1) GQ_FAIL("null argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  if (!gq_stage_prescale_ok(c, dim_ks(c, l), l, 1)) GQ_FAIL("pre-scaled residues are not supported for this level");
  const unsigned rs_j = rescale ? rs_fusable(c, l) : 0;
  if (ks_epilogue_dev(c, l, dim_ks(c, l), chat, add0, add1, c0, c1, ncoef, pick(c, stream), rs_j, true)) return 1;
  if (rescale && !rs_j) return rescale_dev(c, l, c0, c1, true, stream, ncoef);
  return 0;
}

extern "C" int gq_stage_rescale_dev(gq_ctx *c, unsigned l, uint64_t *c0, uint64_t *c1, int divide,
                                    unsigned ncoef, void *stream) {
  if (!c) GQ_FAIL("null context");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  return rescale_dev(c, l, c0, c1, divide != 0, stream, ncoef);
}

// ---------------------------------------------- fused compute + exchange of the limb-sharded pipeline
// The same stages as gq_stage_*_dev with the exchange folded into the kernels (gq_p2p.h): `send` makes the producing
// kernel store its results straight into the peers' receive slabs over NVLink and publish the epoch from its last
// block; `wait` makes the consuming kernel poll this rank's flags before it reads its slab.  One launch where the
// unfused plan has scatter + signal (producer) or wait + kernel (consumer).
static int to_peer_out(const gq_ctx *c, const gq_p2p_xchg *x, unsigned nloc, GqPeerOut *po) {
  if (!x || !x->dst || !x->off || !x->flags || !x->epoch_dev || !x->err_dev || !x->ticket_dev) GQ_FAIL("incomplete send descriptor");
  if (x->G == 0 || x->G > GQ_P2P_MAX_RANKS || x->me >= x->G || x->slot >= 8) GQ_FAIL("bad rank / slot in send descriptor");
  if (nloc == 0 || nloc * x->G != c->n) GQ_FAIL("coefficient slice does not match the rank count");
  for (unsigned s = 0; s < GQ_P2P_MAX_RANKS; s++) {
    po->dst.base[s] = s < x->G ? (uint64_t *)x->dst[s] : nullptr;
    po->flags[s] = s < x->G ? (uint64_t *)x->flags[s] : nullptr;
  }
  for (unsigned s = 0; s <= GQ_P2P_MAX_RANKS; s++) po->dst.off[s] = x->off[s <= x->G ? s : x->G];
  po->G = (int)x->G; po->me = (int)x->me; po->nloc = (int)nloc; po->n = (int)c->n; po->D = (int)x->D;
  if (nloc & (nloc - 1)) GQ_FAIL("coefficient slice is not a power of two");
  po->lg_nloc = 0;
  while ((1u << po->lg_nloc) < nloc) po->lg_nloc++;
  po->slot = (int)x->slot; po->epoch = x->epoch_dev; po->err = x->err_dev; po->ticket = x->ticket_dev;
  return 0;
}
static int to_wait(const gq_p2p_xchg *x, GqWaitIn *w) {
  if (!x || !x->flags_local || !x->epoch_dev || !x->err_dev) GQ_FAIL("incomplete wait descriptor");
  if (x->G == 0 || x->G > GQ_P2P_MAX_RANKS || x->slot >= 8) GQ_FAIL("bad rank count / slot in wait descriptor");
  w->flags = x->flags_local; w->slot = (int)x->slot; w->G = (int)x->G; w->epoch = x->epoch_dev; w->err = x->err_dev;
  w->spin_limit = 20000000ull;          // 4 s of 200 ns polls; GPQHE_B200_P2P_TIMEOUT_MS overrides
  if (const char *e = getenv("GPQHE_B200_P2P_TIMEOUT_MS")) {
    const long long ms = atoll(e);
    if (ms > 0) w->spin_limit = (unsigned long long)ms * 5000ull;
  }
  return 0;
}

// 1 when every stage of he_mul at level l has its fused form for this context (two-kernel transforms, 28-bit digit
// decomposition, fast CRT kernels, pseudo-Mersenne pointwise loaders), else 0: the caller then uses the unfused plan
extern "C" int gq_p2p_fused_ok(gq_ctx *c, unsigned l) {
  if (!c || !c->has_he || l > c->L) return 0;
  const unsigned D1 = dim_tensor(c, l), D2 = dim_ks(c, l);
  if (c->logn <= 12 || c->W > 7 || !c->fast_crt || getenv("GPQHE_B200_DEC32")) return 0;
  if (D1 > c->dimub || D2 > c->dimub || D2 > c->dimevk) return 0;
  if (!fused_inv_ok(c, 0, D1) || !fused_inv_ok(c, 0, D2)) return 0;
  for (unsigned d = 0; d < D2; d++) if (c->primes[d].bits < 33 || c->primes[d].bits > 60) return 0;
  const GqModulus &q = c->qmod[l].dev;
  if (!q.is_pow2 || q.k < 1 || !c->basis[D1 - 1].dev.F || !c->basis[D2 - 1].dev.F) return 0;
  if (c->basis[c->dim - 1].dev.WT > c->basis[D2 - 1].dev.WT) return 0;
  return 1;
}

extern "C" int gq_p2p_decompose_dev(gq_ctx *c, unsigned nlimbs, unsigned npolys, const uint64_t *const *in, unsigned W,
                                    unsigned ncoef, const gq_p2p_xchg *send, void *stream) {
  if (!c || !in || npolys == 0 || npolys > 4 || nlimbs == 0) GQ_FAIL("bad argument");
  GqPeerOut po;
  if (to_peer_out(c, send, ncoef, &po)) return 1;
  GqPtrs4 p = ptrs(in[0], npolys > 1 ? in[1] : nullptr, npolys > 2 ? in[2] : nullptr, npolys > 3 ? in[3] : nullptr);
  return decompose_dev(c, 0, nlimbs, p, W, npolys, nullptr, pick(c, stream), ncoef, &po);
}

extern "C" int gq_p2p_ntt_fwd_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, unsigned npolys, uint64_t *data,
                                  const gq_p2p_xchg *wait, void *stream) {
  if (!c || !data || npolys == 0) GQ_FAIL("bad argument");
  GqWaitIn w;
  if (to_wait(wait, &w)) return 1;
  return ntt_internal(c, limb0, nlimbs, npolys, data, false, pick(c, stream), 0, 0, &w);
}

extern "C" int gq_p2p_tensor_inv_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y_scratch, const uint64_t *x,
                                     const gq_p2p_xchg *send, void *stream) {
  if (!c || !y_scratch || !x) GQ_FAIL("bad argument");
  if (!fused_inv_ok(c, limb0, nlimbs)) GQ_FAIL("no fused tensor + inverse NTT for these primes");
  GqPeerOut po;
  if (to_peer_out(c, send, c->n / (send ? send->G : 1), &po)) return 1;
  if (send && send->pre && (send->D > c->dimub || !c->basis[send->D - 1].fin_s)) GQ_FAIL("no pre-scale tables for this basis");
  return ntt_inv_fused(c, GQ_NTTF_TENSOR, limb0, nlimbs, x, y_scratch, nullptr, nullptr, 1, pick(c, stream),
                       (send && send->pre) ? send->D : 0, &po);
}

extern "C" int gq_p2p_mul_evk_inv_dev(gq_ctx *c, unsigned limb0, unsigned nlimbs, uint64_t *y_scratch, const uint64_t *x,
                                      const gq_evk *evk, const gq_p2p_xchg *send, void *stream) {
  if (!c || !y_scratch || !x || !evk) GQ_FAIL("bad argument");
  if (limb0 + nlimbs > c->dimevk || !fused_inv_ok(c, limb0, nlimbs)) GQ_FAIL("limb range invalid or no fused form for these primes");
  GqPeerOut po;
  if (to_peer_out(c, send, c->n / (send ? send->G : 1), &po)) return 1;
  const size_t off = (size_t)limb0 * c->n;
  if (send && send->pre && (send->D > c->dimub || !c->basis[send->D - 1].fin_s)) GQ_FAIL("no pre-scale tables for this basis");
  return ntt_inv_fused(c, GQ_NTTF_EVK, limb0, nlimbs, x, y_scratch, evk->p0 + off, evk->p1 + off, 1, pick(c, stream),
                       (send && send->pre) ? send->D : 0, &po);
}

extern "C" int gq_p2p_reconstruct_dev(gq_ctx *c, unsigned D, unsigned l, unsigned npolys, const uint64_t *rhat,
                                      uint64_t *const *out, unsigned ncoef, const gq_p2p_xchg *wait, void *stream) {
  if (he_check(c, l)) return 1;
  if (!rhat || !out || npolys == 0 || npolys > 4) GQ_FAIL("bad argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  GqWaitIn w;
  if (to_wait(wait, &w)) return 1;
  GqPtrs4 p = ptrs(out[0], npolys > 1 ? out[1] : nullptr, npolys > 2 ? out[2] : nullptr, npolys > 3 ? out[3] : nullptr);
  const bool pre = wait && wait->pre;
  if (pre && !gq_stage_prescale_ok(c, D, l, 0)) GQ_FAIL("pre-scaled residues are not supported for this basis / level");
  return reconstruct_dev(c, D, rhat, c->qmod[l].dev, p, c->W, npolys, pick(c, stream), ncoef, 0, pre, &w);
}

extern "C" int gq_p2p_ks_epilogue_dev(gq_ctx *c, unsigned l, const uint64_t *chat, const uint64_t *add0,
                                      const uint64_t *add1, uint64_t *c0, uint64_t *c1, unsigned ncoef, int rescale,
                                      const gq_p2p_xchg *wait, void *stream) {
  if (he_check(c, l)) return 1;
  if (rescale && l == 0) GQ_FAIL("cannot go below level 0");
  if (!chat || !c0 || !c1) GQ_FAIL("null argument");
  if (ncoef == 0 || ncoef > c->n) GQ_FAIL("coefficient count out of range");
  GqWaitIn w;
  if (to_wait(wait, &w)) return 1;
  const unsigned rs_j = rescale ? rs_fusable(c, l) : 0;
  const bool pre = wait && wait->pre;
  if (pre && !gq_stage_prescale_ok(c, dim_ks(c, l), l, 1)) GQ_FAIL("pre-scaled residues are not supported for this level");
  if (ks_epilogue_dev(c, l, dim_ks(c, l), chat, add0, add1, c0, c1, ncoef, pick(c, stream), rs_j, pre, &w)) return 1;
  if (rescale && !rs_j) return rescale_dev(c, l, c0, c1, true, stream, ncoef);
  return 0;
}
