/*
 * gpqhe_dropin.c - the reference's hot symbols, re-implemented as thin C host code over
 * the C ABI of the B200 library (include/gpqhe_b200.h).
 *
 * dropin/Makefile compiles the reference's own host sources in place (from $(REF)/src,
 * nothing copied) with the hot definitions renamed to gpqhe_ref_<name>, and links this
 * file in their stead, so that dropin/_build/libgpqhe.so exports the reference's whole
 * symbol table with these entry points running on the GPU:
 *
 *   ntt, invntt                          src/ntt.c:37-73
 *   poly_rns_add, poly_rns_mul           src/poly.c:71-82
 *   rns_decompose                        src/rns.c:37-48
 *   poly_rns2mpi, poly_mul               src/poly.c:84-120
 *   poly_rot, poly_conj                  src/poly.c:263-283
 *   he_mul, he_mulpt                     src/he-mult.c:88-196
 *   he_rs, he_moddown                    src/he-rescale.c:33-70
 *   he_rot, he_conj                      src/he-automorphism.c:87-115
 *   he_add, he_sub, he_addpt, he_subpt, he_neg   src/he-add.c:32-142
 *   he_gemv                              src/he-algo.c:47-90 (device-resident loop)
 *   he_coeff2slot                        src/he-bootstrap.c:359-409 (device-resident composition)
 *   he_exp                               src/he-algo.c:363-458 (device-resident composition)
 *   he_rlsin                             src/he-bootstrap.c:444-475 (device-resident composition)
 *   polyctx_init/exit, hectx_init/exit   src/precomp.c:328-452 (wrappers: the reference
 *                                        builds its tables, we mirror them on the device)
 *   he_free_evk                          src/he-mem.c (wrapper: drops the device copy)
 *
 * Signatures, metadata updates (l, nu, B), aliasing rules and the stderr+abort() error
 * convention are the reference's.  Big integers cross the boundary as word-major
 * two's-complement arrays; the gcry_mpi <-> words marshalling below is the only host
 * arithmetic.  There is no CPU fallback: a failing device call aborts like the reference.
 */
#include "config.h"
#include "types.h"
#include "gpqhe.h"   /* pulls poly.h, params.h */

#include <complex.h>
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <unistd.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "gpqhe_b200.h"

extern struct poly_ctx polyctx;
extern struct he_ctx hectx;

/* the reference's own definitions, renamed by dropin/Makefile */
void gpqhe_ref_polyctx_init(unsigned int logn, MPI q);
void gpqhe_ref_polyctx_exit(void);
void gpqhe_ref_hectx_init(unsigned int logn, MPI q, unsigned int slots, uint64_t Delta);
void gpqhe_ref_hectx_exit(void);
void gpqhe_ref_he_free_evk(he_evk_t *evk);

#define DROPIN_MAX_W 32 /* GQ_MAX_W of the device library */

/* ------------------------------------------------------------------------- state */
struct evk_slot {
  const uint64_t *p0, *p1;  /* host addresses the key was uploaded from */
  uint64_t print;           /* sampled content fingerprint */
  gq_evk *dev;
};

static struct {
  gq_ctx *ctx;
  gq_info info;
  unsigned slots;
  uint64_t Delta;
  uint64_t *stage;          /* pinned staging: 7 big polys of GQ words each */
  size_t stage_words;
  struct evk_slot *evk;
  unsigned nevk, capevk;
  int fast_mpi;             /* direct limb access verified against the public API */
  unsigned long long marshal_in, marshal_out;
} G;

static void die(const char *where)
{
  fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1m%s\033[0m: %s\n", where, gq_last_error());
  errno = EINVAL;
  abort();
}
#define CK(call) do { if ((call) != 0) die(__func__); } while (0)

/* ----------------------------------------------------- gcry_mpi <-> words marshalling */
/* libgcrypt 1.10 (mpi/mpi-internal.h): used only after mpi_layout_selftest() has checked it
 * against the public API on this very library; otherwise the gcry_mpi_print/scan path runs. */
struct mpi_view { int alloced; int nlimbs; int sign; unsigned int flags; unsigned long *d; };

static void mpi_layout_selftest(void)
{
  static const unsigned char be[24] = {0x01, 0x23, 0x45, 0x67, 0x89, 0xab, 0xcd, 0xef, 0xfe, 0xdc, 0xba, 0x98,
                                       0x76, 0x54, 0x32, 0x10, 0x0f, 0x1e, 0x2d, 0x3c, 0x4b, 0x5a, 0x69, 0x78};
  MPI a = NULL;
  G.fast_mpi = 0;
  if (getenv("GPQHE_DROPIN_SLOW_MPI")) return;
  if (gcry_mpi_scan(&a, GCRYMPI_FMT_USG, be, sizeof be, NULL) || !a) return;
  const struct mpi_view *v = (const struct mpi_view *)a;
  int ok = v->nlimbs == 3 && v->alloced >= 3 && v->sign == 0 && v->d &&
           v->d[0] == 0x0f1e2d3c4b5a6978ul && v->d[1] == 0xfedcba9876543210ul && v->d[2] == 0x0123456789abcdeful;
  gcry_mpi_neg(a, a);
  ok = ok && v->sign == 1 && v->nlimbs == 3;
  gcry_mpi_set_ui(a, 0);
  ok = ok && v->nlimbs == 0;
  gcry_mpi_release(a);
  G.fast_mpi = ok;
}

/* a -> W little-endian words, two's complement, written with stride n (word-major poly) */
static void mpi_to_words(uint64_t *out, size_t stride, unsigned W, const MPI a, const char *who)
{
  uint64_t w[DROPIN_MAX_W + 1];
  unsigned nl;
  int neg;
  if (G.fast_mpi) {
    const struct mpi_view *v = (const struct mpi_view *)a;
    nl = (unsigned)v->nlimbs;
    neg = v->sign != 0 && nl != 0;
    if (nl > W) goto too_big;
    for (unsigned i = 0; i < nl; i++) w[i] = v->d[i];
  } else {
    unsigned char buf[8 * DROPIN_MAX_W];
    size_t len = 0;
    if (gcry_mpi_print(GCRYMPI_FMT_USG, buf, sizeof buf, &len, a)) goto too_big;
    nl = (unsigned)((len + 7) / 8);
    if (nl > W) goto too_big;
    for (unsigned i = 0; i < nl; i++) w[i] = 0;
    for (size_t k = 0; k < len; k++) w[k / 8] |= (uint64_t)buf[len - 1 - k] << (8 * (k % 8));
    neg = gcry_mpi_is_neg(a) && len != 0;
  }
  for (unsigned i = nl; i < W; i++) w[i] = 0;
  if (w[W - 1] >> 63) {   /* magnitude uses the sign bit: only -2^(64W-1) is representable */
    int is_min = neg && w[W - 1] == (1ull << 63);
    for (unsigned i = 0; i + 1 < W; i++) is_min = is_min && w[i] == 0;
    if (!is_min) goto too_big;
  }
  if (neg) {
    uint64_t c = 1;
    for (unsigned i = 0; i < W; i++) { uint64_t t = ~w[i] + c; c = (c && t == 0); w[i] = t; }
  }
  for (unsigned i = 0; i < W; i++) out[(size_t)i * stride] = w[i];
  return;
too_big:
  fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1m%s\033[0m: coefficient does not fit %u signed 64-bit words\n", who, W);
  errno = EINVAL;
  abort();
}

static void words_to_mpi(MPI a, const uint64_t *in, size_t stride, unsigned W)
{
  uint64_t w[DROPIN_MAX_W];
  for (unsigned i = 0; i < W; i++) w[i] = in[(size_t)i * stride];
  const int neg = (int)(w[W - 1] >> 63);
  if (neg) {
    uint64_t c = 1;
    for (unsigned i = 0; i < W; i++) { uint64_t t = ~w[i] + c; c = (c && t == 0); w[i] = t; }
  }
  unsigned nl = W;
  while (nl && w[nl - 1] == 0) nl--;
  if (G.fast_mpi) {
    struct mpi_view *v = (struct mpi_view *)a;
    if (v->flags == 0 && v->alloced >= (int)nl && (v->d || nl == 0)) {
      for (unsigned i = 0; i < nl; i++) v->d[i] = w[i];
      v->nlimbs = (int)nl;
      v->sign = neg && nl;
      return;
    }
  }
  if (nl == 0) { gcry_mpi_set_ui(a, 0); return; }
  unsigned char buf[8 * DROPIN_MAX_W];
  for (unsigned i = 0; i < nl; i++)
    for (unsigned k = 0; k < 8; k++) buf[8 * (nl - 1 - i) + (7 - k)] = (unsigned char)(w[i] >> (8 * k));
  MPI t = NULL;
  if (gcry_mpi_scan(&t, GCRYMPI_FMT_USG, buf, 8 * nl, NULL) || !t) { fprintf(stderr, "gcry_mpi_scan failed\n"); abort(); }
  if (neg) gcry_mpi_neg(t, t);
  gcry_mpi_snatch(a, t);
}

static unsigned words_for_bits(unsigned bits) { return bits / 64 + 1; }   /* bits magnitude + sign */

static unsigned poly_words(const MPI *a, unsigned n)
{
  unsigned mx = 0;
  for (unsigned i = 0; i < n; i++) { unsigned b = gcry_mpi_get_nbits(a[i]); if (b > mx) mx = b; }
  return words_for_bits(mx);
}

static unsigned mpi_words_unsigned(uint64_t *out, unsigned cap, const MPI a, const char *who)
{
  unsigned nb = gcry_mpi_get_nbits(a), nw = nb ? (nb + 63) / 64 : 1;
  if (nw + 1 > cap) { fprintf(stderr, "%s: modulus wider than %u words\n", who, cap); errno = EINVAL; abort(); }
  mpi_to_words(out, 1, nw + 1, a, who);
  return nw;
}

static uint64_t *stage(size_t words)
{
  if (G.stage_words < words) {
    if (G.stage) gq_host_free(G.stage);
    void *p = NULL;
    if (gq_host_alloc(&p, words * 8 + 64)) die("stage");
    G.stage = (uint64_t *)p;
    G.stage_words = words;
  }
  return G.stage;
}

/* The MPI <-> words conversion is the drop-in's own host cost (about 17 ns per coefficient: 7.7 ms of an he_mul
 * at N=2^16 whose kernels take 0.4 ms), and every coefficient is an independent object: large polynomials are
 * converted by a few threads (GPQHE_DROPIN_THREADS, default min(8, cores); 1 = the caller's thread only).
 * The reference is single-threaded and non-reentrant; the threads live only inside one conversion. */
struct par_job { void (*fn)(unsigned, unsigned, void *); void *arg; unsigned lo, hi; };
static void *par_tramp(void *p)
{
  struct par_job *j = p;
  j->fn(j->lo, j->hi, j->arg);
  return NULL;
}
static unsigned par_threads(void)
{
  static unsigned T;
  if (!T) {
    const char *e = getenv("GPQHE_DROPIN_THREADS");
    long c = e ? atol(e) : sysconf(_SC_NPROCESSORS_ONLN);
    if (c < 1) c = 1;
    if (c > 8) c = 8;
    T = (unsigned)c;
  }
  return T;
}
static void par_for(unsigned n, void (*fn)(unsigned, unsigned, void *), void *arg)
{
  const unsigned T = par_threads();
  if (T <= 1 || n < 8192) { fn(0, n, arg); return; }
  pthread_t th[8];
  struct par_job job[8];
  unsigned started = 0;
  for (unsigned t = 0; t < T; t++) {
    job[t].fn = fn; job[t].arg = arg;
    job[t].lo = (unsigned)((unsigned long long)n * t / T);
    job[t].hi = (unsigned)((unsigned long long)n * (t + 1) / T);
  }
  for (unsigned t = 1; t < T; t++) {
    if (pthread_create(&th[t], NULL, par_tramp, &job[t])) break;
    started = t;
  }
  fn(job[0].lo, job[0].hi, arg);
  for (unsigned t = started + 1; t < T; t++) fn(job[t].lo, job[t].hi, arg);   /* threads that could not be created */
  for (unsigned t = 1; t <= started; t++) pthread_join(th[t], NULL);
}

struct in_args { uint64_t *dst; const poly_mpi_t *a; unsigned W; const char *who; };
static void poly_in_range(unsigned lo, unsigned hi, void *p)
{
  const struct in_args *x = p;
  const unsigned n = polyctx.n;
  for (unsigned i = lo; i < hi; i++) mpi_to_words(x->dst + i, n, x->W, x->a->coeffs[i], x->who);
}
static void poly_in(uint64_t *dst, const poly_mpi_t *a, unsigned W, const char *who)
{
  struct in_args x = {dst, a, W, who};
  par_for(polyctx.n, poly_in_range, &x);
  G.marshal_in += polyctx.n;
}
struct out_args { poly_mpi_t *r; const uint64_t *src; unsigned W; };
static void poly_out_range(unsigned lo, unsigned hi, void *p)
{
  const struct out_args *x = p;
  const unsigned n = polyctx.n;
  for (unsigned i = lo; i < hi; i++) words_to_mpi(x->r->coeffs[i], x->src + i, n, x->W);
}
static void poly_out(poly_mpi_t *r, const uint64_t *src, unsigned W)
{
  struct out_args x = {r, src, W};
  par_for(polyctx.n, poly_out_range, &x);
  G.marshal_out += polyctx.n;
}

/* ------------------------------------------------------------------- context mirror */
static void ctx_drop(void)
{
  for (unsigned k = 0; k < G.nevk; k++) gq_evk_destroy(G.evk[k].dev);
  G.nevk = 0;
  if (G.ctx) { gq_ctx_destroy(G.ctx); G.ctx = NULL; }
}

/* Mirror the reference's freshly built tables (struct rns_ctx list: p, zetas, zetas_inv in
 * Montgomery form, src/precomp.c:244-264) so both sides use the very same roots. */
static void ctx_build(unsigned slots, uint64_t Delta)
{
  ctx_drop();
  mpi_layout_selftest();
  const unsigned K = polyctx.dimub;
  uint64_t *primes = malloc(K * sizeof *primes);
  const uint64_t **z = malloc(K * sizeof *z), **zi = malloc(K * sizeof *zi);
  const struct rns_ctx *r = polyctx.rns;
  for (unsigned d = 0; d < K; d++, r = r->next) { primes[d] = r->p; z[d] = r->zetas; zi[d] = r->zetas_inv; }
  uint64_t qw[DROPIN_MAX_W + 1];
  unsigned nq = mpi_words_unsigned(qw, DROPIN_MAX_W + 1, polyctx.q, "hectx_init");
  int dev = 0;
  const char *e = getenv("GPQHE_B200_DEVICE");
  if (e) dev = atoi(e);
  if (gq_ctx_create(&G.ctx, polyctx.logn, qw, nq, slots, Delta, GPQHE_LOGP, polyctx.logqub, dev, primes, z, zi))
    die(Delta ? "hectx_init" : "polyctx_init");
  CK(gq_ctx_info(G.ctx, &G.info));
  G.slots = slots; G.Delta = Delta;
  if (Delta && (G.info.L != hectx.L || G.info.dim != hectx.dim || G.info.dimevk != hectx.dimevk)) {
    fprintf(stderr, "gpqhe_dropin: device context disagrees with hectx (L %u/%u dim %u/%u dimevk %u/%u)\n",
            G.info.L, hectx.L, G.info.dim, hectx.dim, G.info.dimevk, hectx.dimevk);
    abort();
  }
  free(primes); free(z); free(zi);
}

void polyctx_init(unsigned int logn, MPI q)
{
  gpqhe_ref_polyctx_init(logn, q);
  ctx_build(0, 0);
}
void polyctx_exit(void)
{
  ctx_drop();
  gpqhe_ref_polyctx_exit();
}
void hectx_init(unsigned int logn, MPI q, unsigned int slots, uint64_t Delta)
{
  gpqhe_ref_hectx_init(logn, q, slots, Delta);   /* calls the reference's polyctx_init itself */
  ctx_build(slots, Delta);
}
void hectx_exit(void)
{
  ctx_drop();
  gpqhe_ref_hectx_exit();
}

static gq_ctx *ctx(void)
{
  if (!G.ctx) {   /* context built by code that bypassed the wrappers (cannot happen via the API) */
    fprintf(stderr, "gpqhe_dropin: no context (call polyctx_init/hectx_init first)\n");
    errno = EINVAL;
    abort();
  }
  return G.ctx;
}

static unsigned limb_of(const struct rns_ctx *rns) { return rns->dim - 1; }   /* src/precomp.c:371 */

/* ------------------------------------------------------------------------ limb layer */
void ntt(uint64_t a[], const struct rns_ctx *rns) { CK(gq_ntt(ctx(), limb_of(rns), a)); }
void invntt(uint64_t a[], const struct rns_ctx *rns) { CK(gq_invntt(ctx(), limb_of(rns), a)); }
void poly_rns_add(uint64_t r[], const uint64_t a[], const uint64_t b[], const struct rns_ctx *rns)
{
  CK(gq_poly_rns_add(ctx(), limb_of(rns), r, a, b));
}
void poly_rns_mul(uint64_t r[], const uint64_t a[], const uint64_t b[], const struct rns_ctx *rns)
{
  CK(gq_poly_rns_mul(ctx(), limb_of(rns), r, a, b));
}

/* -------------------------------------------------------------------- MPI <-> RNS */
void rns_decompose(uint64_t ahat[], const MPI a[], const struct rns_ctx *rns)
{
  const unsigned n = polyctx.n, W = poly_words(a, n);
  if (W > DROPIN_MAX_W) { fprintf(stderr, "rns_decompose: coefficient wider than %u words\n", DROPIN_MAX_W); abort(); }
  uint64_t *s = stage((size_t)W * n);
  for (unsigned i = 0; i < n; i++) mpi_to_words(s + i, n, W, a[i], __func__);
  CK(gq_rns_decompose(ctx(), limb_of(rns), 1, ahat, s, W));
}

void poly_rns2mpi(poly_mpi_t *r, const poly_rns_t *rhat, const struct rns_ctx *rns, const MPI q)
{
  const unsigned n = polyctx.n;
  uint64_t qw[DROPIN_MAX_W + 1];
  const unsigned nq = mpi_words_unsigned(qw, DROPIN_MAX_W + 1, q, __func__);
  const unsigned Wr = words_for_bits(gcry_mpi_get_nbits(q));
  uint64_t *s = stage((size_t)Wr * n);
  CK(gq_poly_rns2mpi(ctx(), rns->dim, s, Wr, rhat->coeffs, qw, nq));
  poly_out(r, s, Wr);
}

void poly_mul(poly_mpi_t *r, const poly_mpi_t *a, const poly_mpi_t *b, const unsigned int dim, const MPI q)
{
  const unsigned n = polyctx.n;
  unsigned W = poly_words(a->coeffs, n), Wb = poly_words(b->coeffs, n);
  if (Wb > W) W = Wb;
  if (W > DROPIN_MAX_W) { fprintf(stderr, "poly_mul: coefficient wider than %u words\n", DROPIN_MAX_W); abort(); }
  uint64_t qw[DROPIN_MAX_W + 1];
  const unsigned nq = mpi_words_unsigned(qw, DROPIN_MAX_W + 1, q, __func__);
  const unsigned Wr = words_for_bits(gcry_mpi_get_nbits(q));
  uint64_t *s = stage((size_t)(2 * W + Wr) * n);
  uint64_t *sa = s, *sb = s + (size_t)W * n, *sr = s + 2 * (size_t)W * n;
  poly_in(sa, a, W, __func__);
  poly_in(sb, b, W, __func__);
  CK(gq_poly_mul(ctx(), sr, Wr, sa, sb, W, dim, qw, nq));
  poly_out(r, sr, Wr);
}

static void automorph(poly_mpi_t *r, const poly_mpi_t *a, int conj, size_t rot, const char *who)
{
  const unsigned n = polyctx.n, W = poly_words(a->coeffs, n);
  if (W > DROPIN_MAX_W) { fprintf(stderr, "%s: coefficient wider than %u words\n", who, DROPIN_MAX_W); abort(); }
  uint64_t *s = stage((size_t)2 * W * n);
  poly_in(s, a, W, who);
  if (conj) CK(gq_poly_conj(ctx(), s + (size_t)W * n, s, W));
  else CK(gq_poly_rot(ctx(), s + (size_t)W * n, s, W, (uint64_t)rot));
  poly_out(r, s + (size_t)W * n, W);
}
void poly_rot(poly_mpi_t *r, const poly_mpi_t *a, size_t rot) { automorph(r, a, 0, rot, __func__); }
void poly_conj(poly_mpi_t *r, const poly_mpi_t *a) { automorph(r, a, 1, 0, __func__); }

/* ---------------------------------------------------------------- evaluation keys */
/* he_genswk writes keys with the CPU ntt() into plain malloc buffers (src/he-kem.c:103-110);
 * the device copy is made on first use and found again by address + sampled fingerprint, so a
 * buffer that was freed and reused for another key is re-uploaded. */
static uint64_t evk_print(const he_evk_t *k)
{
  const size_t words = (size_t)hectx.dimevk * polyctx.n;
  uint64_t h = 0x9E3779B97F4A7C15ull, x = 0x243F6A8885A308D3ull;
  for (unsigned s = 0; s < 2048; s++) {
    x = x * 6364136223846793005ull + 1442695040888963407ull;
    size_t i = (size_t)((x >> 20) % words);
    h = (h ^ k->p0.coeffs[i]) * 0x100000001B3ull;
    h = (h ^ k->p1.coeffs[i]) * 0x100000001B3ull;
  }
  h ^= k->p0.coeffs[0] ^ (k->p1.coeffs[words - 1] << 1);
  return h;
}

static const gq_evk *evk_dev(const he_evk_t *k)
{
  const uint64_t pr = evk_print(k);
  for (unsigned i = 0; i < G.nevk; i++) {
    struct evk_slot *s = &G.evk[i];
    if (s->p0 == k->p0.coeffs && s->p1 == k->p1.coeffs) {
      if (s->print == pr) return s->dev;
      gq_evk_destroy(s->dev);
      CK(gq_evk_create(ctx(), &s->dev, k->p0.coeffs, k->p1.coeffs));
      s->print = pr;
      return s->dev;
    }
  }
  if (G.nevk == G.capevk) {
    G.capevk = G.capevk ? 2 * G.capevk : 16;
    G.evk = realloc(G.evk, G.capevk * sizeof *G.evk);
  }
  struct evk_slot *s = &G.evk[G.nevk++];
  s->p0 = k->p0.coeffs; s->p1 = k->p1.coeffs; s->print = pr; s->dev = NULL;
  CK(gq_evk_create(ctx(), &s->dev, k->p0.coeffs, k->p1.coeffs));
  return s->dev;
}

void he_free_evk(he_evk_t *evk)
{
  for (unsigned i = 0; i < G.nevk; i++)
    if (G.evk[i].p0 == evk->p0.coeffs && G.evk[i].p1 == evk->p1.coeffs) {
      gq_evk_destroy(G.evk[i].dev);
      G.evk[i] = G.evk[--G.nevk];
      break;
    }
  gpqhe_ref_he_free_evk(evk);
}

/* ------------------------------------------------------------------ HE primitives */
void he_mul(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2, const he_evk_t *rlk)
{
  assert(ct1->l == ct2->l);
  /* metadata first and in the reference's statement order (he-mult.c:92-95): with ct aliasing
   * ct1/ct2 the later statements read the values the earlier ones wrote */
  ct->l = ct1->l;
  ct->nu = ct1->nu * ct2->nu;
  ct->B = (ct1->nu) * (ct2->B) + (ct2->nu) * (ct1->B) + (ct1->B) * (ct2->B) + hectx.bnd.Bmult[ct->l];
  const unsigned l = ct->l, n = polyctx.n, W = G.info.W;
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage(6 * poly);
  poly_in(s, &ct1->c0, W, __func__);
  poly_in(s + poly, &ct1->c1, W, __func__);
  if (ct2 == ct1) {
    CK(gq_he_mul(ctx(), l, s + 4 * poly, s + 5 * poly, s, s + poly, s, s + poly, evk_dev(rlk)));
  } else {
    poly_in(s + 2 * poly, &ct2->c0, W, __func__);
    poly_in(s + 3 * poly, &ct2->c1, W, __func__);
    CK(gq_he_mul(ctx(), l, s + 4 * poly, s + 5 * poly, s, s + poly, s + 2 * poly, s + 3 * poly, evk_dev(rlk)));
  }
  poly_out(&ct->c0, s + 4 * poly, W);
  poly_out(&ct->c1, s + 5 * poly, W);
}

void he_mulpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  dest->l = src->l;                                                     /* he-mult.c:162-164 */
  dest->nu = src->nu * pt->nu;
  dest->B = src->B * pt->nu;
  const unsigned l = dest->l, n = polyctx.n, W = G.info.W;
  /* limb count in double arithmetic, truncated, exactly as he-mult.c:168 */
  const unsigned dim = (gcry_mpi_get_nbits(hectx.q[l]) + log2(pt->nu) + polyctx.logn) / GPQHE_LOGP + 1;
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage(5 * poly);
  poly_in(s, &src->c0, W, __func__);
  poly_in(s + poly, &src->c1, W, __func__);
  poly_in(s + 2 * poly, &pt->m, W, __func__);
  CK(gq_he_mulpt(ctx(), l, dim, s + 3 * poly, s + 4 * poly, s, s + poly, s + 2 * poly));
  poly_out(&dest->c0, s + 3 * poly, W);
  poly_out(&dest->c1, s + 4 * poly, W);
}

static void ct_inplace(he_ct_t *ct, int op, unsigned rot, const he_evk_t *k, const char *who)
{
  const unsigned n = polyctx.n, W = G.info.W;
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage(2 * poly);
  poly_in(s, &ct->c0, W, who);
  poly_in(s + poly, &ct->c1, W, who);
  switch (op) {
    case 0: CK(gq_he_rs(ctx(), ct->l, s, s + poly)); break;
    case 1: CK(gq_he_moddown(ctx(), ct->l, s, s + poly)); break;
    case 2: CK(gq_he_rot(ctx(), ct->l, rot, s, s + poly, evk_dev(k))); break;
    case 3: CK(gq_he_conj(ctx(), ct->l, s, s + poly, evk_dev(k))); break;
  }
  poly_out(&ct->c0, s, W);
  poly_out(&ct->c1, s + poly, W);
}

void he_rs(struct he_ct *ct)
{
  ct_inplace(ct, 0, 0, NULL, __func__);
  ct->l -= 1;                                                           /* he-rescale.c:36-38 */
  ct->nu /= hectx.Delta;
  ct->B = ct->B / hectx.Delta + hectx.bnd.Brs;
}
void he_moddown(he_ct_t *ct)
{
  ct_inplace(ct, 1, 0, NULL, __func__);
  ct->l -= 1;                                                           /* he-rescale.c:59 */
}
void he_rot(he_ct_t *ct, const int rot, const he_evk_t *rk) { ct_inplace(ct, 2, (unsigned)rot, &rk[rot], __func__); }
void he_conj(he_ct_t *ct, const he_evk_t *ck) { ct_inplace(ct, 3, 0, ck, __func__); }

/* ---- coefficient-local ops: src/he-add.c:32-146 ------------------------------- */
static void ct_linear(he_ct_t *ct, const he_ct_t *a, const poly_mpi_t *b0, const poly_mpi_t *b1, int op,
                      unsigned l, const char *who)
{
  const unsigned n = polyctx.n, W = G.info.W;
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage(6 * poly);
  poly_in(s, &a->c0, W, who);
  poly_in(s + poly, &a->c1, W, who);
  if (b0) poly_in(s + 2 * poly, b0, W, who);
  if (b1) poly_in(s + 3 * poly, b1, W, who);
  uint64_t *o0 = s + 4 * poly, *o1 = s + 5 * poly;
  switch (op) {
    case 0: CK(gq_he_add(ctx(), l, o0, o1, s, s + poly, s + 2 * poly, s + 3 * poly)); break;
    case 1: CK(gq_he_sub(ctx(), l, o0, o1, s, s + poly, s + 2 * poly, s + 3 * poly)); break;
    case 2: CK(gq_he_addpt(ctx(), l, o0, o1, s, s + poly, s + 2 * poly)); break;
    case 3: CK(gq_he_subpt(ctx(), l, o0, o1, s, s + poly, s + 2 * poly)); break;
    case 4: memcpy(o0, s, 2 * poly * sizeof(uint64_t)); CK(gq_he_neg(ctx(), l, o0, o1)); break;
  }
  poly_out(&ct->c0, o0, W);
  poly_out(&ct->c1, o1, W);
}

void he_add(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2)
{
  assert(ct1->l == ct2->l);
  ct->l = ct1->l;                                                       /* he-add.c:36-38, in order */
  ct->nu = (ct1->nu >= ct2->nu) ? ct1->nu : ct2->nu;
  ct->B = ct1->B + ct2->B;
  ct_linear(ct, ct1, &ct2->c0, &ct2->c1, 0, ct->l, __func__);
}
void he_sub(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2)
{
  assert(ct1->l == ct2->l);
  ct->l = ct1->l;                                                       /* he-add.c:61-63 */
  ct->nu = (ct1->nu >= ct2->nu) ? ct1->nu : ct2->nu;
  ct->B = ct1->B + ct2->B;
  ct_linear(ct, ct1, &ct2->c0, &ct2->c1, 1, ct->l, __func__);
}
void he_addpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  dest->l = src->l;                                                     /* he-add.c:85-87 */
  dest->nu = (src->nu >= pt->nu) ? src->nu : pt->nu;
  dest->B = src->B;
  ct_linear(dest, src, &pt->m, NULL, 2, dest->l, __func__);
}
void he_subpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  dest->l = src->l;                                                     /* he-add.c:109-111 */
  dest->nu = (src->nu >= pt->nu) ? src->nu : pt->nu;
  dest->B = src->B;
  ct_linear(dest, src, &pt->m, NULL, 3, dest->l, __func__);
}
void he_neg(he_ct_t *ct) { ct_linear(ct, ct, NULL, NULL, 4, ct->l, __func__); }

/* ---- he_gemv (src/he-algo.c:47-90), device resident ------------------------------------------
 * The reference walks a baby-step/giant-step loop of he_rot / he_mulpt / he_add on host ciphertexts.
 * Here only the encoding of the matrix diagonals (he_ecd: floating point, host) stays on the host;
 * the ciphertext crosses once in and once out and gq_he_gemv runs the same operations in the same
 * order on the device (baby rotations hoisted and computed once). */
static void gemv_split(unsigned *n1, unsigned *n2)
{
  const unsigned m = hectx.slots;
  *n1 = sqrt(m);                                                        /* he-algo.c:51-54 */
  if (m != *n1 * *n1) *n1 = sqrt(2 * m);
  *n2 = m / *n1;
}

/* diagonal k = i*n1+j of A, pre-rotated by -i*n1 slots, encoded at scale Delta (he-algo.c:70-72), written
 * as word-major polys at dst + k*poly; pt keeps the metadata of the last he_ecd */
static void gemv_diagonals(const _Complex double *A, unsigned n1, uint64_t *dst, const uint64_t **pts,
                           he_pt_t *pt, const char *who)
{
  const unsigned m = hectx.slots, W = G.info.W;
  const size_t poly = (size_t)W * polyctx.n;
  _Complex double *d = malloc(sizeof(_Complex double) * m);
  if (!d) die(who);
  for (unsigned k = 0; k < m; k++) {
    const unsigned shift = (k / n1) * n1;
    for (unsigned t = 0; t < m; t++) {
      const unsigned u = (t + m - shift % m) % m;                       /* slot after rotating by -shift */
      d[t] = A[(size_t)u * m + (k + u) % m];
    }
    he_ecd(pt, d);
    poly_in(dst + (size_t)k * poly, &pt->m, W, who);
    pts[k] = dst + (size_t)k * poly;
  }
  free(d);
}

/* metadata of he_gemv: the reference's sequence of he_mulpt (he-mult.c:162-164), he_add (he-add.c:36-38)
 * and he_rs (he-rescale.c:36-38) updates, in its loop order */
static void gemv_meta(double nu, double B, double ptnu, unsigned n1, unsigned n2, double *nu_out, double *B_out)
{
  double nu_o = 0, B_o = 0;
  for (unsigned i = 0; i < n2; i++) {
    double nu_i = 0, B_i = 0;
    for (unsigned j = 0; j < n1; j++) {
      const double nu_t = nu * ptnu, B_t = B * ptnu;
      if (!j) { nu_i = nu_t; B_i = B_t; }
      else { nu_i = (nu_i >= nu_t) ? nu_i : nu_t; B_i = B_i + B_t; }
    }
    if (!i) { nu_o = nu_i; B_o = B_i; }
    else { nu_o = (nu_o >= nu_i) ? nu_o : nu_i; B_o = B_o + B_i; }
  }
  *nu_out = nu_o / hectx.Delta;
  *B_out = B_o / hectx.Delta + hectx.bnd.Brs;
}

static const gq_evk **gemv_keys(const he_evk_t *rk, unsigned n1, const char *who)
{
  const unsigned m = hectx.slots;
  const gq_evk **keys = calloc(m, sizeof(*keys));
  if (!keys) die(who);
  for (unsigned r = 0; r < m; r++)
    if (r < n1 || r % n1 == 0) keys[r] = evk_dev(&rk[r]);
  return keys;
}

void he_gemv(he_ct_t *ct_dest, const _Complex double *A, const he_ct_t *ct, const he_evk_t *rk)
{
  const unsigned m = hectx.slots, n = polyctx.n, W = G.info.W, l = ct->l;
  unsigned n1, n2;
  gemv_split(&n1, &n2);
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage((4 + (size_t)m) * poly);
  poly_in(s, &ct->c0, W, __func__);
  poly_in(s + poly, &ct->c1, W, __func__);
  he_pt_t pt;
  he_alloc_pt(&pt);
  const uint64_t **pts = malloc(sizeof(*pts) * m);
  if (!pts) die(__func__);
  gemv_diagonals(A, n1, s + 4 * poly, pts, &pt, __func__);
  const gq_evk **keys = gemv_keys(rk, n1, __func__);
  const unsigned dim = (gcry_mpi_get_nbits(hectx.q[l]) + log2(pt.nu) + polyctx.logn) / GPQHE_LOGP + 1;
  double nu_o, B_o;
  gemv_meta(ct->nu, ct->B, pt.nu, n1, n2, &nu_o, &B_o);
  CK(gq_he_gemv(ctx(), l, n1, n2, dim, s + 2 * poly, s + 3 * poly, s, s + poly, pts, keys));
  ct_dest->l = l - 1;
  ct_dest->nu = nu_o;
  ct_dest->B = B_o;
  poly_out(&ct_dest->c0, s + 2 * poly, W);
  poly_out(&ct_dest->c1, s + 3 * poly, W);
  free(pts); free(keys);
  he_free_pt(&pt);
}

/* ---- he_coeff2slot (src/he-bootstrap.c:359-409), device resident ----------------------------------
 * he_conj, four he_gemv, he_add, he_mulpt by he_const_pt(1/n) and he_rs for the real and the imaginary
 * half: one crossing in (the ciphertext + the encoded diagonals of the four matrices of bootstrapctx),
 * one crossing out; the reference's own composition stays available as gpqhe_ref_he_coeff2slot. */
extern struct bootstrap_ctx bootstrapctx;
void he_coeff2slot(he_ct_t *ct_real, he_ct_t *ct_imag, const he_ct_t *ct, const he_evk_t *ck, const he_evk_t *rk)
{
  const unsigned m = hectx.slots, n = polyctx.n, W = G.info.W, l = ct->l;
  unsigned n1, n2;
  gemv_split(&n1, &n2);
  const size_t poly = (size_t)W * n;
  uint64_t *s = stage((7 + 4 * (size_t)m) * poly);
  poly_in(s, &ct->c0, W, __func__);
  poly_in(s + poly, &ct->c1, W, __func__);
  he_pt_t pt, cpt;
  he_alloc_pt(&pt);
  he_alloc_pt(&cpt);
  he_const_pt(&cpt, 1 / polyctx.n);                                     /* he-bootstrap.c:385, integer division as there */
  poly_in(s + 2 * poly, &cpt.m, W, __func__);
  const uint64_t **pts = malloc(sizeof(*pts) * 4 * m);
  if (!pts) die(__func__);
  const _Complex double *mats[4] = {bootstrapctx.U0_conjT, bootstrapctx.U0_T, bootstrapctx.U1_conjT, bootstrapctx.U1_T};
  for (unsigned k = 0; k < 4; k++)
    gemv_diagonals(mats[k], n1, s + (7 + (size_t)k * m) * poly, pts + (size_t)k * m, &pt, __func__);
  const gq_evk **keys = gemv_keys(rk, n1, __func__);
  const unsigned dim_gemv = (gcry_mpi_get_nbits(hectx.q[l]) + log2(pt.nu) + polyctx.logn) / GPQHE_LOGP + 1;
  const unsigned dim_pt = (gcry_mpi_get_nbits(hectx.q[l - 1]) + log2(cpt.nu) + polyctx.logn) / GPQHE_LOGP + 1;
  double nu_g, B_g;
  gemv_meta(ct->nu, ct->B, pt.nu, n1, n2, &nu_g, &B_g);                 /* he_conj keeps nu and B */
  uint64_t *o = s + 3 * poly;
  CK(gq_he_coeff2slot(ctx(), l, n1, n2, dim_gemv, dim_pt, o, o + poly, o + 2 * poly, o + 3 * poly, s, s + poly,
                      pts, s + 2 * poly, evk_dev(ck), keys));
  he_ct_t *outs[2] = {ct_real, ct_imag};
  for (unsigned k = 0; k < 2; k++) {
    double nu = nu_g, B = B_g + B_g;                                    /* he_add, he-add.c:36-38 */
    nu = nu * cpt.nu; B = B * cpt.nu;                                   /* he_mulpt, he-mult.c:162-164 */
    outs[k]->l = l - 2;                                                 /* he_rs, he-rescale.c:36-38 */
    outs[k]->nu = nu / hectx.Delta;
    outs[k]->B = B / hectx.Delta + hectx.bnd.Brs;
    poly_out(&outs[k]->c0, o + (size_t)(2 * k) * poly, W);
    poly_out(&outs[k]->c1, o + (size_t)(2 * k + 1) * poly, W);
  }
  free(pts); free(keys);
  he_free_pt(&pt);
  he_free_pt(&cpt);
}

/* ---- he_exp (src/he-algo.c:363-458) and he_rlsin (src/he-bootstrap.c:444-475), device resident -------------
 * The reference evaluates exp(a*ct) as he_mulpt by a/2^iter, a degree-7 Taylor polynomial (7 he_mul, 5 he_mulpt,
 * he_addpt / he_add / he_rs / he_moddown between them) and `iter` squarings, every primitive a separate call on
 * host ciphertexts.  Here only the encoding of the 8 (he_rlsin: 9) plaintexts stays on the host (he_ecd /
 * he_const_pt: floating point); the ciphertext crosses once in and once out and gq_he_exp / gq_he_rlsin run the
 * same primitives in the same order on the device.  Metadata: the reference's own updates, replayed in order. */
struct meta { double nu, B; };
static struct meta m_mul(unsigned l, struct meta a, struct meta b)            /* he-mult.c:93-95 */
{
  struct meta r = {a.nu * b.nu, (a.nu) * (b.B) + (b.nu) * (a.B) + (a.B) * (b.B) + hectx.bnd.Bmult[l]};
  return r;
}
/* he_mul(&x, &y, &x): the destination aliases ct2, so the B statement reads the nu it has just written (he-mult.c:94-95) */
static struct meta m_mul_d2(unsigned l, struct meta a, struct meta b)
{
  const double nu = a.nu * b.nu;
  struct meta r = {nu, (a.nu) * (b.B) + (nu) * (a.B) + (a.B) * (b.B) + hectx.bnd.Bmult[l]};
  return r;
}
/* he_mul(&x, &x, &x) */
static struct meta m_sqr_inplace(unsigned l, struct meta a)
{
  const double nu = a.nu * a.nu;
  struct meta r = {nu, (nu) * (a.B) + (nu) * (a.B) + (a.B) * (a.B) + hectx.bnd.Bmult[l]};
  return r;
}
static struct meta m_rs(struct meta a) { struct meta r = {a.nu / hectx.Delta, a.B / hectx.Delta + hectx.bnd.Brs}; return r; }
static struct meta m_mulpt(struct meta a, double ptnu) { struct meta r = {a.nu * ptnu, a.B * ptnu}; return r; }
static struct meta m_add(struct meta a, struct meta b) { struct meta r = {(a.nu >= b.nu) ? a.nu : b.nu, a.B + b.B}; return r; }
static struct meta m_addpt(struct meta a, double ptnu) { struct meta r = {(a.nu >= ptnu) ? a.nu : ptnu, a.B}; return r; }
static struct meta m_lin(struct meta act, double ptnu) { return m_rs(m_mulpt(m_addpt(act, ptnu), ptnu)); }

static struct meta exp_meta(unsigned l, struct meta in, double ptnu, unsigned iter)
{
  const struct meta act = m_rs(m_mulpt(in, ptnu));
  const unsigned k = l - 1;
  const struct meta ct2 = m_rs(m_mul(k, act, act));
  const struct meta ct4 = m_rs(m_mul(k - 1, ct2, ct2));
  const struct meta ct01 = m_lin(act, ptnu);
  const struct meta ct23 = m_rs(m_mul_d2(k - 1, ct2, m_lin(act, ptnu)));
  const struct meta ct0123 = m_add(ct01, ct23);
  const struct meta ct45 = m_lin(act, ptnu);
  const struct meta ct67 = m_rs(m_mul_d2(k - 1, ct2, m_lin(act, ptnu)));
  const struct meta ct4567 = m_rs(m_mul_d2(k - 2, ct4, m_add(ct45, ct67)));
  struct meta out = m_add(ct0123, ct4567);
  for (unsigned t = 0; t < iter; t++) out = m_rs(m_sqr_inplace(k - 3 - t, out));
  return out;
}

static unsigned mulpt_dim(unsigned l, double ptnu)                            /* he-mult.c:168 */
{
  return (gcry_mpi_get_nbits(hectx.q[l]) + log2(ptnu) + polyctx.logn) / GPQHE_LOGP + 1;
}

/* the 8 plaintexts of he_exp at dst (pts[0] = he_ecd of a/2^iter in every slot, he-algo.c:443-447; pts[1..7] =
 * he_const_pt of the Taylor constants, he-algo.c:384-411); returns pt->nu (the same for all of them) */
static double exp_plaintexts(_Complex double a, unsigned iter, uint64_t *dst, const uint64_t **pts, const char *who)
{
  static const double consts[7] = {1.0, 3.0, 1.0 / 6, 5.0, 1.0 / 120, 7.0, 1.0 / 5040};
  const unsigned m = hectx.slots, W = G.info.W;
  const size_t poly = (size_t)W * polyctx.n;
  he_pt_t pt;
  he_alloc_pt(&pt);
  _Complex double *coeff = malloc(sizeof(_Complex double) * m);
  if (!coeff) die(who);
  a /= (1 << iter);
  for (unsigned i = 0; i < m; i++) coeff[i] = a;
  he_ecd(&pt, coeff);
  poly_in(dst, &pt.m, W, who);
  pts[0] = dst;
  he_pt_t cpt;                        /* he_exp_taylor's own plaintext: he_const_pt writes coefficients 0 and n/2 only */
  he_alloc_pt(&cpt);
  for (unsigned k = 0; k < 7; k++) {
    he_const_pt(&cpt, consts[k]);
    poly_in(dst + (size_t)(k + 1) * poly, &cpt.m, W, who);
    pts[k + 1] = dst + (size_t)(k + 1) * poly;
  }
  const double nu = pt.nu;
  free(coeff);
  he_free_pt(&pt);
  he_free_pt(&cpt);
  return nu;
}

void he_exp(he_ct_t *ct_exp, _Complex double a, const he_ct_t *ct, const he_evk_t *rlk, const unsigned int iter)
{
  const unsigned n = polyctx.n, W = G.info.W, l = ct->l;
  const size_t poly = (size_t)W * n;
  if (l < 4 + iter) { fprintf(stderr, "he_exp: level %u is below the 4+iter levels the evaluation consumes\n", l); errno = EINVAL; abort(); }
  uint64_t *s = stage(12 * poly);
  const uint64_t *pts[8];
  poly_in(s, &ct->c0, W, __func__);
  poly_in(s + poly, &ct->c1, W, __func__);
  const double ptnu = exp_plaintexts(a, iter, s + 4 * poly, pts, __func__);
  const struct meta in = {ct->nu, ct->B}, out = exp_meta(l, in, ptnu, iter);
  CK(gq_he_exp(ctx(), l, iter, mulpt_dim(l, ptnu), mulpt_dim(l - 1, ptnu), s + 2 * poly, s + 3 * poly, s, s + poly,
               pts, evk_dev(rlk)));
  ct_exp->l = l - 4 - iter;
  ct_exp->nu = out.nu;
  ct_exp->B = out.B;
  poly_out(&ct_exp->c0, s + 2 * poly, W);
  poly_out(&ct_exp->c1, s + 3 * poly, W);
}

void he_rlsin(he_ct_t *ct_dest, const double a, const he_ct_t *ct, const he_evk_t *rlk, const he_evk_t *ck,
              const unsigned int iter)
{
  const unsigned m = hectx.slots, n = polyctx.n, W = G.info.W, l = ct->l;
  const size_t poly = (size_t)W * n;
  if (l < 5 + iter) { fprintf(stderr, "he_rlsin: level %u is below the 5+iter levels the evaluation consumes\n", l); errno = EINVAL; abort(); }
  uint64_t *s = stage(13 * poly);
  const uint64_t *pts[9];
  poly_in(s, &ct->c0, W, __func__);
  poly_in(s + poly, &ct->c1, W, __func__);
  const double ptnu = exp_plaintexts(a * I, iter, s + 4 * poly, pts, __func__);   /* he-bootstrap.c:458 */
  he_pt_t pt;
  he_alloc_pt(&pt);
  _Complex double *coeff = malloc(sizeof(_Complex double) * m);
  if (!coeff) die(__func__);
  const _Complex double b = 1 / (2 * I * a);                                      /* he-bootstrap.c:465-468 */
  for (unsigned i = 0; i < m; i++) coeff[i] = b;
  he_ecd(&pt, coeff);
  poly_in(s + 12 * poly, &pt.m, W, __func__);
  pts[8] = s + 12 * poly;
  const struct meta in = {ct->nu, ct->B}, e = exp_meta(l, in, ptnu, iter);
  const struct meta out = m_rs(m_mulpt(m_add(e, e), pt.nu));       /* he_conj keeps nu, B; he_sub updates as he_add */
  const unsigned lf = l - 4 - iter;
  CK(gq_he_rlsin(ctx(), l, iter, mulpt_dim(l, ptnu), mulpt_dim(l - 1, ptnu), mulpt_dim(lf, pt.nu), s + 2 * poly,
                 s + 3 * poly, s, s + poly, pts, evk_dev(rlk), evk_dev(ck)));
  ct_dest->l = lf - 1;
  ct_dest->nu = out.nu;
  ct_dest->B = out.B;
  poly_out(&ct_dest->c0, s + 2 * poly, W);
  poly_out(&ct_dest->c1, s + 3 * poly, W);
  free(coeff);
  he_free_pt(&pt);
}

/* diagnostics for tests: coefficients marshalled in / out since load, and whether the direct
 * limb path is active */
void gpqhe_dropin_stats(unsigned long long *in, unsigned long long *out, int *fast_mpi)
{
  if (in) *in = G.marshal_in;
  if (out) *out = G.marshal_out;
  if (fast_mpi) *fast_mpi = G.fast_mpi;
}
