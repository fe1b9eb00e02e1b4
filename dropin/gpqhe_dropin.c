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
 *   he_free_evk, he_genrlk/ck/rk         src/he-mem.c, src/he-kem.c (wrappers: drop the device copy of the key)
 *   he_enc_pk/sk, he_dec, he_copy_ct,    src/he-encrypt.c, src/he-mem.c (wrappers: ciphertexts live in device
 *   poly_mpi_free                        shadows between calls; these are the points where the host MPIs are
 *                                        written, read, copied or freed)
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
  unsigned long long used;  /* last use (LRU eviction) */
};

/* device shadow of one he_ct_t, keyed by ct->c0.coeffs (see "device-resident ciphertexts" below) */
struct shadow {
  const void *key;
  uint64_t *dev;            /* c0 at dev, c1 at dev + W*n; NULL until first needed */
  unsigned char dev_ok;     /* the device copy is current */
  unsigned char host_ok;    /* the gcry_mpi objects are current */
  unsigned char tomb;       /* deleted slot of the open-addressing table */
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
  unsigned long long evk_clock;
  struct shadow *sh;        /* open-addressing table, power-of-two capacity */
  size_t shcap, shused;
  uint64_t **pool;          /* free 2-poly device buffers */
  unsigned npool, cappool;
  uint64_t *dscratch;       /* device scratch for plaintexts / encoded diagonals */
  size_t dscratch_words;
  uint64_t *ring;           /* pinned staging ring of the asynchronous uploads / downloads */
  unsigned ring_pos;
  unsigned long long ct_uploads, ct_downloads, shadow_allocs;
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

static void words_to_mpi(MPI a, const uint64_t *in, size_t stride, unsigned W);
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
  if (!ok) return;
  /* the WRITE path too: store values through the limb view and read them back with the public API */
  static const uint64_t vals[3][3] = {{0x1122334455667788ull, 0x99aabbccddeeff00ull, 0x0fedcba987654321ull},
                                      {5, 0, 0}, {0, 0, 0}};
  for (int t = 0; t < 6 && G.fast_mpi; t++) {
    const int neg = t & 1;
    uint64_t w[4] = {vals[t / 2][0], vals[t / 2][1], vals[t / 2][2], 0};
    if (neg) { uint64_t c = 1; for (int i = 0; i < 4; i++) { uint64_t x = ~w[i] + c; c = (c && x == 0); w[i] = x; } }
    MPI b = gcry_mpi_new(320);
    gcry_mpi_set_ui(b, 1);
    gcry_mpi_lshift(b, b, 300);            /* make libgcrypt allocate the limbs */
    words_to_mpi(b, w, 1, 4);
    unsigned char buf[64];
    size_t len = 0;
    int good = gcry_mpi_print(GCRYMPI_FMT_USG, buf, sizeof buf, &len, b) == 0;
    uint64_t back[4] = {0, 0, 0, 0};
    for (size_t k = 0; good && k < len; k++) back[k / 8] |= (uint64_t)buf[len - 1 - k] << (8 * (k % 8));
    const int zero = !vals[t / 2][0] && !vals[t / 2][1] && !vals[t / 2][2];
    good = good && back[0] == vals[t / 2][0] && back[1] == vals[t / 2][1] && back[2] == vals[t / 2][2] &&
           (gcry_mpi_is_neg(b) != 0) == (neg && !zero);
    gcry_mpi_release(b);
    if (!good) G.fast_mpi = 0;
  }
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
static void shadows_drop_all(void);
static void diag_drop_all(void);
static void ctx_drop(void)
{
  diag_drop_all();
  shadows_drop_all();
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

/* ---- batched evaluation-key generation -----------------------------------------------------------
 * he_genswk (src/he-kem.c:103-110) ends with, for every limb d < dimevk and both key halves,
 *     rns_decompose(&swk->p[d*n], poly.coeffs, rns_d);  ntt(&swk->p[d*n], rns_d);
 * i.e. 4*dimevk single-limb host<->device crossings per key, each marshalling the same N big integers again.  While
 * one of the key generators runs (he_genrlk / he_genck / he_genrk are wrapped below and open this scope; the
 * samplers and everything else stay the reference's, RNG order untouched), the FIRST rns_decompose of a polynomial
 * marshals it once, computes NTT(decompose(.)) for ALL dimevk limbs in two batched crossings and keeps the result;
 * every rns_decompose of that polynomial then copies its limb - already transformed - and the ntt() call that
 * follows on the same buffer is a no-op.  A cached polynomial dies with poly_mpi_free (he_genswk frees both at its
 * end), so a recycled address can never hit.  Outside the scope nothing changes.  ntt() keeps the reference's exact
 * bit pattern (gq_ntt_batch), so the keys are byte-identical (tests/keydump.c). */
#define KG_POLYS 2
static struct {
  int active;
  const MPI *a[KG_POLYS];        /* cached polynomials (address of the MPI array) */
  uint64_t *hat[KG_POLYS];       /* NTT(decompose(a)) over dimevk limbs, host */
  const uint64_t *pending[4];    /* buffers whose ntt() call is still to come */
  unsigned npending;
  unsigned long long batched, skipped;
} KG;

static void kg_forget(const void *coeffs)
{
  for (int k = 0; k < KG_POLYS; k++) if (KG.a[k] == (const MPI *)coeffs) KG.a[k] = NULL;
}
static void kg_close(const char *who)
{
  if (KG.npending) { fprintf(stderr, "%s: rns_decompose without its ntt() inside key generation\n", who); abort(); }
  KG.active = 0;
  for (int k = 0; k < KG_POLYS; k++) KG.a[k] = NULL;
}
static int kg_decompose(uint64_t ahat[], const MPI a[], const struct rns_ctx *rns)
{
  const unsigned n = polyctx.n, D = hectx.dimevk, d = limb_of(rns);
  if (!KG.active || d >= D) return 0;
  int k = -1;
  for (int i = 0; i < KG_POLYS; i++) if (KG.a[i] == a) k = i;
  if (k < 0) {
    for (int i = 0; i < KG_POLYS; i++) if (!KG.a[i]) { k = i; break; }
    if (k < 0) k = 0;
    const unsigned W = poly_words(a, n);
    if (W > DROPIN_MAX_W) return 0;
    uint64_t *s = stage((size_t)W * n);
    struct in_args x = {s, &(poly_mpi_t){(MPI *)a}, W, "rns_decompose"};
    par_for(n, poly_in_range, &x);
    if (!KG.hat[k]) {
      void *p = NULL;
      if (gq_host_alloc(&p, (size_t)polyctx.dimub * n * 8)) die("keygen batch");
      KG.hat[k] = p;
    }
    CK(gq_rns_decompose(ctx(), 0, D, KG.hat[k], s, W));
    CK(gq_ntt_batch(ctx(), 0, D, 1, KG.hat[k], 0));
    KG.a[k] = a;
    KG.batched++;
  }
  memcpy(ahat, KG.hat[k] + (size_t)d * n, (size_t)n * 8);
  if (KG.npending == 4) { fprintf(stderr, "rns_decompose: too many transforms pending inside key generation\n"); abort(); }
  KG.pending[KG.npending++] = ahat;
  return 1;
}
static int kg_ntt(const uint64_t a[])
{
  for (unsigned i = 0; i < KG.npending; i++)
    if (KG.pending[i] == a) { KG.pending[i] = KG.pending[--KG.npending]; KG.skipped++; return 1; }
  return 0;
}

/* ------------------------------------------------------------------------ limb layer */
void ntt(uint64_t a[], const struct rns_ctx *rns) { if (KG.active && kg_ntt(a)) return; CK(gq_ntt(ctx(), limb_of(rns), a)); }
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
  if (KG.active && kg_decompose(ahat, a, rns)) return;
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
/* he_genswk writes keys with ntt() into plain malloc buffers (src/he-kem.c:103-110).  The device copy is made on
 * first use and found again by the address of the buffers.  Everything inside this library that writes a key
 * (he_genrlk / he_genck / he_genrk, wrapped below) or frees one (he_free_evk) drops the copy; a caller that rewrites a
 * key buffer by hand says so with gpqhe_dropin_invalidate_evk().  A 64-word sampled fingerprint stays as a cheap
 * second line (hashing 2 x 12 MiB per he_mul would cost ten times the kernels).  The cache is bounded
 * (GPQHE_DROPIN_EVK_CACHE_MB, default 16384): the least recently used copies are dropped first. */
void gpqhe_ref_he_genrlk(he_evk_t *rlk, const poly_mpi_t *sk);
void gpqhe_ref_he_genck(he_evk_t *ck, const poly_mpi_t *sk);
void gpqhe_ref_he_genrk(he_evk_t *rk, const poly_mpi_t *sk);

static uint64_t evk_print(const he_evk_t *k)
{
  const size_t words = (size_t)hectx.dimevk * polyctx.n;
  uint64_t h = 0x9E3779B97F4A7C15ull, x = 0x243F6A8885A308D3ull;
  for (unsigned s = 0; s < 64; s++) {
    x = x * 6364136223846793005ull + 1442695040888963407ull;
    size_t i = (size_t)((x >> 20) % words);
    h = (h ^ k->p0.coeffs[i]) * 0x100000001B3ull;
    h = (h ^ k->p1.coeffs[i]) * 0x100000001B3ull;
  }
  h ^= k->p0.coeffs[0] ^ (k->p1.coeffs[words - 1] << 1);
  return h;
}

static size_t evk_cache_cap(void)
{
  static size_t cap;
  if (!cap) {
    const char *e = getenv("GPQHE_DROPIN_EVK_CACHE_MB");
    long long mb = e ? atoll(e) : 16384;
    if (mb < 64) mb = 64;
    cap = (size_t)mb << 20;
  }
  return cap;
}

static void evk_drop(unsigned i)
{
  gq_ctx_sync(G.ctx);                     /* a kernel in flight may still read the key */
  gq_evk_destroy(G.evk[i].dev);
  G.evk[i] = G.evk[--G.nevk];
}

static const gq_evk *evk_dev(const he_evk_t *k)
{
  const uint64_t pr = evk_print(k);
  G.evk_clock++;
  for (unsigned i = 0; i < G.nevk; i++) {
    struct evk_slot *s = &G.evk[i];
    if (s->p0 == k->p0.coeffs && s->p1 == k->p1.coeffs) {
      s->used = G.evk_clock;
      if (s->print == pr) return s->dev;
      gq_ctx_sync(G.ctx);
      gq_evk_destroy(s->dev);
      CK(gq_evk_create(ctx(), &s->dev, k->p0.coeffs, k->p1.coeffs));
      s->print = pr;
      return s->dev;
    }
  }
  const size_t bytes = 2 * (size_t)hectx.dimevk * polyctx.n * 8;
  while (G.nevk && (size_t)(G.nevk + 1) * bytes > evk_cache_cap()) {     /* least recently used first */
    unsigned lru = 0;
    for (unsigned i = 1; i < G.nevk; i++) if (G.evk[i].used < G.evk[lru].used) lru = i;
    /* one call may hold up to `slots` keys + rlk + ck at once (he_gemv, he_coeff2slot, he_rlsin): copies looked up
     * that recently are never dropped, the cap gives way instead */
    if (G.evk_clock - G.evk[lru].used < (unsigned long long)hectx.slots + 4) break;
    evk_drop(lru);
  }
  if (G.nevk == G.capevk) {
    G.capevk = G.capevk ? 2 * G.capevk : 16;
    G.evk = realloc(G.evk, G.capevk * sizeof *G.evk);
  }
  struct evk_slot *s = &G.evk[G.nevk++];
  s->p0 = k->p0.coeffs; s->p1 = k->p1.coeffs; s->print = pr; s->dev = NULL; s->used = G.evk_clock;
  CK(gq_evk_create(ctx(), &s->dev, k->p0.coeffs, k->p1.coeffs));
  return s->dev;
}

void gpqhe_dropin_invalidate_evk(const he_evk_t *evk)
{
  for (unsigned i = 0; i < G.nevk; i++)
    if (G.evk[i].p0 == evk->p0.coeffs && G.evk[i].p1 == evk->p1.coeffs) { evk_drop(i); break; }
}

void he_free_evk(he_evk_t *evk)
{
  gpqhe_dropin_invalidate_evk(evk);
  gpqhe_ref_he_free_evk(evk);
}
/* key generation stays the reference's (RNG order, stdout prints); its rns_decompose / ntt calls bind to the GPU */
static int kg_enabled(void) { return getenv("GPQHE_DROPIN_NO_KEYGEN_BATCH") == NULL; }
void he_genrlk(he_evk_t *rlk, const poly_mpi_t *sk)
{
  gpqhe_dropin_invalidate_evk(rlk);
  KG.active = kg_enabled();
  gpqhe_ref_he_genrlk(rlk, sk);
  kg_close(__func__);
}
void he_genck(he_evk_t *ck, const poly_mpi_t *sk)
{
  gpqhe_dropin_invalidate_evk(ck);
  KG.active = kg_enabled();
  gpqhe_ref_he_genck(ck, sk);
  kg_close(__func__);
}
void he_genrk(he_evk_t *rk, const poly_mpi_t *sk)
{
  for (unsigned r = 0; r < hectx.slots; r++) gpqhe_dropin_invalidate_evk(&rk[r]);
  KG.active = kg_enabled();
  gpqhe_ref_he_genrk(rk, sk);
  kg_close(__func__);
}
void gpqhe_dropin_keygen_stats(unsigned long long *batched, unsigned long long *skipped)
{
  if (batched) *batched = KG.batched;
  if (skipped) *skipped = KG.skipped;
}

/* ------------------------------------------------- device-resident ciphertexts --------
 * he_ct_t is a public host struct holding 2N gcry_mpi objects (src/gpqhe.h:84-91); converting them costs ten times
 * the kernels of an he_mul.  Every he_* entry point below therefore works on a device SHADOW of the ciphertext: a
 * table keyed by ct->c0.coeffs (the address of the MPI array, stable from he_alloc_ct to he_free_ct) holds the two
 * polynomials as word-major device buffers plus two bits, `dev_ok` (the shadow is current) and `host_ok` (the MPIs
 * are current).  An operation uploads an operand only if its shadow is not current, runs on the device, and marks
 * its result device-only; nothing comes back until somebody READS the MPIs.  Inside libgpqhe.so the readers are
 * he_dec (src/he-encrypt.c:113-123) and he_copy_ct from a host-current source; the writers are he_enc_pk /
 * he_enc_sk (:37-98).  All are wrapped here (same -D rename as the hot symbols), as is poly_mpi_free, which ends a
 * shadow's life.  he_show_ct_params prints metadata only (src/he-info.c:95-104), and l / nu / B always live on the
 * host.  A caller that reads or writes ct->c0.coeffs[i] directly - none of the reference's tests or algorithms does -
 * calls gpqhe_dropin_sync_ct() before reading, gpqhe_dropin_touch_ct() after writing, or sets GPQHE_DROPIN_EAGER=1,
 * which writes every result back at once (the behaviour of round 1). */
void gpqhe_ref_he_enc_pk(struct he_ct *ct, const struct he_pt *pt, const struct he_pk *pk);
void gpqhe_ref_he_enc_sk(struct he_ct *ct, const struct he_pt *pt, const poly_mpi_t *sk);
void gpqhe_ref_he_dec(struct he_pt *pt, const struct he_ct *ct, const poly_mpi_t *sk);
void gpqhe_ref_he_copy_ct(struct he_ct *dest, const struct he_ct *src);
void gpqhe_ref_poly_mpi_free(poly_mpi_t *a);

static size_t ct_poly(void) { return (size_t)G.info.W * polyctx.n; }

static size_t sh_hash(const void *key, size_t cap)
{
  return (size_t)((((uintptr_t)key >> 4) * 0x9E3779B97F4A7C15ull) >> 20) & (cap - 1);
}

static struct shadow *sh_find(const void *key)
{
  if (!G.shcap || !key) return NULL;
  for (size_t i = sh_hash(key, G.shcap);; i = (i + 1) & (G.shcap - 1)) {
    struct shadow *s = &G.sh[i];
    if (s->key == key) return s;
    if (!s->key && !s->tomb) return NULL;
  }
}

static void sh_grow(void)
{
  const size_t ocap = G.shcap, ncap = ocap ? 2 * ocap : 64;
  struct shadow *old = G.sh;
  G.sh = calloc(ncap, sizeof *G.sh);
  if (!G.sh) die("shadow table");
  G.shcap = ncap; G.shused = 0;
  for (size_t i = 0; i < ocap; i++)
    if (old[i].key) {
      size_t j = sh_hash(old[i].key, ncap);
      while (G.sh[j].key) j = (j + 1) & (ncap - 1);
      G.sh[j] = old[i];
      G.shused++;
    }
  free(old);
}

/* entry of a ciphertext, created (host-current, no device buffer) on first sight.  The returned pointer is valid
 * until the next sh_get(); the device buffers themselves never move. */
static struct shadow *sh_get(const he_ct_t *ct)
{
  const void *key = ct->c0.coeffs;
  struct shadow *s = sh_find(key);
  if (s) return s;
  if (2 * (G.shused + 1) > G.shcap) sh_grow();
  size_t i = sh_hash(key, G.shcap);
  while (G.sh[i].key) i = (i + 1) & (G.shcap - 1);
  s = &G.sh[i];
  if (!s->tomb) G.shused++;
  s->key = key; s->tomb = 0; s->dev = NULL; s->dev_ok = 0; s->host_ok = 1;
  return s;
}

static uint64_t *pool_take(void)
{
  if (G.npool) return G.pool[--G.npool];
  void *p = NULL;
  CK(gq_dev_alloc(ctx(), 2 * ct_poly() * 8, &p));
  G.shadow_allocs++;
  return p;
}
static void pool_give(uint64_t *p)
{
  if (!p) return;
  if (G.npool == G.cappool) {
    G.cappool = G.cappool ? 2 * G.cappool : 32;
    G.pool = realloc(G.pool, G.cappool * sizeof *G.pool);
  }
  G.pool[G.npool++] = p;       /* stream order makes reuse safe: the next writer is enqueued after the last reader */
}

static void sh_release(const void *key)
{
  struct shadow *s = sh_find(key);
  if (!s) return;
  pool_give(s->dev);
  s->key = NULL; s->dev = NULL; s->tomb = 1;
}

static void shadows_drop_all(void)
{
  if (G.ctx) gq_ctx_sync(G.ctx);
  for (size_t i = 0; i < G.shcap; i++)
    if (G.sh[i].key && G.sh[i].dev && G.ctx) gq_dev_free(G.ctx, G.sh[i].dev);
  free(G.sh); G.sh = NULL; G.shcap = G.shused = 0;
  for (unsigned i = 0; i < G.npool; i++) if (G.ctx) gq_dev_free(G.ctx, G.pool[i]);
  G.npool = 0;
  if (G.dscratch && G.ctx) gq_dev_free(G.ctx, G.dscratch);
  G.dscratch = NULL; G.dscratch_words = 0;
  if (G.ring) gq_host_free(G.ring);
  G.ring = NULL; G.ring_pos = 0;
}

/* pinned staging ring: a slot handed to an asynchronous copy is not reused before the stream has been synchronised */
#define RING_SLOTS 16
static uint64_t *ring_slot(void)
{
  const size_t poly = ct_poly();
  if (!G.ring) {       /* its own pinned buffer: stage() is rewritten synchronously by poly_mul, rns_decompose ... */
    void *p = NULL;
    if (gq_host_alloc(&p, (size_t)RING_SLOTS * poly * 8 + 64)) die("staging ring");
    G.ring = p;
  }
  if (G.ring_pos == RING_SLOTS) { CK(gq_ctx_sync(ctx())); G.ring_pos = 0; }
  return G.ring + (size_t)(G.ring_pos++) * poly;
}
static void ring_sync(void) { CK(gq_ctx_sync(ctx())); G.ring_pos = 0; }

/* upload one MPI polynomial into dst (device) */
static void poly_up(uint64_t *dst, const poly_mpi_t *a, const char *who)
{
  uint64_t *s = ring_slot();
  poly_in(s, a, G.info.W, who);
  CK(gq_dev_upload(ctx(), dst, s, ct_poly() * 8, NULL));
}

/* device buffer holding a ciphertext's current value (c0 at p, c1 at p + poly); uploads it first if the MPIs are
 * the only current copy.  The pointer stays valid whatever happens to the table afterwards. */
static uint64_t *ct_dev_in(const he_ct_t *ct, const char *who)
{
  struct shadow *s = sh_get(ct);
  if (!s->dev) s->dev = pool_take();
  if (!s->dev_ok) {
    poly_up(s->dev, &ct->c0, who);
    poly_up(s->dev + ct_poly(), &ct->c1, who);
    s->dev_ok = 1;
    G.ct_uploads++;
  }
  return s->dev;
}
/* device buffer that will receive a ciphertext's new value; ct_written() once the operation has been enqueued */
static uint64_t *ct_dev_out(he_ct_t *ct)
{
  struct shadow *s = sh_get(ct);
  if (!s->dev) s->dev = pool_take();
  return s->dev;
}
static void ct_flush(const he_ct_t *ct)
{
  struct shadow *s = sh_find(ct->c0.coeffs);
  if (!s || s->host_ok || !s->dev_ok) return;
  const size_t poly = ct_poly();
  const uint64_t *d = s->dev;
  ring_sync();
  uint64_t *a = ring_slot(), *b = ring_slot();
  CK(gq_dev_download(ctx(), a, d, poly * 8, NULL));
  CK(gq_dev_download(ctx(), b, d + poly, poly * 8, NULL));
  ring_sync();
  poly_out((poly_mpi_t *)&ct->c0, a, G.info.W);
  poly_out((poly_mpi_t *)&ct->c1, b, G.info.W);
  s->host_ok = 1;
  G.ct_downloads++;
}
static int eager(void)
{
  static int e = -1;
  if (e < 0) e = getenv("GPQHE_DROPIN_EAGER") != NULL;
  return e;
}
static void ct_written(he_ct_t *ct)
{
  struct shadow *s = sh_get(ct);
  s->dev_ok = 1; s->host_ok = 0;
  if (eager()) ct_flush(ct);
}

/* public hooks for callers that touch ct->c0.coeffs / c1.coeffs directly; gpqhe_dropin_wait() only waits for the
 * device to finish what has been enqueued (timing), nothing is downloaded */
void gpqhe_dropin_wait(void) { if (G.ctx) CK(gq_ctx_sync(G.ctx)); }
void gpqhe_dropin_sync_ct(const he_ct_t *ct) { ct_flush(ct); }
void gpqhe_dropin_touch_ct(const he_ct_t *ct)
{
  struct shadow *s = sh_find(ct->c0.coeffs);
  if (s) { s->dev_ok = 0; s->host_ok = 1; }
}

/* growable device scratch for plaintexts / encoded diagonals (words) */
static uint64_t *dscratch(size_t words)
{
  if (G.dscratch_words < words) {
    if (G.dscratch) gq_dev_free(ctx(), G.dscratch);
    void *p = NULL;
    CK(gq_dev_alloc(ctx(), words * 8 + 64, &p));
    G.dscratch = p; G.dscratch_words = words;
  }
  return G.dscratch;
}

/* ---- wrapped readers / writers of the host MPIs ---------------------------------------------------- */
void he_enc_pk(struct he_ct *ct, const struct he_pt *pt, const struct he_pk *pk)
{
  gpqhe_ref_he_enc_pk(ct, pt, pk);
  gpqhe_dropin_touch_ct(ct);
}
void he_enc_sk(struct he_ct *ct, const struct he_pt *pt, const poly_mpi_t *sk)
{
  gpqhe_ref_he_enc_sk(ct, pt, sk);
  gpqhe_dropin_touch_ct(ct);
}
void he_dec(struct he_pt *pt, const struct he_ct *ct, const poly_mpi_t *sk)
{
  ct_flush(ct);
  gpqhe_ref_he_dec(pt, ct, sk);
}
void he_copy_ct(struct he_ct *dest, const struct he_ct *src)
{
  if (dest == src || dest->c0.coeffs == src->c0.coeffs) return;
  struct shadow *s = sh_find(src->c0.coeffs);
  const int dev_ok = s && s->dev_ok, host_ok = !s || s->host_ok;
  const uint64_t *sd = s ? s->dev : NULL;
  if (!host_ok) {                            /* the current value lives on the device only: copy it there */
    uint64_t *d = ct_dev_out(dest);
    CK(gq_dev_copy(ctx(), d, sd, 2 * ct_poly() * 8, NULL));
    dest->l = src->l; dest->nu = src->nu; dest->B = src->B;        /* he-mem.c:90-92 */
    ct_written(dest);
    return;
  }
  gpqhe_ref_he_copy_ct(dest, src);
  gpqhe_dropin_touch_ct(dest);
  if (dev_ok) {                              /* both copies current: keep the device one too (no marshalling later) */
    uint64_t *d = ct_dev_out(dest);
    CK(gq_dev_copy(ctx(), d, sd, 2 * ct_poly() * 8, NULL));
    struct shadow *t = sh_get(dest);
    t->dev_ok = 1; t->host_ok = 1;
  }
}
void poly_mpi_free(poly_mpi_t *a)
{
  kg_forget(a->coeffs);
  sh_release(a->coeffs);
  gpqhe_ref_poly_mpi_free(a);
}

/* ------------------------------------------------------------------ HE primitives */
void he_mul(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2, const he_evk_t *rlk)
{
  assert(ct1->l == ct2->l);
  const size_t poly = ct_poly();
  const uint64_t *a = ct_dev_in(ct1, __func__), *b = (ct2 == ct1) ? a : ct_dev_in(ct2, __func__);
  /* metadata in the reference's statement order (he-mult.c:92-95): with ct aliasing ct1/ct2 the later statements
   * read the values the earlier ones wrote */
  ct->l = ct1->l;
  ct->nu = ct1->nu * ct2->nu;
  ct->B = (ct1->nu) * (ct2->B) + (ct2->nu) * (ct1->B) + (ct1->B) * (ct2->B) + hectx.bnd.Bmult[ct->l];
  uint64_t *o = ct_dev_out(ct);
  CK(gq_he_mul_dev(ctx(), ct->l, o, o + poly, a, a + poly, b, b + poly, evk_dev(rlk), NULL));
  ct_written(ct);
}

static const uint64_t *pt_dev(const he_pt_t *pt, unsigned slot, const char *who)
{
  uint64_t *d = dscratch((size_t)4 * ct_poly()) + (size_t)slot * ct_poly();
  poly_up(d, &pt->m, who);
  return d;
}

void he_mulpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  const size_t poly = ct_poly();
  const uint64_t *a = ct_dev_in(src, __func__);
  dest->l = src->l;                                                     /* he-mult.c:162-164 */
  dest->nu = src->nu * pt->nu;
  dest->B = src->B * pt->nu;
  const unsigned l = dest->l;
  /* limb count in double arithmetic, truncated, exactly as he-mult.c:168 */
  const unsigned dim = (gcry_mpi_get_nbits(hectx.q[l]) + log2(pt->nu) + polyctx.logn) / GPQHE_LOGP + 1;
  const uint64_t *m = pt_dev(pt, 0, __func__);
  uint64_t *o = ct_dev_out(dest);
  CK(gq_he_mulpt_dev(ctx(), l, dim, o, o + poly, a, a + poly, m, NULL));
  ct_written(dest);
}

static void ct_inplace(he_ct_t *ct, int op, unsigned rot, const he_evk_t *k, const char *who)
{
  const size_t poly = ct_poly();
  uint64_t *s = ct_dev_in(ct, who);
  switch (op) {
    case 0: CK(gq_he_rs_dev(ctx(), ct->l, s, s + poly, NULL)); break;
    case 1: CK(gq_he_moddown_dev(ctx(), ct->l, s, s + poly, NULL)); break;
    case 2: CK(gq_he_rot_dev(ctx(), ct->l, rot, s, s + poly, evk_dev(k), NULL)); break;
    case 3: CK(gq_he_conj_dev(ctx(), ct->l, s, s + poly, evk_dev(k), NULL)); break;
  }
  ct_written(ct);
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
/* op: 0 add, 1 sub (b = second ciphertext); 2 addpt, 3 subpt (pt); 4 neg */
static void ct_linear(he_ct_t *ct, const he_ct_t *a, const he_ct_t *b, const he_pt_t *pt, int op, unsigned l,
                      const char *who)
{
  const size_t poly = ct_poly();
  const uint64_t *pa = ct_dev_in(a, who), *pb = NULL;
  if (b) pb = (b == a) ? pa : ct_dev_in(b, who);
  if (pt) pb = pt_dev(pt, 0, who);
  uint64_t *o = ct_dev_out(ct);
  switch (op) {
    case 0: CK(gq_he_add_dev(ctx(), l, o, o + poly, pa, pa + poly, pb, pb + poly, NULL)); break;
    case 1: CK(gq_he_sub_dev(ctx(), l, o, o + poly, pa, pa + poly, pb, pb + poly, NULL)); break;
    case 2: CK(gq_he_addpt_dev(ctx(), l, o, o + poly, pa, pa + poly, pb, NULL)); break;
    case 3: CK(gq_he_subpt_dev(ctx(), l, o, o + poly, pa, pa + poly, pb, NULL)); break;
    case 4:
      if (o != pa) CK(gq_dev_copy(ctx(), o, pa, 2 * poly * 8, NULL));
      CK(gq_he_neg_dev(ctx(), l, o, o + poly, NULL));
      break;
  }
  ct_written(ct);
}

void he_add(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2)
{
  assert(ct1->l == ct2->l);
  const unsigned l = ct1->l;
  const double nu = (ct1->nu >= ct2->nu) ? ct1->nu : ct2->nu, B = ct1->B + ct2->B;    /* he-add.c:36-38 */
  ct_linear(ct, ct1, ct2, NULL, 0, l, __func__);
  ct->l = l; ct->nu = nu; ct->B = B;
}
void he_sub(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2)
{
  assert(ct1->l == ct2->l);
  const unsigned l = ct1->l;
  const double nu = (ct1->nu >= ct2->nu) ? ct1->nu : ct2->nu, B = ct1->B + ct2->B;    /* he-add.c:61-63 */
  ct_linear(ct, ct1, ct2, NULL, 1, l, __func__);
  ct->l = l; ct->nu = nu; ct->B = B;
}
void he_addpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  const unsigned l = src->l;
  const double nu = (src->nu >= pt->nu) ? src->nu : pt->nu, B = src->B;                /* he-add.c:85-87 */
  ct_linear(dest, src, NULL, pt, 2, l, __func__);
  dest->l = l; dest->nu = nu; dest->B = B;
}
void he_subpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt)
{
  const unsigned l = src->l;
  const double nu = (src->nu >= pt->nu) ? src->nu : pt->nu, B = src->B;                /* he-add.c:109-111 */
  ct_linear(dest, src, NULL, pt, 3, l, __func__);
  dest->l = l; dest->nu = nu; dest->B = B;
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

/* The encoded diagonals of a matrix, device resident and cached.  The reference re-encodes every diagonal on every
 * he_gemv call (he_ecd at src/he-algo.c:70-72: an inverse canonical embedding plus N double_to_mpi per diagonal,
 * 1-2 ms each at N = 2^14 - for he_coeff2slot's 4 x slots diagonals that is 10 times the device work).  he_ecd is a
 * pure function of the matrix (and of the context), so the encodings are kept on the device keyed by the EXACT
 * contents of A (a copy is compared, no hashing risk) and reused as long as the context lives; the matrices of
 * bootstrapctx never change after he_bootstrapctx_init.  Diagonal k = i*n1+j of A is pre-rotated by -i*n1 slots and
 * encoded at scale Delta (he-algo.c:70-72), stored as word-major polys at dev + k*poly.  `bits` bounds the encoded
 * coefficients (|m| < 2^bits), measured while marshalling: it lets the device accumulate a giant step in the NTT
 * domain (gq_ctx_set_plaintext_bits).  Bounded by GPQHE_DROPIN_DIAG_CACHE_MB (default 4096), least recently used out. */
struct diag_entry {
  _Complex double *A;       /* private copy of the matrix, m*m entries */
  unsigned m, n1;
  uint64_t *dev;            /* m polys */
  double nu;                /* pt.nu of he_ecd (the same for every diagonal: he-encode.c) */
  unsigned bits;
  unsigned long long used;
};
static struct diag_entry *G_diag;
static unsigned G_ndiag, G_capdiag;
static unsigned long long G_diag_clock, G_diag_hits, G_diag_misses;

static void diag_drop(unsigned i)
{
  gq_dev_free(ctx(), G_diag[i].dev);        /* waits for the stream first */
  free(G_diag[i].A);
  G_diag[i] = G_diag[--G_ndiag];
}
static void diag_drop_all(void)
{
  while (G_ndiag) { if (G.ctx) gq_dev_free(G.ctx, G_diag[G_ndiag - 1].dev); free(G_diag[G_ndiag - 1].A); G_ndiag--; }
}

static const struct diag_entry *diag_get(const _Complex double *A, unsigned n1, const char *who)
{
  const unsigned m = hectx.slots;
  const size_t poly = ct_poly(), abytes = sizeof(_Complex double) * m * m;
  G_diag_clock++;
  for (unsigned i = 0; i < G_ndiag; i++)
    if (G_diag[i].m == m && G_diag[i].n1 == n1 && !memcmp(G_diag[i].A, A, abytes)) {
      G_diag[i].used = G_diag_clock;
      G_diag_hits++;
      return &G_diag[i];
    }
  G_diag_misses++;
  static size_t cap;
  if (!cap) {
    const char *e = getenv("GPQHE_DROPIN_DIAG_CACHE_MB");
    long long mb = e ? atoll(e) : 4096;
    cap = (size_t)(mb < 0 ? 0 : mb) << 20;
  }
  size_t held = 0;
  for (unsigned i = 0; i < G_ndiag; i++) held += (size_t)G_diag[i].m * poly * 8;
  while (G_ndiag && held + (size_t)m * poly * 8 > cap) {
    unsigned lru = 0;
    for (unsigned i = 1; i < G_ndiag; i++) if (G_diag[i].used < G_diag[lru].used) lru = i;
    if (G_diag_clock - G_diag[lru].used < 8) break;      /* one call holds up to 4 matrices (he_coeff2slot) */
    held -= (size_t)G_diag[lru].m * poly * 8;
    diag_drop(lru);
  }
  if (G_ndiag == G_capdiag) {
    G_capdiag = G_capdiag ? 2 * G_capdiag : 8;
    G_diag = realloc(G_diag, G_capdiag * sizeof *G_diag);
  }
  struct diag_entry *e = &G_diag[G_ndiag++];
  e->m = m; e->n1 = n1; e->used = G_diag_clock; e->bits = 1;
  e->A = malloc(abytes);
  if (!e->A) die(who);
  memcpy(e->A, A, abytes);
  void *dv = NULL;
  CK(gq_dev_alloc(ctx(), (size_t)m * poly * 8, &dv));
  e->dev = dv;
  he_pt_t pt;
  he_alloc_pt(&pt);
  _Complex double *d = malloc(sizeof(_Complex double) * m);
  if (!d) die(who);
  for (unsigned k = 0; k < m; k++) {
    const unsigned shift = (k / n1) * n1;
    for (unsigned t = 0; t < m; t++) {
      const unsigned u = (t + m - shift % m) % m;                       /* slot after rotating by -shift */
      d[t] = A[(size_t)u * m + (k + u) % m];
    }
    he_ecd(&pt, d);
    for (unsigned i = 0; i < polyctx.n; i++) {
      const unsigned nb = gcry_mpi_get_nbits(pt.m.coeffs[i]) + 1;       /* |m_i| < 2^nbits */
      if (nb > e->bits) e->bits = nb;
    }
    poly_up(e->dev + (size_t)k * poly, &pt.m, who);
  }
  e->nu = pt.nu;
  free(d);
  he_free_pt(&pt);
  return e;
}
void gpqhe_dropin_diag_stats(unsigned long long *hits, unsigned long long *misses)
{
  if (hits) *hits = G_diag_hits;
  if (misses) *misses = G_diag_misses;
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
  const unsigned m = hectx.slots, l = ct->l;
  unsigned n1, n2;
  gemv_split(&n1, &n2);
  const size_t poly = ct_poly();
  const uint64_t *a = ct_dev_in(ct, __func__);
  const struct diag_entry *dg = diag_get(A, n1, __func__);
  const uint64_t **pts = malloc(sizeof(*pts) * m);
  if (!pts) die(__func__);
  for (unsigned k = 0; k < m; k++) pts[k] = dg->dev + (size_t)k * poly;
  const gq_evk **keys = gemv_keys(rk, n1, __func__);
  const unsigned dim = (gcry_mpi_get_nbits(hectx.q[l]) + log2(dg->nu) + polyctx.logn) / GPQHE_LOGP + 1;
  double nu_o, B_o;
  gemv_meta(ct->nu, ct->B, dg->nu, n1, n2, &nu_o, &B_o);
  uint64_t *o = ct_dev_out(ct_dest);
  CK(gq_ctx_set_plaintext_bits(ctx(), dg->bits));
  CK(gq_he_gemv_dev(ctx(), l, n1, n2, dim, o, o + poly, a, a + poly, pts, keys, NULL));
  ct_dest->l = l - 1;
  ct_dest->nu = nu_o;
  ct_dest->B = B_o;
  ct_written(ct_dest);
  free(pts); free(keys);
}

/* ---- he_coeff2slot (src/he-bootstrap.c:359-409), device resident ----------------------------------
 * he_conj, four he_gemv, he_add, he_mulpt by he_const_pt(1/n) and he_rs for the real and the imaginary
 * half: one crossing in (the ciphertext + the encoded diagonals of the four matrices of bootstrapctx),
 * one crossing out; the reference's own composition stays available as gpqhe_ref_he_coeff2slot. */
extern struct bootstrap_ctx bootstrapctx;
void he_coeff2slot(he_ct_t *ct_real, he_ct_t *ct_imag, const he_ct_t *ct, const he_evk_t *ck, const he_evk_t *rk)
{
  const unsigned m = hectx.slots, l = ct->l;
  unsigned n1, n2;
  gemv_split(&n1, &n2);
  const size_t poly = ct_poly();
  if (ct_real == ct_imag) { fprintf(stderr, "he_coeff2slot: the two outputs must be distinct ciphertexts\n"); errno = EINVAL; abort(); }
  const uint64_t *a = ct_dev_in(ct, __func__);
  he_pt_t cpt;
  he_alloc_pt(&cpt);
  he_const_pt(&cpt, 1 / polyctx.n);                                     /* he-bootstrap.c:385, integer division as there */
  const uint64_t *dcpt = pt_dev(&cpt, 1, __func__);
  const uint64_t **pts = malloc(sizeof(*pts) * 4 * m);
  if (!pts) die(__func__);
  const _Complex double *mats[4] = {bootstrapctx.U0_conjT, bootstrapctx.U0_T, bootstrapctx.U1_conjT, bootstrapctx.U1_T};
  double ptnu = 0;
  unsigned ptbits = 1;
  for (unsigned k = 0; k < 4; k++) {
    const struct diag_entry *dg = diag_get(mats[k], n1, __func__);
    uint64_t *dv = dg->dev;                  /* the entry may move when the next matrix is inserted; its buffer does not */
    ptnu = dg->nu;
    if (dg->bits > ptbits) ptbits = dg->bits;
    for (unsigned t = 0; t < m; t++) pts[(size_t)k * m + t] = dv + (size_t)t * poly;
  }
  const gq_evk **keys = gemv_keys(rk, n1, __func__);
  CK(gq_ctx_set_plaintext_bits(ctx(), ptbits));
  const unsigned dim_gemv = (gcry_mpi_get_nbits(hectx.q[l]) + log2(ptnu) + polyctx.logn) / GPQHE_LOGP + 1;
  const unsigned dim_pt = (gcry_mpi_get_nbits(hectx.q[l - 1]) + log2(cpt.nu) + polyctx.logn) / GPQHE_LOGP + 1;
  double nu_g, B_g;
  gemv_meta(ct->nu, ct->B, ptnu, n1, n2, &nu_g, &B_g);                  /* he_conj keeps nu and B */
  /* an output may alias the input (the device call reads `a` up to its last he_gemv): such an output is computed
   * into a spare buffer and copied */
  uint64_t *o_re = ct_dev_out(ct_real), *o_im = ct_dev_out(ct_imag);
  uint64_t *tmp = (o_re == a || o_im == a) ? pool_take() : NULL;
  uint64_t *w_re = (o_re == a) ? tmp : o_re, *w_im = (o_im == a) ? tmp : o_im;
  CK(gq_he_coeff2slot_dev(ctx(), l, n1, n2, dim_gemv, dim_pt, w_re, w_re + poly, w_im, w_im + poly, a, a + poly,
                          pts, dcpt, evk_dev(ck), keys, NULL));
  if (w_re != o_re) CK(gq_dev_copy(ctx(), o_re, w_re, 2 * poly * 8, NULL));
  if (w_im != o_im) CK(gq_dev_copy(ctx(), o_im, w_im, 2 * poly * 8, NULL));
  pool_give(tmp);
  he_ct_t *outs[2] = {ct_real, ct_imag};
  for (unsigned k = 0; k < 2; k++) {
    double nu = nu_g, B = B_g + B_g;                                    /* he_add, he-add.c:36-38 */
    nu = nu * cpt.nu; B = B * cpt.nu;                                   /* he_mulpt, he-mult.c:162-164 */
    outs[k]->l = l - 2;                                                 /* he_rs, he-rescale.c:36-38 */
    outs[k]->nu = nu / hectx.Delta;
    outs[k]->B = B / hectx.Delta + hectx.bnd.Brs;
    ct_written(outs[k]);
  }
  free(pts); free(keys);
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

/* the 8 plaintexts of he_exp uploaded to the DEVICE buffer dst (pts[0] = he_ecd of a/2^iter in every slot, he-algo.c:443-447; pts[1..7] =
 * he_const_pt of the Taylor constants, he-algo.c:384-411); returns pt->nu (the same for all of them) */
static double exp_plaintexts(_Complex double a, unsigned iter, uint64_t *dst, const uint64_t **pts, const char *who)
{
  static const double consts[7] = {1.0, 3.0, 1.0 / 6, 5.0, 1.0 / 120, 7.0, 1.0 / 5040};
  const unsigned m = hectx.slots;
  const size_t poly = ct_poly();
  he_pt_t pt;
  he_alloc_pt(&pt);
  _Complex double *coeff = malloc(sizeof(_Complex double) * m);
  if (!coeff) die(who);
  a /= (1 << iter);
  for (unsigned i = 0; i < m; i++) coeff[i] = a;
  he_ecd(&pt, coeff);
  poly_up(dst, &pt.m, who);
  pts[0] = dst;
  he_pt_t cpt;                        /* he_exp_taylor's own plaintext: he_const_pt writes coefficients 0 and n/2 only */
  he_alloc_pt(&cpt);
  for (unsigned k = 0; k < 7; k++) {
    he_const_pt(&cpt, consts[k]);
    poly_up(dst + (size_t)(k + 1) * poly, &cpt.m, who);
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
  const unsigned l = ct->l;
  const size_t poly = ct_poly();
  if (l < 4 + iter) { fprintf(stderr, "he_exp: level %u is below the 4+iter levels the evaluation consumes\n", l); errno = EINVAL; abort(); }
  const uint64_t *pts[8];
  const uint64_t *x = ct_dev_in(ct, __func__);
  uint64_t *dp = dscratch((4 + 9) * poly) + 4 * poly;
  const double ptnu = exp_plaintexts(a, iter, dp, pts, __func__);
  const struct meta in = {ct->nu, ct->B}, out = exp_meta(l, in, ptnu, iter);
  uint64_t *o = ct_dev_out(ct_exp);
  CK(gq_he_exp_dev(ctx(), l, iter, mulpt_dim(l, ptnu), mulpt_dim(l - 1, ptnu), o, o + poly, x, x + poly,
                   pts, evk_dev(rlk), NULL));
  ct_exp->l = l - 4 - iter;
  ct_exp->nu = out.nu;
  ct_exp->B = out.B;
  ct_written(ct_exp);
}

void he_rlsin(he_ct_t *ct_dest, const double a, const he_ct_t *ct, const he_evk_t *rlk, const he_evk_t *ck,
              const unsigned int iter)
{
  const unsigned m = hectx.slots, l = ct->l;
  const size_t poly = ct_poly();
  if (l < 5 + iter) { fprintf(stderr, "he_rlsin: level %u is below the 5+iter levels the evaluation consumes\n", l); errno = EINVAL; abort(); }
  const uint64_t *pts[9];
  const uint64_t *x = ct_dev_in(ct, __func__);
  uint64_t *dp = dscratch((4 + 9) * poly) + 4 * poly;
  const double ptnu = exp_plaintexts(a * I, iter, dp, pts, __func__);   /* he-bootstrap.c:458 */
  he_pt_t pt;
  he_alloc_pt(&pt);
  _Complex double *coeff = malloc(sizeof(_Complex double) * m);
  if (!coeff) die(__func__);
  const _Complex double b = 1 / (2 * I * a);                                      /* he-bootstrap.c:465-468 */
  for (unsigned i = 0; i < m; i++) coeff[i] = b;
  he_ecd(&pt, coeff);
  poly_up(dp + 8 * poly, &pt.m, __func__);
  pts[8] = dp + 8 * poly;
  const struct meta in = {ct->nu, ct->B}, e = exp_meta(l, in, ptnu, iter);
  const struct meta out = m_rs(m_mulpt(m_add(e, e), pt.nu));       /* he_conj keeps nu, B; he_sub updates as he_add */
  const unsigned lf = l - 4 - iter;
  uint64_t *o = ct_dev_out(ct_dest);
  CK(gq_he_rlsin_dev(ctx(), l, iter, mulpt_dim(l, ptnu), mulpt_dim(l - 1, ptnu), mulpt_dim(lf, pt.nu), o, o + poly,
                     x, x + poly, pts, evk_dev(rlk), evk_dev(ck), NULL));
  ct_dest->l = lf - 1;
  ct_dest->nu = out.nu;
  ct_dest->B = out.B;
  ct_written(ct_dest);
  free(coeff);
  he_free_pt(&pt);
}

/* diagnostics for tests: coefficients marshalled in / out since load, whether the direct limb path is active, and
 * how many whole ciphertexts crossed (uploads / downloads, device buffers allocated for shadows) */
void gpqhe_dropin_stats(unsigned long long *in, unsigned long long *out, int *fast_mpi)
{
  if (in) *in = G.marshal_in;
  if (out) *out = G.marshal_out;
  if (fast_mpi) *fast_mpi = G.fast_mpi;
}
void gpqhe_dropin_ct_stats(unsigned long long *uploads, unsigned long long *downloads, unsigned long long *allocs)
{
  if (uploads) *uploads = G.ct_uploads;
  if (downloads) *downloads = G.ct_downloads;
  if (allocs) *allocs = G.shadow_allocs;
}
