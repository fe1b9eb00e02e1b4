/*
 * dropin_diff.c - differential driver for dropin/_build/libgpqhe.so (TEST TOOL).
 *
 * That library holds, side by side, the GPU implementation of every hot symbol under the
 * reference's own name and the reference's unmodified CPU definition under gpqhe_ref_<name>
 * (dropin/Makefile).  This program builds one context with the reference's API, makes keys
 * and ciphertexts under the deterministic -DSUPERCOP stream, runs each hot entry point both
 * ways on the same inputs and compares every output bit for bit (uint64_t limbs, MPI values,
 * ct metadata).
 *
 *   dropin_diff <logn> <logq> <slots> <logDelta> [nrot]
 * prints one line per case and "PASS k/k" (exit 0) or "FAIL" (exit 1).
 */
#include "config.h"
#include "types.h"
#include "gpqhe.h"

#include <complex.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

extern struct poly_ctx polyctx;
extern struct he_ctx hectx;
extern void ntt(uint64_t a[], const struct rns_ctx *rns);
extern void invntt(uint64_t a[], const struct rns_ctx *rns);
extern void rns_decompose(uint64_t ahat[], const MPI a[], const struct rns_ctx *rns);
extern void sample_z01vec(_Complex double *vec, const size_t m);

/* the reference's CPU definitions kept in the same library */
void gpqhe_ref_ntt(uint64_t a[], const struct rns_ctx *rns);
void gpqhe_ref_invntt(uint64_t a[], const struct rns_ctx *rns);
void gpqhe_ref_poly_rns_add(uint64_t r[], const uint64_t a[], const uint64_t b[], const struct rns_ctx *rns);
void gpqhe_ref_poly_rns_mul(uint64_t r[], const uint64_t a[], const uint64_t b[], const struct rns_ctx *rns);
void gpqhe_ref_rns_decompose(uint64_t ahat[], const MPI a[], const struct rns_ctx *rns);
void gpqhe_ref_poly_rns2mpi(poly_mpi_t *r, const poly_rns_t *rhat, const struct rns_ctx *rns, const MPI q);
void gpqhe_ref_poly_mul(poly_mpi_t *r, const poly_mpi_t *a, const poly_mpi_t *b, const unsigned int dim, const MPI q);
void gpqhe_ref_poly_rot(poly_mpi_t *r, const poly_mpi_t *a, size_t rot);
void gpqhe_ref_poly_conj(poly_mpi_t *r, const poly_mpi_t *a);
void gpqhe_ref_he_mul(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2, const he_evk_t *rlk);
void gpqhe_ref_he_mulpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt);
void gpqhe_ref_he_rs(he_ct_t *ct);
void gpqhe_ref_he_moddown(he_ct_t *ct);
void gpqhe_ref_he_rot(he_ct_t *ct, const int rot, const he_evk_t *rk);
void gpqhe_ref_he_conj(he_ct_t *ct, const he_evk_t *ck);
void gpqhe_ref_he_add(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2);
void gpqhe_ref_he_sub(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2);
void gpqhe_ref_he_addpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt);
void gpqhe_ref_he_subpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt);
void gpqhe_ref_he_neg(he_ct_t *ct);
void gpqhe_ref_he_gemv(he_ct_t *ct_dest, const _Complex double *A, const he_ct_t *ct, const he_evk_t *rk);
void gpqhe_ref_he_coeff2slot(he_ct_t *ct_real, he_ct_t *ct_imag, const he_ct_t *ct, const he_evk_t *ck, const he_evk_t *rk);
void gpqhe_ref_he_exp(he_ct_t *ct_exp, _Complex double a, const he_ct_t *ct, const he_evk_t *rlk, const unsigned int iter);
void gpqhe_ref_he_rlsin(he_ct_t *ct_dest, const double a, const he_ct_t *ct, const he_evk_t *rlk, const he_evk_t *ck,
                        const unsigned int iter);
void gpqhe_dropin_stats(unsigned long long *in, unsigned long long *out, int *fast_mpi);
void gpqhe_dropin_ct_stats(unsigned long long *uploads, unsigned long long *downloads, unsigned long long *allocs);
/* ciphertexts live in device shadows between drop-in calls (dropin/gpqhe_dropin.c): this program reads and writes
 * the MPIs directly (eq_poly, the pure-CPU gpqhe_ref_* bodies), so it uses the two hooks a caller that pokes
 * ct->c0.coeffs is told to use */
void gpqhe_dropin_sync_ct(const he_ct_t *ct);
void gpqhe_dropin_touch_ct(const he_ct_t *ct);
void gpqhe_dropin_wait(void);
#define REFW(ct, call) do { call; gpqhe_dropin_touch_ct(ct); } while (0)   /* a pure-CPU reference body wrote *ct */

#include <time.h>
static double now_ms(void)
{
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}
/* wall time of the drop-in symbol INCLUDING the write-back of its result into the MPIs (gpqhe_dropin_sync_ct: what an
 * isolated call whose result is read at once costs; chained calls skip it, see "chained" below) next to the reference's definition: pure
 * CPU for the primitives; for the compositions marked * the reference's host loop, whose primitive calls bind to the
 * drop-in's GPU symbols (one crossing per primitive) */
#define TIMED2(label, ours, ref) do { double t0 = now_ms(); ours; gpqhe_dropin_sync_ct(&g); double t1 = now_ms(); ref; gpqhe_dropin_sync_ct(&r); double t2 = now_ms(); \
  printf("  time %-28s drop-in %10.3f ms   reference %12.3f ms   x%.1f\n", label, t1 - t0, t2 - t1, (t2 - t1) / (t1 - t0 > 1e-6 ? t1 - t0 : 1e-6)); } while (0)

static unsigned npass, nfail;
static void report(const char *name, int ok)
{
  printf("%-34s %s\n", name, ok ? "ok" : "MISMATCH");
  fflush(stdout);
  if (ok) npass++; else nfail++;
}

static int eq_poly(const poly_mpi_t *a, const poly_mpi_t *b)
{
  for (unsigned i = 0; i < polyctx.n; i++)
    if (gcry_mpi_cmp(a->coeffs[i], b->coeffs[i])) return 0;
  return 1;
}
static int eq_ct(const he_ct_t *a, const he_ct_t *b)
{
  gpqhe_dropin_sync_ct(a);
  gpqhe_dropin_sync_ct(b);
  return a->l == b->l && a->nu == b->nu && a->B == b->B && eq_poly(&a->c0, &b->c0) && eq_poly(&a->c1, &b->c1);
}
static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void)
{
  uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

int main(int argc, char **argv)
{
  if (argc < 5) { fprintf(stderr, "usage: %s logn logq slots logDelta [nrot]\n", argv[0]); return 2; }
  const unsigned logn = atoi(argv[1]), logq = atoi(argv[2]), slots = atoi(argv[3]), logDelta = atoi(argv[4]);
  const unsigned nrot = argc > 5 ? atoi(argv[5]) : 2;
  MPI q = mpi_new(0);
  mpi_set_ui(q, 1);
  mpi_lshift(q, q, logq);
  hectx_init(logn, q, slots, 1ull << logDelta);
  const unsigned n = polyctx.n;
  int fast = 0;
  gpqhe_dropin_stats(NULL, NULL, &fast);
  printf("N=2^%u logq=%u slots=%u logDelta=%u L=%u dim=%u dimevk=%u dimub=%u mpi-path=%s\n", logn, logq, slots, logDelta,
         hectx.L, hectx.dim, hectx.dimevk, polyctx.dimub, fast ? "direct-limb" : "gcry_mpi_print/scan");

  /* ---- limb layer on a few primes */
  {
    uint64_t *a = malloc(n * 8), *b = malloc(n * 8), *x = malloc(n * 8), *y = malloc(n * 8);
    const struct rns_ctx *r = polyctx.rns;
    int ok_f = 1, ok_i = 1, ok_m = 1, ok_a = 1;
    for (unsigned d = 0; d < polyctx.dimub; d++, r = r->next) {
      if (d > 1 && d != polyctx.dimub - 1) continue;
      for (unsigned i = 0; i < n; i++) { a[i] = rnd() % r->p; b[i] = rnd() % r->p; }
      memcpy(x, a, n * 8); memcpy(y, a, n * 8);
      ntt(x, r); gpqhe_ref_ntt(y, r);
      ok_f &= !memcmp(x, y, n * 8);
      memcpy(x, b, n * 8); memcpy(y, b, n * 8);
      invntt(x, r); gpqhe_ref_invntt(y, r);
      ok_i &= !memcmp(x, y, n * 8);
      poly_rns_mul(x, a, b, r); gpqhe_ref_poly_rns_mul(y, a, b, r);
      ok_m &= !memcmp(x, y, n * 8);
      poly_rns_add(x, a, b, r); gpqhe_ref_poly_rns_add(y, a, b, r);
      ok_a &= !memcmp(x, y, n * 8);
    }
    report("ntt", ok_f); report("invntt", ok_i); report("poly_rns_mul", ok_m); report("poly_rns_add", ok_a);
    free(a); free(b); free(x); free(y);
  }

  /* ---- keys and ciphertexts (deterministic stream) */
  poly_mpi_t sk; he_pk_t pk; he_evk_t rlk, ck, *rk;
  he_alloc_sk(&sk); he_alloc_pk(&pk); he_alloc_evk(&rlk); he_alloc_evk(&ck);
  rk = malloc(slots * sizeof *rk);
  for (unsigned i = 0; i < slots; i++) he_alloc_evk(&rk[i]);
  he_keypair(&pk, &sk);
  he_genrlk(&rlk, &sk);
  he_genck(&ck, &sk);
  he_genrk(rk, &sk);
  he_pt_t pt1, pt2; he_ct_t c1, c2, g, r;
  he_alloc_pt(&pt1); he_alloc_pt(&pt2);
  he_alloc_ct(&c1); he_alloc_ct(&c2); he_alloc_ct(&g); he_alloc_ct(&r);
  _Complex double m1[slots], m2[slots];
  sample_z01vec(m1, slots); sample_z01vec(m2, slots);
  he_ecd(&pt1, m1); he_ecd(&pt2, m2);
  he_enc_pk(&c1, &pt1, &pk);
  he_enc_pk(&c2, &pt2, &pk);

  /* ---- MPI <-> RNS */
  {
    uint64_t *x = malloc(n * 8), *y = malloc(n * 8);
    const struct rns_ctx *rn = polyctx.rns;
    int ok = 1;
    for (unsigned d = 0; d < hectx.dimevk; d++, rn = rn->next) {
      rns_decompose(x, c1.c1.coeffs, rn); gpqhe_ref_rns_decompose(y, c1.c1.coeffs, rn);
      ok &= !memcmp(x, y, n * 8);
    }
    report("rns_decompose (all evk limbs)", ok);
    free(x); free(y);
    poly_rns_t hat; poly_rns_alloc(&hat, hectx.dim);
    rn = polyctx.rns;
    for (unsigned d = 0; d < hectx.dim; d++) {
      for (unsigned i = 0; i < n; i++) hat.coeffs[d * n + i] = rnd() % rn->p;
      if (d < hectx.dim - 1) rn = rn->next;
    }
    poly_rns2mpi(&g.c0, &hat, rn, hectx.q[hectx.L]); gpqhe_ref_poly_rns2mpi(&r.c0, &hat, rn, hectx.q[hectx.L]);
    report("poly_rns2mpi", eq_poly(&g.c0, &r.c0));
    poly_rns_free(&hat);
    poly_mul(&g.c0, &c1.c0, &c2.c1, hectx.dim, hectx.q[hectx.L]);
    gpqhe_ref_poly_mul(&r.c0, &c1.c0, &c2.c1, hectx.dim, hectx.q[hectx.L]);
    report("poly_mul", eq_poly(&g.c0, &r.c0));
    poly_rot(&g.c0, &c1.c0, 3); gpqhe_ref_poly_rot(&r.c0, &c1.c0, 3);
    report("poly_rot", eq_poly(&g.c0, &r.c0));
    poly_conj(&g.c0, &c1.c1); gpqhe_ref_poly_conj(&r.c0, &c1.c1);
    report("poly_conj", eq_poly(&g.c0, &r.c0));
  }

  /* ---- HE primitives, walking down the levels like tests/gpqhe.c does */
  he_mul(&g, &c1, &c2, &rlk);                                       /* first use uploads the key */
  TIMED2("he_mul", he_mul(&g, &c1, &c2, &rlk), REFW(&r, gpqhe_ref_he_mul(&r, &c1, &c2, &rlk)));
  report("he_mul", eq_ct(&g, &r));
  TIMED2("he_rs", he_rs(&g), REFW(&r, gpqhe_ref_he_rs(&r)));
  report("he_rs", eq_ct(&g, &r));
  he_mulpt(&g, &g, &pt2); REFW(&r, gpqhe_ref_he_mulpt(&r, &r, &pt2));      /* dest aliases src */
  report("he_mulpt (in place)", eq_ct(&g, &r));
  he_moddown(&g); REFW(&r, gpqhe_ref_he_moddown(&r));
  report("he_moddown", eq_ct(&g, &r));
  he_mul(&g, &g, &g, &rlk); REFW(&r, gpqhe_ref_he_mul(&r, &r, &r, &rlk));   /* he_mul(&x,&x,&x) at a lower level */
  report("he_mul (x,x,x) lower level", eq_ct(&g, &r));
  for (unsigned rot = 0; rot < nrot && rot < slots; rot++) {
    he_copy_ct(&g, &c1); he_copy_ct(&r, &c1);
    he_rot(&g, (int)rot, rk); REFW(&r, gpqhe_ref_he_rot(&r, (int)rot, rk));
    char nm[40]; snprintf(nm, sizeof nm, "he_rot %u", rot);
    report(nm, eq_ct(&g, &r));
  }
  he_copy_ct(&g, &c2); he_copy_ct(&r, &c2);
  he_conj(&g, &ck); REFW(&r, gpqhe_ref_he_conj(&r, &ck));
  report("he_conj", eq_ct(&g, &r));

  /* ---- coefficient-local ops (src/he-add.c), aliased forms included */
  he_add(&g, &c1, &c2); REFW(&r, gpqhe_ref_he_add(&r, &c1, &c2));
  report("he_add", eq_ct(&g, &r));
  he_sub(&g, &g, &c1); REFW(&r, gpqhe_ref_he_sub(&r, &r, &c1));            /* dest aliases ct1 */
  report("he_sub (in place)", eq_ct(&g, &r));
  he_add(&g, &g, &g); REFW(&r, gpqhe_ref_he_add(&r, &r, &r));              /* x + x: wraps past q/2 */
  report("he_add (x,x,x)", eq_ct(&g, &r));
  he_neg(&g); REFW(&r, gpqhe_ref_he_neg(&r));
  report("he_neg", eq_ct(&g, &r));
  he_addpt(&g, &c2, &pt2); REFW(&r, gpqhe_ref_he_addpt(&r, &c2, &pt2));
  report("he_addpt", eq_ct(&g, &r));
  he_subpt(&g, &g, &pt2); REFW(&r, gpqhe_ref_he_subpt(&r, &r, &pt2));
  report("he_subpt (in place)", eq_ct(&g, &r));

  /* ---- he_gemv: device-resident loop vs the reference's host loop over the same primitives */
  if (nrot >= slots) {
    _Complex double *A = malloc(sizeof(_Complex double) * slots * slots);
    for (unsigned i = 0; i < slots * slots; i++)
      A[i] = ((double)(rnd() % 2001) - 1000.0) / 1000.0 + (((double)(rnd() % 2001) - 1000.0) / 1000.0) * I;
    TIMED2("he_gemv *", he_gemv(&g, A, &c1, rk), gpqhe_ref_he_gemv(&r, A, &c1, rk));
    report("he_gemv", eq_ct(&g, &r));
    he_copy_ct(&g, &c2); he_copy_ct(&r, &c2);
    he_gemv(&g, A, &g, rk); gpqhe_ref_he_gemv(&r, A, &r, rk);     /* dest aliases src (he_sum does this) */
    report("he_gemv (in place)", eq_ct(&g, &r));
    free(A);
    /* ---- he_coeff2slot (src/he-bootstrap.c:359-409): one device-resident call vs the reference's own
     * composition (he_conj, 4 he_gemv, he_add, he_mulpt, he_rs as separate calls) */
    he_ct_t gi, ri;
    he_alloc_ct(&gi); he_alloc_ct(&ri);
    he_bootstrapctx_init();
    TIMED2("he_coeff2slot *", he_coeff2slot(&g, &gi, &c1, &ck, rk), gpqhe_ref_he_coeff2slot(&r, &ri, &c1, &ck, rk));
    report("he_coeff2slot (real)", eq_ct(&g, &r));
    report("he_coeff2slot (imag)", eq_ct(&gi, &ri));
    he_bootstrapctx_exit();
    he_free_ct(&gi); he_free_ct(&ri);
  }

  /* ---- he_exp (src/he-algo.c:363-458) and he_rlsin (src/he-bootstrap.c:444-475, EvalMod of he_bootstrap): one
   * device-resident call each vs the reference's own compositions (about 30 primitive calls on host ciphertexts).
   * Needs 5+iter levels; DROPIN_DIFF_NO_EXP=1 skips it (the CPU side takes minutes at N=2^16). */
  if (hectx.L >= 6 && !getenv("DROPIN_DIFF_NO_EXP")) {
    const unsigned iter = hectx.L >= 8 ? 2 : 1;
    const _Complex double a = 0.37 - 0.21 * I;
    TIMED2("he_exp *", he_exp(&g, a, &c1, &rlk, iter), gpqhe_ref_he_exp(&r, a, &c1, &rlk, iter));
    report("he_exp", eq_ct(&g, &r));
    if (!eq_ct(&g, &r)) printf("  l %u/%u nu %.17g/%.17g B %.17g/%.17g c0 %d c1 %d\n", g.l, r.l, g.nu, r.nu, g.B, r.B, eq_poly(&g.c0, &r.c0), eq_poly(&g.c1, &r.c1));
    he_copy_ct(&g, &c2); he_copy_ct(&r, &c2);
    TIMED2("he_rlsin *", he_rlsin(&g, 1.7, &g, &rlk, &ck, iter), gpqhe_ref_he_rlsin(&r, 1.7, &r, &rlk, &ck, iter));   /* in place, as he_bootstrap calls it */
    report("he_rlsin (in place)", eq_ct(&g, &r));
  }

  /* ---- chained use: 9 primitives on one ciphertext with no host access in between - the
   * ciphertext crosses once in and once out, everything else runs on its device shadow */
  if (hectx.L >= 2 && !getenv("DROPIN_DIFF_NO_CHAIN")) {
    he_copy_ct(&g, &c1); he_copy_ct(&r, &c1);
    gpqhe_dropin_sync_ct(&g);
    unsigned long long up0 = 0, dn0 = 0, up1 = 0, dn1 = 0;
    gpqhe_dropin_ct_stats(&up0, &dn0, NULL);
    const double t0 = now_ms();
    he_mul(&g, &g, &g, &rlk); he_rs(&g); he_add(&g, &g, &g); he_rot(&g, 1, rk); he_neg(&g);
    he_mul(&g, &g, &g, &rlk); he_rs(&g); he_conj(&g, &ck); he_sub(&g, &g, &g);
    gpqhe_dropin_wait();
    const double t1 = now_ms();
    gpqhe_dropin_sync_ct(&g);
    const double t2 = now_ms();
    gpqhe_dropin_ct_stats(&up1, &dn1, NULL);
    REFW(&r, gpqhe_ref_he_mul(&r, &r, &r, &rlk)); REFW(&r, gpqhe_ref_he_rs(&r)); REFW(&r, gpqhe_ref_he_add(&r, &r, &r));
    REFW(&r, gpqhe_ref_he_rot(&r, 1, rk)); REFW(&r, gpqhe_ref_he_neg(&r));
    REFW(&r, gpqhe_ref_he_mul(&r, &r, &r, &rlk)); REFW(&r, gpqhe_ref_he_rs(&r)); REFW(&r, gpqhe_ref_he_conj(&r, &ck));
    REFW(&r, gpqhe_ref_he_sub(&r, &r, &r));
    const double t3 = now_ms();
    printf("  time chained x9                   drop-in %10.3f ms (%.3f ms per primitive; + %.3f ms write-back)   reference %12.3f ms   x%.1f\n",
           t1 - t0, (t1 - t0) / 9, t2 - t1, t3 - t2, (t3 - t2) / (t2 - t0 > 1e-6 ? t2 - t0 : 1e-6));
    printf("  chained: ciphertext uploads %llu, downloads %llu\n", up1 - up0, dn1 - dn0);
    const int lazy = getenv("GPQHE_DROPIN_EAGER") == NULL;     /* eager mode writes every result back by design */
    report("chained x9 (device-resident)", eq_ct(&g, &r) && (!lazy || (up1 - up0 <= 1 && dn1 - dn0 == 1)));
  }

  unsigned long long in = 0, out = 0;
  gpqhe_dropin_stats(&in, &out, NULL);
  printf("marshalled coefficients: %llu in, %llu out\n", in, out);
  if (nfail) { printf("FAIL %u/%u\n", nfail, npass + nfail); return 1; }
  printf("PASS %u/%u\n", npass, npass);
  return 0;
}
