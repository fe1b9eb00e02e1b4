/*
 * dropin_bench.c - he_mul + he_rs through the REFERENCE'S OWN SYMBOLS of dropin/_build/libgpqhe.so (what a GPQHE
 * caller links: he_ct_t of gcry_mpi objects, src/gpqhe.h:84-91), timed two ways.  MEASUREMENT TOOL for bench.py's
 * e2e.dropin; built by dropin/Makefile.
 *
 *   isolated  every product takes its two operands from the MPIs (marshalled and uploaded inside the timed region)
 *             and its result is written back into the MPIs: the cost of one he_mul + he_rs for a caller that reads
 *             every coefficient of every result on the host
 *   chained   operands and results stay in their device shadows (the way every algorithm of src/he-algo.c uses the
 *             primitives: nothing reads the MPIs between he_enc and he_dec); the last result is written back once
 *
 *   dropin_bench <logn> <logq> <slots> <logDelta> <ops>      prints one JSON line
 */
#include "config.h"
#include "types.h"
#include "gpqhe.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

extern struct poly_ctx polyctx;
extern struct he_ctx hectx;
void gpqhe_dropin_stats(unsigned long long *in, unsigned long long *out, int *fast_mpi);
void gpqhe_dropin_ct_stats(unsigned long long *uploads, unsigned long long *downloads, unsigned long long *allocs);
void gpqhe_dropin_sync_ct(const he_ct_t *ct);
void gpqhe_dropin_touch_ct(const he_ct_t *ct);
void gpqhe_dropin_wait(void);

static double now_s(void)
{
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec + t.tv_nsec * 1e-9;
}
static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void)
{
  uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
/* uniform on the centred range of q */
static void fill_poly(poly_mpi_t *a, const MPI q, const MPI qh)
{
  const unsigned nb = (gcry_mpi_get_nbits(q) + 7) / 8 + 8;
  unsigned char *buf = malloc(nb);
  for (unsigned i = 0; i < polyctx.n; i++) {
    for (unsigned k = 0; k < nb; k++) buf[k] = (unsigned char)rnd();
    MPI t = NULL;
    gcry_mpi_scan(&t, GCRYMPI_FMT_USG, buf, nb, NULL);
    gcry_mpi_set(a->coeffs[i], t);
    gcry_mpi_release(t);
    mpi_smod(a->coeffs[i], q, qh);
  }
  free(buf);
}

int main(int argc, char **argv)
{
  if (argc < 6) { fprintf(stderr, "usage: %s logn logq slots logDelta ops\n", argv[0]); return 2; }
  const unsigned logn = atoi(argv[1]), logq = atoi(argv[2]), slots = atoi(argv[3]), logDelta = atoi(argv[4]);
  unsigned ops = atoi(argv[5]);
  if (ops < 2) ops = 2;
  FILE *real = stdout;                       /* keygen-free, but hectx_init and friends may print: keep stdout clean */
  MPI q = mpi_new(0);
  mpi_set_ui(q, 1);
  mpi_lshift(q, q, logq);
  hectx_init(logn, q, slots, 1ull << logDelta);
  const unsigned n = polyctx.n, L = hectx.L;
  /* synthetic key: uniform residues in ntt() order, as bench.py's own arm uses */
  he_evk_t rlk;
  he_alloc_evk(&rlk);
  {
    const struct rns_ctx *r = polyctx.rns;
    for (unsigned d = 0; d < hectx.dimevk; d++, r = r->next)
      for (unsigned i = 0; i < n; i++) {
        rlk.p0.coeffs[(size_t)d * n + i] = rnd() % r->p;
        rlk.p1.coeffs[(size_t)d * n + i] = rnd() % r->p;
      }
  }
  enum { POOL = 4 };
  he_ct_t a[POOL], b[POOL], x;
  for (int k = 0; k < POOL; k++) {
    he_alloc_ct(&a[k]); he_alloc_ct(&b[k]);
    a[k].l = b[k].l = L; a[k].nu = b[k].nu = (double)(1ull << logDelta); a[k].B = b[k].B = 1.0;
    fill_poly(&a[k].c0, hectx.q[L], hectx.qh[L]); fill_poly(&a[k].c1, hectx.q[L], hectx.qh[L]);
    fill_poly(&b[k].c0, hectx.q[L], hectx.qh[L]); fill_poly(&b[k].c1, hectx.q[L], hectx.qh[L]);
  }
  he_alloc_ct(&x);
  /* warm-up: key upload, scratch, shadows */
  for (int k = 0; k < 3; k++) { he_mul(&x, &a[k % POOL], &b[k % POOL], &rlk); he_rs(&x); }
  gpqhe_dropin_sync_ct(&x);

  /* isolated: MPIs in, MPIs out, every product */
  double t0 = now_s();
  for (unsigned k = 0; k < ops; k++) {
    gpqhe_dropin_touch_ct(&a[k % POOL]);     /* the host copy is the current one: marshal + upload again */
    gpqhe_dropin_touch_ct(&b[k % POOL]);
    he_mul(&x, &a[k % POOL], &b[k % POOL], &rlk);
    he_rs(&x);
    gpqhe_dropin_sync_ct(&x);
  }
  const double iso = (now_s() - t0) / ops;

  /* chained: device shadows in, device shadow out */
  for (int k = 0; k < POOL; k++) { he_mul(&x, &a[k], &b[k], &rlk); he_rs(&x); }   /* makes every operand device-current */
  gpqhe_dropin_wait();
  unsigned long long up0, dn0, up1, dn1;
  gpqhe_dropin_ct_stats(&up0, &dn0, NULL);
  const unsigned cops = ops * 8;
  t0 = now_s();
  for (unsigned k = 0; k < cops; k++) {
    he_mul(&x, &a[k % POOL], &b[k % POOL], &rlk);
    he_rs(&x);
  }
  gpqhe_dropin_wait();
  const double chained = (now_s() - t0) / cops;
  gpqhe_dropin_sync_ct(&x);
  gpqhe_dropin_ct_stats(&up1, &dn1, NULL);
  int fast = 0;
  gpqhe_dropin_stats(NULL, NULL, &fast);
  fprintf(real,
          "{\"api\": \"he_mul + he_rs through the reference symbols of dropin/_build/libgpqhe.so, he_ct_t of gcry_mpi\", "
          "\"logn\": %u, \"logq\": %u, \"level\": %u, \"mpi_path\": \"%s\", "
          "\"isolated\": {\"ops_per_s\": %.3f, \"ms_per_op\": %.4f, \"ops\": %u, "
          "\"note\": \"operands marshalled from and the result written back into the MPIs for every product\"}, "
          "\"chained\": {\"ops_per_s\": %.3f, \"ms_per_op\": %.4f, \"ops\": %u, \"ct_uploads\": %llu, \"ct_downloads\": %llu, "
          "\"note\": \"operands and results stay in their device shadows; one write-back at the end\"}}\n",
          logn, logq, L, fast ? "direct-limb" : "gcry_mpi_print/scan", 1.0 / iso, iso * 1e3, ops, 1.0 / chained,
          chained * 1e3, cops, up1 - up0, dn1 - dn0);
  return 0;
}
