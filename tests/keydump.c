/*
 * keydump.c - key generation through the reference's API under the deterministic -DSUPERCOP stream: prints one
 * FNV-1a checksum per key buffer and the wall time of each generator.  TEST TOOL, built twice from this one source:
 * dropin/_build/keydump links the drop-in libgpqhe.so (its rns_decompose / ntt / poly_mul run on the GPU),
 * oracle/_ref/keydump links the unmodified reference.  Same seed, same sampler order (the samplers are the
 * reference's own code in both), so every `key` line must be identical: the evaluation keys the GPU path produces
 * are byte-identical to the reference's (src/he-kem.c:74-169).
 *
 *   keydump <logn> <logq> <slots> <logDelta>
 */
#include "config.h"
#include "types.h"
#include "gpqhe.h"

#include <stdio.h>
#include <stdlib.h>
#include <time.h>

extern struct poly_ctx polyctx;
extern struct he_ctx hectx;

static double now_ms(void)
{
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}
static uint64_t fnv(const uint64_t *p, size_t words)
{
  uint64_t h = 0xcbf29ce484222325ull;
  for (size_t i = 0; i < words; i++) { h ^= p[i]; h *= 0x100000001b3ull; }
  return h;
}
static uint64_t fnv_mpi(const poly_mpi_t *a)
{
  uint64_t h = 0xcbf29ce484222325ull;
  unsigned char buf[512];
  for (unsigned i = 0; i < polyctx.n; i++) {
    size_t len = 0;
    gcry_mpi_print(GCRYMPI_FMT_USG, buf, sizeof buf, &len, a->coeffs[i]);
    h ^= (uint64_t)gcry_mpi_is_neg(a->coeffs[i]); h *= 0x100000001b3ull;
    for (size_t k = 0; k < len; k++) { h ^= buf[k]; h *= 0x100000001b3ull; }
  }
  return h;
}

int main(int argc, char **argv)
{
  if (argc < 5) { fprintf(stderr, "usage: %s logn logq slots logDelta\n", argv[0]); return 2; }
  const unsigned logn = atoi(argv[1]), logq = atoi(argv[2]), slots = atoi(argv[3]), logDelta = atoi(argv[4]);
  MPI q = mpi_new(0);
  mpi_set_ui(q, 1);
  mpi_lshift(q, q, logq);
  hectx_init(logn, q, slots, 1ull << logDelta);
  const size_t words = (size_t)hectx.dimevk * polyctx.n;
  poly_mpi_t sk; he_pk_t pk; he_evk_t rlk, ck, *rk;
  he_alloc_sk(&sk); he_alloc_pk(&pk); he_alloc_evk(&rlk); he_alloc_evk(&ck);
  rk = malloc(slots * sizeof *rk);
  for (unsigned i = 0; i < slots; i++) he_alloc_evk(&rk[i]);
  double t0 = now_ms();
  he_keypair(&pk, &sk);
  double t1 = now_ms();
  he_genrlk(&rlk, &sk);
  double t2 = now_ms();
  he_genck(&ck, &sk);
  double t3 = now_ms();
  he_genrk(rk, &sk);
  double t4 = now_ms();
  printf("\nkey sk %016lx\nkey pk %016lx %016lx\n", fnv_mpi(&sk), fnv_mpi(&pk.p0), fnv_mpi(&pk.p1));
  printf("key rlk %016lx %016lx\n", fnv(rlk.p0.coeffs, words), fnv(rlk.p1.coeffs, words));
  printf("key ck %016lx %016lx\n", fnv(ck.p0.coeffs, words), fnv(ck.p1.coeffs, words));
  for (unsigned i = 0; i < slots; i++)
    printf("key rk[%u] %016lx %016lx\n", i, fnv(rk[i].p0.coeffs, words), fnv(rk[i].p1.coeffs, words));
  printf("time he_keypair %.1f ms  he_genrlk %.1f ms  he_genck %.1f ms  he_genrk(%u keys) %.1f ms (%.1f ms per key)\n",
         t1 - t0, t2 - t1, t3 - t2, slots, t4 - t3, (t4 - t3) / slots);
  return 0;
}
