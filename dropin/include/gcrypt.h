/*
 * Declarations-only stand-in for <gcrypt.h> (libgcrypt 1.10.x), product copy used to
 * build dropin/libgpqhe.so (the oracle keeps its own under oracle/shim/).
 *
 * The image ships libgcrypt.so.20 / libgpg-error.so.0 but
 * not the -dev headers, so the reference sources under /root/reference (which
 * `#include <gcrypt.h>`, src/types.h:32) cannot be compiled without this file.
 * It declares exactly the entry points the reference and our drop-in shim call,
 * with the prototypes of the public libgcrypt 1.10 API, plus the `mpi_*` macro
 * aliases gcrypt.h provides when GCRYPT_NO_MPI_MACROS is not defined.  Link with
 * `-l:libgcrypt.so.20`.  Nothing here is copied from the reference repository.
 */
#ifndef GPQHE_B200_DROPIN_GCRYPT_H
#define GPQHE_B200_DROPIN_GCRYPT_H

#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef unsigned int gpg_error_t;
typedef gpg_error_t gcry_error_t;
#define GPG_ERR_NO_ERROR 0

struct gcry_mpi;
typedef struct gcry_mpi *gcry_mpi_t;
struct gcry_sexp;
typedef struct gcry_sexp *gcry_sexp_t;
struct gcry_cipher_handle;
typedef struct gcry_cipher_handle *gcry_cipher_hd_t;

enum gcry_mpi_format {
  GCRYMPI_FMT_NONE = 0,
  GCRYMPI_FMT_STD = 1,
  GCRYMPI_FMT_PGP = 2,
  GCRYMPI_FMT_SSH = 3,
  GCRYMPI_FMT_HEX = 4,
  GCRYMPI_FMT_USG = 5,
  GCRYMPI_FMT_OPAQUE = 8
};

enum gcry_sexp_format {
  GCRYSEXP_FMT_DEFAULT = 0,
  GCRYSEXP_FMT_CANON = 1,
  GCRYSEXP_FMT_BASE64 = 2,
  GCRYSEXP_FMT_ADVANCED = 3
};

typedef enum gcry_random_level {
  GCRY_WEAK_RANDOM = 0,
  GCRY_STRONG_RANDOM = 1,
  GCRY_VERY_STRONG_RANDOM = 2
} gcry_random_level_t;

enum gcry_cipher_algos { GCRY_CIPHER_AES256 = 9 };
enum gcry_cipher_modes { GCRY_CIPHER_MODE_ECB = 1 };

const char *gcry_check_version(const char *req_version);

/* memory */
void *gcry_malloc(size_t n);
void gcry_free(void *a);

/* MPI life cycle */
gcry_mpi_t gcry_mpi_new(unsigned int nbits);
void gcry_mpi_release(gcry_mpi_t a);
gcry_mpi_t gcry_mpi_copy(const gcry_mpi_t a);
gcry_mpi_t gcry_mpi_set(gcry_mpi_t w, const gcry_mpi_t u);
gcry_mpi_t gcry_mpi_set_ui(gcry_mpi_t w, unsigned long u);
void gcry_mpi_snatch(gcry_mpi_t w, gcry_mpi_t u);

/* MPI import / export */
gcry_error_t gcry_mpi_scan(gcry_mpi_t *ret_mpi, enum gcry_mpi_format format,
                           const void *buffer, size_t buflen, size_t *nscanned);
gcry_error_t gcry_mpi_print(enum gcry_mpi_format format, unsigned char *buffer,
                            size_t buflen, size_t *nwritten, const gcry_mpi_t a);

/* MPI arithmetic */
void gcry_mpi_neg(gcry_mpi_t w, gcry_mpi_t u);
int gcry_mpi_cmp(const gcry_mpi_t u, const gcry_mpi_t v);
int gcry_mpi_cmp_ui(const gcry_mpi_t u, unsigned long v);
int gcry_mpi_is_neg(gcry_mpi_t a);
void gcry_mpi_add(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_add_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_addm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_sub(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_sub_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_subm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_mul(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_mul_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_mulm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_div(gcry_mpi_t q, gcry_mpi_t r, gcry_mpi_t dividend,
                  gcry_mpi_t divisor, int round);
void gcry_mpi_mod(gcry_mpi_t r, gcry_mpi_t dividend, gcry_mpi_t divisor);
unsigned int gcry_mpi_get_nbits(gcry_mpi_t a);
int gcry_mpi_test_bit(gcry_mpi_t a, unsigned int n);
void gcry_mpi_set_bit(gcry_mpi_t a, unsigned int n);
void gcry_mpi_lshift(gcry_mpi_t x, gcry_mpi_t a, unsigned int n);
void gcry_mpi_rshift(gcry_mpi_t x, gcry_mpi_t a, unsigned int n);

/* S-expressions (only used by the reference's show_mpi) */
gcry_error_t gcry_sexp_build(gcry_sexp_t *retsexp, size_t *erroff,
                             const char *format, ...);
size_t gcry_sexp_sprint(gcry_sexp_t sexp, int mode, void *buffer, size_t maxlength);
void gcry_sexp_release(gcry_sexp_t sexp);

/* random / cipher (only referenced by rng.c variants that -DSUPERCOP disables) */
void gcry_randomize(void *buffer, size_t length, enum gcry_random_level level);
gcry_error_t gcry_cipher_open(gcry_cipher_hd_t *handle, int algo, int mode,
                              unsigned int flags);
void gcry_cipher_close(gcry_cipher_hd_t h);
gcry_error_t gcry_cipher_setkey(gcry_cipher_hd_t hd, const void *key, size_t keylen);
gcry_error_t gcry_cipher_encrypt(gcry_cipher_hd_t h, void *out, size_t outsize,
                                 const void *in, size_t inlen);
size_t gcry_cipher_get_algo_keylen(int algo);

#ifndef GCRYPT_NO_MPI_MACROS
#define mpi_new(n)          gcry_mpi_new((n))
#define mpi_release(a)      do { gcry_mpi_release((a)); (a) = NULL; } while (0)
#define mpi_copy(a)         gcry_mpi_copy((a))
#define mpi_set(w, u)       gcry_mpi_set((w), (u))
#define mpi_set_ui(w, u)    gcry_mpi_set_ui((w), (u))
#define mpi_cmp(u, v)       gcry_mpi_cmp((u), (v))
#define mpi_cmp_ui(u, v)    gcry_mpi_cmp_ui((u), (v))
#define mpi_is_neg(a)       gcry_mpi_is_neg((a))
#define mpi_neg(w, u)       gcry_mpi_neg((w), (u))
#define mpi_add_ui(w, u, v) gcry_mpi_add_ui((w), (u), (v))
#define mpi_add(w, u, v)    gcry_mpi_add((w), (u), (v))
#define mpi_addm(w, u, v, m) gcry_mpi_addm((w), (u), (v), (m))
#define mpi_sub_ui(w, u, v) gcry_mpi_sub_ui((w), (u), (v))
#define mpi_sub(w, u, v)    gcry_mpi_sub((w), (u), (v))
#define mpi_subm(w, u, v, m) gcry_mpi_subm((w), (u), (v), (m))
#define mpi_mul_ui(w, u, v) gcry_mpi_mul_ui((w), (u), (v))
#define mpi_mul(w, u, v)    gcry_mpi_mul((w), (u), (v))
#define mpi_mulm(w, u, v, m) gcry_mpi_mulm((w), (u), (v), (m))
#define mpi_fdiv(q, r, a, m) gcry_mpi_div((q), (r), (a), (m), -1)
#define mpi_mod(r, a, m)    gcry_mpi_mod((r), (a), (m))
#define mpi_get_nbits(a)    gcry_mpi_get_nbits((a))
#define mpi_test_bit(a, b)  gcry_mpi_test_bit((a), (b))
#define mpi_set_bit(a, b)   gcry_mpi_set_bit((a), (b))
#define mpi_lshift(a, b, c) gcry_mpi_lshift((a), (b), (c))
#define mpi_rshift(a, b, c) gcry_mpi_rshift((a), (b), (c))
#endif /* GCRYPT_NO_MPI_MACROS */

#ifdef __cplusplus
}
#endif

#endif /* GPQHE_B200_DROPIN_GCRYPT_H */
