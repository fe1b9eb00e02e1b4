/*
 * gpqhe_b200.h - C ABI of the B200-native RNS polynomial hot path of GPQHE.
 *
 * This is the drop-in boundary: plain pointers and sizes, no CUDA or torch types.
 * Every entry point names the reference function it replaces (file:line relative
 * to the GPQHE tree, commit 5e70953c).  dropin/ shows the C shim that binds the
 * reference's own symbols (ntt, poly_mul, he_mul, ...) to these calls; the Python
 * mirror in gpqhe_b200/ binds them with ctypes.
 *
 * Data formats
 *   limb-poly   uint64_t[n]          one polynomial modulo one prime p_d, values in
 *                                    [0,p_d] (the reference's ntt() may emit p_d for 0)
 *   poly_rns    uint64_t[dim][n]     limb-major, as struct poly_rns (src/poly.h:74-77)
 *   big poly    uint64_t[W][n]       WORD-MAJOR two's-complement integers: word w of
 *                                    coefficient i is at [w*n + i]; this stands in for
 *                                    struct poly_mpi (src/poly.h:69-72), the MPI <-> words
 *                                    conversion staying on the host side (dropin/)
 *   evk         2 x uint64_t[dimevk][n], already in ntt() output order, exactly what
 *                                    he_genswk stores (src/he-kem.c:103-110)
 *
 * Host entry points take HOST pointers and include the host<->device copies;
 * the *_dev entry points take DEVICE pointers plus a CUDA stream handle
 * (cudaStream_t passed as void*, NULL = the context's own stream) and only
 * enqueue work.  All functions return 0 on success, non-zero on error;
 * gq_last_error() describes the last failure.  The reference aborts on error
 * (SURVEY 8b); the dropin shim turns a non-zero return into the same abort().
 * One context is bound to one device; calls on one context must be serialised
 * by the caller (the reference is single-threaded, process-global state).
 * The scratch buffers of a context (and of each lane) are shared by all its calls: a non-NULL
 * `stream` argument must be ordered after all work enqueued earlier on that context - pass the
 * same stream every time, or make the new stream wait for the old one; independent operations
 * that should overlap belong on different LANES (gq_ctx_set_lanes), which own their scratch.
 */
#ifndef GPQHE_B200_H
#define GPQHE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef struct gq_ctx gq_ctx;
typedef struct gq_evk gq_evk;

typedef struct gq_info {
  unsigned logn, n;
  unsigned logp;      /* GPQHE_LOGP (src/params.h:27), 59 unless TEST_CRT */
  unsigned dimub;     /* polyctx.dimub, src/precomp.c:357 */
  unsigned has_he;    /* 0 when created with Delta == 0 (polyctx_init only) */
  unsigned L;         /* hectx.L,      src/precomp.c:391 */
  unsigned dim;       /* hectx.dim,    src/precomp.c:401 */
  unsigned dimevk;    /* hectx.dimevk, src/precomp.c:407 */
  unsigned W;         /* words per ciphertext coefficient: holds [-q_L/2, q_L/2) */
  unsigned q_is_pow2; /* every q_l and Delta is a power of two (fast path) */
  int device;
} gq_info;

const char *gq_last_error(void);
int gq_abi_version(void);

/* ---- context: polyctx_init + hectx_init (src/precomp.c:328-384, 431-452) ----------
 * q: little-endian words of the top modulus.  Delta == 0 builds the polynomial context
 * only (polyctx_init).  primes/zetas/zetas_inv may be NULL: the library then runs the
 * reference's prime search and root selection itself (src/precomp.c:153-264,370-376).
 * When given, they are the reference's own tables (struct rns_ctx .p, .zetas,
 * .zetas_inv in Montgomery form, one pointer per list node) and are uploaded as is.
 * logqub == 0 applies the reference's security table (src/precomp.c:52-117,339-340). */
int gq_ctx_create(gq_ctx **ctx, unsigned logn, const uint64_t *q, unsigned q_words,
                  unsigned slots, uint64_t Delta, unsigned logp, unsigned logqub,
                  int device, const uint64_t *primes, const uint64_t *const *zetas,
                  const uint64_t *const *zetas_inv);
void gq_ctx_destroy(gq_ctx *ctx);
int gq_ctx_info(const gq_ctx *ctx, gq_info *out);
int gq_ctx_primes(const gq_ctx *ctx, uint64_t *out /* [dimub] */);
/* per-prime scalars of struct rns_ctx: out = {p, pinv_mont, pinv_barr, ninv} */
int gq_ctx_prime_consts(const gq_ctx *ctx, unsigned limb, uint64_t out[4]);
/* limb counts at level l: he-mult.c:99 (tensor) and he-mult.c:51 (key switch) */
int gq_ctx_dims(const gq_ctx *ctx, unsigned l, unsigned *dim_tensor, unsigned *dim_ks);
int gq_ctx_sync(gq_ctx *ctx);
/* Bound on the encoded plaintexts handed to gq_he_gemv / gq_he_coeff2slot / gq_he_slot2coeff (host and _dev forms)
 * from now on: |m| < 2^bits for every coefficient (he_ecd output is about Delta * |value|, src/he-encode.c:53-66).
 * With a bound he_gemv accumulates the n1 products of a giant step in the NTT domain and does ONE inverse NTT + CRT
 * instead of n1 (src/he-algo.c:65-79 does one per he_mulpt) - the same bits, proof in gq_api.cu.  bits = 0 (the
 * default): unknown, every he_mulpt is evaluated on its own. */
int gq_ctx_set_plaintext_bits(gq_ctx *ctx, unsigned bits);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
unsigned long long gq_ctx_launch_count(const gq_ctx *ctx);
void *gq_ctx_stream(gq_ctx *ctx); /* the context's cudaStream_t */
/* coefficients the fast CRT kernels handed to the exact full-width kernel since creation
 * (values within 2^-56 of the +P/2 edge; diagnostics for the edge-case tests), -1 on error */
long long gq_ctx_crt_fixups(gq_ctx *ctx);

/* ---- compute lanes (no reference counterpart: the reference is single-threaded and runs one
 * operation at a time, src/config.h:59 GPQHE_NUM_THREAD is unused) --------------------------
 * One he_mul at N=2^16 does not fill 148 SMs in every stage (kernel tails, < 1 wave in the CRT
 * stages), so independent ciphertext operations are run on `lanes` streams of ONE GPU.  A lane is a
 * context handle that shares every table and key of `ctx` and owns a stream and the scratch, and is
 * accepted by every *_dev entry point (keys created on `ctx` are valid on its lanes).
 *   gq_ctx_set_lanes   lanes >= 1 (1..8); lane 0 is `ctx` itself; grows only
 *   gq_ctx_lane        handle of lane i (NULL if i >= lanes); owned by `ctx`
 *   gq_ctx_lanes_fork  every lane waits for the work enqueued so far on ctx's stream
 *   gq_ctx_lanes_join  ctx's stream waits for the work enqueued so far on every lane
 * gq_he_mul_batch distributes its products over the lanes (3 by default). */
int gq_ctx_set_lanes(gq_ctx *ctx, unsigned lanes);
unsigned gq_ctx_lane_count(const gq_ctx *ctx);
gq_ctx *gq_ctx_lane(gq_ctx *ctx, unsigned i);
int gq_ctx_lanes_fork(gq_ctx *ctx);
int gq_ctx_lanes_join(gq_ctx *ctx);

/* ---- device memory and copies for hosts that do not link the CUDA runtime (dropin/ is plain C): the device-resident
 * shadow of a he_ct_t lives in such buffers (no reference counterpart: struct he_ct keeps N gcry_mpi objects,
 * src/gpqhe.h:84-91).  upload / download / copy only ENQUEUE on the stream (NULL = the context's); the host buffer of
 * an upload must stay untouched, and that of a download unread, until the stream has been synchronised. */
int gq_dev_alloc(gq_ctx *ctx, size_t bytes, void **ptr_dev);
void gq_dev_free(gq_ctx *ctx, void *ptr_dev);
int gq_dev_upload(gq_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes, void *stream);
int gq_dev_download(gq_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes, void *stream);
int gq_dev_copy(gq_ctx *ctx, void *dst_dev, const void *src_dev, size_t bytes, void *stream);

/* pinned host memory for callers that want fast copies */
int gq_host_alloc(void **p, size_t bytes);
void gq_host_free(void *p);

/* ---- limb layer (host pointers) ------------------------------------------------- */
int gq_ntt(gq_ctx *ctx, unsigned limb, uint64_t *a);    /* ntt(),    src/ntt.c:37-52 */
int gq_invntt(gq_ctx *ctx, unsigned limb, uint64_t *a); /* invntt(), src/ntt.c:54-73 */
/* poly_rns_add / poly_rns_mul, src/poly.c:71-82; r may alias a or b */
int gq_poly_rns_add(gq_ctx *ctx, unsigned limb, uint64_t *r, const uint64_t *a, const uint64_t *b);
int gq_poly_rns_mul(gq_ctx *ctx, unsigned limb, uint64_t *r, const uint64_t *a, const uint64_t *b);
/* batched transforms: data[npolys][nlimbs][n], limb d uses prime limb0+d */
int gq_ntt_batch(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys, uint64_t *data,
                 int inverse);

/* ---- MPI <-> RNS (host pointers) --------------------------------------------------- */
/* rns_decompose, src/rns.c:37-48, for limbs limb0..limb0+nlimbs-1 at once */
int gq_rns_decompose(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *ahat,
                     const uint64_t *a, unsigned W);
/* poly_rns2mpi, src/poly.c:109-120: CRT over the first `dim` primes, centre mod P, centre mod q */
int gq_poly_rns2mpi(gq_ctx *ctx, unsigned dim, uint64_t *r, unsigned Wr, const uint64_t *rhat,
                    const uint64_t *q, unsigned q_words);
/* poly_mul, src/poly.c:84-107 */
int gq_poly_mul(gq_ctx *ctx, uint64_t *r, unsigned Wr, const uint64_t *a, const uint64_t *b,
                unsigned W, unsigned dim, const uint64_t *q, unsigned q_words);
/* poly_rot / poly_conj, src/poly.c:263-283 */
int gq_poly_rot(gq_ctx *ctx, uint64_t *r, const uint64_t *a, unsigned W, uint64_t rot);
int gq_poly_conj(gq_ctx *ctx, uint64_t *r, const uint64_t *a, unsigned W);

/* ---- evaluation keys ------------------------------------------------------------- */
/* upload a key laid out as he_evk_t (src/gpqhe.h:78-82): p0, p1 = uint64_t[dimevk][n] */
int gq_evk_create(gq_ctx *ctx, gq_evk **evk, const uint64_t *p0, const uint64_t *p1);
int gq_evk_create_dev(gq_ctx *ctx, gq_evk **evk, const uint64_t *p0_dev, const uint64_t *p1_dev);
/* the device copy of a key back into two host buffers of dimevk*n words each, he_evk_t layout (limb-major, ntt()
   output order): what gpqhe_b200/wire.py serialises (the reference has no key wire format,
   src/precomp.c:455-463 computes kemctx sizes nobody uses) */
int gq_evk_export(const gq_evk *evk, uint64_t *p0, uint64_t *p1);
void gq_evk_destroy(gq_evk *evk);

/* ---- HE primitives, host pointers; every big poly is uint64_t[W][n], W = gq_info.W - */
/* he_mul (mult + relinearise), src/he-mult.c:88-156; `l` = ct1->l = ct2->l; outputs may alias inputs */
int gq_he_mul(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
              const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk);
/* he_mul immediately followed by he_rs on the product (src/he-mult.c:88-156 then
   src/he-rescale.c:33-50) in one host<->device crossing; the result is at level l-1 */
int gq_he_mul_rs(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                 const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk);
/* `count` independent he_mul (+ he_rs when `rescale`) at level l with one relinearisation key - the
   row loop of he_gemv (src/he-linear.c) - pipelined over both copy engines; arrays of `count` poly
   pointers; page-locked host memory (gq_host_alloc) makes the copies overlap the kernels */
int gq_he_mul_batch(gq_ctx *ctx, unsigned l, unsigned count, uint64_t *const *c0, uint64_t *const *c1,
                    const uint64_t *const *a0, const uint64_t *const *a1, const uint64_t *const *b0,
                    const uint64_t *const *b1, const gq_evk *rlk, int rescale);
/* he_mulpt, src/he-mult.c:159-196; `dim` is the limb count the caller derives from pt->nu (:168) */
int gq_he_mulpt(gq_ctx *ctx, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                const uint64_t *a0, const uint64_t *a1, const uint64_t *m);
/* he_rs / he_moddown, src/he-rescale.c:33-70; in place; `l` is the level BEFORE the call */
int gq_he_rs(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1);
int gq_he_moddown(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1);
/* `count` rotations rots[k] of one ciphertext with rns_decompose + ntt of c1 shared (see
   gq_he_rot_hoisted_dev); host pointers, outputs are separate polynomials */
int gq_he_rot_hoisted(gq_ctx *ctx, unsigned l, unsigned count, const unsigned *rots,
                      uint64_t *const *c0_out, uint64_t *const *c1_out, const uint64_t *c0,
                      const uint64_t *c1, const gq_evk *const *rk);
/* the ciphertext arithmetic of he_gemv (src/he-algo.c:47-90): baby-step/giant-step diagonal method,
   slots = n1*n2, with the diagonals already encoded by the caller (pts[i*n1+j] = he_ecd of
   zrotdiag(A, i*n1+j, -i*n1), uint64_t[W][n] each; `dim` = the he_mulpt limb count for their nu,
   src/he-mult.c:168).  rk[r] = key of rotation r (r < n1 and r = i*n1 are used).  Result at level
   l-1 (the closing he_rs is included).  One crossing, every intermediate ciphertext stays in HBM. */
int gq_he_gemv(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0, uint64_t *c1,
               const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
               const gq_evk *const *rk);
/* he_coeff2slot, src/he-bootstrap.c:359-409: ct_conj = he_conj(ct); for part in {real, imag}:
   he_gemv(U_conjT, ct) + he_gemv(U_T, ct_conj) -> he_add -> he_mulpt(constant plaintext) -> he_rs, with every
   intermediate ciphertext on the device.  pts: encoded diagonals of U0_conjT, U0_T, U1_conjT, U1_T (4 sets of
   n1*n2, each as gq_he_gemv wants them); cpt: the encoded constant plaintext; dim_gemv / dim_pt: he_mulpt limb
   counts (he-mult.c:168) at levels l and l-1; rk indexed by rotation amount.  Outputs are at level l-2. */
int gq_he_coeff2slot(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim_gemv, unsigned dim_pt,
                     uint64_t *re0, uint64_t *re1, uint64_t *im0, uint64_t *im1, const uint64_t *a0,
                     const uint64_t *a1, const uint64_t *const *pts, const uint64_t *cpt, const gq_evk *ck,
                     const gq_evk *const *rk);
/* he_exp (src/he-algo.c:435-458 over he_exp_taylor, :364-433): exp(a*ct) as the reference evaluates it - he_mulpt
   by the encoded a/2^iter, he_rs, the degree-7 Taylor polynomial in the reference's evaluation order, `iter`
   squarings - with every intermediate ciphertext on the device.  pts[0] = he_ecd of a/2^iter in every slot,
   pts[1..7] = he_const_pt of 1, 3, 1/6, 5, 1/120, 7, 1/5040 (uint64_t[W][n] each, encoded by the caller: host
   floating point); dim_a / dim_t = he_mulpt limb counts (he-mult.c:168) at levels l and l-1.  Needs
   l >= 4 + iter; result at level l-4-iter. */
int gq_he_exp(gq_ctx *ctx, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, uint64_t *c0, uint64_t *c1,
              const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts, const gq_evk *rlk);
/* he_rlsin (src/he-bootstrap.c:444-475; the EvalMod step of he_bootstrap, :518-520): sin(a*ct)/a =
   he_rs(he_mulpt(he_sub(e, he_conj(e)), pts[8])) with e = he_exp(i*a*ct).  pts[0..7] as gq_he_exp (pts[0]
   encodes i*a/2^iter), pts[8] = he_ecd of 1/(2*i*a); dim_f = he_mulpt limb count at level l-4-iter.  Needs
   l >= 5 + iter; result at level l-5-iter. */
int gq_he_rlsin(gq_ctx *ctx, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, unsigned dim_f,
                uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                const gq_evk *rlk, const gq_evk *ck);
/* he_slot2coeff (src/he-bootstrap.c:414-438): he_gemv(U0, ct0) + he_gemv(U1, ct1); (a0,a1) = ct0, (b0,b1) = ct1,
   pts = the encoded diagonals of U0 then of U1 (2 sets of n1*n2 as gq_he_gemv wants them).  Result at level l-1. */
int gq_he_slot2coeff(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0, uint64_t *c1,
                     const uint64_t *a0, const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                     const uint64_t *const *pts, const gq_evk *const *rk);
/* he_add / he_sub (src/he-add.c:32-76), he_addpt / he_subpt (:79-120: only c0 meets the plaintext,
   c1 is reduced), he_neg (:123-140, in place): out = smod_{q_l}(a op b); outputs may alias inputs */
int gq_he_add(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
              const uint64_t *b0, const uint64_t *b1);
int gq_he_sub(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
              const uint64_t *b0, const uint64_t *b1);
int gq_he_addpt(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                const uint64_t *m);
int gq_he_subpt(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                const uint64_t *m);
int gq_he_neg(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1);
/* he_rot / he_conj, src/he-automorphism.c:87-115; in place; `rk_rot` is &rk[rot] */
int gq_he_rot(gq_ctx *ctx, unsigned l, unsigned rot, uint64_t *c0, uint64_t *c1, const gq_evk *rk_rot);
int gq_he_conj(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const gq_evk *ck);

/* ---- device-resident variants: DEVICE pointers, enqueue only -------------------- */
/* inverse: 0 forward, 1 inverse.  Profiling only: +2 runs just the 4096-point tile kernel,
 * +4 just the strided kernel of a two-kernel transform (the result is then not an NTT). */
int gq_ntt_batch_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys,
                     uint64_t *data_dev, int inverse, void *stream);
int gq_poly_rns_mul_batch_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys,
                              uint64_t *r_dev, const uint64_t *a_dev, const uint64_t *b_dev,
                              void *stream);
int gq_he_mul_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                  const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk,
                  void *stream);
/* he_mul immediately followed by he_rs on the product (src/he-mult.c:88-156, src/he-rescale.c:33-54): when q_l and
   Delta are powers of two the rounded division rides in the key-switch epilogue kernel (the quotient is still in
   registers), otherwise it is the separate rescale kernel; same bits either way.  Result at level l-1. */
int gq_he_mul_rs_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                     const uint64_t *a1, const uint64_t *b0, const uint64_t *b1, const gq_evk *rlk,
                     void *stream);
int gq_he_mulpt_dev(gq_ctx *ctx, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                    const uint64_t *a0, const uint64_t *a1, const uint64_t *m, void *stream);
/* he_mulpt followed by he_rs (the pair he_exp_taylor and he_coeff2slot make, src/he-algo.c:385-388,
   src/he-bootstrap.c:386-388), the division fused behind the CRT as in gq_he_mul_rs_dev */
int gq_he_mulpt_rs_dev(gq_ctx *ctx, unsigned l, unsigned dim, uint64_t *c0, uint64_t *c1,
                       const uint64_t *a0, const uint64_t *a1, const uint64_t *m, void *stream);
int gq_he_rs_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, void *stream);
int gq_he_moddown_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, void *stream);
int gq_he_rot_dev(gq_ctx *ctx, unsigned l, unsigned rot, uint64_t *c0, uint64_t *c1,
                  const gq_evk *rk_rot, void *stream);
int gq_he_conj_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const gq_evk *ck,
                   void *stream);
/* `count` rotations rots[k] of ONE ciphertext (the baby steps of he_gemv, src/he-algo.c:62-66): the same
   bits as `count` he_rot calls, with rns_decompose + ntt of c1 shared (hoisting).  rk[k] = key of
   rotation rots[k]; outputs must not overlap the inputs. */
int gq_he_rot_hoisted_dev(gq_ctx *ctx, unsigned l, unsigned count, const unsigned *rots,
                          uint64_t *const *c0_out, uint64_t *const *c1_out, const uint64_t *c0,
                          const uint64_t *c1, const gq_evk *const *rk, void *stream);
int gq_he_gemv_dev(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                   uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                   const gq_evk *const *rk, void *stream);
int gq_he_coeff2slot_dev(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim_gemv, unsigned dim_pt,
                         uint64_t *re0, uint64_t *re1, uint64_t *im0, uint64_t *im1, const uint64_t *a0,
                         const uint64_t *a1, const uint64_t *const *pts, const uint64_t *cpt, const gq_evk *ck,
                         const gq_evk *const *rk, void *stream);
int gq_he_exp_dev(gq_ctx *ctx, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, uint64_t *c0,
                  uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *const *pts,
                  const gq_evk *rlk, void *stream);
int gq_he_rlsin_dev(gq_ctx *ctx, unsigned l, unsigned iter, unsigned dim_a, unsigned dim_t, unsigned dim_f,
                    uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                    const uint64_t *const *pts, const gq_evk *rlk, const gq_evk *ck, void *stream);
int gq_he_slot2coeff_dev(gq_ctx *ctx, unsigned l, unsigned n1, unsigned n2, unsigned dim, uint64_t *c0,
                         uint64_t *c1, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0,
                         const uint64_t *b1, const uint64_t *const *pts, const gq_evk *const *rk, void *stream);
int gq_he_add_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                  const uint64_t *b0, const uint64_t *b1, void *stream);
int gq_he_sub_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0, const uint64_t *a1,
                  const uint64_t *b0, const uint64_t *b1, void *stream);
int gq_he_addpt_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                    const uint64_t *a1, const uint64_t *m, void *stream);
int gq_he_subpt_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, const uint64_t *a0,
                    const uint64_t *a1, const uint64_t *m, void *stream);
int gq_he_neg_dev(gq_ctx *ctx, unsigned l, uint64_t *c0, uint64_t *c1, void *stream);

/* ---- stage-level entry points (DEVICE pointers, enqueue only) ------------------------
 * One call = one stage of the pipelines above on a limb range [limb0, limb0+nlimbs) (NTT domain,
 * limb-polys of the full ring degree n) or on a slice of `ncoef` coefficients (big-integer domain:
 * big polys are uint64_t[W][ncoef], residues uint64_t[npolys][nlimbs][ncoef]).  The limb-sharded
 * multi-GPU pipeline is built from them with an all-to-all between the two domains.
 *   decompose    rns_decompose (src/rns.c:37-48) of npolys (<= 4) big polys: out[npolys][nlimbs][ncoef]
 *   ntt          the internal transforms (canonical in / out), data[npolys][nlimbs][n]
 *   tensor       he-mult.c:125-136: x[4][nlimbs][n] -> y[3][nlimbs][n]
 *   mul_evk      he-mult.c:61-64:   x[nlimbs][n] -> y[2][nlimbs][n] with the key's limbs limb0..
 *   reconstruct  poly_rns2mpi(q_l) (src/poly.c:109-120) over the first D primes: rhat[npolys][D][ncoef]
 *   ks_epilogue  he-mult.c:67-77 / he-automorphism.c:68-76: chat[2][D2(l)][ncoef], add may be NULL
 *   rescale      he_rs (divide != 0) / he_moddown on a slice; l = level before the call */
int gq_stage_decompose_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys,
                           const uint64_t *const *in_dev, unsigned W, unsigned ncoef, uint64_t *out_dev,
                           void *stream);
int gq_stage_ntt_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys, uint64_t *data_dev,
                     int inverse, void *stream);
int gq_stage_tensor_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_dev,
                        const uint64_t *x_dev, void *stream);
int gq_stage_mul_evk_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_dev,
                         const uint64_t *x_dev, const gq_evk *evk, void *stream);
/* tensor / mul_evk followed by the inverse NTT of the same limb range, the pointwise products formed while the
   inverse transform loads its tiles (one pass instead of two kernels + a round trip): y is in the coefficient
   domain.  ks_epilogue_rs: ks_epilogue followed by he_rs on the slice (one kernel for power-of-two moduli). */
int gq_stage_tensor_inv_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_dev,
                            const uint64_t *x_dev, void *stream);
int gq_stage_mul_evk_inv_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_dev,
                             const uint64_t *x_dev, const gq_evk *evk, void *stream);
int gq_stage_ks_epilogue_rs_dev(gq_ctx *ctx, unsigned l, const uint64_t *chat_dev, const uint64_t *add0_dev,
                                const uint64_t *add1_dev, uint64_t *c0_dev, uint64_t *c1_dev, unsigned ncoef,
                                void *stream);
int gq_stage_reconstruct_dev(gq_ctx *ctx, unsigned D, unsigned l, unsigned npolys, const uint64_t *rhat_dev,
                             uint64_t *const *out_dev, unsigned ncoef, void *stream);
/* gq_stage_reconstruct_dev for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  It has NO COUNTERPART IN THE REFERENCE.
 * Same kernels, same arguments, bit-identical results; the fix-list scratch of the fast CRT kernels comes
 * from a file-scope __device__ array of the library (gq_noalloc.cu: 1 MiB + 16 B per GPU, held until the process
 * exits) instead of the context's cudaMalloc'd buffer, and the call invokes no cudaMalloc / cudaFree / allocating
 * routine.  Limit: npolys * ncoef <= 262 144 (4 polynomials of 65 536 coefficients); larger calls fail with an
 * error.  The array is shared by every call and context on a device and keeps the last call's contents (its header is
 * cleared on `stream` at the start of each call): calls that overlap on the GPU race - order them on one stream. */
int gq_stage_reconstruct_noalloc_dev(gq_ctx *ctx, unsigned D, unsigned l, unsigned npolys, const uint64_t *rhat_dev,
                                     uint64_t *const *out_dev, unsigned ncoef, void *stream);
int gq_stage_ks_epilogue_dev(gq_ctx *ctx, unsigned l, const uint64_t *chat_dev, const uint64_t *add0_dev,
                             const uint64_t *add1_dev, uint64_t *c0_dev, uint64_t *c1_dev, unsigned ncoef,
                             void *stream);
int gq_stage_rescale_dev(gq_ctx *ctx, unsigned l, uint64_t *c0_dev, uint64_t *c1_dev, int divide,
                         unsigned ncoef, void *stream);
/* The same stages on PRE-SCALED residues, the form he_mul / he_rot use internally: the inverse transforms end with the
   first multiplication of rns_reconstruct (src/rns.c:68-69, y_d = a_d (P/p_d)^-1 mod p_d over the basis
   {p_0..p_(D-1)}) folded into their N^-1 scaling, and the CRT stages take such residues - which lets them run as
   tensor-core contractions (gq_crt_tc.cu).  gq_stage_prescale_ok: 1 when these forms exist for the basis D at level l
   (keyswitch != 0: D must be the key-switch basis of l), else use the plain entries above.  `rescale` folds he_rs in. */
int gq_stage_prescale_ok(gq_ctx *ctx, unsigned D, unsigned l, int keyswitch);
int gq_stage_tensor_inv_pre_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned D, uint64_t *y_dev,
                                const uint64_t *x_dev, void *stream);
int gq_stage_mul_evk_inv_pre_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned D, uint64_t *y_dev,
                                 const uint64_t *x_dev, const gq_evk *evk, void *stream);
int gq_stage_reconstruct_pre_dev(gq_ctx *ctx, unsigned D, unsigned l, unsigned npolys, const uint64_t *rhat_dev,
                                 uint64_t *const *out_dev, unsigned ncoef, void *stream);
int gq_stage_ks_epilogue_pre_dev(gq_ctx *ctx, unsigned l, const uint64_t *chat_dev, const uint64_t *add0_dev,
                                 const uint64_t *add1_dev, uint64_t *c0_dev, uint64_t *c1_dev, unsigned ncoef,
                                 int rescale, void *stream);

/* ---- peer-to-peer re-layout for the sharded pipeline (gq_p2p.cu) ----------------------------
 * One receive slab per rank, exported / mapped with CUDA IPC (one process per GPU, same node).
 * scatter_c2l: src[P][D][nloc] -> owner of each limb, scatter_l2c: src[P][Dl][n] -> owner of each
 * coefficient; dst[s] is rank s's destination buffer as mapped in THIS process, off[G+1] the limb
 * ownership.  signal/wait: per-exchange epoch flags in the peers' slabs (slot < 8); the epoch is
 * read from device memory (bump increments it) so the launches can be replayed from a CUDA graph.
 * err_dev is set non-zero when a peer does not show up within GPQHE_B200_P2P_TIMEOUT_MS (default 4000) and is STICKY:
 * a rank whose flag is set signals a poison epoch, so the failure reaches every rank's err_dev within the same op
 * and every later wait returns at once.  A caller must read err_dev before it trusts an output. */
int gq_p2p_alloc(gq_ctx *ctx, size_t bytes, void **ptr_dev, unsigned char handle[64]);
int gq_p2p_open(gq_ctx *ctx, const unsigned char handle[64], void **peer_ptr_dev);
int gq_p2p_close(gq_ctx *ctx, void *peer_ptr_dev);
int gq_p2p_free(gq_ctx *ctx, void *ptr_dev);
int gq_p2p_scatter_c2l_dev(gq_ctx *ctx, const uint64_t *src_dev, unsigned P, unsigned D, unsigned nloc,
                           unsigned G, unsigned me, const uint64_t *const *dst, const int *off, void *stream);
int gq_p2p_scatter_l2c_dev(gq_ctx *ctx, const uint64_t *src_dev, unsigned P, unsigned Dl, unsigned D,
                           unsigned nloc, unsigned G, unsigned me, const uint64_t *const *dst, const int *off,
                           void *stream);
int gq_p2p_bump_dev(gq_ctx *ctx, unsigned long long *epoch_dev, void *stream);
int gq_p2p_signal_dev(gq_ctx *ctx, const uint64_t *const *flags, unsigned slot, unsigned G, unsigned me,
                      const unsigned long long *epoch_dev, const int *err_dev, void *stream);
int gq_p2p_wait_dev(gq_ctx *ctx, const uint64_t *flags_local_dev, unsigned slot, unsigned G,
                    const unsigned long long *epoch_dev, int *err_dev, void *stream);

/* ---- fused compute + exchange (round 2): the stages above with the exchange INSIDE the producing / consuming kernel
 * (gq_p2p.h).  `send`: the kernel stores its results straight into the peers' receive slabs (dst[s] = rank s's
 * buffer of this exchange as mapped here, off[G+1] = limb ownership, D = limbs of the exchange) and its last block
 * publishes *epoch_dev in flag[slot][me] of every rank (flags[s]) - or a poison value when *err_dev is set;
 * ticket_dev is a zeroed device word the kernel uses to find its last block.  `wait`: every block of the kernel polls
 * flag[slot][0..G) of this rank (flags_local) for *epoch_dev before it reads its slab; a timeout
 * (GPQHE_B200_P2P_TIMEOUT_MS) or a poisoned flag sets *err_dev (sticky).  No reference counterpart: the reference has
 * no multi-device path at all.  gq_p2p_fused_ok: 1 when he_mul at level l can run entirely on these entry points.
 *   decompose    in[npolys] (W words x ncoef, my coefficient slice) -> residues of all nlimbs limbs at their owners
 *   ntt_fwd      forward NTT of a receive slab (waits first)
 *   tensor_inv / mul_evk_inv   pointwise stage + inverse NTT of my limbs; y_scratch[3|2][nlimbs][n] is scratch between
 *                              the two passes, the finished coefficients go to their owners
 *   reconstruct / ks_epilogue  CRT stages on a receive slab (wait first); ks_epilogue folds he_rs in when `rescale` */
typedef struct gq_p2p_xchg {
  unsigned G, me, slot, D;
  const uint64_t *const *dst;
  const int *off;
  const uint64_t *const *flags;
  const uint64_t *flags_local;
  const unsigned long long *epoch_dev;
  int *err_dev;
  unsigned *ticket_dev;
  unsigned pre;   /* != 0: the residues of this exchange are pre-scaled for the basis D (gq_stage_prescale_ok) */
} gq_p2p_xchg;
/* Process-wide switch of programmatic dependent launch between the library's kernels (on by default; the environment
   variable GPQHE_B200_NO_PDL=1 forces it off); returns the previous setting.  No reference counterpart: a launch
   detail of this implementation. */
int gq_set_pdl(int on);
int gq_p2p_fused_ok(gq_ctx *ctx, unsigned l);
int gq_p2p_decompose_dev(gq_ctx *ctx, unsigned nlimbs, unsigned npolys, const uint64_t *const *in_dev, unsigned W,
                         unsigned ncoef, const gq_p2p_xchg *send, void *stream);
int gq_p2p_ntt_fwd_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, unsigned npolys, uint64_t *data_dev,
                       const gq_p2p_xchg *wait, void *stream);
int gq_p2p_tensor_inv_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_scratch_dev, const uint64_t *x_dev,
                          const gq_p2p_xchg *send, void *stream);
int gq_p2p_mul_evk_inv_dev(gq_ctx *ctx, unsigned limb0, unsigned nlimbs, uint64_t *y_scratch_dev, const uint64_t *x_dev,
                           const gq_evk *evk, const gq_p2p_xchg *send, void *stream);
int gq_p2p_reconstruct_dev(gq_ctx *ctx, unsigned D, unsigned l, unsigned npolys, const uint64_t *rhat_dev,
                           uint64_t *const *out_dev, unsigned ncoef, const gq_p2p_xchg *wait, void *stream);
int gq_p2p_ks_epilogue_dev(gq_ctx *ctx, unsigned l, const uint64_t *chat_dev, const uint64_t *add0_dev,
                           const uint64_t *add1_dev, uint64_t *c0_dev, uint64_t *c1_dev, unsigned ncoef, int rescale,
                           const gq_p2p_xchg *wait, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* GPQHE_B200_H */
