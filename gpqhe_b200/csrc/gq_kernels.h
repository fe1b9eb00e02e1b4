// gq_kernels.h - host-side launchers of the device kernels (one per pipeline stage).
#pragma once
#include "gq_tables.h"

struct GqPtrs4 { uint64_t *p[4]; };
struct GqPeerOut;   // gq_p2p.h: peer output / flag wait of the sharded pipeline, inside the producing / consuming kernel
struct GqWaitIn;

// K1/K2 (gq_ntt.cu): data[npolys][nlimbs][n] in place; returns kernels launched or <0
int gq_launch_ntt(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                  bool inverse, cudaStream_t st, int only = 0);

// K1/K2 lazy variants for the internal pipelines (gq_ntt_lazy.cu): canonical in, canonical out;
// valid only for primes p = 2^b + c, b <= 59, 16c < 2^b (checked by the caller)
// wt != null (forward, n > 2^12): the strided kernel waits for the peers' flags first (sharded pipeline); -2 = no such form
int gq_launch_ntt_lazy(const GqNttTables &tb, uint64_t *data, int limb0, int nlimbs, int npolys,
                       bool inverse, cudaStream_t st, int only = 0, const GqWaitIn *wt = nullptr);

// pointwise stage fused into the load of the lazy inverse transform (gq_ntt_lazy.cu): x -> y = invntt(product)
//   TENSOR x[4][D][n] -> y[3][D][n];  EVK / EVKPERM x[D][n], e0, e1 -> y[2][D][n];  PT x[3][D][n] -> y[2][D][n];
//   MUL x[2][D][n] -> y[D][n];  GEMV x[T][2][D][n], e0[T][D][n], g = T -> y[2][D][n] = sum_t x[t][k] * e0[t]
enum { GQ_NTTF_TENSOR = 1, GQ_NTTF_EVK = 2, GQ_NTTF_EVKPERM = 3, GQ_NTTF_PT = 4, GQ_NTTF_MUL = 5, GQ_NTTF_GEMV = 6 };
// po != null (TENSOR / EVK, n > 2^12): the last pass stores into the peers' slabs and signals; -2 = no such form
int gq_launch_ntt_lazy_fused(const GqNttTables &tb, int mode, const uint64_t *x, uint64_t *y, const uint64_t *e0,
                             const uint64_t *e1, unsigned g, int limb0, int nlimbs, cudaStream_t st,
                             const GqPeerOut *po = nullptr);

// K3 (gq_pointwise.cu)
void gq_launch_binary(const GqPrime *primes, int n, int limb0, int nlimbs, int npolys, bool mul,
                      uint64_t *r, const uint64_t *a, const uint64_t *b, cudaStream_t st);
// pm: every prime of the range is 2^s + c with c < 2^28, 33 <= s <= 59 (gq_reduce_pm kernels)
void gq_launch_tensor(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                      cudaStream_t st, bool pm = false);
void gq_launch_mul_evk(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                       const uint64_t *e0, const uint64_t *e1, cudaStream_t st, bool pm = false);
// the same with x read through the evaluation-point permutation of the Galois map g (hoisted rotations)
void gq_launch_mul_evk_perm(const GqPrime *primes, int n, int logn, int D, uint64_t *y, const uint64_t *x,
                            const uint64_t *e0, const uint64_t *e1, uint64_t g, cudaStream_t st, bool pm);
void gq_launch_mul_pt(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                      cudaStream_t st, bool pm = false);

// K4/K5/K6 (gq_crt.cu); return the number of kernels launched, <0 when no instantiation fits.
// fix = device scratch {int count (16-byte slot), int list[4n]}: enables the fast low-word kernels
// pre / ratio: the residues come pre-multiplied by (P/p_d)^-1 from the lazy inverse NTT (GqNttTables::fin_s); for the
// key switch `ratio` converts them to the special-modulus basis (GqHostBasis::ratio).  Fast path only.
// rs_j > 0 (power-of-two q and Delta only): he_rs fused behind the CRT, the output is rdiv(., 2^rs_j) smod 2^(k-rs_j)
// digits: 32-bit digit kernel (every prime of the range needs >= 33 and <= 60 bits)
int gq_launch_decompose(const GqDecTables &dt, int n, int limb0, int nlimbs, const GqPtrs4 &in,
                        int W, uint64_t *out, int npolys, cudaStream_t st, int digits = false,
                        const GqPeerOut *po = nullptr);
int gq_launch_reconstruct(const GqCrtBasis &B, const GqPrime *primes, int n, const uint64_t *rhat,
                          const GqModulus &q, const GqPtrs4 &out, int Wout, int npolys,
                          cudaStream_t st, int *fix = nullptr, unsigned rs_j = 0, bool pre = false,
                          const GqWaitIn *wt = nullptr, const uint32_t *tc_tab = nullptr);
// K5 / K6 on the tensor cores (gq_crt_tc.cu): u8 mma.sync contraction of the residue bytes with a constant byte
// table over `slots` limb slots (K: 8 words per slot, u: one estimate word per slot, cw: 8 words)
size_t gq_crt_tc_table_words(int slots);
void gq_crt_tc_table(uint32_t *tab, int slots, const uint64_t *K, const uint64_t *u, const uint64_t *cw);
int gq_launch_reconstruct_tc(const GqCrtBasis &B, const uint32_t *tab, int n, const uint64_t *rhat, unsigned k,
                             const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st, int *fix, unsigned rs_j,
                             const GqWaitIn *wt);
// K4 on the tensor cores: per-limb B fragments (coefficients of W <= 7 words, pseudo-Mersenne primes)
size_t gq_dec_tc_table_words(int nprimes);
void gq_dec_tc_table(unsigned long long *tab, int nprimes, const uint64_t *p, int W);
int gq_launch_decompose_tc(const unsigned long long *tab, const GqPrime *primes, int n, int limb0, int nlimbs,
                           const GqPtrs4 &in, int W, uint64_t *out, int npolys, cudaStream_t st, const GqPeerOut *po);
int gq_launch_ks_epilogue_tc(const GqCrtBasis &B, const GqCrtBasis &BP, const uint32_t *tab, const GqPrime *primes,
                             const ulonglong2 *ratio, int n, const uint64_t *chat, unsigned k, const GqPtrs4 &add,
                             int Wadd, const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st, int *fix,
                             unsigned rs_j, const GqWaitIn *wt);
int gq_launch_ks_epilogue(const GqCrtBasis &B, const GqCrtBasis &BP, const uint64_t *Pinv,
                          const GqPrime *primes, int n, const uint64_t *chat, unsigned k,
                          const GqPtrs4 &add, int Wadd, const GqPtrs4 &out, int Wout, int npolys,
                          cudaStream_t st, int *fix = nullptr, unsigned rs_j = 0, const ulonglong2 *ratio = nullptr,
                          const GqWaitIn *wt = nullptr, const uint32_t *tc_tab = nullptr);
// general-modulus key-switch epilogue (exact long division, slow path)
int gq_launch_ks_epilogue_general(const GqCrtBasis &B, const uint64_t *P, int nP,
                                  const uint64_t *Ph, const GqPrime *primes, int n,
                                  const uint64_t *chat, const GqModulus &q, const GqPtrs4 &add,
                                  int Wadd, const GqPtrs4 &out, int Wout, int npolys,
                                  cudaStream_t st);

// K7/K8 (gq_misc.cu)
// in-place rounded division by 2^j (j == 0: none) then mpi_smod(2^k)
void gq_launch_rescale_pow2(int n, int W, const GqPtrs4 &c, int npolys, unsigned j, unsigned k,
                            cudaStream_t st);
// general Delta / q
void gq_launch_rescale_general(int n, int W, const GqPtrs4 &c, int npolys, uint64_t Delta,
                               const GqModulus &q, cudaStream_t st);
// out[y] = smod_q(a[y] op_y b[y]); op: 0 add, 1 sub, 2 negate a, 3 reduce a  (src/he-add.c:32-142)
void gq_launch_linear(int n, int W, const GqPtrs4 &out, const GqPtrs4 &a, const GqPtrs4 &b, int npolys,
                      int op0, int op1, const GqModulus &q, cudaStream_t st);
// out[(i*g mod 2n) mod n] = +-in[i]   (poly_rot: g = 5^rot mod 2n; poly_conj: g = 2n-1)
void gq_launch_automorph(int n, int W, const GqPtrs4 &out, const GqPtrs4 &in, int npolys,
                         uint64_t g, cudaStream_t st);
