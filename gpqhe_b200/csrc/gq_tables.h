// gq_tables.h - device-resident image of the reference's context tables
// (struct rns_ctx list + struct he_ctx, src/poly.h:28-65, src/gpqhe.h:36-50),
// flattened once per context by gq_context.cu and passed to kernels by value.
#pragma once
#include "gq_arith.h"

#define GQ_MAX_PRIMES 64   // polyctx.dimub is 30 for the largest shipped ring
#define GQ_MAX_W 32        // widest coefficient / modulus (words per big-poly coefficient, q, P): 2048 bits
#define GQ_MAX_WT 48       // widest CRT basis product the exact kernels hold (3072 bits; only the low W words are
                           // touched by the fast kernels, the full width by the rare exact fix-list kernels)
#define GQ_DEC_W 33        // powers 2^(64w) mod p tabulated for w = 0..32

// NTT view: everything kernels K1/K2/K3 need.
struct GqNttTables {
  int logn;
  int n;
  const GqPrime *primes;        // [nprimes]
  const ulonglong2 *tw_fwd;     // [nprimes][n]  {w, floor(w*2^64/p)}, index as zetas[k]
  const ulonglong2 *tw_inv;     // [nprimes][n]  same for zetas_inv[k]
  const ulonglong2 *tw_inv_last;// [nprimes]     zetas_inv[1]*N^-1 mod p, folded final stage
  // final scaling of the LAZY inverse transforms (gq_ntt_lazy.cu), Shoup pairs per prime: fin_s = s, fin_w =
  // zetas_inv[1]*s with s = N^-1 (the plain invntt()), or s = N^-1 * (P/p_d)^-1 mod p_d when the output feeds
  // the CRT over the basis {p_0..p_(D-1)} directly: rns_reconstruct's first multiplication (src/rns.c:68-69)
  // then costs nothing (gq_ctx::ntt_tables(D))
  const ulonglong2 *fin_s;      // [nprimes]
  const ulonglong2 *fin_w;      // [nprimes]
};

// MPI -> RNS view (K4): pow64[d*GQ_DEC_W + w] = 2^(64w) mod p_d; t1_sh[d] is the Shoup
// companion of pow64[d][1].
// pow32[d*GQ_DEC32 + j] = 2^(32j) mod p_d for j = 0..2*GQ_MAX_W-1 (digit form of the same table);
// one_sh[d] = floor(2^64 / p_d), the Shoup companion of the constant 1.
// pow28[d*GQ_DEC28 + j] = 2^(28j) mod p_d for j = 0..15 (k_decompose28: coefficients of at most 7 words).
#define GQ_DEC32 (2 * GQ_MAX_W)
#define GQ_DEC28 16
struct GqDecTables {
  const GqPrime *primes;
  const uint64_t *pow32;
  const uint64_t *pow28;
  const uint64_t *one_sh;
  const uint64_t *pow64;
  const uint64_t *t1_sh;
};

// CRT basis {p_0..p_{D-1}} (the `rns` node with dim == D, src/precomp.c:266-293).
// All multiword constants are zero-padded to WT words, WT = template width.
struct GqCrtBasis {
  int D;
  int WT;                 // multiple of 4, >= words(bits(P)+1)
  unsigned bitsP;
  int sh;                 // bitsP - 58 (>= 0) : window used for the quotient estimate
  uint64_t recip;         // floor((2^120-1) / ((P >> sh) + 1))
  const uint64_t *P;      // [WT]
  const uint64_t *Ph;     // [WT] floor(P/2)
  const uint64_t *phat;   // [D][WT]   P / p_d
  const uint64_t *inv;    // [D]       (P/p_d)^-1 mod p_d          (phat_invmp)
  const uint64_t *inv_sh; // [D]       Shoup companion
  // fixed-point image of 1/p_d and of ceil(P/2)/P with 128 fractional bits: lets the fast CRT
  // kernels find the overflow count floor((S + ceil(P/2))/P) without the high words of S
  // F = 64 + bits(p) - 1 fractional bits (all primes of the basis have the same bit length, else
  // F = 0 and only the exact kernels are used): floor(2^F / p_d) then fits one word
  int F;
  const uint64_t *u;      // [D][2]    floor(2^F / p_d) (second word 0)
  const uint64_t *hc;     // [2]       floor(2^F * ceil(P/2) / P)
};

// a modulus q used by mpi_smod (src/types.c:108): power-of-two fast path or general
struct GqModulus {
  int is_pow2;
  unsigned k;             // q == 2^k when is_pow2
  int nwords;             // significant words of q (general path)
  const uint64_t *q;      // [GQ_MAX_W] zero padded
  const uint64_t *qh;     // [GQ_MAX_W] floor(q/2)
};
