// gq_context.h - host-side context: the flattened, device-resident image of the
// reference's process-global `polyctx` / `hectx` (src/precomp.c:38-47).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "gq_hostbig.h"
#include "gq_kernels.h"

struct GqHostBasis {
  GqCrtBasis dev;   // device pointers
  BigU P;
  void *blob = nullptr;
  std::vector<uint64_t> inv;          // host copy of (P/p_d)^-1 mod p_d
  std::vector<uint64_t> phat;         // host copy of P/p_d, [D][WT] words
  std::vector<uint64_t> u;            // host copy of floor(2^F / p_d), [D]
  // final-scale tables of the lazy inverse NTT with (P/p_d)^-1 folded in (built on first use, gq_api.cu)
  ulonglong2 *fin_s = nullptr, *fin_w = nullptr;
  // key switch over this basis: (P_dim/p_d)^-1 * ((P/p_d)^-1)^-1 mod p_d for d < dim, Shoup pairs; turns the
  // prescaled residues of this basis into those of the special-modulus basis (r = T mod P_dim)
  ulonglong2 *ratio = nullptr;
};

struct GqHostModulus {
  GqModulus dev;
  BigU q;
  void *blob = nullptr;
};

enum { GQ_SLOT_LIMB_A = 0, GQ_SLOT_LIMB_B, GQ_SLOT_BIG, GQ_SLOT_IO, GQ_SLOT_PIPE, GQ_SLOT_GEMV, GQ_SLOT_C2S, GQ_SLOT_EXP, GQ_SLOT_RLSIN, GQ_SLOT_COUNT };

struct gq_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  unsigned logn = 0, n = 0, logp = 59, dimub = 0, logqub = 0;
  unsigned has_he = 0, L = 0, dim = 0, dimevk = 0, W = 0, q_is_pow2 = 0, slots = 0;
  uint64_t Delta = 0;
  unsigned logDelta = 0;      // valid when Delta is a power of two
  BigU q, P, PqL;
  std::vector<BigU> ql, qh;
  std::vector<GqHostModulus> qmod;   // per level
  std::vector<GqPrime> primes;
  std::vector<GqHostBasis> basis;    // basis[D-1] for D = 1..dimub
  // device tables
  GqPrime *d_primes = nullptr;
  ulonglong2 *d_tw_fwd = nullptr, *d_tw_inv = nullptr, *d_tw_inv_last = nullptr;
  ulonglong2 *d_fin_s = nullptr;     // {N^-1, Shoup companion} per prime: the plain final scale of the lazy inverse NTT
  std::vector<uint64_t> h_inv1;      // zetas_inv[1] per prime (plain residue), for the folded final-scale tables
  uint64_t *d_pow64 = nullptr, *d_t1sh = nullptr, *d_pow32 = nullptr, *d_onesh = nullptr, *d_pow28 = nullptr;
  uint64_t *d_Pinv = nullptr;        // P^-1 mod 2^(64*GQ_MAX_W)
  uint64_t *d_Pwords = nullptr;      // P and floor(P/2), GQ_MAX_W words each (general path)
  int *d_fix = nullptr;              // {count, pad x3, list[4n]} of the fast CRT kernels
  bool fast_crt = true;
  uint64_t *d_adhoc = nullptr;       // scratch for an ad-hoc modulus (q, qh)
  // growable device scratch
  void *slot[GQ_SLOT_COUNT] = {};
  size_t slot_bytes[GQ_SLOT_COUNT] = {};
  // copy engines of the batched host entry points (created on first use)
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_pipe[8] = {};       // {in_ready, in_free, out_ready, out_free} x 2 buffers
  unsigned long long launches = 0;
  // compute lanes (gq_ctx_set_lanes): lane 0 is this context; lanes[k-1] (k >= 1) are shadow
  // contexts that share every device table above (parent != null: not owned) and own a stream,
  // the scratch slots, d_fix and d_adhoc, so that independent ciphertext operations overlap
  gq_ctx *parent = nullptr;
  std::vector<gq_ctx *> lanes;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // staging of the batched host entry points when lanes are in use
  std::vector<cudaEvent_t> ev_batch;
  // B-fragment tables of the tensor-core CRT (gq_crt_tc.cu), built on first use per basis D (key D: reconstruct,
  // key 0x10000 + D: key-switch epilogue over the basis D); kept by the ROOT context (lanes look them up there)
  std::map<unsigned, uint32_t *> tc_tabs;
  std::vector<uint64_t> h_Pinv;      // host copy of d_Pinv
  unsigned pt_bits = 0;          // gq_ctx_set_plaintext_bits: |m| < 2^pt_bits for the plaintexts of he_gemv (0 = unknown)
  bool lazy_ntt = true;          // internal pipelines use gq_ntt_lazy.cu when every prime allows it

  // D > 0: inverse transforms end with the CRT pre-multiplication of the basis {p_0..p_(D-1)} folded into N^-1
  // (the tables must have been built: gq_api.cu crt_prescale)
  GqNttTables ntt_tables(unsigned D = 0) const {
    GqNttTables t;
    t.logn = (int)logn; t.n = (int)n; t.primes = d_primes;
    t.tw_fwd = d_tw_fwd; t.tw_inv = d_tw_inv; t.tw_inv_last = d_tw_inv_last;
    t.fin_s = D ? basis[D - 1].fin_s : d_fin_s;
    t.fin_w = D ? basis[D - 1].fin_w : d_tw_inv_last;
    return t;
  }
  GqDecTables dec_tables() const {
    GqDecTables t; t.primes = d_primes; t.pow64 = d_pow64; t.t1_sh = d_t1sh;
    t.pow32 = d_pow32; t.one_sh = d_onesh; t.pow28 = d_pow28; return t;
  }
};

struct gq_evk {
  gq_ctx *ctx = nullptr;
  uint64_t *p0 = nullptr, *p1 = nullptr;   // device, [dimevk][n] each
  bool owned = true;
};

void gq_set_error(const std::string &msg);
