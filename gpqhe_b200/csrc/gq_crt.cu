// gq_crt.cu - the MPI <-> RNS crossings on the device.
//   K4 k_decompose      rns_decompose                      src/rns.c:37-48
//   K5 k_reconstruct    poly_rns2mpi = rns_reconstruct + 2x mpi_smod
//                                                          src/rns.c:60-75, src/poly.c:109-120
//   K6 k_ks_epilogue    poly_rns2mpi(P*q_l) + mpi_rdiv(P) + mpi_addm + mpi_smod
//                                                          src/he-mult.c:67-77, src/he-automorphism.c:68-76
// One thread owns one coefficient; multiword integers live in registers (compile-time
// width WT, fully unrolled), CRT constants are staged in shared memory and read as
// warp broadcasts, limb-poly reads / word-major writes are coalesced across threads.
//
// Exactness notes.
//  * rns_reconstruct returns sum_d a_d*(phat_d*inv_d mod P) mod P in [0,P).  We compute
//    y_d = a_d*inv_d mod p_d, S = sum_d y_d*phat_d (S == that sum mod P, S < D*P), then
//    S mod P exactly: a 64-bit quotient estimate from the top bits that never
//    overshoots, followed by exact compare-and-subtract.
//  * Key switch: the reference reduces T = centred CRT value first to x = T smod (P*q_l),
//    then takes rdiv(x,P) and reduces mod q_l.  x = T + m*P*q_l for an integer m, so
//    floor(x/P) = floor(T/P) + m*q_l and x mod P = T mod P: the smod(P*q_l) step cannot
//    change the final residue mod q_l and is skipped.  floor(T/P) = (T - r)/P with
//    r = T mod P = CRT of the first `dim` residues; the division is exact, so modulo
//    2^(64*WP) it is a multiplication by P^-1 mod 2^(64*WP).
#include "gq_bigint.h"
#include "gq_kernels.h"
#include "gq_p2p.h"

namespace {

template <int WT> struct MaxD { static constexpr int v = WT + WT / 8 + 2; };

template <int WT>
struct CrtShared {
  uint64_t phat[MaxD<WT>::v * WT];
  uint64_t P[WT + 1];
  uint64_t Ph[WT + 1];
  uint64_t inv[MaxD<WT>::v];
  uint64_t inv_sh[MaxD<WT>::v];
  uint64_t p[MaxD<WT>::v];
};

// inv_over != null: Shoup pairs that replace (P/p_d)^-1 - the residues then arrive pre-multiplied for ANOTHER
// basis and the pair converts between the two (GqHostBasis::ratio)
template <int WT>
GQ_DEV void crt_stage(CrtShared<WT> &S, const GqCrtBasis &B, const GqPrime *primes,
                      const ulonglong2 *inv_over = nullptr) {
  for (int j = threadIdx.x; j < B.D * WT; j += blockDim.x) S.phat[j] = B.phat[j];
  for (int j = threadIdx.x; j <= WT; j += blockDim.x) {
    S.P[j] = (j < WT) ? B.P[j] : 0;
    S.Ph[j] = (j < WT) ? B.Ph[j] : 0;
  }
  for (int j = threadIdx.x; j < B.D; j += blockDim.x) {
    S.inv[j] = inv_over ? inv_over[j].x : B.inv[j];
    S.inv_sh[j] = inv_over ? inv_over[j].y : B.inv_sh[j];
    S.p[j] = primes[j].p;
  }
}

// acc[0..WT) <- (sum_d rhat[d][i] * C_d) mod P, in [0,P)
// pre: the residues are already y_d = a_d * (P/p_d)^-1 mod p_d (the inverse NTT folded that constant into its
// final N^-1 scaling, GqNttTables::fin_s), so the first multiplication of rns_reconstruct is skipped
template <int WT>
GQ_DEV void crt_mod_P(uint64_t *acc, const CrtShared<WT> &S, const GqCrtBasis &B,
                      const uint64_t *__restrict__ rhat, int n, int i, bool pre = false) {
#pragma unroll
  for (int w = 0; w <= WT; w++) acc[w] = 0;
  uint64_t a0 = rhat[i], a1 = (B.D > 1) ? rhat[(size_t)n + i] : 0;
  for (int d = 0; d < B.D; d++) {
    const uint64_t a = a0;
    a0 = a1;
    a1 = (d + 2 < B.D) ? rhat[(size_t)(d + 2) * n + i] : 0;
    uint64_t y = pre ? a : gq_mul_shoup(a, S.inv[d], S.inv_sh[d], S.p[d]);
    mw_mac<WT>(acc, y, &S.phat[d * WT]);
  }
  if (B.sh >= 0) {
    const int sw = B.sh >> 6, sb = B.sh & 63;
    uint64_t lo = 0, hi = 0;
#pragma unroll
    for (int w = 0; w <= WT; w++) {   // mask-select: a compare-and-assign chain would be
      lo |= acc[w] & (0 - (uint64_t)(w == sw));        // turned into a dynamic index and
      hi |= acc[w] & (0 - (uint64_t)(w == sw + 1));    // push acc[] out of registers
    }
    uint64_t top = sb ? ((lo >> sb) | (hi << (64 - sb))) : lo;
    uint64_t k = gq_mulhi(top, B.recip) >> 56;
    mw_mulsub<WT + 1>(acc, k, S.P);
  }
  for (int it = 0; it < B.D + 2; it++) {
    if (!mw_geq<WT + 1>(acc, S.P)) break;
    mw_sub<WT + 1>(acc, acc, S.P);
  }
}

// mpi_smod(acc, P, P/2): acc in [0,P) -> [-ceil(P/2), floor(P/2)), two's complement WT words
template <int WT>
GQ_DEV void crt_centre(uint64_t *acc, const CrtShared<WT> &S) {
  if (mw_geq<WT>(acc, S.Ph)) mw_sub<WT>(acc, acc, S.P);
}

// signed acc (WT words) -> mpi_smod(acc, q, qh) for a general modulus (slow, exact)
template <int WT>
GQ_DEV void smod_general(uint64_t *acc, const GqModulus &q) {
  uint64_t u[WT];
  const bool neg = (acc[WT - 1] >> 63) != 0;
#pragma unroll
  for (int w = 0; w < WT; w++) u[w] = acc[w];
  if (neg) mw_neg<WT>(u);
  gq_divrem(u, WT, q.q, q.nwords, nullptr);   // u[0..nwords) = |acc| mod q
  bool zero = true;
  for (int w = 0; w < q.nwords; w++) zero = zero && (u[w] == 0);
  uint64_t qq[WT], hh[WT];
#pragma unroll
  for (int w = 0; w < WT; w++) { qq[w] = (w < q.nwords) ? q.q[w] : 0; hh[w] = (w < q.nwords) ? q.qh[w] : 0; }
  if (neg && !zero) mw_sub<WT>(u, qq, u);      // floor-mod of a negative value
  if (mw_geq<WT>(u, hh)) mw_sub<WT>(u, u, qq);
#pragma unroll
  for (int w = 0; w < WT; w++) acc[w] = u[w];
}

// the four batch pointers are kernel parameters: selecting with a chain keeps them in the constant bank (an
// index computed from blockIdx.y copies the struct to local memory: 64-byte stack frame in the ptxas log)
GQ_DEV uint64_t *ptr4(const GqPtrs4 &P, int b) { return b == 0 ? P.p[0] : b == 1 ? P.p[1] : b == 2 ? P.p[2] : P.p[3]; }

template <int WT>
GQ_DEV void store_words(uint64_t *__restrict__ out, int n, int i, int Wout, const uint64_t *acc) {
  const uint64_t fill = (acc[WT - 1] >> 63) ? ~(uint64_t)0 : 0;
#pragma unroll
  for (int w = 0; w < WT; w++)
    if (w < Wout) out[(size_t)w * n + i] = acc[w];
  for (int w = WT; w < Wout; w++) out[(size_t)w * n + i] = fill;
}

// List mode of the exact kernels runs on GQ_FIX_BLOCKS blocks (an input crafted to sit on the +P/2 edge could put up to
// 4n coefficients on the list: one block would serialise them on one SM).  fix[0] = pending count, fix[1] = lifetime
// total, fix[2] = ticket: the last block to finish empties the list for the next call.
#define GQ_FIX_BLOCKS 148
GQ_DEV void fix_list_done(int *fix_count) {
  __syncthreads();
  if (threadIdx.x == 0) {
#ifdef GQ_EMULATE
    const int t = fix_count[2]++;
#else
    __threadfence();
    const int t = atomicAdd(fix_count + 2, 1);
#endif
    if (t == (int)gridDim.x - 1) { fix_count[0] = 0; fix_count[2] = 0; }
  }
}

// ---- K5 ---------------------------------------------------------------------------
// batch: poly b reads rhat + b*D*n, writes out.p[b]
// rs_j > 0 (POW2 only): he_rs fused behind the CRT - rounded division by 2^rs_j, then smod 2^(k - rs_j).  The
// reference reduces smod 2^k first (src/poly.c:117) and divides afterwards (src/he-rescale.c:45-50); with
// x = T + m*2^k, floor(x / 2^j) = floor(T / 2^j) + m*2^(k-j) and x mod 2^j = T mod 2^j, so dividing the unreduced
// value and reducing once modulo 2^(k-j) gives the same residue, tie rule included.
template <int WT, bool POW2>
GQ_DEV void reconstruct_one(const CrtShared<WT> &S, const GqCrtBasis &B, int n,
                            const uint64_t *__restrict__ rhat, const GqModulus &q, const GqPtrs4 &out,
                            int Wout, int b, int i, unsigned rs_j, bool pre) {
  uint64_t acc[WT + 1];
  crt_mod_P<WT>(acc, S, B, rhat + (size_t)b * B.D * n, n, i, pre);
  crt_centre<WT>(acc, S);
  if (POW2) {
    if (rs_j) mw_rdiv_pow2<WT>(acc, rs_j);
    if (q.k - rs_j < 64u * WT) mw_smod_pow2<WT>(acc, q.k - rs_j);
  } else {
    smod_general<WT>(acc, q);
  }
  store_words<WT>(out.p[b], n, i, Wout, acc);
}

// fix_list == null: one thread per coefficient, blockIdx.y = poly.  Otherwise (one block) the
// coefficients b*n+i that the fast kernel could not decide are taken from the list, which is
// then emptied for the next call.
template <int WT, bool POW2>
__global__ void __launch_bounds__(128)
k_reconstruct(GqCrtBasis B, const GqPrime *__restrict__ primes, int n,
              const uint64_t *__restrict__ rhat, GqModulus q, GqPtrs4 out, int Wout,
              int *fix_count, const int *__restrict__ fix_list, unsigned rs_j, bool pre) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ CrtShared<WT> S;
  if (fix_list) {
    const int cnt = *fix_count;
    if (cnt == 0) return;
    crt_stage<WT>(S, B, primes);
    __syncthreads();
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < cnt; e += gridDim.x * blockDim.x)
      reconstruct_one<WT, POW2>(S, B, n, rhat, q, out, Wout, fix_list[e] / n, fix_list[e] % n, rs_j, pre);
    fix_list_done(fix_count);
    return;
  }
  crt_stage<WT>(S, B, primes);
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  reconstruct_one<WT, POW2>(S, B, n, rhat, q, out, Wout, blockIdx.y, i, rs_j, pre);
}

// ---- K6 (power-of-two q_l fast path) ------------------------------------------------
// B  : key-switch basis (D2 limbs), BP : basis of the special modulus P (`dim` limbs)
// Pinv: P^-1 mod 2^(64*WP).  add.p[b]: additive big poly (Wadd words) or null.
template <int WT, int WP>
GQ_DEV void ks_one(const CrtShared<WT> &S, const CrtShared<WP> &SP, const uint64_t *sPinv,
                   const GqCrtBasis &B, const GqCrtBasis &BP, int n, const uint64_t *__restrict__ chat,
                   unsigned k, const GqPtrs4 &add, int Wadd, const GqPtrs4 &out, int Wout, int b, int i,
                   unsigned rs_j, bool pre) {
  const uint64_t *ch = chat + (size_t)b * B.D * n;
  uint64_t T[WT + 1];
  crt_mod_P<WT>(T, S, B, ch, n, i, pre);
  crt_centre<WT>(T, S);
  uint64_t r[WP + 1];
  crt_mod_P<WP>(r, SP, BP, ch, n, i);            // r = T mod P in [0,P)
  const bool round_up = !mw_geq<WP>(SP.Ph, r);     // r > floor(P/2)
  // low WP words of T - r
  uint64_t x[WP];
  {
    uint64_t tl[WP];
#pragma unroll
    for (int w = 0; w < WP; w++) tl[w] = (w < WT) ? T[w] : ((T[WT - 1] >> 63) ? ~(uint64_t)0 : 0);
    mw_sub<WP>(x, tl, r);
  }
  // Q = x * Pinv mod 2^(64*WP)  (exact quotient modulo 2^(64*WP))
  uint64_t Q[WP];
#pragma unroll
  for (int w = 0; w < WP; w++) Q[w] = 0;
#pragma unroll
  for (int a = 0; a < WP; a++) {
    uint64_t carry = 0;
#pragma unroll
    for (int c = 0; c + a < WP; c++) {
      gq_u128 t = (gq_u128)x[a] * sPinv[c] + Q[a + c] + carry;
      Q[a + c] = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
    }
  }
  // + rounding + additive term (sign-extended to WP words)
  uint64_t ad[WP];
  if (add.p[b]) {
    const uint64_t *ap = add.p[b];
    uint64_t top = 0;
#pragma unroll
    for (int w = 0; w < WP; w++) {
      if (w < Wadd) { ad[w] = ap[(size_t)w * n + i]; top = ad[w]; }
    }
    const uint64_t fill = (top >> 63) ? ~(uint64_t)0 : 0;
#pragma unroll
    for (int w = 0; w < WP; w++) if (w >= Wadd) ad[w] = fill;
  } else {
#pragma unroll
    for (int w = 0; w < WP; w++) ad[w] = 0;
  }
  {
    uint64_t carry = round_up ? 1u : 0u;
#pragma unroll
    for (int w = 0; w < WP; w++) {
      gq_u128 t = (gq_u128)Q[w] + ad[w] + carry;
      Q[w] = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
    }
  }
  if (rs_j) mw_rdiv_pow2<WP>(Q, rs_j);          // he_rs fused (see reconstruct_one)
  mw_smod_pow2<WP>(Q, k - rs_j);
  store_words<WP>(out.p[b], n, i, Wout, Q);
}

template <int WT, int WP>
__global__ void __launch_bounds__(128)
k_ks_epilogue(GqCrtBasis B, GqCrtBasis BP, const uint64_t *__restrict__ Pinv,
              const GqPrime *__restrict__ primes, int n, const uint64_t *__restrict__ chat,
              unsigned k, GqPtrs4 add, int Wadd, GqPtrs4 out, int Wout,
              int *fix_count, const int *__restrict__ fix_list, unsigned rs_j, const ulonglong2 *ratio) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ CrtShared<WT> S;
  __shared__ CrtShared<WP> SP;
  __shared__ uint64_t sPinv[WP];
  if (fix_list && *fix_count == 0) return;
  const bool pre = ratio != nullptr;
  crt_stage<WT>(S, B, primes);
  crt_stage<WP>(SP, BP, primes, ratio);
  for (int j = threadIdx.x; j < WP; j += blockDim.x) sPinv[j] = Pinv[j];
  __syncthreads();
  if (fix_list) {   // one block: the coefficients the fast kernel left undecided
    const int cnt = *fix_count;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < cnt; e += gridDim.x * blockDim.x)
      ks_one<WT, WP>(S, SP, sPinv, B, BP, n, chat, k, add, Wadd, out, Wout, fix_list[e] / n, fix_list[e] % n, rs_j, pre);
    fix_list_done(fix_count);
    return;
  }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ks_one<WT, WP>(S, SP, sPinv, B, BP, n, chat, k, add, Wadd, out, Wout, blockIdx.y, i, rs_j, pre);
}

// ---- K6, general modulus: exact long division by P (slow path) -----------------------
template <int WT>
__global__ void __launch_bounds__(128)
k_ks_epilogue_general(GqCrtBasis B, const uint64_t *__restrict__ Pw, int nP,
                      const uint64_t *__restrict__ Phw, const GqPrime *__restrict__ primes, int n,
                      const uint64_t *__restrict__ chat, GqModulus q, GqPtrs4 add, int Wadd,
                      GqPtrs4 out, int Wout) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ CrtShared<WT> S;
  crt_stage<WT>(S, B, primes);
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.y;
  uint64_t T[WT + 1];
  crt_mod_P<WT>(T, S, B, chat + (size_t)b * B.D * n, n, i);
  crt_centre<WT>(T, S);
  const bool neg = (T[WT - 1] >> 63) != 0;
  uint64_t u[WT], quot[WT + 1], Pl[WT], Phl[WT];
#pragma unroll
  for (int w = 0; w < WT; w++) {
    u[w] = T[w]; quot[w] = 0;
    Pl[w] = (w < nP) ? Pw[w] : 0;
    Phl[w] = (w < nP) ? Phw[w] : 0;
  }
  quot[WT] = 0;
  if (neg) mw_neg<WT>(u);
  gq_divrem(u, WT, Pl, nP, quot);              // |T| = quot*P + u
  for (int w = WT - nP + 1; w < WT; w++) quot[w] = 0;
  bool zero = true;
  for (int w = 0; w < nP; w++) zero = zero && (u[w] == 0);
  if (neg && !zero) {                          // floor: -(quot+1), remainder P - u
    uint64_t cy = 1;
    for (int w = 0; w < WT && cy; w++) { quot[w] += 1; cy = (quot[w] == 0); }
    mw_sub<WT>(u, Pl, u);
  }
  const bool up = !mw_geq<WT>(Phl, u);         // remainder > floor(P/2)
  if (neg) mw_neg<WT>(quot);
  uint64_t ad[WT];
  if (add.p[b]) {
    const uint64_t *ap = add.p[b];
    uint64_t top = 0;
#pragma unroll
    for (int w = 0; w < WT; w++)
      if (w < Wadd) { ad[w] = ap[(size_t)w * n + i]; top = ad[w]; }
    const uint64_t fill = (top >> 63) ? ~(uint64_t)0 : 0;
#pragma unroll
    for (int w = 0; w < WT; w++) if (w >= Wadd) ad[w] = fill;
  } else {
#pragma unroll
    for (int w = 0; w < WT; w++) ad[w] = 0;
  }
  uint64_t carry = up ? 1u : 0u;
#pragma unroll
  for (int w = 0; w < WT; w++) {
    gq_u128 t = (gq_u128)quot[w] + ad[w] + carry;
    quot[w] = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
  }
  if (q.is_pow2) {
    if (q.k < 64u * WT) mw_smod_pow2<WT>(quot, q.k);
  } else {
    smod_general<WT>(quot, q);
  }
  store_words<WT>(out.p[b], n, i, Wout, quot);
}

// ---- fast K5 / K6 (power-of-two q): low words only --------------------------------------
// The results above are only needed modulo q = 2^k (k <= 64*W), and the centred CRT value is
//   T = S - M*P,   S = sum_d y_d * (P/p_d),   M = floor((S + ceil(P/2)) / P)
// (M counts the multiples of P removed by "mod P" plus the one removed by centring).  Modulo
// 2^(64W) only the low W words of every P/p_d and of P enter, so the D x WT multiply-accumulate
// shrinks to D x W - provided M is known exactly.  (S + ceil(P/2))/P = sum_d y_d/p_d + ceil(P/2)/P
// is evaluated in fixed point with F = 64 + bits(p) - 1 fractional bits from below (floor(2^F/p_d)
// fits one word): est <= true < est + D*2^-63.  If the fraction of est is not within 2^-54 of 1
// the integer part IS M.  Otherwise (the value sits on the +P/2 edge; never for honest ciphertexts) the coefficient index is put
// on a list and recomputed by the exact full-width kernel above in list mode.
template <int W>
struct CrtSharedLow {
  uint64_t phat[GQ_MAX_PRIMES * W];
  uint64_t P[W];
  uint64_t inv[GQ_MAX_PRIMES], inv_sh[GQ_MAX_PRIMES], p[GQ_MAX_PRIMES];
  uint64_t u0[GQ_MAX_PRIMES];
  uint64_t hc[2];
};

template <int W>
GQ_DEV void crt_low_stage(CrtSharedLow<W> &S, const GqCrtBasis &B, const GqPrime *primes) {
  for (int j = threadIdx.x; j < B.D * W; j += blockDim.x) {
    const int d = j / W, w = j % W;
    S.phat[j] = (w < B.WT) ? B.phat[d * B.WT + w] : 0;
  }
  for (int j = threadIdx.x; j < W; j += blockDim.x) S.P[j] = (j < B.WT) ? B.P[j] : 0;
  for (int j = threadIdx.x; j < B.D; j += blockDim.x) {
    S.inv[j] = B.inv[j]; S.inv_sh[j] = B.inv_sh[j]; S.p[j] = primes[j].p;
    S.u0[j] = B.u[2 * j];
  }
  if (threadIdx.x < 2) S.hc[threadIdx.x] = B.hc[threadIdx.x];
}

// acc[0..W) = low words of the centred CRT value; false if M could not be decided
template <int W, bool PRE>
GQ_DEV bool crt_low(uint64_t *acc, const CrtSharedLow<W> &S, int D, int fs,
                    const uint64_t *__restrict__ rhat, int n, int i) {
#pragma unroll
  for (int w = 0; w < W; w++) acc[w] = 0;
  uint64_t e0 = S.hc[0], e1 = S.hc[1], e2 = 0;
  // residue loads run two limbs ahead of the arithmetic (long-scoreboard was the top stall)
  uint64_t a0 = rhat[i], a1 = (D > 1) ? rhat[(size_t)n + i] : 0;
  for (int d = 0; d < D; d++) {
    const uint64_t a = a0;
    a0 = a1;
    a1 = (d + 2 < D) ? rhat[(size_t)(d + 2) * n + i] : 0;
    const uint64_t y = PRE ? a : gq_mul_shoup(a, S.inv[d], S.inv_sh[d], S.p[d]);
    mw_mac_low<W>(acc, y, &S.phat[d * W]);
    // est += y * floor(2^F / p_d)   (one 64x64 product, 192-bit accumulator)
    const uint64_t h0 = gq_mulhi(y, S.u0[d]), l0 = y * S.u0[d];
    const uint64_t t = e0 + l0;
    const uint64_t c0 = t < l0;
    e0 = t;
    const uint64_t t1 = e1 + h0;
    uint64_t c1 = t1 < h0;
    e1 = t1 + c0;
    c1 += e1 < c0;
    e2 += c1;
  }
  // est = M * 2^F + fraction; F = 64 + fs, 32 <= fs <= 62
  const uint64_t frac = (e0 >> fs) | (e1 << (64 - fs));
  if (frac >= ~(uint64_t)0 - 1023) return false;
  const uint64_t M = (e1 >> fs) | (e2 << (64 - fs));
  // acc -= M * P  (low words)
  uint64_t carry = 0, borrow = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    gq_u128 t = (gq_u128)M * S.P[w] + carry;
    const uint64_t lo = (uint64_t)t;
    carry = (uint64_t)(t >> 64);
    const uint64_t a = acc[w], dd = a - lo, b1 = a < lo, d2 = dd - borrow, b2 = dd < borrow;
    acc[w] = d2;
    borrow = b1 | b2;
  }
  return true;
}

GQ_DEV void fix_push(int *fix_count, int *fix_list, int entry) {
#ifdef GQ_EMULATE
  fix_list[(*fix_count)++] = entry;
  fix_count[1]++;
#else
  fix_list[atomicAdd(fix_count, 1)] = entry;
  atomicAdd(fix_count + 1, 1);          // lifetime total, read by gq_ctx_crt_fixups()
#endif
}

template <int W, bool PRE>
__global__ void __launch_bounds__(256)
k_reconstruct_fast(GqCrtBasis B, const GqPrime *__restrict__ primes, int n,
                   const uint64_t *__restrict__ rhat, unsigned k, GqPtrs4 out, int Wout,
                   int *fix_count, int *fix_list, unsigned rs_j, GqWaitIn wt) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ CrtSharedLow<W> S;
  crt_low_stage<W>(S, B, primes);
  gq_peer_wait(wt);                   // sharded pipeline: rhat is a receive slab (gq_p2p.h); no-op otherwise
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.y;
  uint64_t acc[W];
  if (!crt_low<W, PRE>(acc, S, B.D, B.F - 64, rhat + (size_t)b * B.D * n, n, i)) { fix_push(fix_count, fix_list, b * n + i); return; }
  if (rs_j) mw_rdiv_pow2<W>(acc, rs_j);         // he_rs fused (see reconstruct_one)
  mw_smod_pow2<W>(acc, k - rs_j);
  store_words<W>(ptr4(out, b), n, i, Wout, acc);
}

// key-switch epilogue: T low WP words by crt_low over the D2 basis, r = T mod P exactly over the
// first `dim` limbs (full width WP, cheap), then the same exact quotient as ks_one.
// WQ <= WP: the result is only needed modulo q = 2^k, k < 64*WQ, so T, T - r and the quotient are carried in
// WQ words (WQ = the context's words per coefficient; 7 against WP = 8 at logq = 438); only r = T mod P needs
// the full width of P.
template <int WP, int WQ, bool PRE>
__global__ void __launch_bounds__(128)
k_ks_epilogue_fast(GqCrtBasis B, GqCrtBasis BP, const uint64_t *__restrict__ Pinv,
                   const GqPrime *__restrict__ primes, int n, const uint64_t *__restrict__ chat,
                   unsigned k, GqPtrs4 add, int Wadd, GqPtrs4 out, int Wout, int *fix_count, int *fix_list,
                   unsigned rs_j, const ulonglong2 *ratio, GqWaitIn wt) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ CrtSharedLow<WQ> S;
  __shared__ CrtShared<WP> SP;
  __shared__ uint64_t sPinv[WQ];
  crt_low_stage<WQ>(S, B, primes);
  crt_stage<WP>(SP, BP, primes, PRE ? ratio : nullptr);
  for (int j = threadIdx.x; j < WQ; j += blockDim.x) sPinv[j] = Pinv[j];
  gq_peer_wait(wt);                   // sharded pipeline: chat is a receive slab (gq_p2p.h); no-op otherwise
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.y;
  const uint64_t *ch = chat + (size_t)b * B.D * n;
  uint64_t x[WQ];
  if (!crt_low<WQ, PRE>(x, S, B.D, B.F - 64, ch, n, i)) { fix_push(fix_count, fix_list, b * n + i); return; }
  bool round_up;
  {
    uint64_t r[WP + 1];
    crt_mod_P<WP>(r, SP, BP, ch, n, i);            // r = T mod P in [0,P)
    round_up = !mw_geq<WP>(SP.Ph, r);              // r > floor(P/2)
    mw_sub<WQ>(x, x, r);                           // low words of T - r
  }
  uint64_t Q[WQ];                                  // (T - r) * P^-1 mod 2^(64*WQ): the exact quotient mod 2^(64*WQ)
#pragma unroll
  for (int w = 0; w < WQ; w++) Q[w] = 0;
#pragma unroll
  for (int a = 0; a < WQ; a++) {
    uint64_t carry = 0;
#pragma unroll
    for (int c = 0; c + a < WQ; c++) {
      gq_u128 t = (gq_u128)x[a] * sPinv[c] + Q[a + c] + carry;
      Q[a + c] = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
    }
  }
  {
    const uint64_t *ap = ptr4(add, b);
    uint64_t carry = round_up ? 1u : 0u, top = 0;
#pragma unroll
    for (int w = 0; w < WQ; w++) {
      uint64_t ad = 0;
      if (ap) { if (w < Wadd) { ad = ap[(size_t)w * n + i]; top = ad; } else ad = (top >> 63) ? ~(uint64_t)0 : 0; }
      gq_u128 t = (gq_u128)Q[w] + ad + carry;
      Q[w] = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
    }
  }
  if (rs_j) mw_rdiv_pow2<WQ>(Q, rs_j);          // he_rs fused into the epilogue: the quotient is in registers
  mw_smod_pow2<WQ>(Q, k - rs_j);
  store_words<WQ>(ptr4(out, b), n, i, Wout, Q);
}

// ---- K4 ---------------------------------------------------------------------------
// (lo,hi) 128-bit value -> canonical residue mod p
GQ_DEV uint64_t reduce128(uint64_t lo, uint64_t hi, const GqPrime &P, uint64_t T1, uint64_t T1sh) {
  if (P.bits >= 32) {
    uint64_t r1 = gq_mul_shoup(hi, T1, T1sh, P.p);
    uint64_t r2 = gq_barrett128(lo, 0, P);
    uint64_t s = r1 + r2;
    return (s >= P.p) ? s - P.p : s;
  }
  gq_u128 v = ((gq_u128)hi << 64) | lo;
  return (uint64_t)(v % P.p);
}

// K4, digit form (primes of at least 33 bits).  The coefficient is read as 2W 32-bit digits x_j;
//   sum_j x_j * (2^(32j) mod p)  =  A0 + 2^32 * A1,   A0 = sum x_j * lo32(U_j) (< 2^70),
//                                                      A1 = sum x_j * hi32(U_j) (< 2^64 for p < 2^61)
// costs two 32x32 multiply-adds per digit instead of a 64x64 product plus a 192-bit add per
// word; the 100-bit total is folded once with two lazy Shoup products (by 2^64 mod p and by 1).
template <int W>
__global__ void __launch_bounds__(256)
k_decompose32(GqDecTables dt, int n, int limb0, int nlimbs, GqPtrs4 in, uint64_t *__restrict__ out) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  __shared__ uint64_t sU[GQ_MAX_PRIMES * 2 * W];
  __shared__ uint64_t sC[GQ_MAX_PRIMES * 6];      // p, T1, T1sh, one_sh, TW, pad
  for (int j = threadIdx.x; j < nlimbs * 2 * W; j += blockDim.x)
    sU[j] = dt.pow32[(size_t)(limb0 + j / (2 * W)) * GQ_DEC32 + j % (2 * W)];
  for (int d = threadIdx.x; d < nlimbs; d += blockDim.x) {
    sC[6 * d + 0] = dt.primes[limb0 + d].p;
    sC[6 * d + 1] = dt.pow64[(size_t)(limb0 + d) * GQ_DEC_W + 1];
    sC[6 * d + 2] = dt.t1_sh[limb0 + d];
    sC[6 * d + 3] = dt.one_sh[limb0 + d];
    sC[6 * d + 4] = dt.pow64[(size_t)(limb0 + d) * GQ_DEC_W + W];
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.y;
  const uint64_t *src = (b == 0) ? in.p[0] : (b == 1) ? in.p[1] : (b == 2) ? in.p[2] : in.p[3];
  uint64_t *dst = out + (size_t)b * nlimbs * n;
  uint32_t x[2 * W];
#pragma unroll
  for (int w = 0; w < W; w++) {
    const uint64_t v = src[(size_t)w * n + i];
    x[2 * w] = (uint32_t)v; x[2 * w + 1] = (uint32_t)(v >> 32);
  }
  const bool neg = (x[2 * W - 1] >> 31) != 0;
  for (int d = 0; d < nlimbs; d++) {
    const uint64_t *U = &sU[d * 2 * W];
    const uint64_t p = sC[6 * d], T1 = sC[6 * d + 1], T1sh = sC[6 * d + 2], osh = sC[6 * d + 3];
    gq_u128 A0 = x[0];
    uint64_t A1 = 0;
#pragma unroll
    for (int j = 1; j < 2 * W; j++) {
      const uint64_t u = U[j];
      A0 += (uint64_t)x[j] * (uint32_t)u;
      A1 += (uint64_t)x[j] * (uint32_t)(u >> 32);
    }
    const gq_u128 V = A0 + ((gq_u128)A1 << 32);
    const uint64_t vlo = (uint64_t)V, vhi = (uint64_t)(V >> 64);
    // vhi*2^64 + vlo  ==  vhi*T1 + vlo*1 (mod p), each lazy product in [0,2p)
    uint64_t r = (vhi * T1 - gq_mulhi(vhi, T1sh) * p) + (vlo - gq_mulhi(vlo, osh) * p);
    r = (r >= 2 * p) ? r - 2 * p : r;
    r = (r >= p) ? r - p : r;
    if (neg) {
      const uint64_t TW = sC[6 * d + 4];
      r = (r >= TW) ? r - TW : r - TW + p;
    }
    dst[(size_t)d * n + i] = r;
  }
}

// K4, 28-bit digit form (coefficients of at most 7 words = 16 digits, primes of 33..60 bits).  With
//   U_j = 2^(28j) mod p = u0_j + 2^32 u1_j   (u1_j < 2^28),
//   sum_j x_j U_j = A0 + 2^32 A1,  A0 = sum x_j u0_j < 16 * 2^28 * 2^32 = 2^64,  A1 = sum x_j u1_j < 2^60:
// both column sums fit 64 bits, so every term is ONE 32x32+64 multiply-add with no carry handling (the 32-bit
// digit form needs a 70-bit A0: three add-with-carry per term; SASS 188 -> about 105 instructions per residue).
// The 93-bit total is folded as in k_decompose32.
// PM: every prime is 2^s + c with 33 <= s <= 59, c < 2^min(s-31,28) (the shape of every GPQHE prime,
// src/precomp.c:358-376: 2^59 + 1 + k*2N).  Then 2^s == -c and the 93-bit total V = Vh*2^s + Vl folds with two
// small products instead of two 64-bit Shoup multiplications:  t = c*Vh < 2^62 = th*2^s + tl,
//   V == Vl - tl + c*th  (mod p),   r = Vl + p - tl + c*th  in (0, 2p + 2^31).
// PEER: the residues are not written to `out` but straight into the receive slabs of the ranks that own their limbs
// (sharded pipeline, gq_p2p.h); the last block to finish publishes the exchange's epoch.
template <int W, bool PM, bool PEER>
__global__ void __launch_bounds__(256)
k_decompose28(GqDecTables dt, int n, int limb0, int nlimbs, GqPtrs4 in, uint64_t *__restrict__ out, GqPeerOut po) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  constexpr int J = (64 * W + 27) / 28;
  static_assert(J <= GQ_DEC28, "at most 16 digits");
  __shared__ uint64_t sU[GQ_MAX_PRIMES * J];
  __shared__ uint64_t sC[GQ_MAX_PRIMES * 6];      // p, T1, T1sh, one_sh, TW, pad
  for (int j = threadIdx.x; j < nlimbs * J; j += blockDim.x)
    sU[j] = dt.pow28[(size_t)(limb0 + j / J) * GQ_DEC28 + j % J];
  for (int d = threadIdx.x; d < nlimbs; d += blockDim.x) {
    sC[6 * d + 0] = dt.primes[limb0 + d].p;
    sC[6 * d + 1] = dt.pow64[(size_t)(limb0 + d) * GQ_DEC_W + 1];
    sC[6 * d + 2] = dt.t1_sh[limb0 + d];
    sC[6 * d + 3] = dt.one_sh[limb0 + d];
    sC[6 * d + 4] = dt.pow64[(size_t)(limb0 + d) * GQ_DEC_W + W];
    sC[6 * d + 5] = dt.primes[limb0 + d].bits - 1;
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  if (!PEER && !live) return;
  const int b = blockIdx.y;
  const uint64_t *src = (b == 0) ? in.p[0] : (b == 1) ? in.p[1] : (b == 2) ? in.p[2] : in.p[3];
  uint64_t *dst = out + (size_t)b * nlimbs * n;
  uint64_t xw[W];
#pragma unroll
  for (int w = 0; w < W; w++) xw[w] = live ? src[(size_t)w * n + i] : 0;
  const bool neg = (xw[W - 1] >> 63) != 0;
  uint32_t x[J];
#pragma unroll
  for (int j = 0; j < J; j++) {
    const int o = 28 * j, w = o >> 6, sft = o & 63;
    uint64_t v = xw[w] >> sft;
    if (sft > 36 && w + 1 < W) v |= xw[w + 1] << (64 - sft);
    x[j] = (uint32_t)v & 0x0fffffffu;
  }
  for (int d = 0; d < nlimbs; d++) {
    const uint64_t *U = &sU[d * J];
    const uint64_t p = sC[6 * d], T1 = sC[6 * d + 1], T1sh = sC[6 * d + 2], osh = sC[6 * d + 3];
    uint64_t A0 = x[0], A1 = 0;
#pragma unroll
    for (int j = 1; j < J; j++) {
      const uint64_t u = U[j];
      A0 = gq_mad_wide(x[j], (uint32_t)u, A0);
      A1 = gq_mad_wide(x[j], (uint32_t)(u >> 32), A1);
    }
    const uint64_t vlo = A0 + (A1 << 32);
    const uint64_t vhi = (A1 >> 32) + (vlo < A0 ? 1u : 0u);
    uint64_t r;
    if (PM) {
      const int sh = (int)sC[6 * d + 5];                       // 33..59
      const uint64_t mask = (((uint64_t)1) << sh) - 1;
      const uint32_t cc = (uint32_t)(p - (mask + 1));
      const uint64_t Vh = (vlo >> sh) | (vhi << (64 - sh));     // < 2^(93-sh) <= 2^60
      const uint64_t t = Vh * cc;                               // < 2^62
      const uint64_t u = gq_mad_wide((uint32_t)(t >> sh), cc, (vlo & mask) + p - (t & mask));
      r = u;
    } else {
      // vhi*2^64 + vlo  ==  vhi*T1 + vlo*1 (mod p), each lazy product in [0,2p)
      r = (vhi * T1 - gq_mulhi(vhi, T1sh) * p) + (vlo - gq_mulhi(vlo, osh) * p);
    }
    r = (r >= 2 * p) ? r - 2 * p : r;
    r = (r >= p) ? r - p : r;
    if (neg) {
      const uint64_t TW = sC[6 * d + 4];
      r = (r >= TW) ? r - TW : r - TW + p;
    }
    if (PEER) { if (live) *gq_peer_c2l(po, b, limb0 + d, i) = r; }
    else dst[(size_t)d * n + i] = r;
  }
  if (PEER) gq_peer_signal(po);
}

template <int W>
__global__ void __launch_bounds__(256)
k_decompose(GqDecTables dt, int n, int limb0, int nlimbs, GqPtrs4 in, uint64_t *__restrict__ out) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.y;
  const uint64_t *src = in.p[b];
  uint64_t *dst = out + (size_t)b * nlimbs * n;
  uint64_t x[W];
#pragma unroll
  for (int w = 0; w < W; w++) x[w] = src[(size_t)w * n + i];
  const bool neg = (x[W - 1] >> 63) != 0;
  for (int d = 0; d < nlimbs; d++) {
    const GqPrime P = dt.primes[limb0 + d];
    const uint64_t *T = dt.pow64 + (size_t)(limb0 + d) * GQ_DEC_W;
    const uint64_t T1 = __ldg(&T[1]), T1sh = __ldg(&dt.t1_sh[limb0 + d]);
    // 192-bit accumulator of sum_w x_w * (2^(64w) mod p)
    uint64_t a0 = x[0], a1 = 0, a2 = 0;
#pragma unroll
    for (int w = 1; w < W; w++) {
      gq_u128 t = (gq_u128)x[w] * __ldg(&T[w]);
      uint64_t tl = (uint64_t)t, th = (uint64_t)(t >> 64);
      a0 += tl;
      uint64_t c = a0 < tl;
      a1 += c;
      uint64_t c2 = a1 < c;
      a1 += th;
      c2 += a1 < th;
      a2 += c2;
    }
    // fold: a0 + a1*T1 + a2*T2  < 2^125
    gq_u128 v = (gq_u128)a1 * T1 + a0;
    if (W > 2) v += (gq_u128)a2 * __ldg(&T[2]);
    uint64_t r = reduce128((uint64_t)v, (uint64_t)(v >> 64), P, T1, T1sh);
    if (neg) {
      const uint64_t TW = __ldg(&T[W]);
      r = (r >= TW) ? r - TW : r - TW + P.p;
    }
    dst[(size_t)d * n + i] = r;
  }
}

}  // namespace

// ---- launchers --------------------------------------------------------------------
// exact (full-width) kernels exist for CRT bases of up to GQ_MAX_WT = 48 words (3 072 bits: the 47-limb key-switch
// basis of the reference's logn = 15, logq = 881 ring needs 2 774); M(WT) is expanded once per width
#define GQ_WT_CASES(WTV, M) \
  switch (WTV) { case 8: M(8); break; case 16: M(16); break; case 24: M(24); break; case 32: M(32); break; \
                 case 40: M(40); break; case 48: M(48); break; default: return -1; }

// fix: {count, list} scratch of the context; non-null selects the fast low-word kernel followed
// by the exact kernel in list mode (power-of-two q only).  Returns kernels launched.
int gq_launch_reconstruct(const GqCrtBasis &B, const GqPrime *primes, int n, const uint64_t *rhat,
                          const GqModulus &q, const GqPtrs4 &out, int Wout, int npolys,
                          cudaStream_t st, int *fix, unsigned rs_j, bool pre, const GqWaitIn *wt,
                          const uint32_t *tc_tab) {
  if (rs_j && (!q.is_pow2 || rs_j > q.k)) return -1;
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  dim3 grid((n + 127) / 128, npolys), block(128);
  int *fc = nullptr, *fl = nullptr;
  int launches = 1;
  if (fix && B.F && q.is_pow2 && q.k >= 1 && q.k <= 64u * GQ_MAX_W) {
    fc = fix; fl = fix + 4;
    const int W = (int)((q.k + 63) / 64);
    dim3 g2((n + 255) / 256, npolys), b2(256);
#define GQ_RF(V) do { if (pre) GQ_LAUNCH((k_reconstruct_fast<V, true>), g2, b2, st, B, primes, n, rhat, q.k, out, Wout, fc, fl, rs_j, w0); \
                      else GQ_LAUNCH((k_reconstruct_fast<V, false>), g2, b2, st, B, primes, n, rhat, q.k, out, Wout, fc, fl, rs_j, w0); } while (0)
    // tensor-core form (gq_crt_tc.cu) when its table was built for this basis: pre-scaled residues, k <= 512
    if (tc_tab && pre && gq_launch_reconstruct_tc(B, tc_tab, n, rhat, q.k, out, Wout, npolys, st, fix, rs_j, wt) == 1) {
    } else
    if (W <= 1) GQ_RF(1); else if (W <= 2) GQ_RF(2); else if (W <= 3) GQ_RF(3); else if (W <= 4) GQ_RF(4);
    else if (W <= 5) GQ_RF(5); else if (W <= 6) GQ_RF(6); else if (W <= 7) GQ_RF(7); else if (W <= 8) GQ_RF(8);
    else if (W <= 12) GQ_RF(12); else if (W <= 16) GQ_RF(16); else if (W <= 24) GQ_RF(24); else GQ_RF(32);
#undef GQ_RF
    grid = dim3(GQ_FIX_BLOCKS, 1);
    launches = 2;
  } else if (pre) {
    return -1;        // pre-scaled residues are only produced for the fast path
  } else if (wt) {
    return -2;        // the flag wait lives in the fast kernel only
  }
#define GQ_REC_T(V) GQ_LAUNCH((k_reconstruct<V, true>), grid, block, st, B, primes, n, rhat, q, out, Wout, fc, fl, rs_j, pre)
#define GQ_REC_F(V) GQ_LAUNCH((k_reconstruct<V, false>), grid, block, st, B, primes, n, rhat, q, out, Wout, fc, fl, rs_j, pre)
  if (q.is_pow2) {
    GQ_WT_CASES(B.WT, GQ_REC_T)
  } else {
    GQ_WT_CASES(B.WT, GQ_REC_F)
  }
#undef GQ_REC_T
#undef GQ_REC_F
  return launches;
}

template <int WT>
static int ks_dispatch_wp(const GqCrtBasis &B, const GqCrtBasis &BP, const uint64_t *Pinv,
                          const GqPrime *primes, int n, const uint64_t *chat, unsigned k,
                          const GqPtrs4 &add, int Wadd, const GqPtrs4 &out, int Wout, dim3 grid,
                          cudaStream_t st, int *fc, int *fl, unsigned rs_j, const ulonglong2 *ratio) {
  dim3 block(128);
  if (BP.WT > WT) return -1;
#define GQ_KS(WPV) GQ_LAUNCH((k_ks_epilogue<WT, (WT >= WPV ? WPV : 8)>), grid, block, st, B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, fc, fl, rs_j, ratio)
  switch (BP.WT) {
    case 8: GQ_KS(8); return 1;
    case 16: if (WT >= 16) { GQ_KS(16); return 1; } return -1;
    case 24: if (WT >= 24) { GQ_KS(24); return 1; } return -1;
    case 32: if (WT >= 32) { GQ_KS(32); return 1; } return -1;
  }
#undef GQ_KS
  return -1;
}

int gq_launch_ks_epilogue(const GqCrtBasis &B, const GqCrtBasis &BP, const uint64_t *Pinv,
                          const GqPrime *primes, int n, const uint64_t *chat, unsigned k,
                          const GqPtrs4 &add, int Wadd, const GqPtrs4 &out, int Wout, int npolys,
                          cudaStream_t st, int *fix, unsigned rs_j, const ulonglong2 *ratio, const GqWaitIn *wt,
                          const uint32_t *tc_tab) {
  if (rs_j > k) return -1;
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  dim3 grid((n + 127) / 128, npolys);
  int *fc = nullptr, *fl = nullptr;
  int launches = 1;
  if (fix && B.F && k >= 1 && BP.WT <= B.WT) {
    fc = fix; fl = fix + 4;
    dim3 block(128);
    // tensor-core form (gq_crt_tc.cu) when its table was built for this pair of bases: pre-scaled residues, k <= 512
    if (tc_tab && ratio &&
        gq_launch_ks_epilogue_tc(B, BP, tc_tab, primes, ratio, n, chat, k, add, Wadd, out, Wout, npolys, st, fix, rs_j, wt) == 1) {
    } else
    switch (BP.WT) {
#define GQ_KFL(WPV, WQV) do { if (ratio) GQ_LAUNCH((k_ks_epilogue_fast<WPV, WQV, true>), grid, block, st, B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, fc, fl, rs_j, ratio, w0); \
                              else GQ_LAUNCH((k_ks_epilogue_fast<WPV, WQV, false>), grid, block, st, B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, fc, fl, rs_j, ratio, w0); } while (0)
#define GQ_KF(V) case V: GQ_KFL(V, V); break;
      case 8:     // the shipped parameter sets (logq = 438: 7 words per coefficient, P of 8 words) and their neighbours
        if (Wout == 7 && Wadd <= 7 && k < 64u * 7) GQ_KFL(8, 7);
        else if (Wout == 6 && Wadd <= 6 && k < 64u * 6) GQ_KFL(8, 6);
        else GQ_KFL(8, 8);
        break;
      GQ_KF(16) GQ_KF(24) GQ_KF(32)
#undef GQ_KF
#undef GQ_KFL
      default: return -1;
    }
    grid = dim3(GQ_FIX_BLOCKS, 1);
    launches = 2;
  } else if (ratio) {
    return -1;        // pre-scaled residues are only produced for the fast path
  } else if (wt) {
    return -2;        // the flag wait lives in the fast kernel only
  }
  int r = -1;
  switch (B.WT) {
    case 8: r = ks_dispatch_wp<8>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
    case 16: r = ks_dispatch_wp<16>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
    case 24: r = ks_dispatch_wp<24>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
    case 32: r = ks_dispatch_wp<32>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
    case 40: r = ks_dispatch_wp<40>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
    case 48: r = ks_dispatch_wp<48>(B, BP, Pinv, primes, n, chat, k, add, Wadd, out, Wout, grid, st, fc, fl, rs_j, ratio); break;
  }
  return r < 0 ? r : launches;
}

template <int W>
static int dec_launch(const GqDecTables &dt, int n, int limb0, int nlimbs, const GqPtrs4 &in,
                      uint64_t *out, int npolys, cudaStream_t st, int digits, const GqPeerOut *po) {
  dim3 grid((n + 255) / 256, npolys), block(256);
  const GqPeerOut none = GqPeerOut();
  if constexpr (W <= 7) {
    if (po) {      // peer output exists for the 28-bit digit kernels only (every shipped parameter set)
      if (digits == 3) { GQ_LAUNCH((k_decompose28<W, true, true>), grid, block, st, dt, n, limb0, nlimbs, in, out, *po); return 1; }
      if (digits == 2) { GQ_LAUNCH((k_decompose28<W, false, true>), grid, block, st, dt, n, limb0, nlimbs, in, out, *po); return 1; }
      return -2;
    }
    if (digits == 3) { GQ_LAUNCH((k_decompose28<W, true, false>), grid, block, st, dt, n, limb0, nlimbs, in, out, none); return 1; }
    if (digits == 2) { GQ_LAUNCH((k_decompose28<W, false, false>), grid, block, st, dt, n, limb0, nlimbs, in, out, none); return 1; }
  }
  if (po) return -2;
  if (digits) GQ_LAUNCH(k_decompose32<W>, grid, block, st, dt, n, limb0, nlimbs, in, out);
  else GQ_LAUNCH(k_decompose<W>, grid, block, st, dt, n, limb0, nlimbs, in, out);
  return 1;
}

// po != null: peer output (sharded pipeline); returns -2 when no peer-output kernel covers these parameters
int gq_launch_decompose(const GqDecTables &dt, int n, int limb0, int nlimbs, const GqPtrs4 &in,
                        int W, uint64_t *out, int npolys, cudaStream_t st, int digits, const GqPeerOut *po) {
  switch (W) {
#define GQ_CASE(V) case V: return dec_launch<V>(dt, n, limb0, nlimbs, in, out, npolys, st, digits, po);
    GQ_CASE(1) GQ_CASE(2) GQ_CASE(3) GQ_CASE(4) GQ_CASE(5) GQ_CASE(6) GQ_CASE(7) GQ_CASE(8)
    GQ_CASE(9) GQ_CASE(10) GQ_CASE(11) GQ_CASE(12) GQ_CASE(13) GQ_CASE(14) GQ_CASE(15) GQ_CASE(16)
    GQ_CASE(17) GQ_CASE(18) GQ_CASE(19) GQ_CASE(20) GQ_CASE(21) GQ_CASE(22) GQ_CASE(23) GQ_CASE(24)
    GQ_CASE(25) GQ_CASE(26) GQ_CASE(27) GQ_CASE(28) GQ_CASE(29) GQ_CASE(30) GQ_CASE(31) GQ_CASE(32)
#undef GQ_CASE
  }
  return -1;
}

int gq_launch_ks_epilogue_general(const GqCrtBasis &B, const uint64_t *P, int nP, const uint64_t *Ph,
                                  const GqPrime *primes, int n, const uint64_t *chat,
                                  const GqModulus &q, const GqPtrs4 &add, int Wadd,
                                  const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st) {
  dim3 grid((n + 127) / 128, npolys), block(128);
#define GQ_KSG(V) GQ_LAUNCH(k_ks_epilogue_general<V>, grid, block, st, B, P, nP, Ph, primes, n, chat, q, add, Wadd, out, Wout)
  GQ_WT_CASES(B.WT, GQ_KSG)
#undef GQ_KSG
  return 1;
}
