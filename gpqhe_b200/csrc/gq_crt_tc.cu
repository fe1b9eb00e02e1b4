// gq_crt_tc.cu - K5 (poly_rns2mpi, src/poly.c:109-120 over rns_reconstruct, src/rns.c:60-75) and K6 (the key-switch
// epilogue of he_mul / he_rot / he_conj: round(T / P) + ct, src/he-mult.c:57-64, src/he-kem.c:24-45) as dense integer
// contractions on the tensor cores.
//
// The integer kernels of gq_crt.cu compute, per coefficient, T = sum_d y_d * (P/p_d) - M * P modulo 2^(64W) with the
// overflow count M from a fixed-point estimate: D x W 64-bit multiply-accumulates with carry chains (about 2 000
// instructions per coefficient; the key-switch epilogue, which adds T mod P over the special limbs and a W x W
// product with P^-1, about 6 000), latency-bound.  Every one of those sums is LINEAR in the bytes of the residues:
//     C[i][j] = sum_{d, a} byte_a(y_d[i]) * byte_{j-a}(K_d)
// is a (coefficients x 8D) * (8D x 64) product in u8 with exact int32 accumulation (at most 8D products of 255^2),
// which mma.sync.m16n8k32 does at 568 T MAC/s on this part (profiles/probe/gqp_imma.cu), with
//     K_d = (P/p_d) mod 2^512                                   for K5,
//     K_d = (PQ/p_d) * P^-1 mod 2^512                           for K6 (the exact quotient (T - r) / P is
//     K'_i = -(P/p_i) * P^-1 mod 2^512 on z_i                   (T - r) * P^-1 modulo a power of two, and both
//                                                               T = sum y_d (PQ/p_d) - M PQ and r = T mod P =
//                                                               sum z_i (P/p_i) - M' P are sums of this kind),
// and 16 more columns per sum for the fixed-point estimates floor(2^F/p_d) of M and M'.  z_i = y_i * ratio_i mod p_i
// (one Shoup product per special limb) are the residues pre-scaled for the basis of P (GqHostBasis::ratio).  What is
// left per coefficient is the carry propagation of the byte columns, -M * C (C = P for K5, PQ * P^-1 = Q for K6; M
// is at most D + 1, so this is a small multiple), + M' + [r > P/2] (decided from the fraction of the M' estimate), the
// ciphertext addend, the fused he_rs, smod 2^k and the store.  The same integers, the same fix-list rule (estimates
// too close to a decision boundary go to the exact kernel in list mode), the same stores as the integer kernels:
// bit-identical (tests: every he_mul / he_rot / poly_mul parity case runs through it, and against the integer kernels
// with GPQHE_B200_NO_TC_CRT).
//
// Fragment layouts (PTX ISA, mma.m16n8k32 .u8; g = lane / 4, t = lane % 4): A: a0..a3 = rows (g, g+8, g, g+8) x
// k-bytes (4t.., 4t.., 16+4t.., 16+4t..); B: b0, b1 = k-bytes (4t.., 16+4t..) x column g; C: c0, c1 = row g, columns
// 2t, 2t+1; c2, c3 = row g+8.  The K and N orders are ours (the table is a constant):
//   * k-slots (4t..4t+3, 16+4t..16+4t+3) of chunk c carry the 8 bytes of the residue of LIMB 4c + t: an A fragment is
//     one 64-bit load per row, 8 consecutive coefficients per quad column (one 64-byte segment);
//   * column 2t + e of n-tile v < 8 is output byte 16t + 2v + e: quad lane t ends up with 16 CONSECUTIVE byte columns =
//     words 2t, 2t+1 of an 8-word result (W <= 8: k <= 512); n-tiles 8, 9 carry estimate bytes 4t + 2(v-8) + e.
// Carries between the four lanes of a quad are small signed numbers (below 2^31) and ripple through three shuffles.
#include "gq_bigint.h"
#include "gq_kernels.h"
#include "gq_p2p.h"

namespace {

GQ_DEV void mma_u8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#if defined(__CUDA_ARCH__)
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
#elif defined(GQ_EMULATE)
  gq_emul_mma_u8(c, a, b0, b1);
#endif
}

GQ_DEV uint64_t shfl64(uint64_t v, int src) {
  const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)v, src), hi = __shfl_sync(0xffffffffu, (uint32_t)(v >> 32), src);
  return ((uint64_t)hi << 32) | lo;
}

GQ_DEV void fix_push_tc(int *fix_count, int *fix_list, int entry) {
#ifdef GQ_EMULATE
  fix_list[(*fix_count)++] = entry;
  fix_count[1]++;
#else
  fix_list[atomicAdd(fix_count, 1)] = entry;
  atomicAdd(fix_count + 1, 1);
#endif
}

GQ_DEV uint64_t *ptr4(const GqPtrs4 &P, int b) { return b == 0 ? P.p[0] : b == 1 ? P.p[1] : b == 2 ? P.p[2] : P.p[3]; }

// three-word accumulator (v2 in two's complement): v += b, v -= b
GQ_DEV void add3(uint64_t &v0, uint64_t &v1, uint64_t &v2, uint64_t b0, uint64_t b1, uint64_t b2) {
#if defined(__CUDA_ARCH__)
  asm("add.cc.u64 %0, %0, %3;\n\taddc.cc.u64 %1, %1, %4;\n\taddc.u64 %2, %2, %5;"
      : "+l"(v0), "+l"(v1), "+l"(v2) : "l"(b0), "l"(b1), "l"(b2));
#else
  const uint64_t s0 = v0 + b0, c0 = s0 < b0;
  const uint64_t s1 = v1 + b1, c1 = s1 < b1, s1b = s1 + c0, c1b = s1b < c0;
  v0 = s0; v1 = s1b; v2 = v2 + b2 + c1 + c1b;
#endif
}
GQ_DEV void sub3(uint64_t &v0, uint64_t &v1, uint64_t &v2, uint64_t b0, uint64_t b1, uint64_t b2) {
#if defined(__CUDA_ARCH__)
  asm("sub.cc.u64 %0, %0, %3;\n\tsubc.cc.u64 %1, %1, %4;\n\tsubc.u64 %2, %2, %5;"
      : "+l"(v0), "+l"(v1), "+l"(v2) : "l"(b0), "l"(b1), "l"(b2));
#else
  const uint64_t d0 = v0 - b0, w0 = v0 < b0;
  const uint64_t d1 = v1 - b1, w1 = v1 < b1, d1b = d1 - w0, w1b = d1 < w0;
  v0 = d0; v1 = d1b; v2 = v2 - b2 - (w1 | w1b);
#endif
}

// four byte columns (each below 2^24) -> c0 + c1 2^8 + c2 2^16 + c3 2^24 (below 2^49): three 32 x 32 + 64 multiply-adds
GQ_DEV uint64_t pack4(int c0, int c1, int c2, int c3) {
  uint64_t r = (uint64_t)(uint32_t)c1 * 0x100u + (uint32_t)c0;
  r = (uint64_t)(uint32_t)c2 * 0x10000u + r;
  return (uint64_t)(uint32_t)c3 * 0x1000000u + r;
}
// P0 + P1 2^32 + P2 2^64 + P3 2^96 (each P below 2^49) as three words
GQ_DEV void sum4(uint64_t P0, uint64_t P1, uint64_t P2, uint64_t P3, uint64_t &w0, uint64_t &w1, uint64_t &w2) {
#if defined(__CUDA_ARCH__)
  uint32_t s1, s2, s3, s4;
  asm("add.cc.u32 %0, %4, %5;\n\taddc.cc.u32 %1, %6, %7;\n\taddc.cc.u32 %2, %8, %9;\n\taddc.u32 %3, %10, 0;"
      : "=r"(s1), "=r"(s2), "=r"(s3), "=r"(s4)
      : "r"((uint32_t)(P0 >> 32)), "r"((uint32_t)P1), "r"((uint32_t)(P1 >> 32)), "r"((uint32_t)P2),
        "r"((uint32_t)(P2 >> 32)), "r"((uint32_t)P3), "r"((uint32_t)(P3 >> 32)));
  w0 = ((uint64_t)s1 << 32) | (uint32_t)P0;
  w1 = ((uint64_t)s3 << 32) | s2;
  w2 = s4;
#else
  const uint64_t a = P1 << 32;
  w0 = P0 + a;
  const uint64_t c0 = w0 < a;
  const uint64_t bsum = (P1 >> 32) + P2 + c0;              // below 2^50: no overflow
  const uint64_t d = P3 << 32;
  w1 = bsum + d;
  w2 = (P3 >> 32) + (w1 < d ? 1u : 0u);
#endif
}
// m * (C1 : C0) for a 32-bit m: five 32-bit limbs through four chained 32 x 32 + 32 multiply-adds
GQ_DEV void mul_small(uint32_t m, uint64_t C0, uint64_t C1, uint64_t &b0, uint64_t &b1, uint64_t &b2) {
  const uint64_t t0 = (uint64_t)m * (uint32_t)C0;
  const uint64_t t1 = (uint64_t)m * (uint32_t)(C0 >> 32) + (t0 >> 32);
  const uint64_t t2 = (uint64_t)m * (uint32_t)C1 + (t1 >> 32);
  const uint64_t t3 = (uint64_t)m * (uint32_t)(C1 >> 32) + (t2 >> 32);
  b0 = (t1 << 32) | (uint32_t)t0;
  b1 = (t3 << 32) | (uint32_t)t2;
  b2 = t3 >> 32;
}

struct TcArgs {
  const unsigned long long *frag;   // B fragments [chunk][10 n-tiles][32 lanes], 64 bits per lane
  const uint64_t *cw;               // [8] multiplier of the overflow count (P for K5, Q = PQ * P^-1 for K6), low 512 bits
  const uint64_t *hc;               // [2] centring constant of the estimate of M (GqCrtBasis::hc)
  const GqPrime *primes;            // K6: the special limbs
  const ulonglong2 *ratio;          // K6: y_i -> z_i (GqHostBasis::ratio)
  int D, ych, DP, zch;              // limbs / chunks of the y part; limbs / chunks of the z part (K6)
  int fs, fsP;                      // fractional bits of the estimates beyond 64 (GqCrtBasis::F - 64)
};

// one chunk (four limbs, one per quad lane) of a 16-coefficient step: ten MMAs, and for a special limb of K6 ten more
// on the residue re-scaled for the basis of P
template <bool KS>
GQ_DEV void tc_chunk(const TcArgs &a, const unsigned long long *sfrag, int c, int lane, int t, uint64_t v0, uint64_t v1,
                     int (&acc)[8][4], int (&eM)[2][4], int (&eP)[2][4]) {
  constexpr int NT = 10;
  if (c >= a.ych) return;
  const uint32_t A[4] = {(uint32_t)v0, (uint32_t)v1, (uint32_t)(v0 >> 32), (uint32_t)(v1 >> 32)};
  const unsigned long long *tb = sfrag + (size_t)c * NT * 32 + lane;
#pragma unroll
  for (int v = 0; v < NT; v++) {
    const unsigned long long bb = tb[v * 32];         // 8 bytes per lane, consecutive lanes: conflict-free
    if (v < 8) mma_u8(acc[v], A, (uint32_t)bb, (uint32_t)(bb >> 32));
    else mma_u8(eM[v - 8], A, (uint32_t)bb, (uint32_t)(bb >> 32));
  }
  if (KS && c < a.zch) {
    const int d = 4 * c + t;
    uint64_t z0 = 0, z1 = 0;
    if (d < a.DP) {
      const ulonglong2 rt = a.ratio[d];
      const uint64_t p = a.primes[d].p;
      z0 = v0 * rt.x - gq_mulhi(v0, rt.y) * p;
      if (z0 >= p) z0 -= p;
      z1 = v1 * rt.x - gq_mulhi(v1, rt.y) * p;
      if (z1 >= p) z1 -= p;
    }
    const uint32_t Az[4] = {(uint32_t)z0, (uint32_t)z1, (uint32_t)(z0 >> 32), (uint32_t)(z1 >> 32)};
    const unsigned long long *tz = sfrag + (size_t)(a.ych + c) * NT * 32 + lane;
#pragma unroll
    for (int v = 0; v < NT; v++) {
      const unsigned long long bb = tz[v * 32];
      if (v < 8) mma_u8(acc[v], Az, (uint32_t)bb, (uint32_t)(bb >> 32));
      else mma_u8(eP[v - 8], Az, (uint32_t)bb, (uint32_t)(bb >> 32));
    }
  }
}

// KS = false: K5, out = smod(round-shift(T)); KS = true: K6, out = smod(round-shift(round(T / P) + add)).
// The residues rhat are PRE-SCALED for the basis (y_d = a_d (P/p_d)^-1 mod p_d, canonical; GqNttTables::fin_s).
// Quad lane t owns words 2t, 2t+1 of an 8-word integer (only the first Wout are stored).
// RCH: chunks whose residues are held in registers and fetched one step ahead (later chunks, if any, are loaded in
// place).  Persistent blocks: the fragment table is staged in shared memory once, each warp walks 16-coefficient steps.
template <bool KS, int RCH>
__global__ void __launch_bounds__(128, 4)
k_crt_tc(TcArgs a, int n, const uint64_t *__restrict__ rhat, unsigned k, GqPtrs4 add, int Wadd, GqPtrs4 out, int Wout,
         int *fix_count, int *fix_list, unsigned rs_j, GqWaitIn wt) {
  GQ_PDL_TRIGGER();
  constexpr int NT = 10;
  GQ_DYN_SMEM(unsigned long long, sfrag);
  {                                                   // 2 560 bytes per chunk: 160 16-byte pieces, loads batched by 5
    const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(a.frag);
    ulonglong2 *dst2 = reinterpret_cast<ulonglong2 *>(sfrag);
    const int pieces = (a.ych + a.zch) * NT * 16;
    for (int j0 = threadIdx.x; j0 < pieces; j0 += 5 * 128) {
      ulonglong2 tmp[5];
#pragma unroll
      for (int u = 0; u < 5; u++) if (j0 + u * 128 < pieces) tmp[u] = src[j0 + u * 128];
#pragma unroll
      for (int u = 0; u < 5; u++) if (j0 + u * 128 < pieces) dst2[j0 + u * 128] = tmp[u];
    }
  }
  GQ_PDL_WAIT();                                      // the table above is constant; the residues are not
  gq_peer_wait(wt);                                   // (ends with the block barrier that also publishes sfrag)
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3, qb = lane & ~3;
  const int b = blockIdx.y, D = a.D;
  const uint64_t *rh = rhat + (size_t)b * D * n;
  uint64_t *dst = ptr4(out, b);
  const uint64_t *ap = KS ? ptr4(add, b) : nullptr;
  const uint64_t C0 = a.cw[2 * t], C1 = a.cw[2 * t + 1];
  const uint64_t hc0 = a.hc[0], hc1 = a.hc[1];
  const unsigned kk = k - rs_j;                       // the result is reduced smod 2^kk
  const unsigned sw = (kk - 1) >> 6, sb = (kk - 1) & 63;
  const uint64_t keep = (sb == 63) ? ~(uint64_t)0 : ((((uint64_t)1) << (sb + 1)) - 1);
  const uint64_t R = rs_j ? (((uint64_t)1) << (rs_j - 1)) - 1 : 0;      // he_rs: floor((x + 2^(j-1) - 1) / 2^j) rounds
                                                                         // to nearest, ties down (mw_rdiv_pow2)
  const unsigned ia = 2 * t, ib = 2 * t + 1;
  const int stride = gridDim.x * 64;
  uint64_t Rs[RCH][2];
  {
    const int i0 = (blockIdx.x * 4 + warp) * 16;
#pragma unroll
    for (int u = 0; u < RCH; u++) {
      const int d = 4 * u + t;
      Rs[u][0] = (d < D && i0 + g < n) ? rh[(size_t)d * n + i0 + g] : 0;
      Rs[u][1] = (d < D && i0 + g + 8 < n) ? rh[(size_t)d * n + i0 + g + 8] : 0;
    }
  }
  for (int i0 = (blockIdx.x * 4 + warp) * 16; i0 < n; i0 += stride) {
    int acc[8][4], eM[2][4], eP[2][4];
#pragma unroll
    for (int v = 0; v < 8; v++)
#pragma unroll
      for (int e = 0; e < 4; e++) acc[v][e] = 0;
#pragma unroll
    for (int v = 0; v < 2; v++)
#pragma unroll
      for (int e = 0; e < 4; e++) eM[v][e] = eP[v][e] = 0;
    const bool r0 = i0 + g < n, r1 = i0 + g + 8 < n;
    uint64_t ad[2][2] = {{0, 0}, {0, 0}};
    if (KS && ap) {                                   // the ciphertext addend: my two words of rows g, g+8
      if (r0 && (int)ia < Wadd) ad[0][0] = ap[(size_t)ia * n + i0 + g];
      if (r0 && (int)ib < Wadd) ad[0][1] = ap[(size_t)ib * n + i0 + g];
      if (r1 && (int)ia < Wadd) ad[1][0] = ap[(size_t)ia * n + i0 + g + 8];
      if (r1 && (int)ib < Wadd) ad[1][1] = ap[(size_t)ib * n + i0 + g + 8];
    }
#pragma unroll
    for (int u = 0; u < RCH; u++) tc_chunk<KS>(a, sfrag, u, lane, t, Rs[u][0], Rs[u][1], acc, eM, eP);
    for (int c = RCH; c < a.ych; c++) {               // bases wider than the register window
      const int d = 4 * c + t;
      const uint64_t v0 = (d < D && r0) ? rh[(size_t)d * n + i0 + g] : 0;
      const uint64_t v1 = (d < D && r1) ? rh[(size_t)d * n + i0 + g + 8] : 0;
      tc_chunk<KS>(a, sfrag, c, lane, t, v0, v1, acc, eM, eP);
    }
    {                                                 // the next step's residues travel while this step's carries resolve
      const int j0 = i0 + stride;
#pragma unroll
      for (int u = 0; u < RCH; u++) {
        const int d = 4 * u + t;
        Rs[u][0] = (d < D && j0 + g < n) ? rh[(size_t)d * n + j0 + g] : 0;
        Rs[u][1] = (d < D && j0 + g + 8 < n) ? rh[(size_t)d * n + j0 + g + 8] : 0;
      }
    }
#pragma unroll
    for (int row = 0; row < 2; row++) {
      const int i = i0 + g + 8 * row, e0 = 2 * row, e1 = 2 * row + 1;
      uint64_t v0, v1, v2;
      sum4(pack4(acc[0][e0], acc[0][e1], acc[1][e0], acc[1][e1]), pack4(acc[2][e0], acc[2][e1], acc[3][e0], acc[3][e1]),
           pack4(acc[4][e0], acc[4][e1], acc[5][e0], acc[5][e1]), pack4(acc[6][e0], acc[6][e1], acc[7][e0], acc[7][e1]),
           v0, v1, v2);
      // fixed-point estimate of the overflow count: every lane gathers the four 4-byte slices
      const uint64_t Em = pack4(eM[0][e0], eM[0][e1], eM[1][e0], eM[1][e1]);
      uint64_t s0, s1, s2;
      sum4(shfl64(Em, qb), shfl64(Em, qb + 1), shfl64(Em, qb + 2), shfl64(Em, qb + 3), s0, s1, s2);
      add3(s0, s1, s2, hc0, hc1, 0);
      const uint64_t frac = (s0 >> a.fs) | (s1 << (64 - a.fs));
      bool undecided = frac >= ~(uint64_t)0 - 1023;                  // the same rule as crt_low (gq_crt.cu)
      const uint32_t M = (uint32_t)((s1 >> a.fs) | (s2 << (64 - a.fs)));   // at most D + 1
      uint64_t b0, b1, b2;
      mul_small(M, C0, C1, b0, b1, b2);
      sub3(v0, v1, v2, b0, b1, b2);
      uint64_t lane0 = R;
      if (KS) {
        const uint64_t Ep = pack4(eP[0][e0], eP[0][e1], eP[1][e0], eP[1][e1]);
        uint64_t p0, p1, p2;
        sum4(shfl64(Ep, qb), shfl64(Ep, qb + 1), shfl64(Ep, qb + 2), shfl64(Ep, qb + 3), p0, p1, p2);
        const uint64_t fracP = (p0 >> a.fsP) | (p1 << (64 - a.fsP));        // (r / P) 2^64, from below, error < 1024
        const uint32_t MP = (uint32_t)((p1 >> a.fsP) | (p2 << (64 - a.fsP)));
        const uint64_t half = ((uint64_t)1) << 63;
        // r = sum z_i (P/p_i) - M' P needs floor() decided; [r > P/2] needs the fraction clear of 1/2 (P is odd:
        // r / P is never 1/2, and the estimate is a lower bound, so fracP >= 1/2 decides "above")
        undecided = undecided || fracP >= ~(uint64_t)0 - 1023 || (fracP < half && fracP >= half - 1024);
        lane0 += (uint64_t)MP + (fracP >= half ? 1u : 0u);           // -r contributes +M' (P * P^-1 = 1); round up
        add3(v0, v1, v2, ad[row][0], ad[row][1], 0);
      }
      add3(v0, v1, v2, t == 0 ? lane0 : 0, 0, 0);
      // ripple the small signed carries across the quad: lane r takes lane r-1's third word
#pragma unroll
      for (int r = 1; r < 4; r++) {
        const int cin = __shfl_sync(0xffffffffu, (int)(uint32_t)v2, qb + r - 1);
        if (t == r) {
          const uint64_t ext = (uint64_t)((int64_t)cin >> 63);
          add3(v0, v1, v2, (uint64_t)(int64_t)cin, ext, ext);
        }
      }
      if (rs_j) {                // floor shift by j < 64: my upper word needs the next lane's lower word
        const uint64_t nxt = shfl64(v0, qb + ((t + 1) & 3));
        const uint64_t top = (t == 3) ? (uint64_t)((int64_t)v1 >> 63) : nxt;
        v0 = (v0 >> rs_j) | (v1 << (64 - rs_j));
        v1 = (v1 >> rs_j) | (top << (64 - rs_j));
      }
      // smod 2^kk: sign-extend from bit kk-1 (word sw, owned by quad lane sw / 2)
      const uint64_t sword = (sw & 1) ? v1 : v0;
      const uint32_t sbit = __shfl_sync(0xffffffffu, (uint32_t)((sword >> sb) & 1u), qb + (int)(sw >> 1));
      const uint64_t fill = 0 - (uint64_t)sbit;
      const uint64_t o0 = (ia < sw) ? v0 : (ia == sw) ? ((v0 & keep) | (fill & ~keep)) : fill;
      const uint64_t o1 = (ib < sw) ? v1 : (ib == sw) ? ((v1 & keep) | (fill & ~keep)) : fill;
      if (i < n) {
        if (undecided) {
          if (t == 0) fix_push_tc(fix_count, fix_list, b * n + i);
        } else {
          if ((int)ia < Wout) dst[(size_t)ia * n + i] = o0;
          if ((int)ib < Wout) dst[(size_t)ib * n + i] = o1;
        }
      }
    }
  }
}


// V = lo4 + hi4 2^32 (lo4, hi4 < 2^47: V < 2^79) -> canonical residue mod p = 2^s + cc, s = 32 + s32 in 33..59,
// cc < 2^min(s - 31, 28): 2^s == -cc, so with V = Vh 2^s + Vl and t = cc Vh = th 2^s + tl (t < 2^48)
//   V == Vl - tl + cc th,   r = Vl + p - tl + cc th  in (0, 2p + 2^31)      (the fold of k_decompose28, gq_crt.cu)
GQ_DEV uint64_t fold_pm(uint64_t lo4, uint64_t hi4, uint64_t p, uint32_t cc, uint32_t s32) {
  const uint32_t m1 = (1u << s32) - 1;
  const uint64_t mid = (lo4 >> 32) + (uint32_t)hi4;           // limb 1 with its carry: below 2^33
  const uint32_t v0 = (uint32_t)lo4, v1 = (uint32_t)mid, v2 = (uint32_t)(hi4 >> 32) + (uint32_t)(mid >> 32);
#if defined(__CUDA_ARCH__)
  const uint32_t vh_lo = __funnelshift_r(v1, v2, s32);
#else
  const uint32_t vh_lo = (uint32_t)(((((uint64_t)v2) << 32) | v1) >> s32);
#endif
  const uint32_t vh_hi = v2 >> s32;                           // Vh < 2^46; vh_hi * cc < 2^16
  const uint64_t tt = gq_mad_wide(vh_lo, cc, (uint64_t)(vh_hi * cc) << 32);
  const uint32_t th = (uint32_t)(tt >> 32) >> s32;
  const uint64_t tl = tt & ((((uint64_t)m1) << 32) | 0xffffffffu);
  const uint64_t vl = (((uint64_t)(v1 & m1)) << 32) | v0;
  uint64_t r = gq_mad_wide(th, cc, vl + p - tl);
  r = (r >= 2 * p) ? r - 2 * p : r;
  r = (r >= p) ? r - p : r;
  return r;
}

// ---- K4 (poly_mpi2rns, src/poly.c:96-107 over rns_decompose, src/rns.c:37-48) as the same kind of contraction --------
// residue_d = sum_{w, a} byte_a(x_w) * (2^(64w + 8a) mod p_d)  -  sign * (2^(64W) mod p_d)        (mod p_d)
// K = the 8 bytes of each of the W <= 7 coefficient words (word slot 7 carries the sign as a 0/1 byte against the
// constant p - 2^(64W) mod p), N = the 8 bytes of the table entry of each limb: per group of four limbs two k-steps x
// four n-tiles; column 2t + e of n-tile v is byte 2v + e of LIMB base + t, so lane (g, t) ends with all eight byte
// columns of (row g, limb base + t) and (row g + 8, limb base + t): the recombination is thread-local, then one
// pseudo-Mersenne fold (every prime 2^s + c, the shape pm_ok checks; same fold as k_decompose28 of gq_crt.cu).
// Table: per limb 72 64-bit words (64 used: [k-chunk 2][n-tile 4][e 2][t 4]; the stride of 576 bytes keeps the four
// limbs of a B fragment in different banks).
#define GQ_DTC_STRIDE 72
template <bool PEER>
__global__ void __launch_bounds__(128)
k_decompose_tc(const unsigned long long *__restrict__ tab, const GqPrime *__restrict__ primes, int n, int limb0,
               int nlimbs, int W, GqPtrs4 in, uint64_t *__restrict__ out, GqPeerOut po) {
  GQ_PDL_TRIGGER();
  GQ_DYN_SMEM(unsigned long long, sT);                // [groups * 4][72], then the primes
  const int groups = (nlimbs + 3) / 4;
  unsigned long long *sP = sT + (size_t)groups * 4 * GQ_DTC_STRIDE, *sQ = sP + groups * 4;
  uint64_t **sDst = reinterpret_cast<uint64_t **>(sQ + groups * 4);     // PEER: where limb d of my slice goes
  {
    const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(tab + (size_t)limb0 * GQ_DTC_STRIDE);
    ulonglong2 *dst2 = reinterpret_cast<ulonglong2 *>(sT);
    const int pieces = nlimbs * (GQ_DTC_STRIDE / 2), all = groups * 4 * (GQ_DTC_STRIDE / 2);
    for (int j0 = threadIdx.x; j0 < all; j0 += 4 * 128) {
      ulonglong2 tmp[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int j = j0 + u * 128;
        tmp[u] = make_ulonglong2(0, 0);
        if (j < pieces) tmp[u] = src[j];
      }
#pragma unroll
      for (int u = 0; u < 4; u++) if (j0 + u * 128 < all) dst2[j0 + u * 128] = tmp[u];
    }
    for (int d = threadIdx.x; d < groups * 4; d += blockDim.x) {
      const uint64_t pd = d < nlimbs ? primes[limb0 + d].p : 0;
      const unsigned sh = d < nlimbs ? primes[limb0 + d].bits - 1 : 33;
      sP[d] = pd;
      sQ[d] = ((unsigned long long)(sh - 32) << 32) | (uint32_t)(pd - (((uint64_t)1) << sh));
      if (PEER) sDst[d] = d < nlimbs ? gq_peer_c2l(po, blockIdx.y, limb0 + d, 0) : nullptr;
    }
  }
  GQ_PDL_WAIT();                                      // tables and primes are constants; the coefficients are not
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3, qb = lane & ~3;
  const int b = blockIdx.y;
  const uint64_t *src = (b == 0) ? in.p[0] : (b == 1) ? in.p[1] : (b == 2) ? in.p[2] : in.p[3];
  uint64_t *dst = out + (size_t)b * nlimbs * n;
  const int stride = gridDim.x * 64;
  const int cs = (W - 1) >> 2, ts = (W - 1) & 3;      // the top word lives in chunk cs, quad lane ts
  // my B-fragment words: limb base + (g >> 1), e = g & 1, k-slot t
  const unsigned long long *myB = sT + (size_t)(g >> 1) * GQ_DTC_STRIDE + (g & 1) * 4 + t;
  uint64_t X[2][2];
  {
    const int i0 = (blockIdx.x * 4 + warp) * 16;
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const int w = 4 * c + t;
      X[c][0] = (w < W && i0 + g < n) ? src[(size_t)w * n + i0 + g] : 0;
      X[c][1] = (w < W && i0 + g + 8 < n) ? src[(size_t)w * n + i0 + g + 8] : 0;
    }
  }
  for (int i0 = (blockIdx.x * 4 + warp) * 16; i0 < n; i0 += stride) {
    uint32_t A[2][4];
    {
      const uint64_t top0 = cs ? X[1][0] : X[0][0], top1 = cs ? X[1][1] : X[0][1];
      const uint32_t s0 = __shfl_sync(0xffffffffu, (uint32_t)(top0 >> 63), qb + ts);
      const uint32_t s1 = __shfl_sync(0xffffffffu, (uint32_t)(top1 >> 63), qb + ts);
#pragma unroll
      for (int c = 0; c < 2; c++) {
        A[c][0] = (uint32_t)X[c][0]; A[c][1] = (uint32_t)X[c][1];
        A[c][2] = (uint32_t)(X[c][0] >> 32); A[c][3] = (uint32_t)(X[c][1] >> 32);
      }
      if (t == 3) { A[1][0] = s0; A[1][1] = s1; A[1][2] = 0; A[1][3] = 0; }      // word slot 7: the sign
    }
    {                                                 // the next step's words travel while this step's limbs are folded
      const int j0 = i0 + stride;
#pragma unroll
      for (int c = 0; c < 2; c++) {
        const int w = 4 * c + t;
        X[c][0] = (w < W && j0 + g < n) ? src[(size_t)w * n + j0 + g] : 0;
        X[c][1] = (w < W && j0 + g + 8 < n) ? src[(size_t)w * n + j0 + g + 8] : 0;
      }
    }
    const bool live0 = i0 + g < n, live1 = i0 + g + 8 < n;
#pragma unroll 2
    for (int G = 0; G < groups; G++) {
      int acc[4][4];
#pragma unroll
      for (int v = 0; v < 4; v++)
#pragma unroll
        for (int e = 0; e < 4; e++) acc[v][e] = 0;
      const unsigned long long *gb = myB + (size_t)G * 4 * GQ_DTC_STRIDE;
#pragma unroll
      for (int c = 0; c < 2; c++)
#pragma unroll
        for (int v = 0; v < 4; v++) {
          const unsigned long long bb = gb[(c * 4 + v) * 8];
          mma_u8(acc[v], A[c], (uint32_t)bb, (uint32_t)(bb >> 32));
        }
      const int d = 4 * G + t;
      const uint64_t p = sP[d];
      if (d < nlimbs) {
        const uint64_t pq = sQ[d];
        const uint32_t cc = (uint32_t)pq, s32 = (uint32_t)(pq >> 32);       // p = 2^(32 + s32) + cc
#pragma unroll
        for (int row = 0; row < 2; row++) {
          const int e0 = 2 * row;
          const uint64_t lo4 = pack4(acc[0][e0], acc[0][e0 + 1], acc[1][e0], acc[1][e0 + 1]);     // bytes 0..3: < 2^47
          const uint64_t hi4 = pack4(acc[2][e0], acc[2][e0 + 1], acc[3][e0], acc[3][e0 + 1]);     // bytes 4..7
          const uint64_t r = fold_pm(lo4, hi4, p, cc, s32);
          const int i = i0 + g + 8 * row;
          if (row ? live1 : live0) {
            if (PEER) sDst[d][i] = r;
            else dst[(size_t)d * n + i] = r;
          }
        }
      }
    }
  }
  if (PEER) gq_peer_signal(po);
}

}  // namespace

// Words (32-bit) of the table for `slots` limb slots (a multiple of 4): fragments, then the 8 words of the multiplier.
size_t gq_crt_tc_table_words(int slots) { return (size_t)(slots / 4) * 10 * 64 + 16; }

// Host side: the B fragments.  K[s*8 + w]: the 8 words of the constant of slot s (zero for padding slots);
// u[s] = the estimate constant floor(2^F / p) of slot s (one word); cw[8]: the multiplier of the overflow count.
void gq_crt_tc_table(uint32_t *tab, int slots, const uint64_t *K, const uint64_t *u, const uint64_t *cw) {
  const int NT = 10, nchunks = slots / 4;
  for (int c = 0; c < nchunks; c++)
    for (int v = 0; v < NT; v++)
      for (int lane = 0; lane < 32; lane++) {
        const int g = lane >> 2, t = lane & 3;        // B fragment: column g of the n-tile, k-bytes 4t.. and 16+4t..
        const int tc = g >> 1, e = g & 1;             // column g belongs to quad lane tc of the C fragment
        const int s = 4 * c + t;
        uint32_t reg[2] = {0, 0};
        for (int half = 0; half < 2; half++)
          for (int kb = 0; kb < 4; kb++) {
            const int a = 4 * half + kb;              // byte a of the residue in slot s
            int val = 0;
            if (v < 8) {
              const int j = 16 * tc + 2 * v + e;      // output byte position (8 words)
              const int src = j - a;                  // byte of K_s
              if (src >= 0 && src < 64) val = (int)((K[(size_t)s * 8 + src / 8] >> (8 * (src % 8))) & 0xff);
            } else {
              const int j = 4 * tc + 2 * (v - 8) + e;
              const int src = j - a;                  // byte of u_s (64 bits)
              if (src >= 0 && src < 8) val = (int)((u[s] >> (8 * src)) & 0xff);
            }
            reg[half] |= (uint32_t)val << (8 * kb);
          }
        tab[(((size_t)c * NT + v) * 32 + lane) * 2 + 0] = reg[0];
        tab[(((size_t)c * NT + v) * 32 + lane) * 2 + 1] = reg[1];
      }
  uint32_t *tail = tab + (size_t)nchunks * NT * 64;
  for (int w = 0; w < 8; w++) { tail[2 * w] = (uint32_t)cw[w]; tail[2 * w + 1] = (uint32_t)(cw[w] >> 32); }
}

// persistent grid: four blocks of four warps per SM (up to 128 registers per thread; five blocks at 96 registers
// measured slower, 19.7 / 26.8 us against 18.1 / 24.5 us), shared between the polynomials
static dim3 tc_grid(int n, int npolys) {
  int gx = (n + 63) / 64, cap = (GQ_NUM_SMS * 4) / npolys;       // never a second, nearly empty wave
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  return dim3(gx, npolys);
}

#define GQ_TC_LAUNCH(KSV, RV, ...) GQ_LAUNCH_DYN((k_crt_tc<KSV, RV>), grid, block, smem, st, __VA_ARGS__)

// K5.  Returns 1 (one launch; the caller follows with the exact list-mode kernel), -1 when this form does not apply.
int gq_launch_reconstruct_tc(const GqCrtBasis &B, const uint32_t *tab, int n, const uint64_t *rhat, unsigned k,
                             const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st, int *fix, unsigned rs_j,
                             const GqWaitIn *wt) {
  if (!tab || !fix || k < 1 || rs_j >= k || rs_j >= 64 || B.F < 64 + 32 || B.F > 64 + 62) return -1;
  if (k > 512 || Wout > 8 || Wout < (int)((k + 63) / 64)) return -1;
  TcArgs a = TcArgs();
  a.D = B.D; a.ych = (B.D + 3) / 4;
  a.frag = (const unsigned long long *)tab;
  a.cw = (const uint64_t *)(tab + (size_t)a.ych * 10 * 64);
  a.hc = B.hc; a.fs = B.F - 64; a.fsP = a.fs;
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  const dim3 grid = tc_grid(n, npolys), block(128);
  const size_t smem = (size_t)a.ych * 10 * 32 * 8;
  if (smem > 48 * 1024) return -1;
  GqPtrs4 none = GqPtrs4();
  if (a.ych <= 2) GQ_TC_LAUNCH(false, 2, a, n, rhat, k, none, 0, out, Wout, fix, fix + 4, rs_j, w0);
  else if (a.ych <= 4) GQ_TC_LAUNCH(false, 4, a, n, rhat, k, none, 0, out, Wout, fix, fix + 4, rs_j, w0);
  else GQ_TC_LAUNCH(false, 6, a, n, rhat, k, none, 0, out, Wout, fix, fix + 4, rs_j, w0);
  return 1;
}

// K6.  tab: the table of gq_crt_tc_table over 4*ceil(D2/4) + 4*ceil(dim/4) slots (y part, then z part).
int gq_launch_ks_epilogue_tc(const GqCrtBasis &B, const GqCrtBasis &BP, const uint32_t *tab, const GqPrime *primes,
                             const ulonglong2 *ratio, int n, const uint64_t *chat, unsigned k, const GqPtrs4 &add,
                             int Wadd, const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st, int *fix,
                             unsigned rs_j, const GqWaitIn *wt) {
  if (!tab || !fix || !ratio || k < 1 || rs_j >= k || rs_j >= 64) return -1;
  if (B.F < 64 + 32 || B.F > 64 + 62 || BP.F < 64 + 32 || BP.F > 64 + 62 || BP.D > B.D) return -1;
  if (k > 512 || Wout > 8 || Wout < (int)((k + 63) / 64)) return -1;
  bool has_add = false;
  for (int b = 0; b < npolys && b < 4; b++) has_add = has_add || add.p[b];
  if (has_add && (Wadd > 8 || 64u * (unsigned)Wadd < k)) return -1;     // the addend's sign words would matter
  TcArgs a = TcArgs();
  a.D = B.D; a.ych = (B.D + 3) / 4; a.DP = BP.D; a.zch = (BP.D + 3) / 4;
  a.frag = (const unsigned long long *)tab;
  a.cw = (const uint64_t *)(tab + (size_t)(a.ych + a.zch) * 10 * 64);
  a.hc = B.hc; a.fs = B.F - 64; a.fsP = BP.F - 64;
  a.primes = primes; a.ratio = ratio;
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  const dim3 grid = tc_grid(n, npolys), block(128);
  const size_t smem = (size_t)(a.ych + a.zch) * 10 * 32 * 8;
  if (smem > 48 * 1024) return -1;
  if (a.ych <= 4) GQ_TC_LAUNCH(true, 4, a, n, chat, k, add, Wadd, out, Wout, fix, fix + 4, rs_j, w0);
  else GQ_TC_LAUNCH(true, 6, a, n, chat, k, add, Wadd, out, Wout, fix, fix + 4, rs_j, w0);
  return 1;
}

// ---- K4 host side: the per-limb B fragments for coefficients of W <= 7 words, nprimes x GQ_DTC_STRIDE 64-bit words
size_t gq_dec_tc_table_words(int nprimes) { return (size_t)nprimes * GQ_DTC_STRIDE; }
void gq_dec_tc_table(unsigned long long *tab, int nprimes, const uint64_t *p, int W) {
  for (int d = 0; d < nprimes; d++) {
    uint64_t U[8][8];                                // U[w][a] = 2^(64w + 8a) mod p; slot 7: the sign constant
    uint64_t x = 1 % p[d];
    for (int w = 0; w < 8; w++)
      for (int a = 0; a < 8; a++) {
        U[w][a] = x;
        x = (uint64_t)(((gq_u128)x * 256u) % p[d]);
      }
    const uint64_t tw = U[W < 8 ? W : 7][0];         // 2^(64W) mod p
    for (int a = 0; a < 8; a++) U[7][a] = 0;
    U[7][0] = tw ? p[d] - tw : 0;
    for (int w = W; w < 7; w++)
      for (int a = 0; a < 8; a++) U[w][a] = 0;
    unsigned long long *T = tab + (size_t)d * GQ_DTC_STRIDE;
    for (int j = 0; j < GQ_DTC_STRIDE; j++) T[j] = 0;
    for (int c = 0; c < 2; c++)
      for (int v = 0; v < 4; v++)
        for (int e = 0; e < 2; e++)
          for (int t = 0; t < 4; t++) {
            const int w = 4 * c + t, j = 2 * v + e;  // word slot, output byte
            uint32_t reg[2] = {0, 0};
            for (int half = 0; half < 2; half++)
              for (int kb = 0; kb < 4; kb++)
                reg[half] |= (uint32_t)((U[w][4 * half + kb] >> (8 * j)) & 0xff) << (8 * kb);
            T[((c * 4 + v) * 2 + e) * 4 + t] = ((unsigned long long)reg[1] << 32) | reg[0];
          }
  }
}

// Returns 1, or -1 when this form does not apply (the caller uses the integer kernels)
int gq_launch_decompose_tc(const unsigned long long *tab, const GqPrime *primes, int n, int limb0, int nlimbs,
                           const GqPtrs4 &in, int W, uint64_t *out, int npolys, cudaStream_t st, const GqPeerOut *po) {
  if (!tab || W < 1 || W > 7 || nlimbs < 1 || npolys < 1 || npolys > 4) return -1;
  const int groups = (nlimbs + 3) / 4;
  const size_t smem = (size_t)groups * 4 * (GQ_DTC_STRIDE + 3) * 8;
  if (smem > 48 * 1024) return -1;
  int gx = (n + 63) / 64, cap = (GQ_NUM_SMS * 8) / npolys;
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  const dim3 grid(gx, npolys), block(128);
  const GqPeerOut none = GqPeerOut();
  if (po) GQ_LAUNCH_DYN((k_decompose_tc<true>), grid, block, smem, st, tab, primes, n, limb0, nlimbs, W, in, out, *po);
  else GQ_LAUNCH_DYN((k_decompose_tc<false>), grid, block, smem, st, tab, primes, n, limb0, nlimbs, W, in, out, none);
  return 1;
}
