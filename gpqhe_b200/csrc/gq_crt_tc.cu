// gq_crt_tc.cu - K5 (poly_rns2mpi, src/poly.c:109-120 over rns_reconstruct, src/rns.c:60-75) as a dense integer
// contraction on the tensor cores.
//
// The fast CRT of gq_crt.cu computes, per coefficient, T = sum_d y_d * (P/p_d) - M * P modulo 2^(64W) with
// M = floor(sum_d y_d * floor(2^F/p_d) / 2^F + ...) from a fixed-point estimate: D x W 64-bit multiply-accumulates with
// carry chains, about 2 100 integer instructions per coefficient, latency-bound (ncu: 54 % issue, 46 % warps active).
// Both sums are contractions of the residue BYTES with constant byte tables:
//     C[i][j] = sum_{d, a} byte_a(y_d[i]) * byte_{j-a}(K_d),     K_d = (P/p_d) mod 2^(64W)   (j < 8W)
//                                                                 K_d = floor(2^F/p_d)         (16 more columns)
// i.e. (coefficients x 8D) * (8D x (8W + 16)) in u8 with exact int32 accumulation (at most 8D products of 255^2:
// below 2^24), which mma.sync.m16n8k32 does at 568 T MAC/s on this part (profiles/probe/gqp_imma.cu).  What is left
// per coefficient is the carry propagation of 8W + 16 byte columns - about 200 instructions - and the tail of the
// integer kernel (subtract M * P, optional fused he_rs, smod 2^k, store).  Same integers, same fix-list rule, same
// stores as k_reconstruct_fast: bit-identical (tests: every he_mul / he_mulpt / poly_mul parity case runs through it).
//
// Fragment layouts (PTX ISA, mma.m16n8k32 .u8; g = lane / 4, t = lane % 4): A: a0..a3 = rows (g, g+8, g, g+8) x
// k-bytes (4t.., 4t.., 16+4t.., 16+4t..); B: b0, b1 = k-bytes (4t.., 16+4t..) x column g; C: c0, c1 = row g, columns
// 2t, 2t+1; c2, c3 = row g+8.  The K and N orders are ours (the table is a constant):
//   * k-slots (4t..4t+3, 16+4t..16+4t+3) of chunk c carry the 8 bytes of the residue of LIMB 4c + t: an A fragment is
//     one 64-bit load per row, 8 consecutive coefficients per quad column (one 64-byte segment);
//   * column 2t + e of n-tile v < W is output byte 2W*t + 2v + e: quad lane t ends up with 2W CONSECUTIVE byte columns
//     (its slice of T is one 128-bit partial sum, W <= 7); n-tiles W, W+1 carry estimate bytes 4t + 2(v-W) + e.
// The four partial sums of a row meet through 24 64-bit shuffles; lanes t = 0, 1 finish rows g, g+8.
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

// T[0..W) += (hi:lo) << off (mod 2^(64W)); off is a compile-time constant after unrolling
template <int W>
GQ_DEV void add_shifted(uint64_t *T, uint64_t lo, uint64_t hi, int off) {
  const int w0 = off >> 6, sb = off & 63;
  const uint64_t p0 = lo << sb;
  const uint64_t p1 = sb ? ((lo >> (64 - sb)) | (hi << sb)) : hi;
  const uint64_t p2 = sb ? (hi >> (64 - sb)) : 0;
  uint64_t carry = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    const uint64_t add = (w == w0) ? p0 : (w == w0 + 1) ? p1 : (w == w0 + 2) ? p2 : 0;
    if (w >= w0) {
      const uint64_t s = T[w] + add, c1 = s < add, s2 = s + carry, c2 = s2 < carry;
      T[w] = s2;
      carry = c1 | c2;
    }
  }
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

// tab: B fragments [chunk][n-tile][lane][2], built by gq_crt_tc_table; the residues rhat are PRE-SCALED
// (y_d = a_d (P/p_d)^-1 mod p_d, folded into the inverse NTT: GqNttTables::fin_s)
template <int W>
__global__ void __launch_bounds__(128)
k_reconstruct_tc(GqCrtBasis B, const uint32_t *__restrict__ tab, int nchunks, int n, const uint64_t *__restrict__ rhat,
                 unsigned k, GqPtrs4 out, int Wout, int *fix_count, int *fix_list, unsigned rs_j, GqWaitIn wt) {
  constexpr int NT = W + 2;
  GQ_DYN_SMEM(uint32_t, sB);
  __shared__ uint64_t sP[W];
  __shared__ uint64_t sHc[2];
  for (int j = threadIdx.x; j < nchunks * NT * 64; j += blockDim.x) sB[j] = tab[j];
  for (int j = threadIdx.x; j < W; j += blockDim.x) sP[j] = (j < B.WT) ? B.P[j] : 0;
  if (threadIdx.x < 2) sHc[threadIdx.x] = B.hc[threadIdx.x];
  gq_peer_wait(wt);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int b = blockIdx.y, D = B.D, fs = B.F - 64;
  const uint64_t *rh = rhat + (size_t)b * D * n;
  uint64_t *dst = ptr4(out, b);
  for (int i0 = (blockIdx.x * 4 + warp) * 16; i0 < n; i0 += gridDim.x * 64) {
    int acc[NT][4];
#pragma unroll
    for (int v = 0; v < NT; v++)
#pragma unroll
      for (int e = 0; e < 4; e++) acc[v][e] = 0;
    const bool r0 = i0 + g < n, r1 = i0 + g + 8 < n;
    for (int c = 0; c < nchunks; c++) {
      const int d = 4 * c + t;
      const uint64_t v0 = (d < D && r0) ? rh[(size_t)d * n + i0 + g] : 0;
      const uint64_t v1 = (d < D && r1) ? rh[(size_t)d * n + i0 + g + 8] : 0;
      const uint32_t A[4] = {(uint32_t)v0, (uint32_t)v1, (uint32_t)(v0 >> 32), (uint32_t)(v1 >> 32)};
#pragma unroll
      for (int v = 0; v < NT; v++) {
        const uint64_t bb = *reinterpret_cast<const uint64_t *>(&sB[((c * NT + v) * 32 + lane) * 2]);
        mma_u8(acc[v], A, (uint32_t)bb, (uint32_t)(bb >> 32));
      }
    }
    // my 2W byte columns of T and my 4 estimate bytes -> 128-bit / 64-bit partial sums, for both of my rows
    uint64_t Sl[2], Sh[2], E[2];
#pragma unroll
    for (int row = 0; row < 2; row++) {
      uint64_t lo = 0, hi = 0;
#pragma unroll
      for (int m = 0; m < 2 * W; m++) {
        const uint64_t cv = (uint32_t)acc[m >> 1][2 * row + (m & 1)];      // < 2^24
        const int sh = 8 * m;
        if (sh < 64) {
          const uint64_t add = cv << sh;
          lo += add;
          hi += (lo < add ? 1u : 0u) + ((sh > 40) ? (cv >> (64 - sh)) : 0);
        } else {
          hi += cv << (sh - 64);
        }
      }
      Sl[row] = lo; Sh[row] = hi;
      uint64_t e = 0;
#pragma unroll
      for (int m = 0; m < 4; m++) e += (uint64_t)(uint32_t)acc[W + (m >> 1)][2 * row + (m & 1)] << (8 * m);   // < 2^49
      E[row] = e;
    }
    // quad lanes 0 and 1 collect the four partial sums of rows g and g+8
    const int qb = lane & ~3, mine = t & 1;
    uint64_t T[W];
#pragma unroll
    for (int w = 0; w < W; w++) T[w] = 0;
    uint64_t e0 = sHc[0], e1 = sHc[1], e2 = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const uint64_t l0 = shfl64(Sl[0], qb + q), l1 = shfl64(Sl[1], qb + q);
      const uint64_t h0 = shfl64(Sh[0], qb + q), h1 = shfl64(Sh[1], qb + q);
      const uint64_t x0 = shfl64(E[0], qb + q), x1 = shfl64(E[1], qb + q);
      add_shifted<W>(T, mine ? l1 : l0, mine ? h1 : h0, 16 * W * q);
      // estimate += E_q << 32q  (192-bit accumulator e2:e1:e0)
      const uint64_t x = mine ? x1 : x0;
      const int off = 32 * q;
      const uint64_t a0 = (off < 64) ? (x << off) : 0;
      const uint64_t a1 = (off == 0) ? 0 : (off < 64) ? (x >> (64 - off)) : (x << (off - 64));
      const uint64_t a2 = (off > 64) ? (x >> (128 - off)) : 0;
      const uint64_t s0 = e0 + a0, c0 = s0 < a0;
      e0 = s0;
      const uint64_t s1 = e1 + a1, c1 = s1 < a1, s1b = s1 + c0, c1b = s1b < c0;
      e1 = s1b;
      e2 += a2 + c1 + c1b;
    }
    const int i = i0 + g + 8 * mine;
    if (t < 2 && i < n) {
      // est = M * 2^F + fraction, F = 64 + fs (the same decision rule as crt_low, gq_crt.cu)
      const uint64_t frac = (e0 >> fs) | (e1 << (64 - fs));
      if (frac >= ~(uint64_t)0 - 1023) {
        fix_push_tc(fix_count, fix_list, b * n + i);
      } else {
        const uint64_t M = (e1 >> fs) | (e2 << (64 - fs));
        uint64_t carry = 0, borrow = 0;
#pragma unroll
        for (int w = 0; w < W; w++) {          // T -= M * P (low words)
          const gq_u128 pr = (gq_u128)M * sP[w] + carry;
          const uint64_t lo = (uint64_t)pr;
          carry = (uint64_t)(pr >> 64);
          const uint64_t a = T[w], dd = a - lo, b1 = a < lo, d2 = dd - borrow, b2 = dd < borrow;
          T[w] = d2;
          borrow = b1 | b2;
        }
        if (rs_j) mw_rdiv_pow2<W>(T, rs_j);    // he_rs fused (see reconstruct_one, gq_crt.cu)
        mw_smod_pow2<W>(T, k - rs_j);
        const uint64_t fill = (T[W - 1] >> 63) ? ~(uint64_t)0 : 0;
#pragma unroll
        for (int w = 0; w < W; w++)
          if (w < Wout) dst[(size_t)w * n + i] = T[w];
        for (int w = W; w < Wout; w++) dst[(size_t)w * n + i] = fill;
      }
    }
  }
}

}  // namespace

// Words of the B-fragment table for a basis of D limbs and W result words.
size_t gq_crt_tc_table_words(int D, int W) { return (size_t)((D + 3) / 4) * (W + 2) * 64; }

// Host side: the B fragments.  phat_low[d*W + w]: low W words of P/p_d; u[d] = floor(2^F / p_d) (one word).
void gq_crt_tc_table(uint32_t *tab, int D, int W, const uint64_t *phat_low, const uint64_t *u) {
  const int NT = W + 2, nchunks = (D + 3) / 4;
  for (int c = 0; c < nchunks; c++)
    for (int v = 0; v < NT; v++)
      for (int lane = 0; lane < 32; lane++) {
        const int g = lane >> 2, t = lane & 3;        // B fragment: column g of the n-tile, k-bytes 4t.. and 16+4t..
        // column g belongs to quad lane tc = g / 2 of the C fragment, e = g % 2
        const int tc = g >> 1, e = g & 1;
        const int d = 4 * c + t;
        uint32_t reg[2] = {0, 0};
        for (int half = 0; half < 2; half++)
          for (int kb = 0; kb < 4; kb++) {
            const int a = 4 * half + kb;              // byte a of the residue of limb d
            int val = 0;
            if (d < D) {
              if (v < W) {
                const int j = 2 * W * tc + 2 * v + e; // output byte position
                const int src = j - a;                // byte of K_d = phat_d mod 2^(64W)
                if (src >= 0 && src < 8 * W) val = (int)((phat_low[(size_t)d * W + src / 8] >> (8 * (src % 8))) & 0xff);
              } else {
                const int j = 4 * tc + 2 * (v - W) + e;
                const int src = j - a;                // byte of u_d (64 bits)
                if (src >= 0 && src < 8) val = (int)((u[d] >> (8 * src)) & 0xff);
              }
            }
            reg[half] |= (uint32_t)val << (8 * kb);
          }
        tab[(((size_t)c * NT + v) * 32 + lane) * 2 + 0] = reg[0];
        tab[(((size_t)c * NT + v) * 32 + lane) * 2 + 1] = reg[1];
      }
}

// Returns 1 (one launch; the caller follows with the exact list-mode kernel), -1 when no instantiation fits.
int gq_launch_reconstruct_tc(const GqCrtBasis &B, const uint32_t *tab, int n, const uint64_t *rhat, unsigned k,
                             const GqPtrs4 &out, int Wout, int npolys, cudaStream_t st, int *fix, unsigned rs_j,
                             const GqWaitIn *wt) {
  if (!tab || !fix || k < 1 || rs_j > k || B.F < 64 + 32 || B.F > 64 + 62) return -1;
  const int W = (int)((k + 63) / 64), nchunks = (B.D + 3) / 4;
  if (W > 7 || W > Wout) return -1;
  GqWaitIn w0 = GqWaitIn();
  if (wt) w0 = *wt;
  int gx = (n + 255) / 256;           // 256 coefficients per block: the 9-14 KB table is staged once per block
  if (gx < 1) gx = 1;
  const dim3 grid(gx, npolys), block(128);
  const size_t smem = gq_crt_tc_table_words(B.D, W) * 4;
  int *fc = fix, *fl = fix + 4;
#ifdef GQ_EMULATE
#define GQ_TC(V) gq_emul::launch_dyn(grid, block, smem, [&]() { k_reconstruct_tc<V>(B, tab, nchunks, n, rhat, k, out, Wout, fc, fl, rs_j, w0); })
#else
#define GQ_TC(V) k_reconstruct_tc<V><<<grid, block, smem, st>>>(B, tab, nchunks, n, rhat, k, out, Wout, fc, fl, rs_j, w0)
#endif
  switch (W) {
    case 1: GQ_TC(1); break;
    case 2: GQ_TC(2); break;
    case 3: GQ_TC(3); break;
    case 4: GQ_TC(4); break;
    case 5: GQ_TC(5); break;
    case 6: GQ_TC(6); break;
    case 7: GQ_TC(7); break;
    default: return -1;
  }
#undef GQ_TC
  return 1;
}
