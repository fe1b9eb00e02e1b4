// gq_misc.cu - coefficient-local big-integer kernels on word-major polys.
//   K7 k_rescale_pow2 / k_rescale_general   he_rs, he_moddown   src/he-rescale.c:33-70
//                                           (mpi_rdiv + mpi_smod, src/types.c:108-128)
//   K8 k_automorph                          poly_rot, poly_conj src/poly.c:263-283
//   K10 k_linear_pow2 / k_linear_general    he_add, he_sub, he_addpt, he_subpt, he_neg
//                                           src/he-add.c:32-142 (mpi_addm/subm/neg + mpi_smod)
// HBM-bound: one coalesced read and one write of W words per coefficient.
#include "gq_bigint.h"
#include "gq_kernels.h"

namespace {

__global__ void __launch_bounds__(256)
k_rescale_pow2(int n, int W, GqPtrs4 c, unsigned j, unsigned k) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t *x = c.p[blockIdx.y] + i;
  const uint64_t sign = (x[(size_t)(W - 1) * n] >> 63) ? ~(uint64_t)0 : 0;
  const int jw = j >> 6, jb = j & 63;
  uint64_t carry = 0;
  if (j > 0) {   // remainder > 2^(j-1)  <=>  bit j-1 set and a lower bit set
    const unsigned hb = j - 1;
    const int hw = hb >> 6, hbb = hb & 63;
    uint64_t lower = 0;
    for (int w = 0; w < hw; w++) lower |= x[(size_t)w * n];
    const uint64_t xw = x[(size_t)hw * n];
    const uint64_t half = (xw >> hbb) & 1u;
    if (hbb) lower |= xw & ((((uint64_t)1) << hbb) - 1);
    carry = (half && lower) ? 1u : 0u;
  }
  if (k == 0) {   // q == 1: mpi_smod yields -1 for every input (0 >= qh == 0, minus q)
    for (int w = 0; w < W; w++) x[(size_t)w * n] = ~(uint64_t)0;
    return;
  }
  const int sw = (k - 1) >> 6, sb = (k - 1) & 63;
  uint64_t fill = 0;
  for (int w = 0; w < W; w++) {
    uint64_t val;
    if (w <= sw) {
      const uint64_t lo = (w + jw < W) ? x[(size_t)(w + jw) * n] : sign;
      const uint64_t hi = (w + jw + 1 < W) ? x[(size_t)(w + jw + 1) * n] : sign;
      val = jb ? ((lo >> jb) | (hi << (64 - jb))) : lo;
      val += carry;
      carry = (val < carry) ? 1u : 0u;
      if (w == sw) {
        fill = ((val >> sb) & 1u) ? ~(uint64_t)0 : 0;
        const uint64_t keep = (sb == 63) ? ~(uint64_t)0 : ((((uint64_t)1) << (sb + 1)) - 1);
        val = (val & keep) | (fill & ~keep);
      }
    } else {
      val = fill;
    }
    x[(size_t)w * n] = val;
  }
}

// exact slow path: arbitrary Delta (64-bit) and arbitrary q
__global__ void __launch_bounds__(128)
k_rescale_general(int n, int W, GqPtrs4 c, uint64_t Delta, GqModulus q) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t *x = c.p[blockIdx.y] + i;
  uint64_t u[GQ_MAX_W];
  for (int w = 0; w < W; w++) u[w] = x[(size_t)w * n];
  bool neg = (u[W - 1] >> 63) != 0;
  if (neg) {   // magnitude
    uint64_t cy = 1;
    for (int w = 0; w < W; w++) { uint64_t t = ~u[w] + cy; cy = (cy && t == 0) ? 1u : 0u; u[w] = t; }
  }
  if (Delta > 1) {
    // |x| = qm*Delta + rm ; floor(x/Delta) = -qm - (rm != 0), remainder = Delta - rm for negative x
    uint64_t rem = 0;
    for (int w = W - 1; w >= 0; w--) u[w] = gq_div128(rem, u[w], Delta, &rem);
    uint64_t r = rem;
    uint64_t inc = 0;
    if (neg && rem) { inc = 1; r = Delta - rem; }     // magnitude of the floor quotient grows by one
    // rounding: +1 on the signed quotient iff r > floor(Delta/2)
    const bool up = r > (Delta >> 1);
    // signed quotient = neg ? -(u+inc) + up : u + up  -> work on the magnitude
    if (neg) {
      // value = -(u + inc) + up
      if (inc && up) { /* cancels */ }
      else if (inc) { uint64_t cy = 1; for (int w = 0; w < W && cy; w++) { u[w] += 1; cy = (u[w] == 0); } }
      else if (up) {   // -(u) + 1 = -(u - 1); u > 0 here unless u == 0
        bool zero = true;
        for (int w = 0; w < W; w++) zero = zero && (u[w] == 0);
        if (zero) { neg = false; u[0] = 1; }
        else { uint64_t bw = 1; for (int w = 0; w < W && bw; w++) { bw = (u[w] == 0); u[w] -= 1; } }
      }
    } else if (up) {
      uint64_t cy = 1; for (int w = 0; w < W && cy; w++) { u[w] += 1; cy = (u[w] == 0); }
    }
  }
  // mpi_smod(value, q, qh) with value = neg ? -u : u
  if (q.is_pow2) {
    if (neg) { uint64_t cy = 1; for (int w = 0; w < W; w++) { uint64_t t = ~u[w] + cy; cy = (cy && t == 0) ? 1u : 0u; u[w] = t; } }
    const int sw = (q.k - 1) >> 6, sb = (q.k - 1) & 63;
    if (q.k == 0) {
      for (int w = 0; w < W; w++) u[w] = ~(uint64_t)0;
    } else if (sw < W) {
      const uint64_t fill = ((u[sw] >> sb) & 1u) ? ~(uint64_t)0 : 0;
      const uint64_t keep = (sb == 63) ? ~(uint64_t)0 : ((((uint64_t)1) << (sb + 1)) - 1);
      u[sw] = (u[sw] & keep) | (fill & ~keep);
      for (int w = sw + 1; w < W; w++) u[w] = fill;
    }
  } else {
    gq_divrem(u, W, q.q, q.nwords, nullptr);
    bool zero = true;
    for (int w = 0; w < q.nwords; w++) zero = zero && (u[w] == 0);
    if (neg && !zero) {   // u = q - u
      uint64_t bw = 0;
      for (int w = 0; w < W; w++) {
        uint64_t a = (w < q.nwords) ? q.q[w] : 0, b = u[w];
        uint64_t d = a - b, b1 = a < b, d2 = d - bw, b2 = d < bw;
        u[w] = d2; bw = b1 | b2;
      }
    }
    // if u >= qh: u -= q
    bool ge = true;
    for (int w = W - 1; w >= 0; w--) {
      uint64_t h = (w < q.nwords) ? q.qh[w] : 0;
      if (u[w] != h) { ge = u[w] > h; break; }
    }
    if (ge) {
      uint64_t bw = 0;
      for (int w = 0; w < W; w++) {
        uint64_t a = u[w], b = (w < q.nwords) ? q.q[w] : 0;
        uint64_t d = a - b, b1 = a < b, d2 = d - bw, b2 = d < bw;
        u[w] = d2; bw = b1 | b2;
      }
    }
  }
  for (int w = 0; w < W; w++) x[(size_t)w * n] = u[w];
}

__global__ void __launch_bounds__(256)
k_automorph(int n, int W, GqPtrs4 out, GqPtrs4 in, uint64_t g) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t kk = ((uint64_t)i * g) & (uint64_t)(2 * n - 1);
  const bool neg = kk >= (uint64_t)n;
  const int dst = (int)(kk & (uint64_t)(n - 1));
  const uint64_t *src = in.p[blockIdx.y] + i;
  uint64_t *o = out.p[blockIdx.y] + dst;
  uint64_t carry = 1;
  for (int w = 0; w < W; w++) {
    uint64_t x = src[(size_t)w * n];
    if (neg) {
      x = ~x + carry;
      carry = (carry && x == 0) ? 1u : 0u;
    }
    o[(size_t)w * n] = x;
  }
}

// out = smod_q(a op b); op: 0 a+b, 1 a-b, 2 -a, 3 a (reduce only); q = 2^k.  Wrapping W-word
// arithmetic is exact here because 2^k divides 2^(64W).
__global__ void __launch_bounds__(256)
k_linear_pow2(int n, int W, GqPtrs4 out, GqPtrs4 a, GqPtrs4 b, int op0, int op1, unsigned k) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int op = blockIdx.y ? op1 : op0;
  const uint64_t *pa = a.p[blockIdx.y] + i, *pb = (op < 2) ? b.p[blockIdx.y] + i : nullptr;
  uint64_t *po = out.p[blockIdx.y] + i;
  if (k == 0) {   // q == 1: mpi_smod gives -1 (0 >= qh == 0, minus q)
    for (int w = 0; w < W; w++) po[(size_t)w * n] = ~(uint64_t)0;
    return;
  }
  const int sw = (int)((k - 1) >> 6), sb = (int)((k - 1) & 63);
  uint64_t cy = (op == 1 || op == 2) ? 1u : 0u, fill = 0;   // a + ~b + 1, or ~a + 1
  for (int w = 0; w < W; w++) {
    uint64_t val;
    if (w <= sw) {
      const uint64_t x = pa[(size_t)w * n];
      if (op == 3) {
        val = x;
      } else if (op == 2) {
        val = ~x + cy;
        cy = (cy && val == 0) ? 1u : 0u;
      } else {
        uint64_t y = pb[(size_t)w * n];
        if (op == 1) y = ~y;
        const uint64_t t = x + y;
        val = t + cy;
        cy = (t < x) | (val < t);
      }
      if (w == sw) {
        fill = ((val >> sb) & 1u) ? ~(uint64_t)0 : 0;
        const uint64_t keep = (sb == 63) ? ~(uint64_t)0 : ((((uint64_t)1) << (sb + 1)) - 1);
        val = (val & keep) | (fill & ~keep);
      }
    } else {
      val = fill;
    }
    po[(size_t)w * n] = val;
  }
}

// arbitrary q: exact (W+1)-word signed result, magnitude mod q, then the centring of mpi_smod
__global__ void __launch_bounds__(128)
k_linear_general(int n, int W, GqPtrs4 out, GqPtrs4 a, GqPtrs4 b, int op0, int op1, GqModulus q) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int op = blockIdx.y ? op1 : op0;
  const uint64_t *pa = a.p[blockIdx.y] + i, *pb = (op < 2) ? b.p[blockIdx.y] + i : nullptr;
  uint64_t *po = out.p[blockIdx.y] + i;
  uint64_t u[GQ_MAX_W + 1];
  const int WU = W + 1;
  {
    uint64_t cy = (op == 1 || op == 2) ? 1u : 0u;
    const uint64_t sa = (pa[(size_t)(W - 1) * n] >> 63) ? ~(uint64_t)0 : 0;
    const uint64_t sb = pb ? ((pb[(size_t)(W - 1) * n] >> 63) ? ~(uint64_t)0 : 0) : 0;
    for (int w = 0; w < WU; w++) {
      uint64_t x = (w < W) ? pa[(size_t)w * n] : sa;
      uint64_t y = pb ? ((w < W) ? pb[(size_t)w * n] : sb) : 0;
      if (op == 2) { y = ~x; x = 0; }
      if (op == 1) y = ~y;
      const uint64_t t = x + y;
      const uint64_t v = t + cy;
      cy = (t < x) | (v < t);
      u[w] = v;
    }
  }
  const bool neg = (u[WU - 1] >> 63) != 0;
  if (neg) { uint64_t cy = 1; for (int w = 0; w < WU; w++) { uint64_t t = ~u[w] + cy; cy = (cy && t == 0) ? 1u : 0u; u[w] = t; } }
  gq_divrem(u, WU, q.q, q.nwords, nullptr);
  bool zero = true;
  for (int w = 0; w < q.nwords; w++) zero = zero && (u[w] == 0);
  if (neg && !zero) {   // u = q - u
    uint64_t bw = 0;
    for (int w = 0; w < WU; w++) {
      uint64_t x = (w < q.nwords) ? q.q[w] : 0, y = u[w];
      uint64_t d = x - y, b1 = x < y, d2 = d - bw, b2 = d < bw;
      u[w] = d2; bw = b1 | b2;
    }
  }
  bool ge = true;   // u >= qh: u -= q
  for (int w = WU - 1; w >= 0; w--) {
    uint64_t h = (w < q.nwords) ? q.qh[w] : 0;
    if (u[w] != h) { ge = u[w] > h; break; }
  }
  if (ge) {
    uint64_t bw = 0;
    for (int w = 0; w < WU; w++) {
      uint64_t x = u[w], y = (w < q.nwords) ? q.q[w] : 0;
      uint64_t d = x - y, b1 = x < y, d2 = d - bw, b2 = d < bw;
      u[w] = d2; bw = b1 | b2;
    }
  }
  for (int w = 0; w < W; w++) po[(size_t)w * n] = u[w];
}

}  // namespace

void gq_launch_linear(int n, int W, const GqPtrs4 &out, const GqPtrs4 &a, const GqPtrs4 &b, int npolys,
                      int op0, int op1, const GqModulus &q, cudaStream_t st) {
  if (q.is_pow2)
    GQ_LAUNCH(k_linear_pow2, dim3((n + 255) / 256, npolys), dim3(256), st, n, W, out, a, b, op0, op1, (unsigned)q.k);
  else
    GQ_LAUNCH(k_linear_general, dim3((n + 127) / 128, npolys), dim3(128), st, n, W, out, a, b, op0, op1, q);
}

void gq_launch_rescale_pow2(int n, int W, const GqPtrs4 &c, int npolys, unsigned j, unsigned k,
                            cudaStream_t st) {
  GQ_LAUNCH(k_rescale_pow2, dim3((n + 255) / 256, npolys), dim3(256), st, n, W, c, j, k);
}

void gq_launch_rescale_general(int n, int W, const GqPtrs4 &c, int npolys, uint64_t Delta,
                               const GqModulus &q, cudaStream_t st) {
  GQ_LAUNCH(k_rescale_general, dim3((n + 127) / 128, npolys), dim3(128), st, n, W, c, Delta, q);
}

void gq_launch_automorph(int n, int W, const GqPtrs4 &out, const GqPtrs4 &in, int npolys,
                         uint64_t g, cudaStream_t st) {
  GQ_LAUNCH(k_automorph, dim3((n + 255) / 256, npolys), dim3(256), st, n, W, out, in, g);
}
