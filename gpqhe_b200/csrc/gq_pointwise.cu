// gq_pointwise.cu - K3: elementwise limb-poly kernels, 128-bit vectorised, grid-stride,
// grid sized in multiples of the SM count.  Bit-exact with poly_rns_mul / poly_rns_add
// (src/poly.c:71-82): same Barrett step, same canonical output.
//
// Batch layout: limb-poly m of a batch is at base + m*n, its prime is limb0 + m % nlimbs.
#include "gq_kernels.h"

namespace {

GQ_DEV ulonglong2 ld2(const uint64_t *p) { return *reinterpret_cast<const ulonglong2 *>(p); }
GQ_DEV void st2(uint64_t *p, ulonglong2 v) { *reinterpret_cast<ulonglong2 *>(p) = v; }

// r = a (*|+) b
template <bool MUL>
__global__ void __launch_bounds__(256)
k_binary(const GqPrime *__restrict__ primes, int n, int limb0, int nlimbs, int npolys_total,
         uint64_t *r, const uint64_t *a, const uint64_t *b) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long total = (long long)npolys_total * half;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(i / half);
    const GqPrime P = primes[limb0 + m % nlimbs];
    ulonglong2 x = ld2(a + 2 * i), y = ld2(b + 2 * i), z;
    if (MUL) {
      z.x = gq_mulmod_barrett(x.x, y.x, P);
      z.y = gq_mulmod_barrett(x.y, y.y, P);
    } else {
      z.x = gq_addmod_barrett(x.x, y.x, P);
      z.y = gq_addmod_barrett(x.y, y.y, P);
    }
    st2(r + 2 * i, z);
  }
}

// Tensor product in the NTT domain (he_mul hot loop A, src/he-mult.c:125-136):
//   in : x[4][D][n] = NTT(ct1.c0), NTT(ct1.c1), NTT(ct2.c0), NTT(ct2.c1)
//   out: y[3][D][n] = a0*b0, a0*b1 + a1*b0, a1*b1
// The reference adds the two cross terms after their inverse NTTs (:136); the inverse
// NTT is linear mod p and its outputs are canonical, so adding before it gives the same
// canonical limb-poly.
__global__ void __launch_bounds__(256)
k_tensor(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
         const uint64_t *__restrict__ x) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i / half);
    const GqPrime P = primes[d];
    ulonglong2 a0 = ld2(x + 2 * i), a1 = ld2(x + stride + 2 * i);
    ulonglong2 b0 = ld2(x + 2 * stride + 2 * i), b1 = ld2(x + 3 * stride + 2 * i);
    ulonglong2 d0, d1, d2;
    d0.x = gq_mulmod_barrett(a0.x, b0.x, P);
    d0.y = gq_mulmod_barrett(a0.y, b0.y, P);
    d2.x = gq_mulmod_barrett(a1.x, b1.x, P);
    d2.y = gq_mulmod_barrett(a1.y, b1.y, P);
    d1.x = gq_addmod_barrett(gq_mulmod_barrett(a0.x, b1.x, P), gq_mulmod_barrett(a1.x, b0.x, P), P);
    d1.y = gq_addmod_barrett(gq_mulmod_barrett(a0.y, b1.y, P), gq_mulmod_barrett(a1.y, b0.y, P), P);
    st2(y + 2 * i, d0);
    st2(y + stride + 2 * i, d1);
    st2(y + 2 * stride + 2 * i, d2);
  }
}

// Key-switch inner product (he_relin / he_swk hot loop C, src/he-mult.c:61-64):
//   y[0][D][n] = x * evk.p0,  y[1][D][n] = x * evk.p1     (x = NTT(decompose(d)))
// ymul2: same with a plaintext: y[k] = x[k] * m for two inputs (he_mulpt, :186-189).
__global__ void __launch_bounds__(256)
k_mul_evk(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
          const uint64_t *__restrict__ x, const uint64_t *__restrict__ e0,
          const uint64_t *__restrict__ e1) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const GqPrime P = primes[(int)(i / half)];
    ulonglong2 a = ld2(x + 2 * i), k0 = ld2(e0 + 2 * i), k1 = ld2(e1 + 2 * i), r0, r1;
    r0.x = gq_mulmod_barrett(a.x, k0.x, P);
    r0.y = gq_mulmod_barrett(a.y, k0.y, P);
    r1.x = gq_mulmod_barrett(a.x, k1.x, P);
    r1.y = gq_mulmod_barrett(a.y, k1.y, P);
    st2(y + 2 * i, r0);
    st2(y + stride + 2 * i, r1);
  }
}

__global__ void __launch_bounds__(256)
k_mul_pt(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
         const uint64_t *__restrict__ x) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  // x[3][D][n] = NTT(c0), NTT(c1), NTT(m);  y[2][D][n] = c0*m, c1*m
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const GqPrime P = primes[(int)(i / half)];
    ulonglong2 c0 = ld2(x + 2 * i), c1 = ld2(x + stride + 2 * i), mm = ld2(x + 2 * stride + 2 * i);
    ulonglong2 r0, r1;
    r0.x = gq_mulmod_barrett(c0.x, mm.x, P);
    r0.y = gq_mulmod_barrett(c0.y, mm.y, P);
    r1.x = gq_mulmod_barrett(c1.x, mm.x, P);
    r1.y = gq_mulmod_barrett(c1.y, mm.y, P);
    st2(y + 2 * i, r0);
    st2(y + stride + 2 * i, r1);
  }
}

// ---- the same three internal kernels for primes 2^s + c, c < 2^28 (gq_reduce_pm) -------------
typedef GqPM PM;
GQ_DEV PM load_pm(const GqPrime &P) { return gq_load_pm(P); }
GQ_DEV uint64_t mulm_pm(uint64_t a, uint64_t b, const PM &m) { return gq_mulm_pm(a, b, m); }
GQ_DEV uint64_t mul2add_pm(uint64_t a, uint64_t b, uint64_t c2, uint64_t d, const PM &m) { return gq_mul2add_pm(a, b, c2, d, m); }

__global__ void __launch_bounds__(256)
k_tensor_pm(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
            const uint64_t *__restrict__ x) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const PM m = load_pm(primes[(int)(i / half)]);
    const ulonglong2 a0 = ld2(x + 2 * i), a1 = ld2(x + stride + 2 * i);
    const ulonglong2 b0 = ld2(x + 2 * stride + 2 * i), b1 = ld2(x + 3 * stride + 2 * i);
    ulonglong2 d0, d1, d2;
    d0.x = mulm_pm(a0.x, b0.x, m); d0.y = mulm_pm(a0.y, b0.y, m);
    d2.x = mulm_pm(a1.x, b1.x, m); d2.y = mulm_pm(a1.y, b1.y, m);
    d1.x = mul2add_pm(a0.x, b1.x, a1.x, b0.x, m);
    d1.y = mul2add_pm(a0.y, b1.y, a1.y, b0.y, m);
    st2(y + 2 * i, d0);
    st2(y + stride + 2 * i, d1);
    st2(y + 2 * stride + 2 * i, d2);
  }
}

__global__ void __launch_bounds__(256)
k_mul_evk_pm(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
             const uint64_t *__restrict__ x, const uint64_t *__restrict__ e0,
             const uint64_t *__restrict__ e1) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const PM m = load_pm(primes[(int)(i / half)]);
    const ulonglong2 a = ld2(x + 2 * i), k0 = ld2(e0 + 2 * i), k1 = ld2(e1 + 2 * i);
    ulonglong2 r0, r1;
    r0.x = mulm_pm(a.x, k0.x, m); r0.y = mulm_pm(a.y, k0.y, m);
    r1.x = mulm_pm(a.x, k1.x, m); r1.y = mulm_pm(a.y, k1.y, m);
    st2(y + 2 * i, r0);
    st2(y + stride + 2 * i, r1);
  }
}

__global__ void __launch_bounds__(256)
k_mul_pt_pm(const GqPrime *__restrict__ primes, int n, int D, uint64_t *__restrict__ y,
            const uint64_t *__restrict__ x) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const int half = n >> 1;
  const long long per = (long long)D * half;
  const size_t stride = (size_t)D * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const PM m = load_pm(primes[(int)(i / half)]);
    const ulonglong2 c0 = ld2(x + 2 * i), c1 = ld2(x + stride + 2 * i), mm = ld2(x + 2 * stride + 2 * i);
    ulonglong2 r0, r1;
    r0.x = mulm_pm(c0.x, mm.x, m); r0.y = mulm_pm(c0.y, mm.y, m);
    r1.x = mulm_pm(c1.x, mm.x, m); r1.y = mulm_pm(c1.y, mm.y, m);
    st2(y + 2 * i, r0);
    st2(y + stride + 2 * i, r1);
  }
}

// Hoisted rotations: the key switch of sigma_g(d) from xhat = NTT(decompose(d)) computed ONCE.
// sigma_g commutes with the coefficient-wise floor-mod (it only moves and negates coefficients) and
// acts on the evaluation vector as a pure permutation: slot j of ntt()'s output holds the value at
// psi^(2*brv(j)+1), and (sigma_g a)(psi^e) = a(psi^(e*g)).  So
//   NTT(decompose(sigma_g d))[j] = xhat[pi(j)],  2*brv(pi(j))+1 = (2*brv(j)+1)*g mod 2n,
// exactly (canonical residues on both sides), and y_k[j] = xhat[pi(j)] * evk.p_k[j] as in k_mul_evk.
GQ_DEV unsigned gq_brev(unsigned x, int bits) { return __brev(x) >> (32 - bits); }

template <bool PMODE>
__global__ void __launch_bounds__(256)
k_mul_evk_perm(const GqPrime *__restrict__ primes, int n, int logn, int D, uint64_t *__restrict__ y,
               const uint64_t *__restrict__ x, const uint64_t *__restrict__ e0,
               const uint64_t *__restrict__ e1, unsigned g) {
  GQ_PDL_TRIGGER();
  GQ_PDL_WAIT();
  const long long per = (long long)D * n;
  const size_t stride = (size_t)D * n;
  const unsigned mask2n = 2u * (unsigned)n - 1u;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < per;
       i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i >> logn);
    const unsigned j = (unsigned)(i & (n - 1));
    const unsigned e = ((2u * gq_brev(j, logn) + 1u) * g) & mask2n;
    const unsigned src = gq_brev((e - 1u) >> 1, logn);
    const uint64_t a = x[(size_t)d * n + src], k0 = e0[i], k1 = e1[i];
    if (PMODE) {
      const PM m = load_pm(primes[d]);
      y[i] = mulm_pm(a, k0, m);
      y[stride + i] = mulm_pm(a, k1, m);
    } else {
      const GqPrime P = primes[d];
      y[i] = gq_mulmod_barrett(a, k0, P);
      y[stride + i] = gq_mulmod_barrett(a, k1, P);
    }
  }
}

inline int grid_for(long long work_items, int block) {
  long long blocks = (work_items + block - 1) / block;
  long long cap = (long long)GQ_NUM_SMS * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

void gq_launch_binary(const GqPrime *primes, int n, int limb0, int nlimbs, int npolys, bool mul,
                      uint64_t *r, const uint64_t *a, const uint64_t *b, cudaStream_t st) {
  const int total = nlimbs * npolys;
  const int grid = grid_for((long long)total * (n >> 1), 256);
  if (mul)
    GQ_LAUNCH(k_binary<true>, dim3(grid), dim3(256), st, primes, n, limb0, nlimbs, total, r, a, b);
  else
    GQ_LAUNCH(k_binary<false>, dim3(grid), dim3(256), st, primes, n, limb0, nlimbs, total, r, a, b);
}

void gq_launch_tensor(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                      cudaStream_t st, bool pm) {
  if (pm) { GQ_LAUNCH(k_tensor_pm, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x); return; }
  GQ_LAUNCH(k_tensor, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x);
}

void gq_launch_mul_evk(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                       const uint64_t *e0, const uint64_t *e1, cudaStream_t st, bool pm) {
  if (pm) { GQ_LAUNCH(k_mul_evk_pm, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x, e0, e1); return; }
  GQ_LAUNCH(k_mul_evk, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x,
            e0, e1);
}

void gq_launch_mul_evk_perm(const GqPrime *primes, int n, int logn, int D, uint64_t *y, const uint64_t *x,
                            const uint64_t *e0, const uint64_t *e1, uint64_t g, cudaStream_t st, bool pm) {
  const dim3 grid(grid_for((long long)D * n, 256));
  if (pm) GQ_LAUNCH(k_mul_evk_perm<true>, grid, dim3(256), st, primes, n, logn, D, y, x, e0, e1, (unsigned)g);
  else GQ_LAUNCH(k_mul_evk_perm<false>, grid, dim3(256), st, primes, n, logn, D, y, x, e0, e1, (unsigned)g);
}

void gq_launch_mul_pt(const GqPrime *primes, int n, int D, uint64_t *y, const uint64_t *x,
                      cudaStream_t st, bool pm) {
  if (pm) { GQ_LAUNCH(k_mul_pt_pm, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x); return; }
  GQ_LAUNCH(k_mul_pt, dim3(grid_for((long long)D * (n >> 1), 256)), dim3(256), st, primes, n, D, y, x);
}
