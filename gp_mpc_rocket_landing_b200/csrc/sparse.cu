// Inducing-point (sparse) GP glue kernels: the element-wise and reduction steps of SparseGP.fit that sit between the
// Gram / Cholesky / TRSM / GEMM kernels (reference src/gp/sparse_gp.py:150-253).  The heavy steps reuse gram.cu and
// linalg.cu:  K_uu, K_uf (gram_kernel) -> L_uu (potrf) -> A = L_uu^{-1} K_uf (trsm_small, fused q_ff = sum A^2)
//   FITC: fitc_scale (Lambda, A Lambda^{-1/2}, y / Lambda) -> B = I + As As' (dgemm_nt) -> L_B (potrf) -> alpha (potrs)
//   VFE : B = K_uu + K_uf K_fu / s2 (dgemm_nt) + jitter I (add_diag) -> L_B -> alpha
// and sparse_lml computes the FITC log marginal likelihood / VFE bound.  Compiled with -fmad=false so the element-wise
// expressions round like the numpy reference.
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

__global__ void add_diag_kernel(double* __restrict__ A, int n, long lda, double v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) A[(long)i * lda + i] += v;
}

// sparse_gp.py:194-203,208: Lambda = max(k_ff - q_ff + noise, floor); As = A * (1/sqrt(Lambda)); Ylam = y / Lambda.
// One thread per training point (column of A), coalesced across the row-major M x n_pad matrices.
__global__ void __launch_bounds__(128) fitc_scale_kernel(const double* __restrict__ A, long lda, int M, int n, int n_pad,
                                                         const double* __restrict__ qff, double prior, double noise,
                                                         double floor_, const double* __restrict__ Yn, int P,
                                                         double* __restrict__ Lam, double* __restrict__ As, long ldas,
                                                         double* __restrict__ Ylam) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  if (i >= n) {  // padding columns: neutral values
    Lam[i] = 1.0;
    for (int m = 0; m < M; ++m) As[(long)m * ldas + i] = 0.0;
    for (int p = 0; p < P; ++p) Ylam[(long)p * n_pad + i] = 0.0;
    return;
  }
  double lam = (prior - qff[i]) + noise;
  lam = fmax(lam, floor_);
  Lam[i] = lam;
  const double inv_sqrt = 1.0 / sqrt(lam);
  for (int m = 0; m < M; ++m) As[(long)m * ldas + i] = A[(long)m * lda + i] * inv_sqrt;
  for (int p = 0; p < P; ++p) Ylam[(long)p * n_pad + i] = Yn[(long)p * n_pad + i] / lam;
}

__device__ double block_sum_256(double v, double* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double r = 0.0;
  for (int w = 0; w < 8; ++w) r += red[w];
  return r;
}

// One CTA (256 threads) per output p.  method 0 = FITC (sparse_gp.py:212-218), 1 = VFE bound (:234-251).
__global__ void __launch_bounds__(256) sparse_lml_kernel(int method, int M, int n, int n_pad, int M_pad,
                                                         const double* __restrict__ Kuf, const double* __restrict__ Kuu,
                                                         const double* __restrict__ LB, const double* __restrict__ Luu,
                                                         const double* __restrict__ Yn, const double* __restrict__ alpha,
                                                         const double* __restrict__ c, const double* __restrict__ Lam,
                                                         const double* __restrict__ qff, double sv, double noise,
                                                         double* __restrict__ lml) {
  __shared__ double red[8];
  const int p = blockIdx.x, tid = threadIdx.x;
  const double* y = Yn + (long)p * n_pad;
  const double* a = alpha + (long)p * M_pad;
  const double kLog2Pi = 1.8378770664093453;  // np.log(2*np.pi)
  double logLB = 0.0, logLuu = 0.0;
  for (int m = tid; m < M; m += 256) {
    logLB += log(LB[(long)m * M_pad + m]);
    logLuu += log(Luu[(long)m * M_pad + m]);
  }
  logLB = block_sum_256(logLB, red);
  logLuu = block_sum_256(logLuu, red);
  if (method == 0) {
    const double* cp = c + (long)p * M_pad;
    double yly = 0.0, logLam = 0.0, ca = 0.0;
    for (int i = tid; i < n; i += 256) {
      yly += (y[i] * y[i]) / Lam[i];
      logLam += log(Lam[i]);
    }
    for (int m = tid; m < M; m += 256) ca += cp[m] * a[m];
    yly = block_sum_256(yly, red);
    logLam = block_sum_256(logLam, red);
    ca = block_sum_256(ca, red);
    if (tid == 0) {
      const double data_fit = -0.5 * (yly - ca);
      const double complexity = -logLB - 0.5 * logLam;
      lml[p] = (data_fit + complexity) + (-0.5 * n * kLog2Pi);
    }
  } else {
    const double s2 = noise;
    double yy = 0.0, yKa = 0.0, aKa = 0.0, sumA2 = 0.0;
    for (int i = tid; i < n; i += 256) {
      yy += y[i] * y[i];
      double s = 0.0;
      for (int m = 0; m < M; ++m) s += Kuf[(long)m * n_pad + i] * a[m];   // (K_uf' alpha)_i, coalesced over i
      yKa += y[i] * s;
      sumA2 += qff[i];
    }
    for (int m = tid; m < M; m += 256) {
      double s = 0.0;
      for (int k = 0; k < M; ++k) s += Kuu[(long)m * M_pad + k] * a[k];
      aKa += a[m] * s;
    }
    yy = block_sum_256(yy, red);
    yKa = block_sum_256(yKa, red);
    aKa = block_sum_256(aKa, red);
    sumA2 = block_sum_256(sumA2, red);
    if (tid == 0) {
      const double trace_term = (n * sv - sumA2) / s2;
      const double data_fit = ((-0.5 / s2) * yy + (1.0 / s2) * yKa) - 0.5 * aKa;
      const double complexity = (-logLB + logLuu) - (0.5 * n) * log(s2);
      lml[p] = ((data_fit - 0.5 * trace_term) + complexity) - (0.5 * n) * kLog2Pi;
    }
  }
}

}  // namespace

extern "C" int gpmpc_add_diag(double* A, int n, long lda, double value, void* stream) {
  GPMPC_REQUIRE(n >= 0, "gpmpc_add_diag: n < 0");
  if (n == 0) return GPMPC_OK;
  add_diag_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(A, n, lda, value);
  return gpmpc_check_launch("add_diag");
}

extern "C" int gpmpc_fitc_scale(const double* A, long lda, int M, int n, int n_pad, const double* qff, double prior,
                                double noise, double floor_, const double* Yn, int P, double* Lam, double* As, long ldas,
                                double* Ylam, void* stream) {
  GPMPC_REQUIRE(M > 0 && n > 0 && n_pad >= n && P >= 0, "gpmpc_fitc_scale: bad sizes");
  fitc_scale_kernel<<<(n_pad + 127) / 128, 128, 0, (cudaStream_t)stream>>>(A, lda, M, n, n_pad, qff, prior, noise, floor_,
                                                                          Yn, P, Lam, As, ldas, Ylam);
  return gpmpc_check_launch("fitc_scale");
}

extern "C" int gpmpc_sparse_lml(int method, int M, int n, int n_pad, int M_pad, const double* Kuf, const double* Kuu,
                                const double* LB, const double* Luu, const double* Yn, const double* alpha,
                                const double* c, const double* Lam, const double* qff, double sv, double noise, int P,
                                double* lml, void* stream) {
  GPMPC_REQUIRE(method == 0 || method == 1, "gpmpc_sparse_lml: method must be 0 (FITC) or 1 (VFE)");
  GPMPC_REQUIRE(P >= 1 && M >= 1 && n >= 1, "gpmpc_sparse_lml: bad sizes");
  sparse_lml_kernel<<<P, 256, 0, (cudaStream_t)stream>>>(method, M, n, n_pad, M_pad, Kuf, Kuu, LB, Luu, Yn, alpha, c, Lam,
                                                         qff, sv, noise, lml);
  return gpmpc_check_launch("sparse_lml");
}
