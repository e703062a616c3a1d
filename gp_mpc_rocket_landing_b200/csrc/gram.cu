// Gram-matrix build (SE-ARD / Matern-3/2 / Matern-5/2), feature extraction and the small predict epilogues.
//   reference: src/gp/kernels.py:205-262 (SE-ARD), :516-545 (Matern32), :610-638 (Matern52), :264-277 (diagonal),
//              :279-318 (SE-ARD hyper-gradients); src/gp/features.py:194-258, :305-354, :403-444.
// The kernel follows the reference operation order exactly:
//   Xs = X / l (division) ; s = numpy-pairwise row sums of Xs^2 ; d2 = max((s1 + s2) - 2 * <Xs1, Xs2>, 0)
// with the <.,.> cross term on FP64 DMMA (K = padded feature dim) and exp / Matern fused in the epilogue, so a
// Gram entry costs its 8-byte store and nothing else in HBM traffic.
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

// ------------------------------------------------------------------------------------------------ input scaling
__global__ void scale_inputs_kernel(const double* __restrict__ X, int n, int D, long ldx,
                                    const double* __restrict__ ls, double* __restrict__ Xs, int n_pad, int DP,
                                    double* __restrict__ sq) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  double sqv[64];
  double* out = Xs + (long)i * DP;
  if (i < n) {
    for (int d = 0; d < D; ++d) {
      double v = X[(long)i * ldx + d] / ls[d];
      out[d] = v;
      sqv[d] = v * v;
    }
    for (int d = D; d < DP; ++d) out[d] = 0.0;
    sq[i] = numpy_pairwise_sum_small(sqv, D);
  } else {
    for (int d = 0; d < DP; ++d) out[d] = 0.0;
    sq[i] = 0.0;
  }
}

extern "C" int gpmpc_scale_inputs(const double* X, int n, int D, long ldx, const double* lengthscales, double* Xs,
                                  int n_pad, int DP, double* sq, void* stream) {
  GPMPC_REQUIRE(D >= 1 && D <= 64 && DP >= D && DP % 4 == 0 && DP <= 64, "gpmpc_scale_inputs: need 1<=D<=DP<=64, DP%4==0");
  GPMPC_REQUIRE(n_pad >= n, "gpmpc_scale_inputs: n_pad < n");
  scale_inputs_kernel<<<(n_pad + 127) / 128, 128, 0, (cudaStream_t)stream>>>(X, n, D, ldx, lengthscales, Xs, n_pad, DP, sq);
  return gpmpc_check_launch("scale_inputs");
}

// ------------------------------------------------------------------------------------------------ kernel functions
// 64 x 128 output tile per CTA, 8 warps (2 x 4), warp tile 32 x 32 = 4 x 4 DMMA tiles, K = DP (<= 64).  (Was 128 x 128 with
// 64 values per thread: 172 registers, one CTA per SM, and the exp epilogue ran latency-bound at 8 warps per SM -- 203 us per
// 4096 x 4096 lower triangle; the smaller tile keeps two CTAs resident.)
constexpr int GRAM_BM = 64;
template <int MODE>  // 0: Gram ; 1: SE-ARD d/dlog(l_dim) = K * (x1_d - x2_d)^2 / l_d^2 with x/l already scaled
__global__ void __launch_bounds__(256, 2) gram_kernel(int kind, const double* __restrict__ Xs1,
                                                   const double* __restrict__ sq1, int n1_valid,
                                                   const double* __restrict__ Xs2, const double* __restrict__ sq2,
                                                   int n2_valid, int DP, double sv, double diag_add, int pad_identity,
                                                   int lower_only, int grad_dim, double* __restrict__ K, long ldk) {
  extern __shared__ double sm[];
  const int pitch = ((DP + 15) / 16) * 16 + 4;
  double* A = sm;                     // GRAM_BM x pitch
  double* B = sm + GRAM_BM * pitch;   // 128 x pitch
  const int m0 = blockIdx.y * GRAM_BM, n0 = blockIdx.x * 128;
  if (lower_only && n0 > m0 + GRAM_BM - 1) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int wm = warp >> 2, wn = warp & 3;
  for (int idx = tid; idx < 128 * DP; idx += 256) {
    int r = idx / DP, d = idx - r * DP;
    if (r < GRAM_BM) A[r * pitch + d] = Xs1[(long)(m0 + r) * DP + d];
    B[r * pitch + d] = Xs2[(long)(n0 + r) * DP + d];
  }
  __syncthreads();
  constexpr int MT = GRAM_BM / 16;
  double acc[MT][4][2];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  const double* as = A + (wm * (GRAM_BM / 2) + g) * pitch + t;
  const double* bs = B + (wn * 32 + g) * pitch + t;
  for (int kk = 0; kk < DP; kk += 4) {
    double af[MT], bf[4];
#pragma unroll
    for (int i = 0; i < MT; ++i) af[i] = as[i * 8 * pitch + kk];
#pragma unroll
    for (int j = 0; j < 4; ++j) bf[j] = bs[j * 8 * pitch + kk];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
  }
#pragma unroll
  for (int i = 0; i < MT; ++i) {
    const int ml = wm * (GRAM_BM / 2) + i * 8 + g, m = m0 + ml;
    const double s1 = sq1[m];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int nl = wn * 32 + j * 8 + 2 * t, n = n0 + nl;
      double out[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int nn = n + e;
        double v;
        if (m < n1_valid && nn < n2_valid) {
          const double d2 = (s1 + sq2[nn]) - 2.0 * acc[i][j][e];
          v = kernel_value(kind, d2, sv);
          if (MODE == 1) {
            const double df = A[ml * pitch + grad_dim] - B[(nl + e) * pitch + grad_dim];
            v = v * (df * df);
          }
          if (m == nn) v += diag_add;
        } else {
          v = (pad_identity && m == nn) ? 1.0 : 0.0;
        }
        out[e] = v;
      }
      *reinterpret_cast<double2*>(K + (long)m * ldk + n) = make_double2(out[0], out[1]);
    }
  }
}

static int launch_gram(int mode, int kind, const double* Xs1, const double* sq1, int n1_valid, int n1_pad,
                       const double* Xs2, const double* sq2, int n2_valid, int n2_pad, int DP, double sv,
                       double diag_add, int pad_identity, int lower_only, int grad_dim, double* K, long ldk,
                       cudaStream_t stream) {
  GPMPC_REQUIRE(n1_pad % 128 == 0 && n2_pad % 128 == 0, "gpmpc_gram: padded sizes must be multiples of 128");
  GPMPC_REQUIRE(DP % 4 == 0 && DP >= 4 && DP <= 64 && (ldk & 1) == 0, "gpmpc_gram: DP%4==0, 4<=DP<=64, even ldk");
  GPMPC_REQUIRE(kind >= 0 && kind <= 3, "gpmpc_gram: unknown kernel kind");
  if (n1_pad == 0 || n2_pad == 0) return GPMPC_OK;
  const int pitch = ((DP + 15) / 16) * 16 + 4;
  size_t smem = sizeof(double) * (GRAM_BM + 128) * pitch;
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(gram_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(gram_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    configured.here() = 1;
  }
  dim3 grid(n2_pad / 128, n1_pad / GRAM_BM);
  if (mode == 0)
    gram_kernel<0><<<grid, 256, smem, stream>>>(kind, Xs1, sq1, n1_valid, Xs2, sq2, n2_valid, DP, sv, diag_add,
                                                pad_identity, lower_only, grad_dim, K, ldk);
  else
    gram_kernel<1><<<grid, 256, smem, stream>>>(kind, Xs1, sq1, n1_valid, Xs2, sq2, n2_valid, DP, sv, diag_add,
                                                pad_identity, lower_only, grad_dim, K, ldk);
  return gpmpc_check_launch("gram");
}

extern "C" int gpmpc_gram(int kind, const double* Xs1, const double* sq1, int n1_valid, int n1_pad, const double* Xs2,
                          const double* sq2, int n2_valid, int n2_pad, int DP, double signal_var, double diag_add,
                          int pad_identity, int lower_only, double* K, long ldk, void* stream) {
  return launch_gram(0, kind, Xs1, sq1, n1_valid, n1_pad, Xs2, sq2, n2_valid, n2_pad, DP, signal_var, diag_add,
                     pad_identity, lower_only, 0, K, ldk, (cudaStream_t)stream);
}

extern "C" int gpmpc_gram_grad_lengthscale(const double* Xs1, const double* sq1, int n1_valid, int n1_pad,
                                           const double* Xs2, const double* sq2, int n2_valid, int n2_pad, int DP,
                                           double signal_var, int dim, double* G, long ldg, void* stream) {
  GPMPC_REQUIRE(dim >= 0 && dim < DP, "gpmpc_gram_grad_lengthscale: dim out of range");
  return launch_gram(1, GPMPC_KERNEL_SE, Xs1, sq1, n1_valid, n1_pad, Xs2, sq2, n2_valid, n2_pad, DP, signal_var, 0.0,
                     0, 0, dim, G, ldg, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------ predict epilogues
// mean[q,p] = (sum_j Ks[q,j] alpha[p,j]) * ystd[p] + ymean[p]   (exact_gp.py:240-245) -- one warp per query row.
__global__ void __launch_bounds__(256) predict_mean_kernel(const double* __restrict__ Ks, long ldk, int Q, int n,
                                                           const double* __restrict__ alpha, int n_pad, int P,
                                                           const double* __restrict__ ymean,
                                                           const double* __restrict__ ystd, double* __restrict__ mean,
                                                           long ldm, int col0) {
  const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= Q) return;
  const double* row = Ks + (long)q * ldk;
  for (int p0 = 0; p0 < P; p0 += 4) {
    double s[4] = {0, 0, 0, 0};
    for (int j = lane; j < n; j += 32) {
      const double k = row[j];
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (p0 + e < P) s[e] += k * alpha[(long)(p0 + e) * n_pad + j];
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      double v = warp_sum(s[e]);
      if (lane == 0 && p0 + e < P) mean[(long)q * ldm + col0 + p0 + e] = v * ystd[p0 + e] + ymean[p0 + e];
    }
  }
}

extern "C" int gpmpc_predict_mean(const double* Ks, long ldk, int Q, int n, const double* alpha, int n_pad, int P,
                                  const double* ymean, const double* ystd, double* mean, long ldm, int col0,
                                  void* stream) {
  if (Q == 0) return GPMPC_OK;
  predict_mean_kernel<<<(Q + 7) / 8, 256, 0, (cudaStream_t)stream>>>(Ks, ldk, Q, n, alpha, n_pad, P, ymean, ystd, mean,
                                                                    ldm, col0);
  return gpmpc_check_launch("predict_mean");
}

// var[q,p] = max(prior - sumsq[q] (+ plus[q]), floor) * ystd[p]^2   (exact_gp.py:259-263, sparse_gp.py:299-300,
// fast_gp.py:199-200 with floor 0 and ystd 1)
__global__ void predict_var_kernel(const double* __restrict__ sumsq, const double* __restrict__ plus, int Q, int P,
                                   double prior, double floor_, const double* __restrict__ ystd,
                                   double* __restrict__ var, long ldv, int col0) {
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  double v = prior - sumsq[q];
  if (plus) v += plus[q];
  v = fmax(v, floor_);
  for (int p = 0; p < P; ++p) {
    const double s = ystd ? ystd[p] : 1.0;
    var[(long)q * ldv + col0 + p] = v * (s * s);
  }
}

extern "C" int gpmpc_predict_var(const double* sumsq, const double* plus, int Q, int P, double prior, double floor_,
                                 const double* ystd, double* var, long ldv, int col0, void* stream) {
  if (Q == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(predict_var_kernel);
  predict_var_kernel<<<(Q + 255) / 256, 256, 0, (cudaStream_t)stream>>>(sumsq, plus, Q, P, prior, floor_, ystd, var, ldv, col0);
  return gpmpc_check_launch("predict_var");
}

// ------------------------------------------------------------------------------------------------ features
__global__ void features_kernel(int which, const double* __restrict__ X, long ldx, const double* __restrict__ U,
                                long ldu, int B, double* __restrict__ Z, long ldz) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const double* x = X + (long)i * ldx;
  const double* u = U + (long)i * ldu;
  double* z = Z + (long)i * ldz;
  if (which == GPMPC_FEATURES_3DOF) {
    double xl[7], ul[3], zl[11];
    for (int d = 0; d < 7; ++d) xl[d] = x[d];
    for (int d = 0; d < 3; ++d) ul[d] = u[d];
    features_3dof_dev(xl, ul, zl);
    for (int d = 0; d < 11; ++d) z[d] = zl[d];
    return;
  }
  double xl[14], ul[3], zl[13];
  for (int d = 0; d < 14; ++d) xl[d] = x[d];
  for (int d = 0; d < 3; ++d) ul[d] = u[d];
  const int trans = (which == GPMPC_FEATURES_6DOF_TRANS);
  features_6dof_dev(trans, xl, ul, zl);
  for (int d = 0; d < (trans ? 13 : 12); ++d) z[d] = zl[d];
}

extern "C" int gpmpc_features(int which, const double* X, long ldx, const double* U, long ldu, int B, double* Z,
                              long ldz, void* stream) {
  GPMPC_REQUIRE(which >= 0 && which <= 2, "gpmpc_features: unknown extractor");
  if (B == 0) return GPMPC_OK;
  features_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(which, X, ldx, U, ldu, B, Z, ldz);
  return gpmpc_check_launch("features");
}
