// Analytic gradient of the exact-GP log marginal likelihood w.r.t. the SE-ARD log hyper-parameters.
//
// The reference optimises hyper-parameters with L-BFGS-B on finite differences of the LML (exact_gp.py:357-421: one full
// refit per parameter per gradient).  The kernel derivatives it would need are stated in kernels.py:279-318
// (dK/dlog s2 = K, dK/dlog l_d = K (x_d - x'_d)^2 / l_d^2); with alpha = K^{-1} y and W = alpha alpha' - K^{-1}
//     dLML/dtheta = 1/2 sum_ij W_ij (dK/dtheta)_ij                         (Rasmussen & Williams eq. 5.9)
// so all D + 2 components come from ONE pass over the n x n pairs: the kernel value is recomputed from the scaled inputs
// (never stored), the lower triangle of K^{-1} is read once (n^2 / 2 doubles: the algorithmic HBM traffic of the pass).
#include "common.cuh"

namespace {

// grid (n_pad/128, n_pad/128, P), 128 threads: thread t owns row i of the tile and walks its 128 columns.
// part[((p * gy + by) * gx + bx) * (DP + 2) + c]: c = 0 signal variance, 1..D lengthscales, DP + 1 noise.
template <int DP>
__global__ void __launch_bounds__(128) lml_grad_se_kernel(const double* __restrict__ Xs, int n, double sv, double noise,
                                                          const double* __restrict__ Kinv, long ldk,
                                                          const double* __restrict__ alpha, long lda,
                                                          double* __restrict__ part) {
  __shared__ double xj[128 * DP];
  __shared__ double aj[128];
  __shared__ double red[4][DP + 2];
  const int t = threadIdx.x, p = blockIdx.z;
  const int i = blockIdx.y * 128 + t, j0 = blockIdx.x * 128;
  // W and dK/dtheta are symmetric: only the tiles with column tile >= row tile are visited (they read the LOWER triangle
  // of K^{-1}, entry (j, i) with j >= i tile-wise) and the off-diagonal ones count twice
  const long blk = ((long)p * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (blockIdx.x < blockIdx.y) {
    if (t < DP + 2) part[blk * (DP + 2) + t] = 0.0;
    return;
  }
  const double sym = (blockIdx.x > blockIdx.y) ? 2.0 : 1.0;
  for (int idx = t; idx < 128 * DP; idx += 128) xj[idx] = Xs[(long)j0 * DP + idx];
  aj[t] = alpha[(long)p * lda + j0 + t];
  double xi[DP], acc[DP + 2];
#pragma unroll
  for (int d = 0; d < DP; ++d) xi[d] = Xs[(long)i * DP + d];
#pragma unroll
  for (int d = 0; d < DP + 2; ++d) acc[d] = 0.0;
  const double ai = alpha[(long)p * lda + i];
  __syncthreads();
  if (i < n) {
    const int jn = min(128, n - j0);
    for (int c = 0; c < jn; ++c) {
      double d2[DP], s = 0.0;
#pragma unroll
      for (int d = 0; d < DP; ++d) {
        const double df = xi[d] - xj[c * DP + d];
        d2[d] = df * df;
        s += d2[d];
      }
      // K^{-1} is symmetric: read (j, i) so that consecutive threads read consecutive addresses
      const double w = sym * (ai * aj[c] - Kinv[(long)(j0 + c) * ldk + i]) * (sv * exp(-0.5 * s));
      acc[0] += w;
#pragma unroll
      for (int d = 0; d < DP; ++d) acc[1 + d] += w * d2[d];
    }
    if (blockIdx.x == blockIdx.y) acc[DP + 1] = noise * (ai * ai - Kinv[(long)i * ldk + i]);
  }
#pragma unroll
  for (int d = 0; d < DP + 2; ++d) {
    double v = acc[d];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((t & 31) == 0) red[t >> 5][d] = v;
  }
  __syncthreads();
  if (t < DP + 2) part[blk * (DP + 2) + t] = (red[0][t] + red[1][t]) + (red[2][t] + red[3][t]);
}

// grad[p * (D + 2) + c] = 1/2 * sum over tiles, in a fixed order (bit-reproducible)
__global__ void lml_grad_reduce_kernel(const double* __restrict__ part, int tiles, int DP, int D, double* __restrict__ grad) {
  const int p = blockIdx.x, c = threadIdx.x;
  if (c >= D + 2) return;
  const int src = (c == D + 1) ? DP + 1 : c;
  double s = 0.0;
  for (int b = 0; b < tiles; ++b) s += part[((long)p * tiles + b) * (DP + 2) + src];
  grad[p * (D + 2) + c] = 0.5 * s;
}

}  // namespace

extern "C" size_t gpmpc_lml_grad_se_work_doubles(int n_pad, int DP, int P) {
  const size_t g = (size_t)(n_pad / 128);
  return g * g * (size_t)P * (size_t)(DP + 2);
}

extern "C" int gpmpc_lml_grad_se(const double* Xs, int n, int n_pad, int DP, int D, double signal_var, double noise_var,
                                 const double* Kinv, long ldk, const double* alpha, long lda, int P, double* work,
                                 double* grad, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0 && n >= 1 && n <= n_pad, "gpmpc_lml_grad_se: n_pad must be a multiple of 128 and >= n");
  GPMPC_REQUIRE(DP % 4 == 0 && DP >= 4 && DP <= 32 && D >= 1 && D <= DP, "gpmpc_lml_grad_se: need DP % 4 == 0, D <= DP <= 32");
  GPMPC_REQUIRE(P >= 1 && P <= 65535, "gpmpc_lml_grad_se: bad number of outputs");
  const int g = n_pad / 128;
  dim3 grid(g, g, P);
#define GPMPC_LAUNCH_GRAD(dp) \
  case dp: lml_grad_se_kernel<dp><<<grid, 128, 0, stream>>>(Xs, n, signal_var, noise_var, Kinv, ldk, alpha, lda, work); break;
  switch (DP) {
    GPMPC_LAUNCH_GRAD(4) GPMPC_LAUNCH_GRAD(8) GPMPC_LAUNCH_GRAD(12) GPMPC_LAUNCH_GRAD(16)
    GPMPC_LAUNCH_GRAD(20) GPMPC_LAUNCH_GRAD(24) GPMPC_LAUNCH_GRAD(28) GPMPC_LAUNCH_GRAD(32)
  }
#undef GPMPC_LAUNCH_GRAD
  lml_grad_reduce_kernel<<<P, 64, 0, stream>>>(work, g * g, DP, D, grad);
  return gpmpc_check_launch("lml_grad_se");
}
