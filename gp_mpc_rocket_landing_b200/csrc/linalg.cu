// Batched FP64 dense linear algebra for the GP fit / predict path on sm_100a:
//   * blocked right-looking Cholesky (64-wide panels; panel TRSM and trailing SYRK on DMMA via dgemm.cuh)
//     -- replaces np.linalg.cholesky at reference exact_gp.py:164,170 / sparse_gp.py:187,205,228 / fast_gp.py:153
//   * cho_solve for a few right-hand sides (exact_gp.py:179, sparse_gp.py:210,232)
//   * blocked lower-triangular solve with MANY right-hand sides + fused column sum-of-squares
//     -- replaces solve_triangular + np.sum(v**2, axis=0) at exact_gp.py:260-261 (never materialises a separate v)
//   * small-M triangular solves for the sparse GP (sparse_gp.py:190,293,296)
// Layout: row-major, lower triangle, dimensions padded to a multiple of 128 with a unit diagonal in the padding.
#include <math.h>
#include <stdio.h>

#include <stdlib.h>
#include "dgemm.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

// ------------------------------------------------------------------------------------------------ error state
static thread_local char g_err[512] = "";
extern "C" void gpmpc_set_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }
extern "C" const char* gpmpc_last_error(void) { return g_err; }
int gpmpc_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return GPMPC_ERR_CUDA;
  }
  return GPMPC_OK;
}
extern "C" int gpmpc_version(void) { return 100; }

// ------------------------------------------------------------------------------------------------ potrf diag 64
// One CTA factors a 64x64 diagonal block in shared memory and also forms its inverse W = L11^{-1}; W turns the
// panel solve and every later triangular solve against this block into a tensor-core GEMM.
__global__ void __launch_bounds__(256) potrf_diag64_kernel(double* __restrict__ Abase, long lda, long strideA,
                                                           double* __restrict__ Wbase, long strideW,
                                                           int* __restrict__ info, int k0) {
  extern __shared__ double potrf_sm[];
  double(*S)[65] = reinterpret_cast<double(*)[65]>(potrf_sm);
  double(*Wi)[65] = reinterpret_cast<double(*)[65]>(potrf_sm + 64 * 65);
  const int b = blockIdx.x, tid = threadIdx.x;
  double* A = Abase + b * strideA + (long)k0 * lda + k0;
  double* W = Wbase + b * strideW;
  for (int idx = tid; idx < 4096; idx += 256) {
    int i = idx >> 6, k = idx & 63;
    S[i][k] = A[(long)i * lda + k];
    Wi[i][k] = 0.0;
  }
  __syncthreads();
  // Right-looking Cholesky, four threads per row; thread (i, q) keeps its 16 entries S[i][4 m + q] in REGISTERS.  Per column
  // j: the (unscaled) column is published to shared memory by its owners at the end of step j - 1, every thread reads the
  // pivot and the entries it needs from there, updates its registers -- (a_ij rd)(a_kj rd), the arithmetic and its order are
  // those of the shared-memory version this replaces (75 -> 65 us per block: the row updates no longer go through dependent
  // LDS / STS pairs; what remains is 64 x (barrier + DSQRT + DDIV) and the FP64 issue rate of eight warps) -- and the owners of column j + 1 publish it.  ONE barrier per column, double-buffered column.
  __shared__ double colbuf[2][64];
  {
    const int i = tid >> 2, q = tid & 3;
    double r[16];
#pragma unroll
    for (int m = 0; m < 16; ++m) r[m] = S[i][4 * m + q];
    if (q == 0) colbuf[0][i] = r[0];
    __syncthreads();
    // fully unrolled: every register index below is a compile-time constant (a run-time index sends r[] to local memory)
#pragma unroll
    for (int j = 0; j < 64; ++j) {
      const double* col = colbuf[j & 1];
      double d = col[j];
      if (!(d > 0.0)) {  // uniform: every thread reads the same value
        if (tid == 0 && info[b] == 0) info[b] = k0 + j + 1;  // LAPACK: order of the first non-PD leading minor
        d = 1.0;  // keep going on finite numbers; the host checks info and retries with jitter
      }
      d = sqrt(d);
      const double rd = 1.0 / d;
      if (i > j) {
        const double lij = col[i] * rd;
#pragma unroll
        for (int m = 0; m < 16; ++m) {
          const int k = 4 * m + q;
          if (k > j && k <= i) r[m] -= lij * (col[k] * rd);
        }
        if (q == (j & 3)) {
#pragma unroll
          for (int m = 0; m < 16; ++m)
            if (m == (j >> 2)) r[m] = lij;
        }
      } else if (i == j && q == (j & 3)) {
#pragma unroll
        for (int m = 0; m < 16; ++m)
          if (m == (j >> 2)) r[m] = d;
      }
      if (j + 1 < 64 && q == ((j + 1) & 3)) {   // column j + 1 is final: publish it for the next step
        double v = 0.0;
#pragma unroll
        for (int m = 0; m < 16; ++m)
          if (m == ((j + 1) >> 2)) v = r[m];
        colbuf[(j + 1) & 1][i] = v;
      }
      __syncthreads();
    }
#pragma unroll
    for (int m = 0; m < 16; ++m) S[i][4 * m + q] = r[m];
    __syncthreads();
  }
  // W = L^{-1}: four threads per column j split every dot product (k = q, q+4, ...) and combine it by two shuffles; the
  // column of W lives in registers (w[m] = W[4 m + q][j]); after the butterfly all four lanes hold the same sum, so each
  // computes the new entry and the lane that owns row i keeps it.
  {
    const int j = tid >> 2, q = tid & 3;
    double w[16];
#pragma unroll
    for (int m = 0; m < 16; ++m) w[m] = 0.0;
#pragma unroll
    for (int i = 0; i < 64; ++i) {   // uniform trip count: the shuffles below need the whole warp; unrolled as above
      double sacc = 0.0;
#pragma unroll
      for (int m = 0; m < 16; ++m) {
        const int k = 4 * m + q;
        if (k >= j && k < i) sacc = fma(S[i][k], w[m], sacc);
      }
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
      if (i >= j && q == (i & 3)) {
        const double val = (i == j) ? 1.0 / S[j][j] : -sacc / S[i][i];
#pragma unroll
        for (int m = 0; m < 16; ++m)
          if (m == (i >> 2)) w[m] = val;
      }
    }
#pragma unroll
    for (int m = 0; m < 16; ++m) Wi[4 * m + q][j] = w[m];
  }
  __syncthreads();
  for (int idx = tid; idx < 4096; idx += 256) {
    int i = idx >> 6, k = idx & 63;
    A[(long)i * lda + k] = (k <= i) ? S[i][k] : 0.0;
    W[idx] = Wi[i][k];
  }
}

// W = L_bb^{-1} for 64x64 diagonal blocks of an already factored L (the second half of potrf_diag64_kernel on its own):
// used after rows were appended to a factor, when only the diagonal blocks that received new rows change.
__global__ void __launch_bounds__(256) trinv64_kernel(const double* __restrict__ L, long lda, double* __restrict__ Wbase,
                                                      int b0) {
  extern __shared__ double trinv_sm[];
  double(*S)[65] = reinterpret_cast<double(*)[65]>(trinv_sm);
  double(*Wi)[65] = reinterpret_cast<double(*)[65]>(trinv_sm + 64 * 65);
  const int blk = b0 + blockIdx.x, tid = threadIdx.x;
  const double* A = L + (long)blk * 64 * lda + (long)blk * 64;
  for (int idx = tid; idx < 4096; idx += 256) {
    int i = idx >> 6, k = idx & 63;
    S[i][k] = A[(long)i * lda + k];
    Wi[i][k] = 0.0;
  }
  __syncthreads();
  {
    const int j = tid >> 2, q = tid & 3;
    for (int i = 0; i < 64; ++i) {
      double sacc = 0.0;
      for (int k = j + ((q - j) & 3); k < i; k += 4) sacc = fma(S[i][k], Wi[k][j], sacc);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
      if (q == 0 && i >= j) Wi[i][j] = (i == j) ? 1.0 / S[j][j] : -sacc / S[i][i];
      __syncwarp();
    }
  }
  __syncthreads();
  double* W = Wbase + (long)blk * 4096;
  for (int idx = tid; idx < 4096; idx += 256) W[idx] = Wi[idx >> 6][idx & 63];
}

extern "C" int gpmpc_trinv64_blocks(const double* L, int n_pad, long lda, double* Winv, int block_first, int block_last,
                                    void* stream) {
  GPMPC_REQUIRE(n_pad % 64 == 0 && block_first >= 0 && block_last < n_pad / 64 && block_first <= block_last,
                "gpmpc_trinv64_blocks: bad block range");
  const size_t smem = sizeof(double) * 2 * 64 * 65;
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(trinv64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured.here() = 1;
  }
  trinv64_kernel<<<block_last - block_first + 1, 256, smem, (cudaStream_t)stream>>>(L, lda, Winv, block_first);
  return gpmpc_check_launch("trinv64");
}

__global__ void zero_upper_kernel(double* __restrict__ A, int n, long lda, long strideA) {
  double* a = A + blockIdx.z * strideA;
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  int i = blockIdx.y;
  if (j < n && j > i) a[(long)i * lda + j] = 0.0;
}

extern "C" int gpmpc_potrf(double* A, int n_pad, long lda, long strideA, int batch, double* Winv, int* info,
                           void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad > 0 && n_pad % 128 == 0, "gpmpc_potrf: n_pad must be a positive multiple of 128");
  GPMPC_REQUIRE((lda & 1) == 0 && lda >= n_pad, "gpmpc_potrf: lda must be even and >= n_pad");
  const long strideW = (long)(n_pad / 64) * 4096;
  cudaMemsetAsync(info, 0, sizeof(int) * batch, stream);
  const size_t diag_smem = sizeof(double) * 2 * 64 * 65;
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(potrf_diag64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)diag_smem);
    configured.here() = 1;
  }
  // Two-level blocking: 64-wide right-looking steps INSIDE a 256-wide outer panel (their rank-64 updates touch the panel
  // columns only), then ONE rank-256 update of the trailing matrix per outer panel.  The trailing matrix is streamed
  // through HBM n/256 times instead of n/64 times and its GEMM runs a 16-step k-loop instead of 4.
  constexpr int NB = 512;   // measured at n = 8000, 16 matrices: NB 256 / 512 / 1024 -> 129.8 / 123.6 / 126.6 ms
  for (int K0 = 0; K0 < n_pad; K0 += NB) {
    const int Kend = (K0 + NB < n_pad) ? K0 + NB : n_pad;
    for (int k0 = K0; k0 < Kend; k0 += 64) {
      double* W = Winv + (long)(k0 / 64) * 4096;
      potrf_diag64_kernel<<<batch, 256, diag_smem, stream>>>(A, lda, strideA, W, strideW, info, k0);
      int rc = gpmpc_check_launch("potrf_diag64");
      if (rc) return rc;
      if (k0 + 64 >= n_pad) break;
      const int o = ((k0 + 64) / 128) * 128;  // 128-aligned tile origin; rows/cols below k0+64 are masked
      GemmParams p{};
      p.A = A + (long)o * lda + k0; p.lda = lda; p.strideA = strideA;
      p.B = W; p.ldb = 64; p.strideB = strideW;
      p.C = A + (long)o * lda + k0; p.ldc = lda; p.strideC = strideA;
      p.M = n_pad - o; p.N = 64; p.K = 64; p.alpha = 1.0; p.beta = 0.0; p.epi = EPI_STORE;
      p.row_min = k0 + 64 - o;
      rc = launch_dgemm_nt<128, 64, 4, 2>(p, batch, stream);  // L21 = A21 * W^T  (in place, all rows below)
      if (rc) return rc;
      if (k0 + 64 >= Kend) break;
      GemmParams s{};                                          // panel columns [k0+64, Kend): rank-64 update
      s.A = A + (long)o * lda + k0; s.lda = lda; s.strideA = strideA;
      s.B = s.A; s.ldb = lda; s.strideB = strideA;
      s.C = A + (long)o * lda + o; s.ldc = lda; s.strideC = strideA;
      s.M = n_pad - o; s.N = Kend - o; s.K = 64; s.alpha = -1.0; s.beta = 1.0; s.epi = EPI_STORE;
      s.lower_only = 1; s.row_min = k0 + 64 - o; s.col_min = k0 + 64 - o;
      rc = launch_dgemm_nt<64, 128, 2, 4>(s, batch, stream);
      if (rc) return rc;
    }
    if (Kend >= n_pad) break;
    GemmParams t{};                                            // trailing matrix: A22 -= L21 * L21^T, K = 256
    t.A = A + (long)Kend * lda + K0; t.lda = lda; t.strideA = strideA;
    t.B = t.A; t.ldb = lda; t.strideB = strideA;
    t.C = A + (long)Kend * lda + Kend; t.ldc = lda; t.strideC = strideA;
    t.M = n_pad - Kend; t.N = n_pad - Kend; t.K = Kend - K0; t.alpha = -1.0; t.beta = 1.0; t.epi = EPI_STORE;
    t.lower_only = 1;
    // 64 x 128 tiles, two CTAs per SM: 9 % faster than 128 x 128 with one (refit sweep 16 x n = 4000: 27.8 -> 25.5 ms,
    // n = 8000: 139.1 -> 125.7 ms)
    int rc = launch_dgemm_nt<64, 128, 2, 4>(t, batch, stream);
    if (rc) return rc;
  }
  dim3 grid((n_pad + 255) / 256, n_pad, batch);
  zero_upper_kernel<<<grid, 256, 0, stream>>>(A, n_pad, lda, strideA);
  return gpmpc_check_launch("zero_upper");
}

// ------------------------------------------------------------------------------------------------ potrs (few RHS)
// One CTA per (factor, rhs): y = L^{-1} b by 64-blocks (GEMV against the already-solved part, then the 64x64
// inverse block), then a = L^{-T} y.  The solution vector lives in shared memory.
__global__ void __launch_bounds__(256) potrs_kernel(const double* __restrict__ Lbase, int n_pad, long lda,
                                                    long strideL, const double* __restrict__ Wbase,
                                                    double* __restrict__ Bbase, int nrhs) {
  extern __shared__ double sh[];
  double* y = sh;               // n_pad
  double* r = sh + n_pad;       // 64
  double* part = r + 64;        // 4*64
  const int f = blockIdx.x / nrhs;
  const double* L = Lbase + f * strideL;
  const double* Wf = Wbase + (long)f * (n_pad / 64) * 4096;
  double* b = Bbase + (long)blockIdx.x * n_pad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nb = n_pad / 64;
  for (int i = tid; i < n_pad; i += 256) y[i] = b[i];
  __syncthreads();
  // forward
  for (int ib = 0; ib < nb; ++ib) {
    const int i0 = ib * 64;
    for (int rr = warp; rr < 64; rr += 8) {
      const double* row = L + (long)(i0 + rr) * lda;
      double s = 0.0;
      for (int k = lane; k < i0; k += 32) s += row[k] * y[k];
      s = warp_sum(s);
      if (lane == 0) r[rr] = y[i0 + rr] - s;
    }
    __syncthreads();
    if (tid < 64) {
      const double* w = Wf + (long)ib * 4096 + tid * 64;
      double s = 0.0;
      for (int k = 0; k <= tid; ++k) s += w[k] * r[k];
      y[i0 + tid] = s;
    }
    __syncthreads();
  }
  // backward: a_i = W_i^T (y_i - sum_{k > block i} L[k, block i]^T a_k)
  for (int ib = nb - 1; ib >= 0; --ib) {
    const int i0 = ib * 64;
    const int c = tid & 63, kp = tid >> 6;
    double s = 0.0;
    for (int k = i0 + 64 + kp; k < n_pad; k += 4) s += L[(long)k * lda + i0 + c] * y[k];
    part[kp * 64 + c] = s;
    __syncthreads();
    if (tid < 64) r[tid] = y[i0 + tid] - (part[tid] + part[64 + tid] + part[128 + tid] + part[192 + tid]);
    __syncthreads();
    if (tid < 64) {
      const double* w = Wf + (long)ib * 4096;
      double s2 = 0.0;
      for (int k = tid; k < 64; ++k) s2 += w[k * 64 + tid] * r[k];
      y[i0 + tid] = s2;
    }
    __syncthreads();
  }
  for (int i = tid; i < n_pad; i += 256) b[i] = y[i];
}

// Large n, few right-hand sides: the single-CTA solve above streams the whole factor through one SM (42 ms at
// n = 8000).  Blocked right-looking substitution instead, one launch per 64-block: every CTA recomputes the small
// solved block (64 x 64 product with the cached inverse), CTA 0 stores it, and the CTAs together subtract its
// contribution from the not-yet-solved part -- a tall GEMV spread over all SMs.  The launch boundary is the dependency.
__global__ void __launch_bounds__(256) potrs_fwd_step_kernel(const double* __restrict__ Lbase, int n_pad, long lda,
                                                             long strideL, const double* __restrict__ Wbase,
                                                             double* __restrict__ Bbase, double* __restrict__ Ybase,
                                                             int nrhs, int j0) {
  __shared__ double bj[64], yj[64], part[4][64];
  const int f = blockIdx.y / nrhs;
  const double* L = Lbase + f * strideL;
  const double* W = Wbase + ((long)f * (n_pad / 64) + j0 / 64) * 4096;
  double* b = Bbase + (long)blockIdx.y * n_pad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid < 64) bj[tid] = b[j0 + tid];
  __syncthreads();
  {
    const int i = tid & 63, q = tid >> 6;
    double s = 0.0;
    for (int k = q; k <= i; k += 4) s += W[i * 64 + k] * bj[k];
    part[q][i] = s;
  }
  __syncthreads();
  if (tid < 64) {
    const double v = (part[0][tid] + part[1][tid]) + (part[2][tid] + part[3][tid]);
    yj[tid] = v;
    if (blockIdx.x == 0) Ybase[(long)blockIdx.y * n_pad + j0 + tid] = v;
  }
  __syncthreads();
  const double y0 = yj[2 * lane], y1 = yj[2 * lane + 1];
  // 64 rows per CTA (n_pad - j0 - 64 is a multiple of 64), 8 per warp, 4 in flight: loads and reductions overlap
  if (j0 + 64 + ((int)blockIdx.x + 1) * 64 > n_pad) return;   // last block: nothing left to update
  const int r0 = j0 + 64 + blockIdx.x * 64 + warp * 8;
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    double sv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const double2 l2 = *reinterpret_cast<const double2*>(L + (long)(r0 + g * 4 + u) * lda + j0 + 2 * lane);
      sv[u] = l2.x * y0 + l2.y * y1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < 4; ++u) sv[u] += __shfl_xor_sync(0xffffffffu, sv[u], o);
    }
    if (lane < 4) {
      const double v = lane == 0 ? sv[0] : (lane == 1 ? sv[1] : (lane == 2 ? sv[2] : sv[3]));
      b[r0 + g * 4 + lane] -= v;
    }
  }
}

__global__ void __launch_bounds__(256) potrs_bwd_step_kernel(const double* __restrict__ Lbase, int n_pad, long lda,
                                                             long strideL, const double* __restrict__ Wbase,
                                                             double* __restrict__ Ybase, double* __restrict__ Xbase,
                                                             int nrhs, int j0) {
  __shared__ double rj[64], xj[64], part[4][64];
  const int f = blockIdx.y / nrhs;
  const double* L = Lbase + f * strideL;
  const double* W = Wbase + ((long)f * (n_pad / 64) + j0 / 64) * 4096;
  double* y = Ybase + (long)blockIdx.y * n_pad;
  const int tid = threadIdx.x;
  if (tid < 64) rj[tid] = y[j0 + tid];
  __syncthreads();
  {
    const int c = tid & 63, q = tid >> 6;
    double s = 0.0;
    for (int k = c + q; k < 64; k += 4) s += W[k * 64 + c] * rj[k];   // x_j = W_j' r_j
    part[q][c] = s;
  }
  __syncthreads();
  if (tid < 64) {
    const double v = (part[0][tid] + part[1][tid]) + (part[2][tid] + part[3][tid]);
    xj[tid] = v;
    if (blockIdx.x == 0) Xbase[(long)blockIdx.y * n_pad + j0 + tid] = v;
  }
  __syncthreads();
  // 64 columns per CTA, the 64 rows of the block split over 4 thread groups (16 rows each, coalesced across columns)
  const int cl = tid & 63, q = tid >> 6;
  const int c = blockIdx.x * 64 + cl;
  double s0 = 0.0, s1 = 0.0;
  if (c < j0) {
    const double* col = L + (long)(j0 + q * 16) * lda + c;
#pragma unroll
    for (int r = 0; r < 16; r += 2) {
      s0 += col[(long)r * lda] * xj[q * 16 + r];
      s1 += col[(long)(r + 1) * lda] * xj[q * 16 + r + 1];
    }
  }
  __syncthreads();            // part[] was read above to form xj
  part[q][cl] = s0 + s1;
  __syncthreads();
  if (tid < 64 && c < j0) y[c] -= (part[0][tid] + part[1][tid]) + (part[2][tid] + part[3][tid]);
}

static int potrs_blocked_min() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GPMPC_POTRS_BLOCKED_MIN");   // measurement knob; 0 disables the blocked path
    v = e ? atoi(e) : 2048;
    if (v <= 0) v = 1 << 30;
  }
  return v;
}

static int potrs_blocked(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* B, int nrhs,
                         int batch, cudaStream_t stream) {
  double* Y = nullptr;
  if (cudaMallocAsync(&Y, sizeof(double) * (size_t)batch * nrhs * n_pad, stream) != cudaSuccess) {
    cudaGetLastError();
    gpmpc_set_error("gpmpc_potrs: scratch allocation failed");
    return GPMPC_ERR_CUDA;
  }
  const int nv = batch * nrhs;
  for (int j0 = 0; j0 < n_pad; j0 += 64) {
    const int rest = n_pad - j0 - 64;
    dim3 grid(rest > 0 ? rest / 64 : 1, nv);
    potrs_fwd_step_kernel<<<grid, 256, 0, stream>>>(L, n_pad, lda, strideL, Winv, B, Y, nrhs, j0);
  }
  for (int j0 = n_pad - 64; j0 >= 0; j0 -= 64) {
    dim3 grid(j0 > 0 ? j0 / 64 : 1, nv);
    potrs_bwd_step_kernel<<<grid, 256, 0, stream>>>(L, n_pad, lda, strideL, Winv, Y, B, nrhs, j0);
  }
  cudaFreeAsync(Y, stream);
  return gpmpc_check_launch("potrs_blocked");
}

extern "C" int gpmpc_potrs(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* B,
                           int nrhs, int batch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0 && n_pad > 0, "gpmpc_potrs: n_pad must be a multiple of 128");
  if (n_pad >= potrs_blocked_min() && (long)batch * nrhs <= 64 && (lda & 1) == 0)
    return potrs_blocked(L, n_pad, lda, strideL, Winv, B, nrhs, batch, stream);
  size_t smem = sizeof(double) * (n_pad + 64 + 256);
  GPMPC_REQUIRE(smem <= 200 * 1024, "gpmpc_potrs: n_pad too large for the shared-memory solve");
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(potrs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    configured.here() = 1;
  }
  potrs_kernel<<<batch * nrhs, 256, smem, stream>>>(L, n_pad, lda, strideL, Winv, B, nrhs);
  return gpmpc_check_launch("potrs");
}

// ------------------------------------------------------------------------------------------------ targets / LML
// exact_gp.py:141-146: y_mean = mean(y), y_std = std(y) (population), y_std<1e-10 -> 1; Yn[p][i] = (y - mean)/std.
__global__ void __launch_bounds__(256) normalize_targets_kernel(const double* __restrict__ Y, int n, int P, long ldy,
                                                                int n_pad, int normalize, double* __restrict__ Yn,
                                                                double* __restrict__ ymean, double* __restrict__ ystd) {
  __shared__ double red[8];
  __shared__ double s_val;
  const int p = blockIdx.x, tid = threadIdx.x;
  auto block_sum = [&](double v) {
    v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    if (tid == 0) {
      double t = 0.0;
      for (int w = 0; w < 8; ++w) t += red[w];
      s_val = t;
    }
    __syncthreads();
    return s_val;
  };
  double mean = 0.0, sd = 1.0;
  if (normalize) {
    double s = 0.0;
    for (int i = tid; i < n; i += 256) s += Y[(long)i * ldy + p];
    mean = block_sum(s) / n;
    double q = 0.0;
    for (int i = tid; i < n; i += 256) {
      double d = Y[(long)i * ldy + p] - mean;
      q += d * d;
    }
    sd = sqrt(block_sum(q) / n);
    if (sd < 1e-10) sd = 1.0;
  }
  for (int i = tid; i < n_pad; i += 256)
    Yn[(long)p * n_pad + i] = (i < n) ? (Y[(long)i * ldy + p] - mean) / sd : 0.0;
  if (tid == 0) {
    ymean[p] = mean;
    ystd[p] = sd;
  }
}

extern "C" int gpmpc_normalize_targets(const double* Y, int n, int P, long ldy, int n_pad, int normalize, double* Yn,
                                       double* ymean, double* ystd, void* stream) {
  normalize_targets_kernel<<<P, 256, 0, (cudaStream_t)stream>>>(Y, n, P, ldy, n_pad, normalize, Yn, ymean, ystd);
  return gpmpc_check_launch("normalize_targets");
}

// exact_gp.py:186-204: lml = -0.5 y^T alpha - sum log diag(L) - n/2 log(2 pi), one CTA per (factor, rhs).
__global__ void __launch_bounds__(256) lml_kernel(const double* __restrict__ Lbase, int n, int n_pad, long lda,
                                                  long strideL, const double* __restrict__ Yn,
                                                  const double* __restrict__ alpha, int nrhs,
                                                  double* __restrict__ lml) {
  __shared__ double red[8];
  const int f = blockIdx.x / nrhs, tid = threadIdx.x;
  const double* L = Lbase + f * strideL;
  const double* y = Yn + (long)blockIdx.x * n_pad;
  const double* a = alpha + (long)blockIdx.x * n_pad;
  double s = 0.0;
  for (int i = tid; i < n; i += 256) s += -0.5 * y[i] * a[i] - log(L[(long)i * lda + i]);
  s = warp_sum(s);
  if ((tid & 31) == 0) red[tid >> 5] = s;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    lml[blockIdx.x] = t - 0.5 * n * log(2.0 * M_PI);
  }
}

extern "C" int gpmpc_lml(const double* L, int n, int n_pad, long lda, long strideL, const double* Yn,
                         const double* alpha, int nrhs, int batch, double* lml, void* stream) {
  lml_kernel<<<batch * nrhs, 256, 0, (cudaStream_t)stream>>>(L, n, n_pad, lda, strideL, Yn, alpha, nrhs, lml);
  return gpmpc_check_launch("lml");
}

// ------------------------------------------------------------------------------------------------ TRSM, many RHS
// V = L^{-1} Ks^T for Q right-hand sides stored as rows of Ks (Q_pad x n_pad, row-major) -- solved IN PLACE, block
// row by block row (64 wide):   R_i = Ks_i - L[i, :i] V[:i]   (DMMA GEMM, K = 64 i)
//                               V_i = W_i R_i                (DMMA GEMM, K = 64) + sumsq[q] += sum_rows V_i^2
// Same arithmetic as a blocked dtrsm; the fused column sum of squares is exact_gp.py:261.
extern "C" int gpmpc_trsm_sumsq_batched(const double* L, int n_pad, long lda, long strideL, const double* Winv,
                                        double* Ks, int q_pad, long ldk, long strideK, double* sumsq, int batch,
                                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0 && q_pad % 128 == 0, "gpmpc_trsm_sumsq: n_pad and q_pad must be multiples of 128");
  GPMPC_REQUIRE(batch >= 1, "gpmpc_trsm_sumsq: batch < 1");
  const long strideW = (long)(n_pad / 64) * 4096;
  cudaMemsetAsync(sumsq, 0, sizeof(double) * (size_t)q_pad * batch, stream);
  for (int ib = 0; ib < n_pad / 64; ++ib) {
    const int i0 = ib * 64;
    if (ib > 0) {
      GemmParams p{};
      p.A = L + (long)i0 * lda; p.lda = lda; p.strideA = strideL;
      p.B = Ks; p.ldb = ldk; p.strideB = strideK;
      p.C = Ks + i0; p.ldc = ldk; p.strideC = strideK; p.c_trans = 1;
      p.M = 64; p.N = q_pad; p.K = i0; p.alpha = -1.0; p.beta = 1.0; p.epi = EPI_STORE;
      int rc = launch_dgemm_nt<64, 128, 2, 4>(p, batch, stream);
      if (rc) return rc;
    }
    GemmParams s{};
    s.A = Winv + (long)ib * 4096; s.lda = 64; s.strideA = strideW;
    s.B = Ks + i0; s.ldb = ldk; s.strideB = strideK;
    s.C = Ks + i0; s.ldc = ldk; s.strideC = strideK; s.c_trans = 1;
    s.M = 64; s.N = q_pad; s.K = 64; s.alpha = 1.0; s.beta = 0.0; s.epi = EPI_STORE_SUMSQ;
    s.sumsq = sumsq; s.strideS = q_pad;
    int rc = launch_dgemm_nt<64, 128, 2, 4>(s, batch, stream);
    if (rc) return rc;
  }
  return GPMPC_OK;
}

extern "C" int gpmpc_trsm_sumsq(const double* L, int n_pad, long lda, const double* Winv, double* Ks, int q_pad,
                                long ldk, double* sumsq, void* stream_) {
  return gpmpc_trsm_sumsq_batched(L, n_pad, lda, 0, Winv, Ks, q_pad, ldk, 0, sumsq, 1, stream_);
}

// ------------------------------------------------------------------------------- explicit inverse factor, many RHS
// The horizon pipeline evaluates the predictive variance (exact_gp.py:259-264) for B*N queries against ONE factor per
// control step.  With W = L^{-1} formed once per fit, sum_i v_i^2 = |W k*|^2 is a single triangular DMMA GEMM launch
// (same n^2 flops per query as the blocked solve above, no 2(n/64)-1 dependent launches, V never written back).
__global__ void set_identity_kernel(double* __restrict__ A, int n, long lda) {
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long)n * n) return;
  const int r = (int)(idx / n), c = (int)(idx % n);
  A[(long)r * lda + c] = (r == c) ? 1.0 : 0.0;
}
__global__ void __launch_bounds__(256) transpose_inplace_kernel(double* __restrict__ A, int n, long lda) {
  __shared__ double t0[32][33], t1[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj > bi) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    t0[r][tx] = A[(long)(bi * 32 + r) * lda + bj * 32 + tx];
    t1[r][tx] = A[(long)(bj * 32 + r) * lda + bi * 32 + tx];
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    A[(long)(bi * 32 + r) * lda + bj * 32 + tx] = t1[tx][r];
    if (bi != bj) A[(long)(bj * 32 + r) * lda + bi * 32 + tx] = t0[tx][r];
  }
}

extern "C" int gpmpc_tri_inverse(const double* L, int n_pad, long lda, const double* Winv, double* Linv, long ldi,
                                 void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0, "gpmpc_tri_inverse: n_pad must be a multiple of 128");
  const long tot = (long)n_pad * n_pad;
  set_identity_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(Linv, n_pad, ldi);
  // rows of the identity solved in place: row q <- (L^{-1} e_q)' = column q of L^{-1}, i.e. the result is W'.
  // Right-looking form: its rank-64 updates cover all remaining columns at once (2x faster here than the left-looking
  // solve that `predict` uses for many query rows).
  int rc = gpmpc_trsm_right_batched(L, n_pad, lda, 0, Winv, Linv, n_pad, ldi, 0, 1, stream_);
  if (rc) return rc;
  transpose_inplace_kernel<<<dim3(n_pad / 32, n_pad / 32), 256, 0, stream>>>(Linv, n_pad, ldi);
  return gpmpc_check_launch("tri_inverse");
}

__global__ void sum_parts_kernel(const double* __restrict__ part, int nparts, long stride, int q, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= q) return;
  double s = 0.0;
  for (int t = 0; t < nparts; ++t) s += part[(long)t * stride + i];  // fixed order: bit-reproducible
  out[i] = s;
}

// The triangular GEMM alone: per-tile partial sums work[t * q_pad + q], t < n_pad / 64 (the fused step reduces them inside
// its variance / covariance kernel, in the same fixed order as sum_parts_kernel).
extern "C" int gpmpc_linv_sumsq_parts_n(const double* Linv, int n, int n_pad, long ldi, const double* Ks, int q_pad, long ldk,
                                        double* work, void* stream_) {
  GPMPC_REQUIRE(n_pad % 128 == 0 && q_pad % 128 == 0, "gpmpc_linv_sumsq_parts: n_pad and q_pad must be multiples of 128");
  GPMPC_REQUIRE(Linv && Ks && work, "gpmpc_linv_sumsq_parts: NULL buffer");
  GPMPC_REQUIRE(n >= 0 && n <= n_pad, "gpmpc_linv_sumsq_parts: n out of range");
  GemmParams p{};
  p.A = Linv; p.lda = ldi;
  p.B = Ks; p.ldb = ldk;
  p.M = n_pad; p.N = q_pad; p.K = n_pad; p.alpha = 1.0; p.beta = 0.0;
  p.k_tri = 1; p.m_fast = 1; p.epi = EPI_SUMSQ_PART; p.sumsq = work; p.part_stride = q_pad;
  // Linv comes from gpmpc_tri_inverse: exact zeros above the diagonal, identity on the padding rows n..n_pad-1 (whose K*
  // columns are zero): skip the DMMA strips that only multiply those zeros (2.82 -> 2.62 ms at n = 1000, n_pad = 1024,
  // 81 920 queries; 128 x 128 tiles are slower: 2.86 ms)
  static const bool no_skip = getenv("GPMPC_NO_TRI_SKIP") != nullptr;   // A/B switch for tools/bench_trigemm.py
  p.tri_skip = no_skip ? 0 : 1;
  p.m_valid = no_skip ? 0 : n;
  return launch_dgemm_nt<64, 128, 2, 4>(p, 1, (cudaStream_t)stream_);
}

extern "C" int gpmpc_linv_sumsq_parts(const double* Linv, int n_pad, long ldi, const double* Ks, int q_pad, long ldk,
                                      double* work, void* stream_) {
  return gpmpc_linv_sumsq_parts_n(Linv, 0, n_pad, ldi, Ks, q_pad, ldk, work, stream_);
}

extern "C" int gpmpc_linv_sumsq(const double* Linv, int n_pad, long ldi, const double* Ks, int q_pad, long ldk,
                                double* work, double* sumsq, int one_cta_per_sm, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0 && q_pad % 128 == 0, "gpmpc_linv_sumsq: n_pad and q_pad must be multiples of 128");
  GemmParams p{};
  p.A = Linv; p.lda = ldi;
  p.B = Ks; p.ldb = ldk;
  p.M = n_pad; p.N = q_pad; p.K = n_pad; p.alpha = 1.0; p.beta = 0.0;
  p.k_tri = 1; p.m_fast = 1; p.epi = EPI_SUMSQ_PART; p.sumsq = work; p.part_stride = q_pad;
  p.tri_skip = 1;
  // one_cta_per_sm: a persistent grid of one CTA per SM that asks for 114.5 KB of shared memory, so that exactly one CTA is
  // resident per SM and 112 KB stay free -- room for two QPs of the ADMM kernel running concurrently on another stream
  // (DMMA runs on the tensor pipe, the ADMM chains on the FP64 FMA pipe: separate units).
  int rc;
  if (one_cta_per_sm) {
    static GpmpcPerDevice sms;
    if (!sms.here()) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms.here(), cudaDevAttrMultiProcessorCount, dev);
    }
    rc = launch_dgemm_nt_persistent<64, 128, 2, 4>(p, sms.here(), 117248, stream);
  } else {
    rc = launch_dgemm_nt<64, 128, 2, 4>(p, 1, stream);
  }
  if (rc) return rc;
  GPMPC_PREFER_SHARED(sum_parts_kernel);
  sum_parts_kernel<<<(q_pad + 255) / 256, 256, 0, stream>>>(work, n_pad / 64, q_pad, q_pad, sumsq);
  return gpmpc_check_launch("linv_sumsq");
}

// Right-looking form of the row solve V <- (L^{-1} V')' for FEW rows (q_pad = 128..512): per 64-block one in-place product
// with the cached inverse block and one rank-64 update of all columns to the right, which spreads over n/128 CTAs.  The
// left-looking gpmpc_trsm_sumsq is built for many query rows (its k-loop grows with the solved part and its grid with
// the rows: with 128 rows that is one long serial CTA per step -- 43 ms at n = 8000 against 2 ms here).
extern "C" int gpmpc_trsm_right_batched(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* V,
                                        int q_pad, long ldv, long strideV, int batch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GPMPC_REQUIRE(n_pad % 128 == 0 && q_pad % 128 == 0 && q_pad > 0 && batch >= 1,
                "gpmpc_trsm_right: n_pad, q_pad must be multiples of 128");
  GPMPC_REQUIRE((lda & 1) == 0 && (ldv & 1) == 0 && ldv >= n_pad, "gpmpc_trsm_right: leading dimensions must be even");
  const long strideW = (long)(n_pad / 64) * 4096;
  for (int j0 = 0; j0 < n_pad; j0 += 64) {
    GemmParams a{};                                   // V[:, j0:j0+64] <- V[:, j0:j0+64] W_j'   (in place: K = N = 64)
    a.A = V + j0; a.lda = ldv; a.strideA = strideV;
    a.B = Winv + (long)(j0 / 64) * 4096; a.ldb = 64; a.strideB = strideW;
    a.C = V + j0; a.ldc = ldv; a.strideC = strideV;
    a.M = q_pad; a.N = 64; a.K = 64; a.alpha = 1.0; a.beta = 0.0; a.epi = EPI_STORE;
    int rc = launch_dgemm_nt<128, 64, 4, 2>(a, batch, stream);
    if (rc) return rc;
    if (j0 + 64 >= n_pad) break;
    const int o = ((j0 + 64) / 128) * 128;            // 128-aligned origin; columns below j0 + 64 are masked
    GemmParams u{};                                   // V[:, j0+64:] -= V[:, j0:j0+64] L[j0+64:, j0:j0+64]'
    u.A = V + j0; u.lda = ldv; u.strideA = strideV;
    u.B = L + (long)o * lda + j0; u.ldb = lda; u.strideB = strideL;
    u.C = V + o; u.ldc = ldv; u.strideC = strideV;
    u.M = q_pad; u.N = n_pad - o; u.K = 64; u.alpha = -1.0; u.beta = 1.0; u.epi = EPI_STORE;
    u.col_min = j0 + 64 - o;
    rc = launch_dgemm_nt<64, 128, 2, 4>(u, batch, stream);
    if (rc) return rc;
  }
  return GPMPC_OK;
}

extern "C" int gpmpc_trsm_right(const double* L, int n_pad, long lda, const double* Winv, double* V, int q_pad, long ldv,
                                void* stream_) {
  return gpmpc_trsm_right_batched(L, n_pad, lda, 0, Winv, V, q_pad, ldv, 0, 1, stream_);
}

// ------------------------------------------------------------------------------------------------ small-M TRSM
// X = L^{-1} B for a small lower-triangular L (M <= 128, e.g. the sparse GP's L_uu / L_B, sparse_gp.py:190,293,296)
// and many columns: L in shared memory, one column per thread (broadcast reads), optional column sum of squares.
__global__ void __launch_bounds__(128) trsm_small_kernel(const double* __restrict__ L, int M, long ldl,
                                                         double* __restrict__ B, long ldb, int ncols, int b_colmajor,
                                                         double* __restrict__ sumsq, double sumsq_sign) {
  extern __shared__ double sL[];  // M*(M+1)
  for (int idx = threadIdx.x; idx < M * M; idx += blockDim.x) sL[(idx / M) * (M + 1) + idx % M] = L[(long)(idx / M) * ldl + idx % M];
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncols) return;
  // element (i, c): column-major storage B[c*ldb + i] (rows of a (ncols x M) matrix) or row-major B[i*ldb + c]
  auto at = [&](int i) -> double& { return b_colmajor ? B[(long)c * ldb + i] : B[(long)i * ldb + c]; };
  double ss = 0.0;
  for (int i = 0; i < M; ++i) {
    double s = at(i);
    for (int k = 0; k < i; ++k) s -= sL[i * (M + 1) + k] * at(k);
    s /= sL[i * (M + 1) + i];
    at(i) = s;
    ss += s * s;
  }
  if (sumsq) sumsq[c] += sumsq_sign * ss;
}

extern "C" int gpmpc_trsm_small(const double* L, int M, long ldl, double* B, long ldb, int ncols, int b_colmajor,
                                double* sumsq, double sumsq_sign, void* stream) {
  GPMPC_REQUIRE(M > 0 && M <= 144, "gpmpc_trsm_small: M must be in 1..144");
  size_t smem = sizeof(double) * M * (M + 1);
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(trsm_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 180 * 1024);
    configured.here() = 1;
  }
  trsm_small_kernel<<<(ncols + 127) / 128, 128, smem, (cudaStream_t)stream>>>(L, M, ldl, B, ldb, ncols, b_colmajor,
                                                                            sumsq, sumsq_sign);
  return gpmpc_check_launch("trsm_small");
}

// ------------------------------------------------------------------------------------------------ plain GEMM entry
extern "C" int gpmpc_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc, int M, int N,
                              int K, double alpha, double beta, int lower_only, void* stream) {
  GemmParams p{};
  p.A = A; p.lda = lda; p.B = B; p.ldb = ldb; p.C = C; p.ldc = ldc;
  p.M = M; p.N = N; p.K = K; p.alpha = alpha; p.beta = beta; p.epi = EPI_STORE; p.lower_only = lower_only;
  return launch_dgemm_nt<64, 128, 2, 4>(p, 1, (cudaStream_t)stream);
}
