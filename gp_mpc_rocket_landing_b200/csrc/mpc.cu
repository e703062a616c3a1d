// MPC-side kernels of the 3-DoF GP-MPC hot path (7 states [m, r(3), v(3)], 3 controls):
//   * GP-augmented horizon rollout            reference gp_mpc.py:277-281 + _predict_with_gp :139-175
//   * linear covariance propagation           reference uncertainty_prop.py:117-177
//   * QP assembly in the OSQP-RTI layout      reference osqp_rti.py:203-372, linearisation :374-401 / :656-710
//   * SQP / RTI bookkeeping after the QP      reference gp_mpc.py:336-345, osqp_rti.py:532-552
// The dynamics are the in-repo Euler statement (nominal_mpc.py:586-605); `simdyn` is third-party and absent.
// This file is compiled with -fmad=false so that the element-wise model arithmetic rounds like the numpy
// restatement; the one hot loop (the GP cross-covariance row) uses explicit fma().
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

constexpr int NX = 7, NU = 3, NS = 10;

// Euler step, nominal_mpc.py:586-605:  m+ = m - dt*alpha*|u| ; r+ = r + dt v ; v+ = v + dt (u/m + g)
__device__ __forceinline__ void step3(const double* x, const double* u, const GpmpcDyn3& p, double* xn) {
  const double T = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
  xn[0] = x[0] + p.dt * (-p.alpha * T);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    xn[1 + i] = x[1 + i] + p.dt * x[4 + i];
    xn[4 + i] = x[4 + i] + p.dt * (u[i] / x[0] + p.g[i]);
  }
}

// Analytic discrete Jacobians, osqp_rti.py:682-708 (T_mag carries the reference's +1e-10).
__device__ __forceinline__ void jac3_analytic(const double* x, const double* u, const GpmpcDyn3& p, double A[NX][NX],
                                              double Bm[NX][NU]) {
  const double m = x[0];
  const double T = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]) + 1e-10;
#pragma unroll
  for (int i = 0; i < NX; ++i) {
#pragma unroll
    for (int j = 0; j < NX; ++j) A[i][j] = (i == j) ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < NU; ++j) Bm[i][j] = 0.0;
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    A[1 + i][4 + i] = p.dt;
    A[4 + i][0] = -u[i] / (m * m) * p.dt;
    Bm[0][i] = -p.alpha * u[i] / T * p.dt;
    Bm[4 + i][i] = p.dt / m;
  }
}

// Forward finite differences with eps = 1e-6, osqp_rti.py:374-401.
__device__ __forceinline__ void jac3_fd(const double* x, const double* u, const GpmpcDyn3& p, double A[NX][NX],
                                        double Bm[NX][NU]) {
  const double eps = 1e-6;
  double f0[NX], f1[NX], xp[NX], up[NU];
  step3(x, u, p, f0);
  for (int i = 0; i < NX; ++i) {
    for (int d = 0; d < NX; ++d) xp[d] = x[d];
    xp[i] += eps;
    step3(xp, u, p, f1);
    for (int r = 0; r < NX; ++r) A[r][i] = (f1[r] - f0[r]) / eps;
  }
  for (int i = 0; i < NU; ++i) {
    for (int d = 0; d < NU; ++d) up[d] = u[d];
    up[i] += eps;
    step3(x, up, p, f1);
    for (int r = 0; r < NX; ++r) Bm[r][i] = (f1[r] - f0[r]) / eps;
  }
}

// ------------------------------------------------------------------------------------------------ rollout
// One warp per scenario; the horizon is a dependency chain (x_{k+1} needs the GP mean at x_k), so the parallelism
// is scenarios x training points: each lane evaluates n/32 cross-covariances per stage, writes them as one
// coalesced row of Ks (kept for the variance TRSM) and accumulates the P posterior means on the fly.
constexpr int ROLL_WARPS = 8;
constexpr int MAXP = 6, MAXDP = 64;

// GP posterior means of one (x,u) query evaluated by one warp: features (features.py:403-444), scaling and row norm
// (kernels.py:223-229), cross-covariance row against the n training points (each lane n/32 of them, written as one
// coalesced row of Ks when requested) and the P means accumulated on the fly (exact_gp.py:237-245).
template <int DP_T>  // DP_T > 0: compile-time padded feature dimension (registers); 0: runtime (shared memory)
__device__ __forceinline__ void gp_query_warp(const GpmpcExactModel& M, const double* x, const double* u, double* zs,
                                              int lane, double* __restrict__ ks_row, double* dm) {
  const int DP = DP_T > 0 ? DP_T : M.DP;
  double z[11], sqv[11];
  features_3dof_dev(x, u, z);
#pragma unroll
  for (int d = 0; d < 11; ++d) {
    z[d] = z[d] / M.lengthscales[d];
    sqv[d] = z[d] * z[d];
  }
  const double sqz = numpy_pairwise_sum_small(sqv, 11);
  __syncwarp();
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < 11; ++d) zs[d] = z[d];
    for (int d = 11; d < DP; ++d) zs[d] = 0.0;
  }
  __syncwarp();
  double zr[DP_T > 0 ? DP_T : 1];
  if (DP_T > 0) {
#pragma unroll
    for (int d = 0; d < DP_T; ++d) zr[d] = zs[d];
  }
  double acc[MAXP];
#pragma unroll
  for (int p = 0; p < MAXP; ++p) acc[p] = 0.0;
  for (int j = lane; j < M.n_pad; j += 32) {
    double kv = 0.0;
    if (j < M.n) {
      const double* xr = M.Xs + (long)j * DP;
      double dot = 0.0;
      if (DP_T > 0) {
#pragma unroll
        for (int d = 0; d < DP_T; d += 4) {
          const double4 t = *reinterpret_cast<const double4*>(xr + d);
          dot = fma(zr[d], t.x, dot);
          dot = fma(zr[d + 1], t.y, dot);
          dot = fma(zr[d + 2], t.z, dot);
          dot = fma(zr[d + 3], t.w, dot);
        }
      } else {
        for (int d = 0; d < DP; ++d) dot = fma(zs[d], __ldg(xr + d), dot);
      }
      const double d2 = (sqz + __ldg(M.sq + j)) - 2.0 * dot;
      kv = kernel_value(M.kind, d2, M.signal_var);
#pragma unroll
      for (int p = 0; p < MAXP; ++p)
        if (p < M.P) acc[p] = fma(kv, __ldg(M.alpha + (long)p * M.n_pad + j), acc[p]);
    }
    if (ks_row) ks_row[j] = kv;
  }
#pragma unroll
  for (int p = 0; p < MAXP; ++p)
    if (p < M.P) dm[p] = warp_sum(acc[p]) * M.ystd[p] + M.ymean[p];  // exact_gp.py:240-245
}

template <int DP_T>
__global__ void __launch_bounds__(ROLL_WARPS * 32) rollout3_kernel(GpmpcExactModel M, int has_model, GpmpcDyn3 dyn,
                                                                   int nbatch, int N, const double* __restrict__ x0,
                                                                   const double* __restrict__ U,
                                                                   double* __restrict__ X, double* __restrict__ dmean,
                                                                   double* __restrict__ Ks, long ldk,
                                                                   const int* __restrict__ active) {
  __shared__ double zs_sm[ROLL_WARPS][MAXDP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * ROLL_WARPS + warp;
  if (b >= nbatch) return;
  const bool act = (active == nullptr) || active[b] != 0;
  double* zs = zs_sm[warp];
  double x[NX], xn[NX], u[NU];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = x0[(long)b * NX + d];
  if (act && lane < NX) X[(long)b * (N + 1) * NX + lane] = x[lane];
  for (int k = 0; k < N; ++k) {
    const long row = (long)b * N + k;
    if (!act) {  // keep the rows of skipped scenarios finite for the variance TRSM that follows
      if (has_model && Ks)
        for (int j = lane; j < M.n_pad; j += 32) Ks[row * ldk + j] = 0.0;
      continue;
    }
#pragma unroll
    for (int d = 0; d < NU; ++d) u[d] = U[row * NU + d];
    double dm[MAXP];
#pragma unroll
    for (int p = 0; p < MAXP; ++p) dm[p] = 0.0;
    if (has_model) gp_query_warp<DP_T>(M, x, u, zs, lane, Ks ? Ks + row * ldk : nullptr, dm);
    step3(x, u, dyn, xn);
#pragma unroll
    for (int i = 0; i < 3; ++i) xn[4 + i] = xn[4 + i] + dm[i] * dyn.dt;  // gp_mpc.py:165-167
    if (lane < NX) X[((long)b * (N + 1) + k + 1) * NX + lane] = xn[lane];
    if (dmean && lane < 3) dmean[row * 3 + lane] = dm[lane];
#pragma unroll
    for (int d = 0; d < NX; ++d) x[d] = xn[d];
  }
}

// Shared-memory variant (the one the 1000-point model takes): the whole model -- scaled training inputs (feature-major, so
// lane j reads a conflict-free column), row norms and the P alpha vectors, 16 n_pad (DP+1+P) bytes = 131 kB at n_pad = 1024
// -- is staged once per CTA, and every warp rolls out TWO scenarios at once, so each training point read from shared
// memory serves two queries and the two dot-product chains overlap.  The grid is sized so that all scenario pairs run in
// ONE wave (one CTA per SM, W warps each, W * #SM >= B/2).  Same per-element arithmetic as gp_query_warp.
constexpr int ROLLS_MAX_WARPS = 16;

template <int DP_T, int P_T>
__global__ void __launch_bounds__(ROLLS_MAX_WARPS * 32, 1)
    rollout3_smem_kernel(GpmpcExactModel M, GpmpcDyn3 dyn, int nbatch, int N, const double* __restrict__ x0,
                         const double* __restrict__ U, double* __restrict__ X, double* __restrict__ dmean,
                         double* __restrict__ Ks, long ldk, const int* __restrict__ active) {
  extern __shared__ __align__(16) double rsm[];
  const int n_pad = M.n_pad, n = M.n;
  double* XsT = rsm;                          // DP_T x n_pad
  double* sqs = XsT + (size_t)DP_T * n_pad;   // n_pad
  double* als = sqs + n_pad;                  // P_T x n_pad
  for (int e = threadIdx.x; e < n_pad * DP_T; e += blockDim.x) {
    const int j = e / DP_T, d = e - j * DP_T;
    XsT[(size_t)d * n_pad + j] = M.Xs[e];
  }
  for (int e = threadIdx.x; e < n_pad; e += blockDim.x) sqs[e] = M.sq[e];
  for (int e = threadIdx.x; e < n_pad * P_T; e += blockDim.x) als[e] = M.alpha[e];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
  const int ntask = (nbatch + 1) / 2;
  double ls[11];
#pragma unroll
  for (int d = 0; d < 11; ++d) ls[d] = M.lengthscales[d];
  for (int task = blockIdx.x * W + warp; task < ntask; task += gridDim.x * W) {
    const int b0 = 2 * task, b1 = min(b0 + 1, nbatch - 1);
    const bool has1 = b0 + 1 < nbatch;
    const bool act0 = (active == nullptr) || active[b0] != 0;
    const bool act1 = has1 && ((active == nullptr) || active[b1] != 0);
    double xa[NX], xb[NX], xn[NX], ua[NU], ub[NU];
#pragma unroll
    for (int d = 0; d < NX; ++d) {
      xa[d] = x0[(long)b0 * NX + d];
      xb[d] = x0[(long)b1 * NX + d];
    }
    if (lane < NX) {
      if (act0) X[(long)b0 * (N + 1) * NX + lane] = xa[lane];
      if (act1) X[(long)b1 * (N + 1) * NX + lane] = xb[lane];
    }
    for (int k = 0; k < N; ++k) {
      const long r0 = (long)b0 * N + k, r1 = (long)b1 * N + k;
      double* k0 = Ks ? Ks + r0 * ldk : nullptr;
      double* k1 = (Ks && has1) ? Ks + r1 * ldk : nullptr;
      if (!act0 && !act1) {  // keep the rows of skipped scenarios finite for the variance TRSM that follows
        for (int j = lane; j < n_pad; j += 32) {
          if (k0) k0[j] = 0.0;
          if (k1) k1[j] = 0.0;
        }
        continue;
      }
#pragma unroll
      for (int d = 0; d < NU; ++d) {
        ua[d] = U[r0 * NU + d];
        ub[d] = U[r1 * NU + d];
      }
      double za[DP_T], zb[DP_T], sqv[11];
      features_3dof_dev(xa, ua, za);
#pragma unroll
      for (int d = 0; d < 11; ++d) {
        za[d] = za[d] / ls[d];
        sqv[d] = za[d] * za[d];
      }
      const double sqa = numpy_pairwise_sum_small(sqv, 11);
      features_3dof_dev(xb, ub, zb);
#pragma unroll
      for (int d = 0; d < 11; ++d) {
        zb[d] = zb[d] / ls[d];
        sqv[d] = zb[d] * zb[d];
      }
      const double sqb = numpy_pairwise_sum_small(sqv, 11);
#pragma unroll
      for (int d = 11; d < DP_T; ++d) za[d] = zb[d] = 0.0;
      double acca[P_T], accb[P_T];
#pragma unroll
      for (int p = 0; p < P_T; ++p) acca[p] = accb[p] = 0.0;
#pragma unroll 2
      for (int j = lane; j < n_pad; j += 32) {
        double kva = 0.0, kvb = 0.0;
        if (j < n) {
          double da = 0.0, db = 0.0;
#pragma unroll
          for (int d = 0; d < DP_T; ++d) {
            const double t = XsT[(size_t)d * n_pad + j];
            da = fma(za[d], t, da);
            db = fma(zb[d], t, db);
          }
          const double sj = sqs[j];
          kva = kernel_value(M.kind, (sqa + sj) - 2.0 * da, M.signal_var);
          kvb = kernel_value(M.kind, (sqb + sj) - 2.0 * db, M.signal_var);
#pragma unroll
          for (int p = 0; p < P_T; ++p) {
            const double a = als[(size_t)p * n_pad + j];
            acca[p] = fma(kva, a, acca[p]);
            accb[p] = fma(kvb, a, accb[p]);
          }
        }
        if (k0) k0[j] = act0 ? kva : 0.0;
        if (k1) k1[j] = act1 ? kvb : 0.0;
      }
      double dma[3], dmb[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        dma[p] = warp_sum(acca[p]) * M.ystd[p] + M.ymean[p];  // exact_gp.py:240-245
        dmb[p] = warp_sum(accb[p]) * M.ystd[p] + M.ymean[p];
      }
      if (act0) {
        step3(xa, ua, dyn, xn);
#pragma unroll
        for (int i = 0; i < 3; ++i) xn[4 + i] = xn[4 + i] + dma[i] * dyn.dt;  // gp_mpc.py:165-167
        if (lane < NX) X[((long)b0 * (N + 1) + k + 1) * NX + lane] = xn[lane];
        if (dmean && lane < 3) dmean[r0 * 3 + lane] = dma[lane];
#pragma unroll
        for (int d = 0; d < NX; ++d) xa[d] = xn[d];
      }
      if (act1) {
        step3(xb, ub, dyn, xn);
#pragma unroll
        for (int i = 0; i < 3; ++i) xn[4 + i] = xn[4 + i] + dmb[i] * dyn.dt;
        if (lane < NX) X[((long)b1 * (N + 1) + k + 1) * NX + lane] = xn[lane];
        if (dmean && lane < 3) dmean[r1 * 3 + lane] = dmb[lane];
#pragma unroll
        for (int d = 0; d < NX; ++d) xb[d] = xn[d];
      }
    }
  }
}

// GP means at given trajectory points (X[b,k], U[b,k]), k < N: the affine terms c_k of one SQP iteration
// (gp_mpc.py:309-317).  One warp per point.
template <int DP_T>
__global__ void __launch_bounds__(ROLL_WARPS * 32) gp_points3_kernel(GpmpcExactModel M, int nbatch, int N,
                                                                     const double* __restrict__ X,
                                                                     const double* __restrict__ U,
                                                                     double* __restrict__ dmean,
                                                                     double* __restrict__ Ks, long ldk,
                                                                     const int* __restrict__ active) {
  __shared__ double zs_sm[ROLL_WARPS][MAXDP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long pt = (long)blockIdx.x * ROLL_WARPS + warp;
  if (pt >= (long)nbatch * N) return;
  const int b = (int)(pt / N), k = (int)(pt - (long)b * N);
  if (active && !active[b]) {
    if (Ks)
      for (int j = lane; j < M.n_pad; j += 32) Ks[pt * ldk + j] = 0.0;
    return;
  }
  double x[NX], u[NU], dm[MAXP];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = X[((long)b * (N + 1) + k) * NX + d];
#pragma unroll
  for (int d = 0; d < NU; ++d) u[d] = U[pt * NU + d];
#pragma unroll
  for (int p = 0; p < MAXP; ++p) dm[p] = 0.0;
  gp_query_warp<DP_T>(M, x, u, zs_sm[warp], lane, Ks ? Ks + pt * ldk : nullptr, dm);
  if (lane < 3) dmean[pt * 3 + lane] = dm[lane];
}

// z = [x_0,u_0,...,x_N] from (X, U): osqp_rti.py:601-614 (_solution_to_vector), the primal warm start.
__global__ void __launch_bounds__(256) pack_traj3_kernel(int nbatch, int N, const double* __restrict__ X,
                                                         const double* __restrict__ U, double* __restrict__ z,
                                                         const int* __restrict__ active) {
  const int n = N * NS + NX;
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long)nbatch * n) return;
  const int b = (int)(idx / n), j = (int)(idx - (long)b * n);
  if (active && !active[b]) return;
  const int k = j / NS, i = j - k * NS;
  z[idx] = (i < NX) ? X[((long)b * (N + 1) + k) * NX + i] : U[((long)b * N + k) * NU + (i - NX)];
}

// ------------------------------------------------------------------------------------------------ covariance
// One thread per scenario, Sigma (7x7) in registers; A_k = I + E_k with 6 off-diagonal entries, so A Sigma A' is two
// sparse row/column updates (same products as the dense form of uncertainty_prop.py:166, zeros skipped).
__global__ void __launch_bounds__(128) propagate_cov3_kernel(GpmpcDyn3 dyn, int nbatch, int N,
                                                             const double* __restrict__ X,
                                                             const double* __restrict__ U,
                                                             const double* __restrict__ dvar, double sigma0,
                                                             double* __restrict__ covs,
                                                             const int* __restrict__ active) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  double S[NX][NX], Mx[NX][NX];
#pragma unroll
  for (int i = 0; i < NX; ++i)
#pragma unroll
    for (int j = 0; j < NX; ++j) S[i][j] = (i == j) ? sigma0 : 0.0;
  double* out = covs + (long)b * (N + 1) * NX * NX;
#pragma unroll
  for (int i = 0; i < NX; ++i)
#pragma unroll
    for (int j = 0; j < NX; ++j) out[i * NX + j] = S[i][j];
  for (int k = 0; k < N; ++k) {
    const double* x = X + ((long)b * (N + 1) + k) * NX;
    const double* u = U + ((long)b * N + k) * NU;
    const double m = x[0];
    double a[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) a[i] = -u[i] / (m * m) * dyn.dt;
    // Mx = A S
#pragma unroll
    for (int j = 0; j < NX; ++j) {
      Mx[0][j] = S[0][j];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        Mx[1 + i][j] = S[1 + i][j] + dyn.dt * S[4 + i][j];
        Mx[4 + i][j] = a[i] * S[0][j] + S[4 + i][j];
      }
    }
    // S = Mx A'
#pragma unroll
    for (int r = 0; r < NX; ++r) {
      S[r][0] = Mx[r][0];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        S[r][1 + i] = Mx[r][1 + i] + Mx[r][4 + i] * dyn.dt;
        S[r][4 + i] = Mx[r][0] * a[i] + Mx[r][4 + i];
      }
    }
    if (dvar) {
      const double* v = dvar + ((long)b * N + k) * 3;
#pragma unroll
      for (int i = 0; i < 3; ++i) S[4 + i][4 + i] = S[4 + i][4 + i] + v[i] * (dyn.dt * dyn.dt);  // :157-160
    }
    double* o = out + (long)(k + 1) * NX * NX;
#pragma unroll
    for (int i = 0; i < NX; ++i)
#pragma unroll
      for (int j = 0; j < NX; ++j) o[i * NX + j] = S[i][j];
  }
}


// ------------------------------------------------------------------------------------------------ unscented transform
// uncertainty_prop.py:179-264 for the 7-state model.  The reference class carries 14 states: n = 14 fixes lambda and the
// weights, and of its 29 sigma points the 14 that perturb only the 7 inert states propagate exactly like the centre
// (their block of Sigma never couples with the first 7).  Per stage: ut_sigma3_kernel (7x7 Cholesky of
// (n + lambda) Sigma + 1e-10 I, 15 distinct points), the N = 1 rollout kernel images them (step + GP mean, and the
// cross-covariance row of the centre for the variance), ut_combine3_kernel recombines in the reference's term order.
__global__ void __launch_bounds__(128) ut_sigma3_kernel(int nbatch, double scale, const double* __restrict__ xk,
                                                        const double* __restrict__ Sk, const double* __restrict__ U,
                                                        long ldu, double* __restrict__ SP, double* __restrict__ Usp,
                                                        int* __restrict__ info) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbatch) return;
  double L[NX][NX], x[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) {
    x[i] = xk[(long)b * NX + i];
#pragma unroll
    for (int j = 0; j < NX; ++j) L[i][j] = scale * Sk[((long)b * NX + i) * NX + j] + ((i == j) ? 1e-10 : 0.0);
  }
#pragma unroll
  for (int j = 0; j < NX; ++j) {
    double d = L[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
    if (!(d > 0.0)) {
      if (info[0] == 0) info[0] = b + 1;   // np.linalg.cholesky raises LinAlgError; the host checks this flag
      d = 1.0;
    }
    const double dj = sqrt(d);
    L[j][j] = dj;
#pragma unroll
    for (int i = j + 1; i < NX; ++i) {
      double v = L[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) v -= L[i][k] * L[j][k];
      L[i][j] = v / dj;
    }
#pragma unroll
    for (int i = 0; i < j; ++i) L[i][j] = 0.0;
  }
  double* sp = SP + (long)b * 15 * NX;
  double* us = Usp + (long)b * 15 * NU;
#pragma unroll
  for (int i = 0; i < NX; ++i) sp[i] = x[i];
#pragma unroll
  for (int c = 0; c < NX; ++c) {
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      sp[(1 + c) * NX + i] = x[i] + L[i][c];
      sp[(1 + NX + c) * NX + i] = x[i] - L[i][c];
    }
  }
  const double u0 = U[(long)b * ldu], u1 = U[(long)b * ldu + 1], u2 = U[(long)b * ldu + 2];
  for (int s = 0; s < 15; ++s) {
    us[s * NU] = u0; us[s * NU + 1] = u1; us[s * NU + 2] = u2;
  }
}

// Xn: images of the sigma points, (nbatch * 15) x 2 x 7 (the N = 1 rollout layout: row 1 of each pair).
__global__ void __launch_bounds__(128) ut_combine3_kernel(int nbatch, int n_inert, double wm0, double wc0, double wi,
                                                          const double* __restrict__ Xn, const double* __restrict__ var,
                                                          double dt, double* __restrict__ xnext,
                                                          double* __restrict__ Snext) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbatch) return;
  const double* base = Xn + (long)b * 15 * 2 * NX + NX;   // image of point s at base + s * 2 * NX
  double c[NX], xn[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) {
    c[i] = base[i];
    xn[i] = wm0 * c[i];
  }
  // rows 0..28 of the reference's sigma-point array: centre, +real (7), +inert, -real (7), -inert
  for (int s = 1; s <= NX; ++s)
#pragma unroll
    for (int i = 0; i < NX; ++i) xn[i] = xn[i] + wi * base[(long)s * 2 * NX + i];
  for (int s = 0; s < n_inert; ++s)
#pragma unroll
    for (int i = 0; i < NX; ++i) xn[i] = xn[i] + wi * c[i];
  for (int s = NX + 1; s <= 2 * NX; ++s)
#pragma unroll
    for (int i = 0; i < NX; ++i) xn[i] = xn[i] + wi * base[(long)s * 2 * NX + i];
  for (int s = 0; s < n_inert; ++s)
#pragma unroll
    for (int i = 0; i < NX; ++i) xn[i] = xn[i] + wi * c[i];
  double S[NX][NX];
#pragma unroll
  for (int i = 0; i < NX; ++i)
#pragma unroll
    for (int j = 0; j < NX; ++j) S[i][j] = 0.0;
  auto add = [&](const double* q, double w) {
    double df[NX];
#pragma unroll
    for (int i = 0; i < NX; ++i) df[i] = q[i] - xn[i];
#pragma unroll
    for (int i = 0; i < NX; ++i)
#pragma unroll
      for (int j = 0; j < NX; ++j) S[i][j] = S[i][j] + w * (df[i] * df[j]);
  };
  add(c, wc0);
  for (int s = 1; s <= NX; ++s) add(base + (long)s * 2 * NX, wi);
  for (int s = 0; s < n_inert; ++s) add(c, wi);
  for (int s = NX + 1; s <= 2 * NX; ++s) add(base + (long)s * 2 * NX, wi);
  for (int s = 0; s < n_inert; ++s) add(c, wi);
  if (var) {
#pragma unroll
    for (int i = 0; i < 3; ++i) S[4 + i][4 + i] += var[(long)b * 3 + i] * (dt * dt);
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) {
    xnext[(long)b * NX + i] = xn[i];
#pragma unroll
    for (int j = 0; j < NX; ++j) Snext[((long)b * NX + i) * NX + j] = S[i][j];
  }
}

// ------------------------------------------------------------------------------------------------ QP assembly
// One thread per (scenario, stage); stage N writes the initial-state rows and the terminal cost vector.
__global__ void __launch_bounds__(128) qp_assemble3_kernel(GpmpcDyn3 dyn, int nbatch, int N, int lin_mode, int rhs_mode,
                                                           const double* __restrict__ X_lin,
                                                           const double* __restrict__ U_lin,
                                                           const double* __restrict__ x_init,
                                                           const double* __restrict__ x_ref, long ldxref,
                                                           const double* __restrict__ c_gp,
                                                           const double* __restrict__ Qd,
                                                           const double* __restrict__ Qfd, double* __restrict__ AB,
                                                           double* __restrict__ beq, double* __restrict__ q,
                                                           const int* __restrict__ active,
                                                           double* __restrict__ zpack) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbatch * (N + 1)) return;
  const int b = idx / (N + 1), k = idx - b * (N + 1);
  if (active && !active[b]) return;
  const int n = N * NS + NX, neq = NX * (N + 1);
  const double* xr = x_ref + (long)b * ldxref + (long)k * NX;
  double* qb = q + (long)b * n + (long)k * NS;
  if (zpack) {  // primal warm start z = [x_0,u_0,...,x_N] (osqp_rti.py:601-614), fused: one launch less per SQP iteration
    double* zb = zpack + (long)b * n + (long)k * NS;
#pragma unroll
    for (int i = 0; i < NX; ++i) zb[i] = X_lin[((long)b * (N + 1) + k) * NX + i];
    if (k < N)
#pragma unroll
      for (int i = 0; i < NU; ++i) zb[NX + i] = U_lin[((long)b * N + k) * NU + i];
  }
  if (k == N) {
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      qb[i] = -(Qfd[i] * xr[i]);                       // osqp_rti.py:254-256
      beq[(long)b * neq + i] = x_init[(long)b * NX + i];  // :286-288
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) qb[i] = -(Qd[i] * xr[i]);  // :249-252
#pragma unroll
  for (int i = 0; i < NU; ++i) qb[NX + i] = 0.0;
  double x[NX], u[NU], A[NX][NX], Bm[NX][NU];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = X_lin[((long)b * (N + 1) + k) * NX + d];
#pragma unroll
  for (int d = 0; d < NU; ++d) u[d] = U_lin[((long)b * N + k) * NU + d];
  if (lin_mode == 0) jac3_analytic(x, u, dyn, A, Bm);
  else jac3_fd(x, u, dyn, A, Bm);
  // affine term
  double c[NX];
  if (rhs_mode == 0) {
    step3(x, u, dyn, c);  // c_k = f(x_lin,u_lin) - A x_lin - B u_lin   (:338-344, sign as the reference writes it)
  } else {
#pragma unroll
    for (int d = 0; d < NX; ++d) c[d] = X_lin[((long)b * (N + 1) + k + 1) * NX + d];  // gp_mpc.py:402-411
  }
#pragma unroll
  for (int r = 0; r < NX; ++r) {
    double ax = 0.0, bu = 0.0;
#pragma unroll
    for (int j = 0; j < NX; ++j) ax += A[r][j] * x[j];
#pragma unroll
    for (int j = 0; j < NU; ++j) bu += Bm[r][j] * u[j];
    c[r] = (c[r] - ax) - bu;
  }
  if (c_gp) {
#pragma unroll
    for (int i = 0; i < 3; ++i) c[4 + i] = c[4 + i] + c_gp[((long)b * N + k) * 3 + i] * dyn.dt;
  }
  double* be = beq + (long)b * neq + NX + (long)k * NX;
#pragma unroll
  for (int r = 0; r < NX; ++r) be[r] = (rhs_mode == 0) ? c[r] : -c[r];
  double* ab = AB + ((long)b * N + k) * NX * NS;
#pragma unroll
  for (int r = 0; r < NX; ++r) {
#pragma unroll
    for (int j = 0; j < NX; ++j) ab[r * NS + j] = (fabs(A[r][j]) > 1e-10) ? A[r][j] : 0.0;  // :301
#pragma unroll
    for (int j = 0; j < NU; ++j) ab[r * NS + NX + j] = (fabs(Bm[r][j]) > 1e-10) ? Bm[r][j] : 0.0;  // :309
  }
}

// ------------------------------------------------------------------------------------------------ convex rows
// Polyhedral thrust / glide-slope rows as 7 extra stage "states" (include/gpmpc_b200.h, GpmpcCone3).  One thread per
// (scenario, stage); stage N writes the initial rows and the terminal entries.
constexpr int NXC = 2 * NX, NSC = NXC + NU;

// the three normals of the inscribed triangle with a vertex in the radial direction of (y, z)
__device__ __forceinline__ void glide_normals(double y, double z, double (&ny)[3], double (&nz)[3]) {
  const double rho = sqrt(y * y + z * z);
  double cy = 1.0, cz = 0.0;
  if (rho > 1e-9) {
    cy = y / rho;
    cz = z / rho;
  }
  const double c60 = 0.5, s60 = 0.8660254037844386;
  ny[0] = c60 * cy - s60 * cz;
  nz[0] = s60 * cy + c60 * cz;
  ny[1] = c60 * cy + s60 * cz;
  nz[1] = -s60 * cy + c60 * cz;
  ny[2] = -cy;
  nz[2] = -cz;
}

// xi at a nominal point: the four thrust facets of u_prev and the three glide facets of x (oriented by x itself)
__device__ __forceinline__ void cone_values(const GpmpcCone3& cone, const double* x, const double* u_prev, double* xi) {
  xi[0] = u_prev[0] + cone.facet_b * u_prev[1];
  xi[1] = u_prev[0] - cone.facet_b * u_prev[1];
  xi[2] = u_prev[0] + cone.facet_b * u_prev[2];
  xi[3] = u_prev[0] - cone.facet_b * u_prev[2];
  double ny[3], nz[3];
  glide_normals(x[2], x[3], ny, nz);
#pragma unroll
  for (int i = 0; i < 3; ++i) xi[4 + i] = (ny[i] * x[2] + nz[i] * x[3]) - cone.half_tan_gs * x[1];
}

__global__ void __launch_bounds__(128) qp_augment_cone3_kernel(GpmpcCone3 cone, int nbatch, int N,
                                                               const double* __restrict__ AB,
                                                               const double* __restrict__ beq,
                                                               const double* __restrict__ q,
                                                               const double* __restrict__ X_nom,
                                                               const double* __restrict__ U_nom,
                                                               double* __restrict__ ABc, double* __restrict__ beqc,
                                                               double* __restrict__ qc, double* __restrict__ zc,
                                                               const int* __restrict__ active) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbatch * (N + 1)) return;
  const int b = idx / (N + 1), k = idx - b * (N + 1);
  if (active && !active[b]) return;
  const int n = N * NS + NX, neq = NX * (N + 1), nc = N * NSC + NXC, neqc = NXC * (N + 1);
  const double* qb = q + (long)b * n + (long)k * NS;
  double* qcb = qc + (long)b * nc + (long)k * NSC;
#pragma unroll
  for (int i = 0; i < NX; ++i) {
    qcb[i] = qb[i];
    qcb[NX + i] = 0.0;
  }
  const double* xk = X_nom + ((long)b * (N + 1) + k) * NX;
  if (zc) {
    double* zb = zc + (long)b * nc + (long)k * NSC;
#pragma unroll
    for (int i = 0; i < NX; ++i) zb[i] = xk[i];
    if (k == 0) {
#pragma unroll
      for (int i = 0; i < NX; ++i) zb[NX + i] = 0.0;
    } else {
      cone_values(cone, xk, U_nom + ((long)b * N + k - 1) * NU, zb + NX);
    }
    if (k < N)
#pragma unroll
      for (int i = 0; i < NU; ++i) zb[NXC + i] = U_nom[((long)b * N + k) * NU + i];
  }
  if (k == N) {
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      beqc[(long)b * neqc + i] = beq[(long)b * neq + i];
      beqc[(long)b * neqc + NX + i] = 0.0;
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < NU; ++i) qcb[NXC + i] = qb[NX + i];
  const double* ab = AB + ((long)b * N + k) * NX * NS;
  const double* be = beq + (long)b * neq + NX + (long)k * NX;
  double* abc = ABc + ((long)b * N + k) * NXC * NSC;
  double* bec = beqc + (long)b * neqc + NXC + (long)k * NXC;
  for (int i = 0; i < NXC * NSC; ++i) abc[i] = 0.0;
#pragma unroll
  for (int r = 0; r < NX; ++r) {
#pragma unroll
    for (int j = 0; j < NX; ++j) abc[r * NSC + j] = ab[r * NS + j];
#pragma unroll
    for (int j = 0; j < NU; ++j) abc[r * NSC + NXC + j] = ab[r * NS + NX + j];
    bec[r] = be[r];
  }
  // thrust facets on u_k
  const double fb = cone.facet_b;
  abc[(NX + 0) * NSC + NXC + 0] = 1.0; abc[(NX + 0) * NSC + NXC + 1] = fb;
  abc[(NX + 1) * NSC + NXC + 0] = 1.0; abc[(NX + 1) * NSC + NXC + 1] = -fb;
  abc[(NX + 2) * NSC + NXC + 0] = 1.0; abc[(NX + 2) * NSC + NXC + 2] = fb;
  abc[(NX + 3) * NSC + NXC + 0] = 1.0; abc[(NX + 3) * NSC + NXC + 2] = -fb;
#pragma unroll
  for (int i = 0; i < 4; ++i) bec[NX + i] = 0.0;
  // glide facets on x_{k+1} = A x_k + B u_k - beq_k, oriented by the nominal x_{k+1}
  const double* xn = X_nom + ((long)b * (N + 1) + k + 1) * NX;
  double ny[3], nz[3];
  glide_normals(xn[2], xn[3], ny, nz);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double* row = abc + (NX + 4 + i) * NSC;
#pragma unroll
    for (int j = 0; j < NX; ++j)
      row[j] = (ny[i] * ab[2 * NS + j] + nz[i] * ab[3 * NS + j]) - cone.half_tan_gs * ab[1 * NS + j];
#pragma unroll
    for (int j = 0; j < NU; ++j)
      row[NXC + j] = (ny[i] * ab[2 * NS + NX + j] + nz[i] * ab[3 * NS + NX + j]) - cone.half_tan_gs * ab[1 * NS + NX + j];
    bec[NX + 4 + i] = (ny[i] * be[2] + nz[i] * be[3]) - cone.half_tan_gs * be[1];
  }
}

__global__ void __launch_bounds__(128) qp_extract_cone3_kernel(int nbatch, int N, const double* __restrict__ zc,
                                                               double* __restrict__ z, const int* __restrict__ active) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbatch * (N + 1)) return;
  const int b = idx / (N + 1), k = idx - b * (N + 1);
  if (active && !active[b]) return;
  const double* zb = zc + (long)b * (N * NSC + NXC) + (long)k * NSC;
  double* o = z + (long)b * (N * NS + NX) + (long)k * NS;
#pragma unroll
  for (int i = 0; i < NX; ++i) o[i] = zb[i];
  if (k < N)
#pragma unroll
    for (int i = 0; i < NU; ++i) o[NX + i] = zb[NXC + i];
}

// Discrete Jacobians at a list of points (dynamics.linearize protocol: osqp_rti.py:374-401 / :656-710), unfiltered.
__global__ void __launch_bounds__(128) linearize3_kernel(GpmpcDyn3 dyn, int npts, int lin_mode,
                                                         const double* __restrict__ X, const double* __restrict__ U,
                                                         double* __restrict__ AB) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npts) return;
  double x[NX], u[NU], A[NX][NX], Bm[NX][NU];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = X[(long)i * NX + d];
#pragma unroll
  for (int d = 0; d < NU; ++d) u[d] = U[(long)i * NU + d];
  if (lin_mode == 0) jac3_analytic(x, u, dyn, A, Bm);
  else jac3_fd(x, u, dyn, A, Bm);
  double* ab = AB + (long)i * NX * NS;
#pragma unroll
  for (int r = 0; r < NX; ++r) {
#pragma unroll
    for (int j = 0; j < NX; ++j) ab[r * NS + j] = A[r][j];
#pragma unroll
    for (int j = 0; j < NU; ++j) ab[r * NS + NX + j] = Bm[r][j];
  }
}

// ------------------------------------------------------------------------------------------------ Monte Carlo step
// One closed-loop simulation step per scenario (experiments/monte_carlo.py:455-549): plant step x <- f(x,u) (+ the
// drag-like "true" residual of the synthetic workload, cf. experiments/dispersion.py:349-360), then the termination
// rules the reference applies at the top of its next loop iteration (:460-488) and LandingConstraints.check_landing
// (:54-104, defaults).  Outcome codes follow LandingOutcome (:24-33): 1 success, 2 crash, 3 fuel, 4 violation,
// 5 timeout, 6 divergence; 0 = still running.
__device__ __forceinline__ int classify3(const double* x, double m0) {
  if (x[1] < 0.0) return 2;
  if (x[0] <= 1.0 + 0.01) return 3;
  bool bad = false;
#pragma unroll
  for (int d = 0; d < NX; ++d) bad = bad || !(fabs(x[d]) <= 1e6);
  if (bad) return 6;
  if (x[1] < 1.0 && fabs(x[4]) < 5.0) {
    const bool ok = fabs(x[1]) <= 1.0 && fabs(x[2]) <= 5.0 && fabs(x[3]) <= 5.0 && fabs(x[4]) <= 2.0 &&
                    fabs(x[5]) <= 1.0 && fabs(x[6]) <= 1.0 && !(1.0 - x[0] / m0 > 1.0 - 0.05);
    return ok ? 1 : 4;
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ post-QP update
// One warp per scenario.  mode 0 = RTI (osqp_rti.py:532-552), mode 1 = GP-MPC SQP iteration (gp_mpc.py:336-345).
__device__ __forceinline__ void mpc_update3_body(int b, int lane, int N, int mode, const double* __restrict__ z,
                                                 const int* __restrict__ qp_status, const double* __restrict__ x_ref,
                                                 long ldxref, const double* __restrict__ Qd,
                                                 const double* __restrict__ Qfd, const double* __restrict__ Rd,
                                                 double* __restrict__ Xb, double* __restrict__ Ub,
                                                 double* __restrict__ X_prev, double* __restrict__ U_prev,
                                                 double* __restrict__ cost, int* __restrict__ converged,
                                                 int* __restrict__ active, int* __restrict__ n_iter,
                                                 const int* __restrict__ qp_it, int* __restrict__ qp_iters) {
  const int n = N * NS + NX;
  const bool solved = qp_status[b] == GPMPC_QP_SOLVED;
  const double* zb = z + (long)b * n;
  double dmaxx = 0.0, dmaxu = 0.0, cs = 0.0;
  if (solved) {
    for (int j = lane; j < n; j += 32) {
      const int k = j / NS, i = j - k * NS;
      const double v = zb[j];
      if (i < NX) {
        double* dst = Xb + k * NX + i;
        dmaxx = fmax(dmaxx, fabs(v - *dst));
        const double e = v - x_ref[(long)b * ldxref + (long)k * NX + i];
        cs += ((k == N) ? Qfd[i] : Qd[i]) * (e * e);
        *dst = v;
        if (X_prev) {  // shifted warm start: X_prev[k-1] = X[k], X_prev[N] = X[N]
          double* xp = X_prev + (long)b * (N + 1) * NX;
          if (k >= 1) xp[(k - 1) * NX + i] = v;
          if (k == N) xp[N * NX + i] = v;
        }
      } else {
        double* dst = Ub + k * NU + (i - NX);
        dmaxu = fmax(dmaxu, fabs(v - *dst));
        cs += Rd[i - NX] * (v * v);
        *dst = v;
        if (U_prev) {
          double* up = U_prev + (long)b * N * NU;
          if (k >= 1) up[(k - 1) * NU + (i - NX)] = v;
          if (k == N - 1) up[(N - 1) * NU + (i - NX)] = v;
        }
      }
    }
  }
  dmaxx = warp_max(dmaxx);
  dmaxu = warp_max(dmaxu);
  cs = warp_sum(cs);
  if (lane == 0) {
    cost[b] = solved ? cs : INFINITY;
    if (qp_iters && qp_it) qp_iters[b] += qp_it[b];
    if (mode == 0) {
      converged[b] = solved ? 1 : 0;
    } else {
      const int conv = (dmaxx < 1e-4 && dmaxu < 1e-4) ? 1 : 0;
      converged[b] = conv;
      if (n_iter) n_iter[b] += 1;
      if (conv && active) active[b] = 0;
    }
  }
}

__global__ void __launch_bounds__(128) mpc_update3_kernel(int nbatch, int N, int mode, const double* __restrict__ z,
                                                          const int* __restrict__ qp_status,
                                                          const double* __restrict__ x_ref, long ldxref,
                                                          const double* __restrict__ Qd,
                                                          const double* __restrict__ Qfd,
                                                          const double* __restrict__ Rd, double* __restrict__ X,
                                                          double* __restrict__ U, double* __restrict__ X_prev,
                                                          double* __restrict__ U_prev, double* __restrict__ cost,
                                                          int* __restrict__ converged, int* __restrict__ active,
                                                          int* __restrict__ n_iter,
                                                          const int* __restrict__ qp_it, int* __restrict__ qp_iters,
                                                          double* __restrict__ U_warm, double* __restrict__ X_warm,
                                                          int* __restrict__ has_warm) {
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= nbatch) return;
  double* Xb = X + (long)b * (N + 1) * NX;
  double* Ub = U + (long)b * N * NU;
  const bool act = !(active && !active[b]);
  if (U_warm) {  // last SQP iteration: the iterate of EVERY scenario becomes its next warm start (gp_mpc.py:355-357)
    if (act) {
      mpc_update3_body(b, lane, N, mode, z, qp_status, x_ref, ldxref, Qd, Qfd, Rd, Xb, Ub, X_prev, U_prev, cost, converged,
                       active, n_iter, qp_it, qp_iters);
      __syncwarp();
    }
    for (int j = lane; j < N * NU; j += 32) U_warm[(long)b * N * NU + j] = Ub[j];
    if (X_warm)
      for (int j = lane; j < (N + 1) * NX; j += 32) X_warm[(long)b * (N + 1) * NX + j] = Xb[j];
    if (has_warm && lane == 0) has_warm[b] = 1;
    return;
  }
  if (!act) return;
  mpc_update3_body(b, lane, N, mode, z, qp_status, x_ref, ldxref, Qd, Qfd, Rd, Xb, Ub, X_prev, U_prev, cost, converged,
                   active, n_iter, qp_it, qp_iters);
}


// ------------------------------------------------------------------------------------------------ fused-step kernels
// Prologue of one batched solve, one thread per (scenario, stage row k <= N): the reference trajectory row (a given X_ref,
// the tiled target of gp_mpc.py:236-245, or the per-step target the Monte Carlo loop derives for `.solve` controllers,
// monte_carlo.py:497-499), the initial controls (warm start | U_ref | hover along axis 0, gp_mpc.py:266-275) and the
// per-scenario flags.
__global__ void __launch_bounds__(128) solve_prologue3_kernel(int nbatch, int N, int target_mode, double g0,
                                                              const double* __restrict__ x0,
                                                              const double* __restrict__ x_target,
                                                              const double* __restrict__ X_ref, long ld_xref_in,
                                                              const double* __restrict__ U_ref, long ld_uref,
                                                              const int* __restrict__ active_in,
                                                              const int* __restrict__ has_warm,
                                                              const double* __restrict__ U_warm,
                                                              double* __restrict__ x_ref, double* __restrict__ U,
                                                              int* __restrict__ active, int* __restrict__ n_iter,
                                                              int* __restrict__ qp_iters, int* __restrict__ converged) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbatch * (N + 1)) return;
  const int b = idx / (N + 1), k = idx - b * (N + 1);
  double* xr = x_ref + ((long)b * (N + 1) + k) * NX;
  if (X_ref) {
#pragma unroll
    for (int i = 0; i < NX; ++i) xr[i] = X_ref[(long)b * ld_xref_in + (long)k * NX + i];
  } else if (target_mode == 0) {
#pragma unroll
    for (int i = 0; i < NX; ++i) xr[i] = x_target[(long)b * NX + i];
  } else {  // monte_carlo.py:497-499: x_target = x; x_target[4:7] = 0; x_target[1] = max(0.5, altitude - 2.0)
    const double* x = x0 + (long)b * NX;
    xr[0] = x[0];
    xr[1] = fmax(0.5, x[1] - 2.0);
    xr[2] = x[2];
    xr[3] = x[3];
    xr[4] = xr[5] = xr[6] = 0.0;
  }
  if (k < N) {
    double* u = U + ((long)b * N + k) * NU;
    if (has_warm && has_warm[b]) {
#pragma unroll
      for (int i = 0; i < NU; ++i) u[i] = U_warm[((long)b * N + k) * NU + i];
    } else if (U_ref) {
#pragma unroll
      for (int i = 0; i < NU; ++i) u[i] = U_ref[(long)b * ld_uref + (long)k * NU + i];
    } else {
      u[0] = x0[(long)b * NX] * g0;
      u[1] = 0.0;
      u[2] = 0.0;
    }
  } else {
    active[b] = active_in ? (active_in[b] != 0) : 1;
    n_iter[b] = 0;
    qp_iters[b] = 0;
    converged[b] = 0;
  }
}

// GP variances from the per-tile partial sums of |L^-1 k*|^2 (fixed summation order, as sum_parts_kernel), the floor and
// the output scale of exact_gp.py:259-263 (as predict_var_kernel), then the covariance chain of propagate_cov3_kernel:
// one launch instead of three, same arithmetic.
__global__ void __launch_bounds__(128) var_cov3_kernel(GpmpcDyn3 dyn, int nbatch, int N, const double* __restrict__ part,
                                                       int nparts, long part_stride, double prior, double floor_,
                                                       const double* __restrict__ ystd, double sigma0,
                                                       const double* __restrict__ X, const double* __restrict__ U,
                                                       double* __restrict__ dvar, double* __restrict__ covs,
                                                       const int* __restrict__ active) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  double S[NX][NX], Mx[NX][NX];
#pragma unroll
  for (int i = 0; i < NX; ++i)
#pragma unroll
    for (int j = 0; j < NX; ++j) S[i][j] = (i == j) ? sigma0 : 0.0;
  double* out = covs + (long)b * (N + 1) * NX * NX;
#pragma unroll
  for (int i = 0; i < NX; ++i)
#pragma unroll
    for (int j = 0; j < NX; ++j) out[i * NX + j] = S[i][j];
  double ys2[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) ys2[i] = ystd[i] * ystd[i];
  for (int k = 0; k < N; ++k) {
    const long qrow = (long)b * N + k;
    double ss = 0.0;
    for (int t = 0; t < nparts; ++t) ss += part[(long)t * part_stride + qrow];
    const double vq = fmax(prior - ss, floor_);
    double v[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      v[i] = vq * ys2[i];
      dvar[qrow * 3 + i] = v[i];
    }
    const double* x = X + ((long)b * (N + 1) + k) * NX;
    const double* u = U + ((long)b * N + k) * NU;
    const double m = x[0];
    double a[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) a[i] = -u[i] / (m * m) * dyn.dt;
#pragma unroll
    for (int j = 0; j < NX; ++j) {
      Mx[0][j] = S[0][j];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        Mx[1 + i][j] = S[1 + i][j] + dyn.dt * S[4 + i][j];
        Mx[4 + i][j] = a[i] * S[0][j] + S[4 + i][j];
      }
    }
#pragma unroll
    for (int r = 0; r < NX; ++r) {
      S[r][0] = Mx[r][0];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        S[r][1 + i] = Mx[r][1 + i] + Mx[r][4 + i] * dyn.dt;
        S[r][4 + i] = Mx[r][0] * a[i] + Mx[r][4 + i];
      }
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) S[4 + i][4 + i] = S[4 + i][4 + i] + v[i] * (dyn.dt * dyn.dt);  // uncertainty_prop.py:157-160
    double* o = out + (long)(k + 1) * NX * NX;
#pragma unroll
    for (int i = 0; i < NX; ++i)
#pragma unroll
      for (int j = 0; j < NX; ++j) o[i * NX + j] = S[i][j];
  }
}

// Closed-loop advance (see GpmpcMcStep3 in the header), one thread per scenario.
__global__ void __launch_bounds__(128) mc_advance3_kernel(GpmpcMcStep3 a) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  if (!a.active[b]) return;
  double xl[NX], xn[NX];
#pragma unroll
  for (int d = 0; d < NX; ++d) xl[d] = a.x[(long)b * NX + d];
  int code = 0;
  bool stepped = false;
  if (a.classify_only) {
    code = classify3(xl, a.m0[b]);
#pragma unroll
    for (int d = 0; d < NX; ++d) xn[d] = xl[d];
  } else {
    const bool ok = a.success == nullptr || a.success[b] != 0;
    if (!ok && a.fail_steps) a.fail_steps[b] += 1;
    if (a.counts) {  // counts[0]: solves consumed by the loop, counts[7]: their ADMM iterations
      atomicAdd(a.counts, 1ULL);
      if (a.qp_iters) atomicAdd(a.counts + 7, (unsigned long long)a.qp_iters[b]);
    }
    if (!ok && a.fail_policy == 1) {  // monte_carlo.py:507-510: "MPC solve failed" -> DIVERGENCE, no plant step
      code = 6;
#pragma unroll
      for (int d = 0; d < NX; ++d) xn[d] = xl[d];
    } else {
      double ul[NU];
#pragma unroll
      for (int d = 0; d < NU; ++d) ul[d] = a.u[(long)b * a.ldu + d];
      step3(xl, ul, a.dyn, xn);
      if (a.drag_cda != 0.0) {
        const double rho = 1.225 * exp(-xl[1] / 8500.0);
        const double sp = sqrt(xl[4] * xl[4] + xl[5] * xl[5] + xl[6] * xl[6]);
#pragma unroll
        for (int i = 0; i < 3; ++i) xn[4 + i] = xn[4 + i] + a.dyn.dt * (-0.5 * rho * a.drag_cda * sp * xl[4 + i] / xl[0]);
      }
      stepped = true;
      const int t = a.t_step[b] + 1;
      a.t_step[b] = t;
      // the reference classifies x_0 .. x_{max_steps-1} at the top of its loop; the state after the last plant step is
      // not inspected: for/else -> TIMEOUT (monte_carlo.py:455,539-542)
      code = (t >= a.max_steps) ? 5 : classify3(xn, a.m0[b]);
    }
  }
  if (stepped)
#pragma unroll
    for (int d = 0; d < NX; ++d) a.x[(long)b * NX + d] = xn[d];
  if (code == 0) return;
  a.outcome[b] = code;
  if (a.x_final)
#pragma unroll
    for (int d = 0; d < NX; ++d) a.x_final[(long)b * NX + d] = xn[d];
  if (a.counts) atomicAdd(a.counts + code, 1ULL);
  if (!a.respawn) {
    a.active[b] = 0;
    return;
  }
  // respawn: the next run of this slot starts from its initial condition with a fresh controller (monte_carlo.py:606-616
  // builds the loop around run_single; GPMPC.reset_warm_start semantics per scenario)
#pragma unroll
  for (int d = 0; d < NX; ++d) a.x[(long)b * NX + d] = a.x_init[(long)b * NX + d];
  a.t_step[b] = 0;
  if (a.has_warm) a.has_warm[b] = 0;
  if (a.rho) a.rho[b] = 0.0;
  if (a.y)
    for (int e = 0; e < a.m; ++e) a.y[(long)b * a.m + e] = 0.0;
}


// ------------------------------------------------------------------------------------------------ predictive safety filter
// src/safety/safety_filter.py:184-361 for the 7-state model, one thread per scenario (see GpmpcSafety3 in the header).
struct SafetyEval {
  const GpmpcSafety3& s;
  __device__ __forceinline__ void backup(const double* x, double* u) const {   // backup_controller.py:231-252
#pragma unroll
    for (int r = 0; r < NU; ++r) {
      double acc = 0.0;
#pragma unroll
      for (int c = 0; c < NX; ++c) acc += s.K[r * NX + c] * (x[c] - s.x_eq[c]);
      double v = s.u_eq[r] - acc;
      if (s.has_u_limits) v = fmin(fmax(v, s.u_min[r]), s.u_max[r]);
      u[r] = v;
    }
  }
  __device__ __forceinline__ bool constraints(const double* x, const double* u) const {   // safety_filter.py:315-346
    const double T = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
    if (T < s.T_min || T > s.T_max) return false;
    if (x[1] > 0.1) {
      if ((x[1] * x[1]) * (s.tan_gamma * s.tan_gamma) < x[2] * x[2] + x[3] * x[3]) return false;
    }
    return true;
  }
  __device__ __forceinline__ double value(const double* x) const {   // invariant_sets.py:178-184
    double dx[NX], t[NX];
#pragma unroll
    for (int c = 0; c < NX; ++c) dx[c] = x[c] - s.x_eq[c];
#pragma unroll
    for (int c = 0; c < NX; ++c) {   // dx' P first (row vector times matrix), then . dx -- numpy's  dx.T @ P @ dx
      double acc = 0.0;
#pragma unroll
      for (int r = 0; r < NX; ++r) acc += dx[r] * s.P[r * NX + c];
      t[c] = acc;
    }
    double v = 0.0;
#pragma unroll
    for (int c = 0; c < NX; ++c) v += t[c] * dx[c];
    return v;
  }
  // `_predict_with_backup` :281-313 fused with the path-constraint scan of `_check_safety` :265-271; returns V(x_N)
  __device__ __forceinline__ double predict(const double* x0, const double* u_first, bool* path_ok) const {
    double x[NX], xn[NX], u[NU];
#pragma unroll
    for (int c = 0; c < NX; ++c) x[c] = x0[c];
    bool ok = true;
    for (int k = 0; k < s.N; ++k) {
      if (k == 0) {
#pragma unroll
        for (int r = 0; r < NU; ++r) u[r] = u_first[r];
      } else {
        backup(x, u);
      }
      if (path_ok) ok = ok && constraints(x, u);
      step3(x, u, s.dyn, xn);
#pragma unroll
      for (int c = 0; c < NX; ++c) x[c] = xn[c];
    }
    if (path_ok) *path_ok = ok;
    return value(x);
  }
};

__global__ void __launch_bounds__(128) safety_filter3_kernel(GpmpcSafety3 s, int nbatch, const double* __restrict__ X,
                                                             const double* __restrict__ U_nom, long ldu,
                                                             double* __restrict__ U_safe, int* __restrict__ flags,
                                                             double* __restrict__ margin, double* __restrict__ backup_value,
                                                             const int* __restrict__ active) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  const SafetyEval ev{s};
  double x[NX], un[NU], u[NU];
#pragma unroll
  for (int c = 0; c < NX; ++c) x[c] = X[(long)b * NX + c];
#pragma unroll
  for (int r = 0; r < NU; ++r) un[r] = u[r] = U_nom[(long)b * ldu + r];
  // `_check_safety` of the nominal control
  bool path_ok;
  double V = ev.predict(x, un, &path_ok);
  bool safe = ev.constraints(x, un) && (s.alpha - V > 0.0) && path_ok;
  int fl = 0;
  if (!safe) {
    fl = 1;
    // `_solve_gradient_qp` :483-531
    for (int it = 0; it < s.n_iters; ++it) {
      const double Vt = ev.predict(x, u, nullptr);
      if (Vt <= s.alpha * s.backup_margin) break;
      const double eps = 1e-4;
      double g[NU];
#pragma unroll
      for (int r = 0; r < NU; ++r) {
        double up[NU] = {u[0], u[1], u[2]};
        up[r] += eps;
        g[r] = (ev.predict(x, up, nullptr) - Vt) / eps;
      }
#pragma unroll
      for (int r = 0; r < NU; ++r) u[r] = u[r] - s.lr * (g[r] + 2.0 * (u[r] - un[r]));
      const double T = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
      if (T < s.T_min) {
        const double d = fmax(T, 1e-6);
#pragma unroll
        for (int r = 0; r < NU; ++r) u[r] = u[r] / d * s.T_min;
      } else if (T > s.T_max) {
#pragma unroll
        for (int r = 0; r < NU; ++r) u[r] = u[r] / T * s.T_max;
      }
    }
    V = ev.predict(x, u, &path_ok);
    safe = ev.constraints(x, u) && (s.alpha - V > 0.0) && path_ok;
  }
  if (safe) fl |= 2;
#pragma unroll
  for (int r = 0; r < NU; ++r) U_safe[(long)b * NU + r] = u[r];
  if (flags) flags[b] = fl;
  if (margin) margin[b] = s.alpha - V;
  if (backup_value) backup_value[b] = V;
}


// ------------------------------------------------------------------------------------------------ Monte Carlo propagation
// uncertainty_prop.py:295-315, one warp per scenario (see gpmpc_mc_combine_3dof in the header).
__global__ void __launch_bounds__(128) mc_combine3_kernel(int nbatch, int S, const double* __restrict__ Xn,
                                                          const double* __restrict__ var, const double* __restrict__ xi,
                                                          double dt, double* __restrict__ particles,
                                                          double* __restrict__ mean, double* __restrict__ cov) {
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= nbatch) return;
  double sum[NX];
#pragma unroll
  for (int c = 0; c < NX; ++c) sum[c] = 0.0;
  for (int i = lane; i < S; i += 32) {
    const long pi = (long)b * S + i;
    const double* img = Xn + pi * 2 * NX + NX;
    double x[NX];
#pragma unroll
    for (int c = 0; c < NX; ++c) x[c] = img[c];
#pragma unroll
    for (int c = 0; c < 3; ++c) x[4 + c] = x[4 + c] + (sqrt(var[pi * 3 + c]) * xi[pi * 3 + c]) * dt;
#pragma unroll
    for (int c = 0; c < NX; ++c) {
      particles[pi * NX + c] = x[c];
      sum[c] += x[c];
    }
  }
  double m[NX];
#pragma unroll
  for (int c = 0; c < NX; ++c) m[c] = warp_sum(sum[c]) / (double)S;
  double acc[NX][NX];
#pragma unroll
  for (int r = 0; r < NX; ++r)
#pragma unroll
    for (int c = 0; c < NX; ++c) acc[r][c] = 0.0;
  __syncwarp();
  for (int i = lane; i < S; i += 32) {
    const double* x = particles + ((long)b * S + i) * NX;
    double d[NX];
#pragma unroll
    for (int c = 0; c < NX; ++c) d[c] = x[c] - m[c];
#pragma unroll
    for (int r = 0; r < NX; ++r)
#pragma unroll
      for (int c = r; c < NX; ++c) acc[r][c] += d[r] * d[c];
  }
#pragma unroll
  for (int r = 0; r < NX; ++r)
#pragma unroll
    for (int c = r; c < NX; ++c) {
      const double v = warp_sum(acc[r][c]) / (double)(S - 1);
      if (lane == 0) {
        cov[((long)b * NX + r) * NX + c] = v;
        cov[((long)b * NX + c) * NX + r] = v;
      }
    }
  if (lane < NX) mean[(long)b * NX + lane] = m[lane];
}

}  // namespace

// ================================================================================================== C ABI
extern "C" int gpmpc_rollout_3dof(const GpmpcExactModel* model, const GpmpcDyn3* dyn, int B, int N, const double* x0,
                                  const double* U, double* X, double* dmean, double* Ks, long ldk, const int* active,
                                  void* stream) {
  GPMPC_REQUIRE(dyn != nullptr && B >= 0 && N >= 1, "gpmpc_rollout_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  GpmpcExactModel M{};
  int has_model = 0;
  if (model) {
    M = *model;
    has_model = 1;
    GPMPC_REQUIRE(M.D == 11 && M.DP >= 12 && M.DP <= MAXDP && M.DP % 4 == 0, "gpmpc_rollout_3dof: model must have D=11, DP%4==0");
    GPMPC_REQUIRE(M.P >= 3 && M.P <= MAXP, "gpmpc_rollout_3dof: model must have 3..6 outputs");
    GPMPC_REQUIRE(Ks == nullptr || ldk >= M.n_pad, "gpmpc_rollout_3dof: ldk < n_pad");
  }
  const int grid = (B + ROLL_WARPS - 1) / ROLL_WARPS;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t smem = sizeof(double) * (size_t)M.n_pad * (12 + 1 + 3);
  if (has_model && M.DP == 12 && M.P == 3 && smem <= 200 * 1024) {
    // one wave: one CTA per SM, W warps each, two scenarios per warp
    static GpmpcPerDevice n_sm_dev;
    if (n_sm_dev.here() == 0) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&n_sm_dev.here(), cudaDevAttrMultiProcessorCount, dev);
      cudaFuncSetAttribute(rollout3_smem_kernel<12, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    const int n_sm = n_sm_dev.here();
    const int ntask = (B + 1) / 2;
    const int passes = (ntask + n_sm * ROLLS_MAX_WARPS - 1) / (n_sm * ROLLS_MAX_WARPS);
    int W = (ntask + n_sm * passes - 1) / (n_sm * passes);
    W = W < 1 ? 1 : (W > ROLLS_MAX_WARPS ? ROLLS_MAX_WARPS : W);
    const int ctas = (ntask + W - 1) / W < n_sm ? (ntask + W - 1) / W : n_sm;
    rollout3_smem_kernel<12, 3><<<ctas, W * 32, smem, s>>>(M, *dyn, B, N, x0, U, X, dmean, Ks, ldk, active);
  } else if (has_model && M.DP == 12)
    rollout3_kernel<12><<<grid, ROLL_WARPS * 32, 0, s>>>(M, has_model, *dyn, B, N, x0, U, X, dmean, Ks, ldk, active);
  else
    rollout3_kernel<0><<<grid, ROLL_WARPS * 32, 0, s>>>(M, has_model, *dyn, B, N, x0, U, X, dmean, Ks, ldk, active);
  return gpmpc_check_launch("rollout_3dof");
}

extern "C" int gpmpc_gp_mean_points_3dof(const GpmpcExactModel* model, int B, int N, const double* X, const double* U,
                                        double* dmean, double* Ks, long ldk, const int* active, void* stream) {
  GPMPC_REQUIRE(model != nullptr && B >= 0 && N >= 1 && dmean != nullptr, "gpmpc_gp_mean_points_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  GpmpcExactModel M = *model;
  GPMPC_REQUIRE(M.D == 11 && M.DP >= 12 && M.DP <= MAXDP && M.DP % 4 == 0, "gpmpc_gp_mean_points_3dof: model must have D=11, DP%4==0");
  GPMPC_REQUIRE(M.P >= 3 && M.P <= MAXP, "gpmpc_gp_mean_points_3dof: model must have 3..6 outputs");
  GPMPC_REQUIRE(Ks == nullptr || ldk >= M.n_pad, "gpmpc_gp_mean_points_3dof: ldk < n_pad");
  const long pts = (long)B * N;
  const int grid = (int)((pts + ROLL_WARPS - 1) / ROLL_WARPS);
  cudaStream_t s = (cudaStream_t)stream;
  if (M.DP == 12) gp_points3_kernel<12><<<grid, ROLL_WARPS * 32, 0, s>>>(M, B, N, X, U, dmean, Ks, ldk, active);
  else gp_points3_kernel<0><<<grid, ROLL_WARPS * 32, 0, s>>>(M, B, N, X, U, dmean, Ks, ldk, active);
  return gpmpc_check_launch("gp_mean_points_3dof");
}

extern "C" int gpmpc_linearize_3dof(const GpmpcDyn3* dyn, int npts, int lin_mode, const double* X, const double* U,
                                    double* AB, void* stream) {
  GPMPC_REQUIRE(dyn != nullptr && npts >= 0 && (lin_mode == 0 || lin_mode == 1), "gpmpc_linearize_3dof: bad arguments");
  if (npts == 0) return GPMPC_OK;
  linearize3_kernel<<<(npts + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*dyn, npts, lin_mode, X, U, AB);
  return gpmpc_check_launch("linearize_3dof");
}

extern "C" int gpmpc_pack_traj_3dof(int B, int N, const double* X, const double* U, double* z, const int* active,
                                    void* stream) {
  GPMPC_REQUIRE(B >= 0 && N >= 1, "gpmpc_pack_traj_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  const long total = (long)B * (N * NS + NX);
  GPMPC_PREFER_SHARED(pack_traj3_kernel);
  pack_traj3_kernel<<<(int)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, N, X, U, z, active);
  return gpmpc_check_launch("pack_traj_3dof");
}

// Per-scenario controller reset (a fresh GPMPC instance for a respawned Monte Carlo run): hover warm start, zero duals,
// rho back to "unset".  One thread per (scenario, element) of the longest array.
__global__ void __launch_bounds__(256) reset_scenarios_kernel(int nbatch, int N, int nu, int m, const int* __restrict__ mask,
                                                              const double* __restrict__ hover0, double* __restrict__ Uw,
                                                              double* __restrict__ y, double* __restrict__ rho,
                                                              int* __restrict__ has_warm) {
  const int per = max(m, N * nu);
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long)nbatch * per) return;
  const int b = (int)(idx / per), e = (int)(idx - (long)b * per);
  if (!mask[b]) return;
  if (Uw && hover0 && e < N * nu) Uw[(long)b * N * nu + e] = (e % nu == 0) ? hover0[b] : 0.0;
  if (e < m) y[(long)b * m + e] = 0.0;
  if (e == 0) {
    rho[b] = 0.0;
    if (has_warm) has_warm[b] = 0;
  }
}

extern "C" int gpmpc_reset_scenarios(int B, int N, int nu, int m, const int* mask, const double* hover0, double* U_warm,
                                     double* y, double* rho, int* has_warm, void* stream) {
  GPMPC_REQUIRE(B >= 0 && N >= 1 && nu >= 1 && m >= 1 && mask && y && rho, "gpmpc_reset_scenarios: bad arguments");
  if (B == 0) return GPMPC_OK;
  const long total = (long)B * (m > N * nu ? m : N * nu);
  reset_scenarios_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, N, nu, m, mask, hover0, U_warm,
                                                                                            y, rho, has_warm);
  return gpmpc_check_launch("reset_scenarios");
}

extern "C" int gpmpc_ut_sigma_3dof(int B, double scale, const double* xk, const double* Sk, const double* U, long ldu,
                                   double* SP, double* Usp, int* info, void* stream) {
  GPMPC_REQUIRE(B >= 0 && xk && Sk && U && SP && Usp && info, "gpmpc_ut_sigma_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  ut_sigma3_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, scale, xk, Sk, U, ldu, SP, Usp, info);
  return gpmpc_check_launch("ut_sigma_3dof");
}

extern "C" int gpmpc_ut_combine_3dof(int B, int n_inert, double wm0, double wc0, double wi, const double* Xn,
                                     const double* var, double dt, double* xnext, double* Snext, void* stream) {
  GPMPC_REQUIRE(B >= 0 && n_inert >= 0 && Xn && xnext && Snext, "gpmpc_ut_combine_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  ut_combine3_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, n_inert, wm0, wc0, wi, Xn, var, dt, xnext,
                                                                        Snext);
  return gpmpc_check_launch("ut_combine_3dof");
}

extern "C" int gpmpc_propagate_cov_3dof(const GpmpcDyn3* dyn, int B, int N, const double* X, const double* U,
                                        const double* dvar, double sigma0, double* covs, const int* active,
                                        void* stream) {
  GPMPC_REQUIRE(dyn != nullptr && B >= 0 && N >= 1, "gpmpc_propagate_cov_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(propagate_cov3_kernel);
  propagate_cov3_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*dyn, B, N, X, U, dvar, sigma0, covs, active);
  return gpmpc_check_launch("propagate_cov_3dof");
}

extern "C" int gpmpc_qp_assemble_3dof(const GpmpcDyn3* dyn, int B, int N, int lin_mode, int rhs_mode,
                                      const double* X_lin, const double* U_lin, const double* x_init,
                                      const double* x_ref, long ldxref, const double* c_gp, const double* Qdiag,
                                      const double* Qfdiag, double* AB, double* beq, double* q, const int* active,
                                      void* stream) {
  GPMPC_REQUIRE(dyn != nullptr && B >= 0 && N >= 1, "gpmpc_qp_assemble_3dof: bad arguments");
  GPMPC_REQUIRE(lin_mode >= 0 && lin_mode <= 1 && rhs_mode >= 0 && rhs_mode <= 1, "gpmpc_qp_assemble_3dof: bad mode");
  if (B == 0) return GPMPC_OK;
  const int total = B * (N + 1);
  GPMPC_PREFER_SHARED(qp_assemble3_kernel);
  qp_assemble3_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      *dyn, B, N, lin_mode, rhs_mode, X_lin, U_lin, x_init, x_ref, ldxref, c_gp, Qdiag, Qfdiag, AB, beq, q, active,
      nullptr);
  return gpmpc_check_launch("qp_assemble_3dof");
}

extern "C" int gpmpc_mpc_update_3dof(int B, int N, int mode, const double* z, const int* qp_status,
                                     const double* x_ref, long ldxref, const double* Qdiag, const double* Qfdiag,
                                     const double* Rdiag, double* X, double* U, double* X_prev, double* U_prev,
                                     double* cost, int* converged, int* active, int* n_iter, void* stream) {
  GPMPC_REQUIRE(B >= 0 && N >= 1 && (mode == 0 || mode == 1), "gpmpc_mpc_update_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(mpc_update3_kernel);
  mpc_update3_kernel<<<(B + 3) / 4, 128, 0, (cudaStream_t)stream>>>(B, N, mode, z, qp_status, x_ref, ldxref, Qdiag,
                                                                     Qfdiag, Rdiag, X, U, X_prev, U_prev, cost,
                                                                     converged, active, n_iter, nullptr, nullptr, nullptr,
                                                                     nullptr, nullptr);
  return gpmpc_check_launch("mpc_update_3dof");
}

// ================================================================================================== fused step (C ABI)
extern "C" int gpmpc_events_create(int n, void** events_out) {
  GPMPC_REQUIRE(n >= 0 && (n == 0 || events_out), "gpmpc_events_create: bad arguments");
  for (int i = 0; i < n; ++i) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) {
      gpmpc_set_error("gpmpc_events_create: cudaEventCreate failed");
      return GPMPC_ERR_CUDA;
    }
    events_out[i] = (void*)e;
  }
  return GPMPC_OK;
}

extern "C" int gpmpc_events_destroy(int n, void** events) {
  for (int i = 0; i < n; ++i)
    if (events && events[i]) cudaEventDestroy((cudaEvent_t)events[i]);
  return GPMPC_OK;
}

extern "C" int gpmpc_events_elapsed_ms(void* a, void* b, float* ms_out) {
  GPMPC_REQUIRE(a && b && ms_out, "gpmpc_events_elapsed_ms: bad arguments");
  if (cudaEventSynchronize((cudaEvent_t)b) != cudaSuccess || cudaEventElapsedTime(ms_out, (cudaEvent_t)a, (cudaEvent_t)b) != cudaSuccess) {
    cudaGetLastError();
    gpmpc_set_error("gpmpc_events_elapsed_ms: event not recorded");
    return GPMPC_ERR_CUDA;
  }
  return GPMPC_OK;
}

extern "C" int gpmpc_mc_combine_3dof(int B, int S, const double* Xn, const double* var, const double* xi, double dt,
                                     double* particles, double* mean, double* cov, void* stream) {
  GPMPC_REQUIRE(B >= 0 && S >= 2 && Xn && var && xi && particles && mean && cov, "gpmpc_mc_combine_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(mc_combine3_kernel);
  mc_combine3_kernel<<<(B + 3) / 4, 128, 0, (cudaStream_t)stream>>>(B, S, Xn, var, xi, dt, particles, mean, cov);
  return gpmpc_check_launch("mc_combine_3dof");
}

extern "C" int gpmpc_safety_filter_3dof(const GpmpcSafety3* setup, int B, const double* X, const double* U_nom, long ldu,
                                        double* U_safe, int* flags, double* margin, double* backup_value, const int* active,
                                        void* stream) {
  GPMPC_REQUIRE(setup != nullptr && B >= 0 && X && U_nom && U_safe && ldu >= NU, "gpmpc_safety_filter_3dof: bad arguments");
  GPMPC_REQUIRE(setup->N >= 1 && setup->n_iters >= 0, "gpmpc_safety_filter_3dof: bad horizon / iteration count");
  if (B == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(safety_filter3_kernel);
  safety_filter3_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*setup, B, X, U_nom, ldu, U_safe, flags, margin,
                                                                            backup_value, active);
  return gpmpc_check_launch("safety_filter_3dof");
}

extern "C" int gpmpc_mc_advance_3dof(const GpmpcMcStep3* a, void* stream) {
  GPMPC_REQUIRE(a != nullptr && a->B >= 0, "gpmpc_mc_advance_3dof: bad arguments");
  GPMPC_REQUIRE(a->x && a->m0 && a->t_step && a->active && a->outcome, "gpmpc_mc_advance_3dof: NULL state buffer");
  GPMPC_REQUIRE(a->classify_only || a->u, "gpmpc_mc_advance_3dof: u is NULL");
  GPMPC_REQUIRE(!a->respawn || a->x_init, "gpmpc_mc_advance_3dof: respawn needs x_init");
  if (a->B == 0) return GPMPC_OK;
  GPMPC_PREFER_SHARED(mc_advance3_kernel);
  mc_advance3_kernel<<<(a->B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*a);
  return gpmpc_check_launch("mc_advance_3dof");
}

namespace {
struct PhaseRec {
  const GpmpcSolve3* p;
  cudaStream_t s;
  void mark(int phase) const {
    if (!p->events || !p->event_count) return;
    const int i = *p->event_count;
    if (i >= p->event_cap) return;
    cudaEventRecord((cudaEvent_t)p->events[i], s);
    p->event_phase[i] = phase;
    *p->event_count = i + 1;
  }
};
}  // namespace

extern "C" int gpmpc_qp_augment_cone_3dof(const GpmpcCone3* cone, int B, int N, const double* AB, const double* beq,
                                          const double* q, const double* X_nom, const double* U_nom, double* AB_c,
                                          double* beq_c, double* q_c, double* z_c, const int* active, void* stream) {
  GPMPC_REQUIRE(cone && B >= 0 && N >= 1 && AB && beq && q && X_nom && U_nom && AB_c && beq_c && q_c,
                "gpmpc_qp_augment_cone_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  const int total = B * (N + 1);
  GPMPC_PREFER_SHARED(qp_augment_cone3_kernel);
  qp_augment_cone3_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*cone, B, N, AB, beq, q, X_nom, U_nom, AB_c,
                                                                                 beq_c, q_c, z_c, active);
  return gpmpc_check_launch("qp_augment_cone_3dof");
}

extern "C" int gpmpc_qp_extract_cone_3dof(int B, int N, const double* z_c, double* z, const int* active, void* stream) {
  GPMPC_REQUIRE(B >= 0 && N >= 1 && z_c && z, "gpmpc_qp_extract_cone_3dof: bad arguments");
  if (B == 0) return GPMPC_OK;
  const int total = B * (N + 1);
  GPMPC_PREFER_SHARED(qp_extract_cone3_kernel);
  qp_extract_cone3_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, N, z_c, z, active);
  return gpmpc_check_launch("qp_extract_cone_3dof");
}

extern "C" int gpmpc_solve_batch_3dof(const GpmpcSolve3* p, int* n_launches, void* stream_) {
  GPMPC_REQUIRE(p != nullptr && p->B >= 0 && p->N >= 1 && p->max_sqp_iter >= 1, "gpmpc_solve_batch_3dof: bad arguments");
  GPMPC_REQUIRE(p->x0 && p->x_ref && p->X && p->U && p->AB && p->beq && p->q && p->z && p->y && p->rho,
                "gpmpc_solve_batch_3dof: NULL state buffer");
  GPMPC_REQUIRE(p->target_mode == 1 || p->x_target || p->X_ref, "gpmpc_solve_batch_3dof: no target");
  cudaStream_t s = (cudaStream_t)stream_;
  const int B = p->B, N = p->N;
  int launches = 0;
  if (n_launches) *n_launches = 0;
  if (B == 0) return GPMPC_OK;
  const bool use_mean = p->has_model && p->use_gp_mean;
  const bool use_unc = p->has_model && p->use_gp_uncertainty;
  GPMPC_REQUIRE(!use_unc || (p->Ks && p->sumsq && p->dvar && p->covs), "gpmpc_solve_batch_3dof: uncertainty buffers are NULL");
  GPMPC_REQUIRE(!use_unc || p->Linv == nullptr || p->var_work, "gpmpc_solve_batch_3dof: var_work is NULL");
  GPMPC_REQUIRE(!p->has_model || p->dmean, "gpmpc_solve_batch_3dof: dmean is NULL");
  const bool tighten = use_unc && p->tighten_kappa > 0.0;
  GPMPC_REQUIRE(!tighten || (p->lb_t && p->ub_t), "gpmpc_solve_batch_3dof: tightened boxes are NULL");
  GPMPC_REQUIRE(!p->cone.enabled || (p->AB_c && p->beq_c && p->q_c && p->z_c && p->y_c && p->Pdiag_c && p->lb_c && p->ub_c),
                "gpmpc_solve_batch_3dof: convex-row buffers are NULL");
  GPMPC_REQUIRE(!p->cone.enabled || !tighten, "gpmpc_solve_batch_3dof: convex rows and per-stage tightening are exclusive");
  PhaseRec rec{p, s};
  const int total = B * (N + 1);
  int rc;
  rec.mark(GPMPC_PHASE_BEGIN);
  GPMPC_PREFER_SHARED(solve_prologue3_kernel);
  solve_prologue3_kernel<<<(total + 127) / 128, 128, 0, s>>>(B, N, p->target_mode, p->g0, p->x0, p->x_target, p->X_ref,
                                                             p->ld_xref_in, p->U_ref, p->ld_uref, p->active_in,
                                                             p->has_warm, p->U_warm, p->x_ref, p->U, p->active, p->n_iter,
                                                             p->qp_iters, p->converged);
  if ((rc = gpmpc_check_launch("solve_prologue_3dof"))) return rc;
  ++launches;
  rec.mark(GPMPC_PHASE_PROLOGUE);
  const GpmpcExactModel* mdl = p->has_model ? &p->model : nullptr;
  // variances of the rows of Ks + covariance chain along (X, U)
  auto uncertainty = [&](const double* Xt) -> int {
    int r;
    const double* part;
    int nparts;
    if (p->Linv) {
      if ((r = gpmpc_linv_sumsq_parts_n(p->Linv, p->model.n, p->model.n_pad, p->ldi, p->Ks, p->q_pad, p->ldk, p->var_work,
                                        (void*)s)))
        return r;
      part = p->var_work;
      nparts = p->model.n_pad / 64;
      ++launches;
    } else {
      if ((r = gpmpc_trsm_sumsq(p->model.L, p->model.n_pad, p->model.n_pad, p->model.Winv, p->Ks, p->q_pad, p->ldk,
                                p->sumsq, (void*)s)))
        return r;
      part = p->sumsq;
      nparts = 1;
      launches += 2 * (p->model.n_pad / 64);
    }
    rec.mark(GPMPC_PHASE_GP_VARIANCE);
    GPMPC_PREFER_SHARED(var_cov3_kernel);
    // one warp per CTA: the kernel is a latency chain per thread (20 sequential 7x7 products), so the 4096 threads are spread
    // over 128 SMs instead of 32
    var_cov3_kernel<<<(B + 31) / 32, 32, 0, s>>>(p->dyn, B, N, part, nparts, (long)p->q_pad, p->model.signal_var, 1e-10,
                                                    p->model.ystd, 1e-6, Xt, p->U, p->dvar, p->covs, p->active);
    if ((r = gpmpc_check_launch("var_cov_3dof"))) return r;
    ++launches;
    rec.mark(GPMPC_PHASE_COV);
    return GPMPC_OK;
  };
  // 1. rollout with the GP mean (gp_mpc.py:277-281)
  rc = gpmpc_rollout_3dof((use_mean || use_unc) ? mdl : nullptr, &p->dyn, B, N, p->x0, p->U, p->X,
                          p->has_model ? p->dmean : nullptr, use_unc ? p->Ks : nullptr, p->ldk, p->active, (void*)s);
  if (rc) return rc;
  ++launches;
  if (p->has_model && !use_mean && use_unc) {  // uncertainty only: the GP mean must not enter the prediction
    if ((rc = gpmpc_rollout_3dof(nullptr, &p->dyn, B, N, p->x0, p->U, p->X, nullptr, nullptr, 0, p->active, (void*)s))) return rc;
    ++launches;
  }
  rec.mark(GPMPC_PHASE_ROLLOUT);
  // 2. uncertainty propagation (gp_mpc.py:284-289)
  if (use_unc && (rc = uncertainty(p->X))) return rc;
  // 3. SQP iterations (gp_mpc.py:296-353)
  GpmpcAdmmSettings S = p->admm;
  S.bound_stride = tighten ? (N * NS + NX) : 0;
  for (int it = 0; it < p->max_sqp_iter; ++it) {
    const bool last = it == p->max_sqp_iter - 1;
    if (tighten) {
      if ((rc = gpmpc_tighten_bounds(NX, NU, B, N, p->lb, p->ub, p->covs, p->tighten_kappa, p->tighten_max_variance,
                                     p->lb_t, p->ub_t, nullptr, (void*)s)))
        return rc;
      ++launches;
    }
    if (use_mean && it > 0) {
      if ((rc = gpmpc_gp_mean_points_3dof(mdl, B, N, p->X, p->U, p->dmean, nullptr, 0, p->active, (void*)s))) return rc;
      ++launches;
      rec.mark(GPMPC_PHASE_GP_MEAN);
    }
    GPMPC_PREFER_SHARED(qp_assemble3_kernel);
    qp_assemble3_kernel<<<(total + 127) / 128, 128, 0, s>>>(p->dyn, B, N, 0, 1, p->X, p->U, p->x0, p->x_ref,
                                                            (long)(N + 1) * NX, use_mean ? p->dmean : nullptr, p->Qd,
                                                            p->Qfd, p->AB, p->beq, p->q, p->active, p->z);
    if ((rc = gpmpc_check_launch("qp_assemble_3dof"))) return rc;
    ++launches;
    if (p->cone.enabled) {  // polyhedral thrust / glide-slope rows: solve in the augmented (14, 3) stage form
      if ((rc = gpmpc_qp_augment_cone_3dof(&p->cone, B, N, p->AB, p->beq, p->q, p->X, p->U, p->AB_c, p->beq_c, p->q_c, p->z_c,
                                           p->active, (void*)s)))
        return rc;
      ++launches;
      rec.mark(GPMPC_PHASE_ASSEMBLE);
      if ((rc = gpmpc_admm_solve(NXC, NU, N, B, p->AB_c, p->beq_c, p->q_c, p->Pdiag_c, p->lb_c, p->ub_c, &S, p->active,
                                 p->z_c, p->y_c, p->rho, p->qp_status, p->qp_it, p->obj, p->pri, p->dua, (void*)s)))
        return rc;
      ++launches;
      rec.mark(GPMPC_PHASE_ADMM);
      if ((rc = gpmpc_qp_extract_cone_3dof(B, N, p->z_c, p->z, p->active, (void*)s))) return rc;
      ++launches;
    } else {
      rec.mark(GPMPC_PHASE_ASSEMBLE);
      if ((rc = gpmpc_admm_solve(NX, NU, N, B, p->AB, p->beq, p->q, p->Pdiag, tighten ? p->lb_t : p->lb,
                                 tighten ? p->ub_t : p->ub, &S, p->active, p->z, p->y, p->rho, p->qp_status, p->qp_it,
                                 p->obj, p->pri, p->dua, (void*)s)))
        return rc;
      ++launches;
      rec.mark(GPMPC_PHASE_ADMM);
    }
    GPMPC_PREFER_SHARED(mpc_update3_kernel);
    mpc_update3_kernel<<<(B + 3) / 4, 128, 0, s>>>(B, N, 1, p->z, p->qp_status, p->x_ref, (long)(N + 1) * NX, p->Qd, p->Qfd,
                                                   p->Rd, p->X, p->U, nullptr, nullptr, p->cost, p->converged, p->active,
                                                   p->n_iter, p->qp_it, p->qp_iters, last ? p->U_warm : nullptr,
                                                   last ? p->X_warm : nullptr, last ? p->has_warm : nullptr);
    if ((rc = gpmpc_check_launch("mpc_update_3dof"))) return rc;
    ++launches;
    rec.mark(GPMPC_PHASE_UPDATE);
    if (last) break;
    if (use_unc && p->repropagate) {  // gp_mpc.py:347-353
      GPMPC_REQUIRE(p->Xs && p->dmean_s, "gpmpc_solve_batch_3dof: re-propagation scratch is NULL");
      if ((rc = gpmpc_rollout_3dof(mdl, &p->dyn, B, N, p->x0, p->U, p->Xs, p->dmean_s, p->Ks, p->ldk, p->active, (void*)s)))
        return rc;
      ++launches;
      rec.mark(GPMPC_PHASE_ROLLOUT);
      if ((rc = uncertainty(p->Xs))) return rc;
    }
  }
  if (n_launches) *n_launches = launches;
  return GPMPC_OK;
}

// Layout probes for the binding tests (tests/test_abi.py): which = 0 GpmpcSolve3, 1 GpmpcMcStep3, 2 GpmpcAdmmSettings,
// 3 GpmpcExactModel; field < 0: sizeof, else the offset of a sentinel field.
#include <stddef.h>
extern "C" long gpmpc_struct_layout(int which, int field) {
  if (which == 0) {
    switch (field) {
      case -1: return (long)sizeof(GpmpcSolve3);
      case 0: return (long)offsetof(GpmpcSolve3, dyn);
      case 1: return (long)offsetof(GpmpcSolve3, admm);
      case 2: return (long)offsetof(GpmpcSolve3, model);
      case 3: return (long)offsetof(GpmpcSolve3, Linv);
      case 4: return (long)offsetof(GpmpcSolve3, x0);
      case 5: return (long)offsetof(GpmpcSolve3, Ks);
      case 6: return (long)offsetof(GpmpcSolve3, q_pad);
      case 7: return (long)offsetof(GpmpcSolve3, var_work);
      case 8: return (long)offsetof(GpmpcSolve3, U_warm);
      case 9: return (long)offsetof(GpmpcSolve3, events);
      case 10: return (long)offsetof(GpmpcSolve3, event_count);
      case 11: return (long)offsetof(GpmpcSolve3, cone);
      case 12: return (long)offsetof(GpmpcSolve3, Pdiag_c);
    }
  } else if (which == 1) {
    switch (field) {
      case -1: return (long)sizeof(GpmpcMcStep3);
      case 0: return (long)offsetof(GpmpcMcStep3, dyn);
      case 1: return (long)offsetof(GpmpcMcStep3, max_steps);
      case 2: return (long)offsetof(GpmpcMcStep3, x);
      case 3: return (long)offsetof(GpmpcMcStep3, counts);
      case 4: return (long)offsetof(GpmpcMcStep3, rho);
    }
  } else if (which == 4) {
    switch (field) {
      case -1: return (long)sizeof(GpmpcSafety3);
      case 0: return (long)offsetof(GpmpcSafety3, lr);
      case 1: return (long)offsetof(GpmpcSafety3, K);
      case 2: return (long)offsetof(GpmpcSafety3, u_max);
    }
  } else if (which == 2 && field < 0) {
    return (long)sizeof(GpmpcAdmmSettings);
  } else if (which == 3 && field < 0) {
    return (long)sizeof(GpmpcExactModel);
  }
  return -1;
}
