// MPC-side kernels of the 6-DoF quaternion GP-MPC step (BASELINE configs[3]: 14 states [m, r_I(3), v_I(3), q_BI(4), w_B(3)],
// 3 controls, N = 30, sparse structured GP):
//   * discrete model: the continuous dynamics stated at nominal_mpc.py:163-203, Euler / the reference's "rk4" (4 Euler
//     sub-steps, :205-236), quaternion renormalisation (rocket_6dof.py:370-387)
//   * forward-difference linearisation, eps = 1e-6 (osqp_rti.py:374-401)
//   * GP-augmented horizon rollout with the two structured residual GPs (gp_mpc.py:139-175,277-281;
//     structured_gp.py:230-305; posterior mean of a sparse GP = K*u alpha, sparse_gp.py:281)
//   * QP assembly in the OSQP-RTI layout for n_x = 14 (osqp_rti.py:168-201,203-372; affine GP term gp_mpc.py:300-317,402-411)
// `simdyn` (the reference's dynamics backend) is third-party and absent: the in-repo CasADi statement is what is
// implemented (oracle/mpc6.py is the CPU restatement the tests compare against).  Compiled with -fmad=false so the
// element-wise model arithmetic rounds like the numpy restatement.
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

constexpr int NX = 14, NU = 3, NS = 17;

__device__ __forceinline__ void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

// x_dot, nominal_mpc.py:163-203
__device__ __forceinline__ void f6(const double* x, const double* u, const GpmpcDyn6& p, double* xd) {
  const double m = x[0];
  const double qw = x[7], qx = x[8], qy = x[9], qz = x[10];
  const double* w = x + 11;
  double C[3][3];
  C[0][0] = 1 - 2 * (qy * qy + qz * qz); C[0][1] = 2 * (qx * qy - qw * qz);     C[0][2] = 2 * (qx * qz + qw * qy);
  C[1][0] = 2 * (qx * qy + qw * qz);     C[1][1] = 1 - 2 * (qx * qx + qz * qz); C[1][2] = 2 * (qy * qz - qw * qx);
  C[2][0] = 2 * (qx * qz - qw * qy);     C[2][1] = 2 * (qy * qz + qw * qx);     C[2][2] = 1 - 2 * (qx * qx + qy * qy);
  const double Tm = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
  xd[0] = -p.alpha * Tm;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    xd[1 + r] = x[4 + r];
    const double TI = C[r][0] * u[0] + C[r][1] * u[1] + C[r][2] * u[2];
    xd[4 + r] = TI / m + p.g[r];
  }
  const double qv[3] = {qx, qy, qz};
  xd[7] = 0.5 * (-(w[0] * qv[0] + w[1] * qv[1] + w[2] * qv[2]));
  double wxq[3];
  cross3(w, qv, wxq);
#pragma unroll
  for (int r = 0; r < 3; ++r) xd[8 + r] = 0.5 * (qw * w[r] + wxq[r]);
  double Jw[3], tq[3], wJw[3], rhs[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) Jw[r] = p.J[r * 3 + 0] * w[0] + p.J[r * 3 + 1] * w[1] + p.J[r * 3 + 2] * w[2];
  cross3(p.rT, u, tq);
  cross3(w, Jw, wJw);
#pragma unroll
  for (int r = 0; r < 3; ++r) rhs[r] = tq[r] - wJw[r];
#pragma unroll
  for (int r = 0; r < 3; ++r)
    xd[11 + r] = p.Jinv[r * 3 + 0] * rhs[0] + p.Jinv[r * 3 + 1] * rhs[1] + p.Jinv[r * 3 + 2] * rhs[2];
}

__device__ __forceinline__ void step6(const double* x, const double* u, const GpmpcDyn6& p, double* xn) {
  double xc[NX], xd[NX];
#pragma unroll
  for (int d = 0; d < NX; ++d) xc[d] = x[d];
  const double h = p.dt / p.substeps;
  for (int s = 0; s < p.substeps; ++s) {
    f6(xc, u, p, xd);
#pragma unroll
    for (int d = 0; d < NX; ++d) xc[d] = xc[d] + h * xd[d];
  }
  if (p.normalize_quat) {
    const double nq = sqrt(xc[7] * xc[7] + xc[8] * xc[8] + xc[9] * xc[9] + xc[10] * xc[10]);
#pragma unroll
    for (int d = 7; d < 11; ++d) xc[d] = xc[d] / nq;
  }
#pragma unroll
  for (int d = 0; d < NX; ++d) xn[d] = xc[d];
}

// Forward finite differences, osqp_rti.py:374-401: column c of [A | B] from one perturbed step.
__device__ __forceinline__ void fd_column(const double* x, const double* u, const GpmpcDyn6& p, const double* f0, int c,
                                          double* col) {
  const double eps = 1e-6;
  double xp[NX], up[NU], f1[NX];
#pragma unroll
  for (int d = 0; d < NX; ++d) xp[d] = x[d];
#pragma unroll
  for (int d = 0; d < NU; ++d) up[d] = u[d];
  if (c < NX) xp[c] += eps;
  else up[c - NX] += eps;
  step6(xp, up, p, f1);
#pragma unroll
  for (int r = 0; r < NX; ++r) col[r] = (f1[r] - f0[r]) / eps;
}

__global__ void __launch_bounds__(128) step6_kernel(GpmpcDyn6 dyn, int npts, const double* __restrict__ X,
                                                    const double* __restrict__ U, double* __restrict__ Xn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npts) return;
  double x[NX], u[NU], xn[NX];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = X[(long)i * NX + d];
#pragma unroll
  for (int d = 0; d < NU; ++d) u[d] = U[(long)i * NU + d];
  step6(x, u, dyn, xn);
#pragma unroll
  for (int d = 0; d < NX; ++d) Xn[(long)i * NX + d] = xn[d];
}

// one thread per (point, column): 17 perturbed steps per point run in parallel
__global__ void __launch_bounds__(128) linearize6_kernel(GpmpcDyn6 dyn, int npts, const double* __restrict__ X,
                                                         const double* __restrict__ U, double* __restrict__ AB) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= npts * NS) return;
  const int i = idx / NS, c = idx - i * NS;
  double x[NX], u[NU], f0[NX], col[NX];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = X[(long)i * NX + d];
#pragma unroll
  for (int d = 0; d < NU; ++d) u[d] = U[(long)i * NU + d];
  step6(x, u, dyn, f0);
  fd_column(x, u, dyn, f0, c, col);
#pragma unroll
  for (int r = 0; r < NX; ++r) AB[((long)i * NX + r) * NS + c] = col[r];
}

// ------------------------------------------------------------------------------------------------ rollout
// One warp per scenario (the horizon is a dependency chain); per stage the warp evaluates both residual GPs: features,
// scaling and row norm (kernels.py:223-229), cross-covariances against the M inducing points (each lane M/32 of them)
// and the 3 + 3 posterior means (sparse_gp.py:281-284).
constexpr int ROLL6_WARPS = 4;
constexpr int MAXDP6 = 16;

__device__ __forceinline__ void gp_mean_warp6(const GpmpcExactModel& M, int trans, const double* x, const double* u,
                                              double* zs, int lane, double* dm) {
  const int D = trans ? 13 : 12;
  double z[13], sqv[13];
  features_6dof_dev(trans, x, u, z);
  for (int d = 0; d < D; ++d) {
    z[d] = z[d] / M.lengthscales[d];
    sqv[d] = z[d] * z[d];
  }
  const double sqz = numpy_pairwise_sum_small(sqv, D);
  __syncwarp();
  if (lane == 0) {
    for (int d = 0; d < D; ++d) zs[d] = z[d];
    for (int d = D; d < M.DP; ++d) zs[d] = 0.0;
  }
  __syncwarp();
  double acc[3] = {0.0, 0.0, 0.0};
  for (int j = lane; j < M.n; j += 32) {
    const double* xr = M.Xs + (long)j * M.DP;
    double dot = 0.0;
    for (int d = 0; d < M.DP; ++d) dot = fma(zs[d], __ldg(xr + d), dot);
    const double d2 = (sqz + __ldg(M.sq + j)) - 2.0 * dot;
    const double kv = kernel_value(M.kind, d2, M.signal_var);
#pragma unroll
    for (int p = 0; p < 3; ++p)
      if (p < M.P) acc[p] = fma(kv, __ldg(M.alpha + (long)p * M.n_pad + j), acc[p]);
  }
#pragma unroll
  for (int p = 0; p < 3; ++p) dm[p] = (p < M.P) ? warp_sum(acc[p]) * M.ystd[p] + M.ymean[p] : 0.0;
}

__global__ void __launch_bounds__(ROLL6_WARPS * 32) rollout6_kernel(GpmpcExactModel Mv, int has_v, GpmpcExactModel Mw,
                                                                    int has_w, GpmpcDyn6 dyn, int nbatch, int N,
                                                                    const double* __restrict__ x0,
                                                                    const double* __restrict__ U, double* __restrict__ X,
                                                                    double* __restrict__ dmean,
                                                                    const int* __restrict__ active) {
  __shared__ double zs_sm[ROLL6_WARPS][MAXDP6];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * ROLL6_WARPS + warp;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  double* zs = zs_sm[warp];
  double x[NX], xn[NX], u[NU];
#pragma unroll
  for (int d = 0; d < NX; ++d) x[d] = x0[(long)b * NX + d];
  if (lane < NX) X[(long)b * (N + 1) * NX + lane] = x[lane];
  for (int k = 0; k < N; ++k) {
    const long row = (long)b * N + k;
#pragma unroll
    for (int d = 0; d < NU; ++d) u[d] = U[row * NU + d];
    double dv[3] = {0.0, 0.0, 0.0}, dw[3] = {0.0, 0.0, 0.0};
    if (has_v) gp_mean_warp6(Mv, 1, x, u, zs, lane, dv);
    if (has_w) gp_mean_warp6(Mw, 0, x, u, zs, lane, dw);
    step6(x, u, dyn, xn);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      xn[4 + i] = xn[4 + i] + dv[i] * dyn.dt;    // gp_mpc.py:165-167
      xn[11 + i] = xn[11 + i] + dw[i] * dyn.dt;  // gp_mpc.py:168
    }
    if (lane < NX) X[((long)b * (N + 1) + k + 1) * NX + lane] = xn[lane];
    if (dmean && lane < 6) dmean[row * 6 + lane] = (lane < 3) ? dv[lane] : dw[lane - 3];
#pragma unroll
    for (int d = 0; d < NX; ++d) x[d] = xn[d];
  }
}

// ------------------------------------------------------------------------------------------------ QP assembly
// PTS (scenario, stage) points per CTA, 17 threads per point: thread c computes column c of [A_k | B_k] by one perturbed
// step; the affine term is then formed per row from the staged block.  Stage N writes the initial-state rows and the
// terminal cost vector.
constexpr int ASM6_PTS = 7;

__global__ void __launch_bounds__(128) qp_assemble6_kernel(GpmpcDyn6 dyn, int nbatch, int N,
                                                           const double* __restrict__ X_lin,
                                                           const double* __restrict__ U_lin,
                                                           const double* __restrict__ x_init,
                                                           const double* __restrict__ x_ref, long ldxref,
                                                           const double* __restrict__ c_gp, const double* __restrict__ Qd,
                                                           const double* __restrict__ Qfd, double* __restrict__ AB,
                                                           double* __restrict__ beq, double* __restrict__ q,
                                                           const int* __restrict__ active) {
  __shared__ double blk[ASM6_PTS][NX][NS + 1];
  __shared__ double xs[ASM6_PTS][NX + NU];
  const int slot = threadIdx.x / NS, c = threadIdx.x - slot * NS;
  const int pt = blockIdx.x * ASM6_PTS + slot;
  const bool valid = slot < ASM6_PTS && pt < nbatch * (N + 1);
  const int b = valid ? pt / (N + 1) : 0, k = valid ? pt - b * (N + 1) : 0;
  const bool act = valid && (!active || active[b]);
  const int n = N * NS + NX, neq = NX * (N + 1);
  const bool dynrow = act && k < N;
  if (dynrow) {
    double x[NX], u[NU], f0[NX], col[NX];
#pragma unroll
    for (int d = 0; d < NX; ++d) x[d] = X_lin[((long)b * (N + 1) + k) * NX + d];
#pragma unroll
    for (int d = 0; d < NU; ++d) u[d] = U_lin[((long)b * N + k) * NU + d];
    step6(x, u, dyn, f0);
    fd_column(x, u, dyn, f0, c, col);
#pragma unroll
    for (int r = 0; r < NX; ++r) blk[slot][r][c] = col[r];
    xs[slot][c] = (c < NX) ? x[c] : u[c - NX];
  }
  __syncthreads();
  if (!act) return;
  const double* xr = x_ref + (long)b * ldxref + (long)k * NX;
  double* qb = q + (long)b * n + (long)k * NS;
  if (k == N) {
    if (c < NX) {
      qb[c] = -(Qfd[c] * xr[c]);                            // osqp_rti.py:254-256
      beq[(long)b * neq + c] = x_init[(long)b * NX + c];    // :286-288
    }
    return;
  }
  qb[c] = (c < NX) ? -(Qd[c] * xr[c]) : 0.0;               // :249-252
  double* ab = AB + ((long)b * N + k) * NX * NS;
#pragma unroll
  for (int r = 0; r < NX; ++r) {
    const double a = blk[slot][r][c];
    ab[r * NS + c] = (fabs(a) > 1e-10) ? a : 0.0;          // :301,309
  }
  if (c < NX) {
    const int r = c;
    double ax = 0.0, bu = 0.0;
#pragma unroll
    for (int j = 0; j < NX; ++j) ax += blk[slot][r][j] * xs[slot][j];
#pragma unroll
    for (int j = 0; j < NU; ++j) bu += blk[slot][r][NX + j] * xs[slot][NX + j];
    double cv = (X_lin[((long)b * (N + 1) + k + 1) * NX + r] - ax) - bu;    // gp_mpc.py:402-411
    if (c_gp) {
      if (r >= 4 && r < 7) cv = cv + c_gp[((long)b * N + k) * 6 + (r - 4)] * dyn.dt;
      if (r >= 11) cv = cv + c_gp[((long)b * N + k) * 6 + 3 + (r - 11)] * dyn.dt;
    }
    beq[(long)b * neq + NX + (long)k * NX + r] = -cv;
  }
}

__global__ void pack_traj_kernel(int nx, int nu, int nbatch, int N, const double* __restrict__ X,
                                 const double* __restrict__ U, double* __restrict__ z, int unpack, double* Xo, double* Uo) {
  const int ns = nx + nu, n = N * ns + nx;
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long)nbatch * n) return;
  const int b = (int)(idx / n), j = (int)(idx - (long)b * n);
  const int k = j / ns, i = j - k * ns;
  if (!unpack) {
    z[idx] = (i < nx) ? X[((long)b * (N + 1) + k) * nx + i] : U[((long)b * N + k) * nu + (i - nx)];
  } else {
    if (i < nx) Xo[((long)b * (N + 1) + k) * nx + i] = z[idx];
    else Uo[((long)b * N + k) * nu + (i - nx)] = z[idx];
  }
}

// Post-QP bookkeeping for any state dimension, one warp per scenario: unpack z, SQP step size, cost, warm-start shift.
// mode 0 = RTI (osqp_rti.py:532-552), mode 1 = GP-MPC SQP iteration (gp_mpc.py:336-345).  Same rules as the 3-DoF kernel.
__global__ void __launch_bounds__(128) mpc_update_kernel(int nx, int nu, int nbatch, int N, int mode,
                                                         const double* __restrict__ z, const int* __restrict__ qp_status,
                                                         const double* __restrict__ x_ref, long ldxref,
                                                         const double* __restrict__ Qd, const double* __restrict__ Qfd,
                                                         const double* __restrict__ Rd, double* __restrict__ X,
                                                         double* __restrict__ U, double* __restrict__ X_prev,
                                                         double* __restrict__ U_prev, double* __restrict__ cost,
                                                         int* __restrict__ converged, int* __restrict__ active,
                                                         int* __restrict__ n_iter) {
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  const int ns = nx + nu, n = N * ns + nx;
  const bool solved = qp_status[b] == GPMPC_QP_SOLVED;
  const double* zb = z + (long)b * n;
  double* Xb = X + (long)b * (N + 1) * nx;
  double* Ub = U + (long)b * N * nu;
  double dmaxx = 0.0, dmaxu = 0.0, cs = 0.0;
  if (solved) {
    for (int j = lane; j < n; j += 32) {
      const int k = j / ns, i = j - k * ns;
      const double v = zb[j];
      if (i < nx) {
        double* dst = Xb + k * nx + i;
        dmaxx = fmax(dmaxx, fabs(v - *dst));
        const double e = v - x_ref[(long)b * ldxref + (long)k * nx + i];
        cs += ((k == N) ? Qfd[i] : Qd[i]) * (e * e);
        *dst = v;
        if (X_prev) {
          double* xp = X_prev + (long)b * (N + 1) * nx;
          if (k >= 1) xp[(k - 1) * nx + i] = v;
          if (k == N) xp[N * nx + i] = v;
        }
      } else {
        double* dst = Ub + k * nu + (i - nx);
        dmaxu = fmax(dmaxu, fabs(v - *dst));
        cs += Rd[i - nx] * (v * v);
        *dst = v;
        if (U_prev) {
          double* up = U_prev + (long)b * N * nu;
          if (k >= 1) up[(k - 1) * nu + (i - nx)] = v;
          if (k == N - 1) up[(N - 1) * nu + (i - nx)] = v;
        }
      }
    }
  }
  dmaxx = warp_max(dmaxx);
  dmaxu = warp_max(dmaxu);
  cs = warp_sum(cs);
  if (lane == 0) {
    cost[b] = solved ? cs : INFINITY;
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

}  // namespace

extern "C" int gpmpc_mpc_update(int nx, int nu, int B, int N, int mode, const double* z, const int* qp_status,
                                const double* x_ref, long ldxref, const double* Qd, const double* Qfd, const double* Rd,
                                double* X, double* U, double* X_prev, double* U_prev, double* cost, int* converged,
                                int* active, int* n_iter, void* stream) {
  if (B == 0) return GPMPC_OK;
  mpc_update_kernel<<<(B + 3) / 4, 128, 0, (cudaStream_t)stream>>>(nx, nu, B, N, mode, z, qp_status, x_ref, ldxref, Qd,
                                                                   Qfd, Rd, X, U, X_prev, U_prev, cost, converged,
                                                                   active, n_iter);
  return gpmpc_check_launch("mpc_update");
}

extern "C" int gpmpc_step_6dof(const GpmpcDyn6* dyn, int npts, const double* X, const double* U, double* Xn,
                               void* stream) {
  GPMPC_REQUIRE(dyn && dyn->substeps >= 1, "gpmpc_step_6dof: bad dynamics parameters");
  if (npts == 0) return GPMPC_OK;
  step6_kernel<<<(npts + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*dyn, npts, X, U, Xn);
  return gpmpc_check_launch("step6");
}

extern "C" int gpmpc_linearize_6dof(const GpmpcDyn6* dyn, int npts, const double* X, const double* U, double* AB,
                                    void* stream) {
  GPMPC_REQUIRE(dyn && dyn->substeps >= 1, "gpmpc_linearize_6dof: bad dynamics parameters");
  if (npts == 0) return GPMPC_OK;
  linearize6_kernel<<<(npts * NS + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*dyn, npts, X, U, AB);
  return gpmpc_check_launch("linearize6");
}

extern "C" int gpmpc_rollout_6dof(const GpmpcExactModel* gp_v, const GpmpcExactModel* gp_w, const GpmpcDyn6* dyn, int B,
                                  int N, const double* x0, const double* U, double* X, double* dmean, const int* active,
                                  void* stream) {
  GPMPC_REQUIRE(dyn && dyn->substeps >= 1, "gpmpc_rollout_6dof: bad dynamics parameters");
  GPMPC_REQUIRE(!gp_v || (gp_v->D == 13 && gp_v->DP <= MAXDP6 && gp_v->P <= 3), "gpmpc_rollout_6dof: gp_v must have 13 features");
  GPMPC_REQUIRE(!gp_w || (gp_w->D == 12 && gp_w->DP <= MAXDP6 && gp_w->P <= 3), "gpmpc_rollout_6dof: gp_w must have 12 features");
  if (B == 0) return GPMPC_OK;
  GpmpcExactModel zero{};
  rollout6_kernel<<<(B + ROLL6_WARPS - 1) / ROLL6_WARPS, ROLL6_WARPS * 32, 0, (cudaStream_t)stream>>>(
      gp_v ? *gp_v : zero, gp_v != nullptr, gp_w ? *gp_w : zero, gp_w != nullptr, *dyn, B, N, x0, U, X, dmean, active);
  return gpmpc_check_launch("rollout6");
}

extern "C" int gpmpc_qp_assemble_6dof(const GpmpcDyn6* dyn, int B, int N, const double* X_lin, const double* U_lin,
                                      const double* x_init, const double* x_ref, long ldxref, const double* c_gp,
                                      const double* Qd, const double* Qfd, double* AB, double* beq, double* q,
                                      const int* active, void* stream) {
  GPMPC_REQUIRE(dyn && dyn->substeps >= 1, "gpmpc_qp_assemble_6dof: bad dynamics parameters");
  if (B == 0) return GPMPC_OK;
  const int pts = B * (N + 1);
  qp_assemble6_kernel<<<(pts + ASM6_PTS - 1) / ASM6_PTS, 128, 0, (cudaStream_t)stream>>>(
      *dyn, B, N, X_lin, U_lin, x_init, x_ref, ldxref, c_gp, Qd, Qfd, AB, beq, q, active);
  return gpmpc_check_launch("qp_assemble6");
}

// ------------------------------------------------------------------------------------------------ covariance chain
// Sigma_{k+1} = A_k Sigma_k A_k' + diag(q_k) (uncertainty_prop.py:117-177), any nx <= 16.  One CTA of 16 x 16 threads per
// scenario: thread (i, j) owns entry (i, j); Sigma and T = A Sigma live in shared memory, A_k is staged per stage.
__global__ void __launch_bounds__(256) propagate_cov_kernel(int nx, int N, const double* __restrict__ A, long ld_row,
                                                            long stage_stride, const double* __restrict__ qdiag,
                                                            double sigma0, double* __restrict__ covs,
                                                            const int* __restrict__ active) {
  __shared__ double Sg[16][17], T[16][17], Ak[16][17];
  const int b = blockIdx.x;
  if (active && !active[b]) return;
  const int i = threadIdx.x >> 4, j = threadIdx.x & 15;
  const bool in = i < nx && j < nx;
  double* out = covs + (long)b * (N + 1) * nx * nx;
  double sij = (i == j) ? sigma0 : 0.0;
  if (in) {
    Sg[i][j] = sij;
    out[i * nx + j] = sij;
  }
  for (int k = 0; k < N; ++k) {
    const double* Ag = A + ((long)b * N + k) * stage_stride;
    if (in) Ak[i][j] = Ag[i * ld_row + j];
    __syncthreads();
    if (in) {
      double t = 0.0;
      for (int l = 0; l < nx; ++l) t += Ak[i][l] * Sg[l][j];
      T[i][j] = t;
    }
    __syncthreads();
    if (in) {
      double v = 0.0;
      for (int l = 0; l < nx; ++l) v += T[i][l] * Ak[j][l];
      if (i == j) v += qdiag[((long)b * N + k) * nx + i];
      sij = v;
      out[(long)(k + 1) * nx * nx + i * nx + j] = v;
    }
    __syncthreads();
    if (in) Sg[i][j] = sij;
  }
}

// Tube recursion e_{k+1} = |A_k| e_k + w_k (uncertainty_prop.py:435-470, safety/tube_mpc.py:104-183), any nx <= 32: one warp
// per scenario, lane i owns component i; e_k is exchanged by shuffles.
__global__ void __launch_bounds__(128) propagate_tube_kernel(int nx, int nbatch, int N, const double* __restrict__ A,
                                                             long ld_row, long stage_stride,
                                                             const double* __restrict__ w, double* __restrict__ tubes,
                                                             const int* __restrict__ active) {
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  double e = 0.0;
  double* out = tubes + (long)b * (N + 1) * nx;
  if (lane < nx) out[lane] = 0.0;
  for (int k = 0; k < N; ++k) {
    const double* Ak = A + ((long)b * N + k) * stage_stride + (long)lane * ld_row;
    double s = 0.0;
    for (int j = 0; j < nx; ++j) {
      const double ej = __shfl_sync(0xffffffffu, e, j);
      if (lane < nx) s += fabs(Ak[j]) * ej;
    }
    if (lane < nx) {
      s += w[((long)b * N + k) * nx + lane];
      out[(long)(k + 1) * nx + lane] = s;
    }
    e = (lane < nx) ? s : 0.0;
  }
}

extern "C" int gpmpc_propagate_tube(int nx, int B, int N, const double* A, long ld_row, long stage_stride, const double* w,
                                    double* tubes, const int* active, void* stream) {
  GPMPC_REQUIRE(nx >= 1 && nx <= 32 && B >= 0 && N >= 1 && A && w && tubes && ld_row >= nx,
                "gpmpc_propagate_tube: need 1 <= nx <= 32, N >= 1, ld_row >= nx");
  if (B == 0) return GPMPC_OK;
  propagate_tube_kernel<<<(B + 3) / 4, 128, 0, (cudaStream_t)stream>>>(nx, B, N, A, ld_row, stage_stride, w, tubes, active);
  return gpmpc_check_launch("propagate_tube");
}

// lb' = lb + kappa sigma, ub' = ub - kappa sigma for the states of stages k >= 1; one thread per (scenario, variable).
__global__ void __launch_bounds__(256) tighten_bounds_kernel(int nx, int nu, int N, long total, const double* __restrict__ lb,
                                                             const double* __restrict__ ub,
                                                             const double* __restrict__ covs, double kappa,
                                                             double max_variance, double* __restrict__ lbo,
                                                             double* __restrict__ ubo, const int* __restrict__ active) {
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ns = nx + nu, n = N * ns + nx;
  const int b = (int)(idx / n), j = (int)(idx - (long)b * n);
  if (active && !active[b]) return;
  const int k = j / ns, i = j - k * ns;
  double l = lb[j], u = ub[j];
  if (k >= 1 && i < nx) {
    const double* Sk = covs + ((long)b * (N + 1) + k) * nx * nx;
    // gp_mpc.py:206-208: no tightening when the stage is too uncertain (there: the mean velocity std; here: any variance)
    bool ok = true;
    for (int d = 0; d < nx; ++d) ok = ok && (Sk[d * nx + d] <= max_variance);
    if (ok) {
      const double back = kappa * sqrt(fmax(Sk[i * nx + i], 0.0));
      const bool lf = l > -1e20, uf = u < 1e20;
      if (lf) l += back;
      if (uf) u -= back;
      if (lf && uf && l > u) l = u = 0.5 * (l + u);
    }
  }
  lbo[idx] = l;
  ubo[idx] = u;
}

extern "C" int gpmpc_tighten_bounds(int nx, int nu, int B, int N, const double* lb, const double* ub, const double* covs,
                                    double kappa, double max_variance, double* lb_out, double* ub_out, const int* active,
                                    void* stream) {
  GPMPC_REQUIRE(nx >= 1 && nu >= 0 && B >= 0 && N >= 1 && lb && ub && covs && lb_out && ub_out,
                "gpmpc_tighten_bounds: bad arguments");
  if (B == 0) return GPMPC_OK;
  const long total = (long)B * (N * (nx + nu) + nx);
  tighten_bounds_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      nx, nu, N, total, lb, ub, covs, kappa, max_variance, lb_out, ub_out, active);
  return gpmpc_check_launch("tighten_bounds");
}

extern "C" int gpmpc_propagate_cov(int nx, int B, int N, const double* A, long ld_row, long stage_stride,
                                   const double* qdiag, double sigma0, double* covs, const int* active, void* stream) {
  GPMPC_REQUIRE(nx >= 1 && nx <= 16 && B >= 0 && N >= 1, "gpmpc_propagate_cov: need 1 <= nx <= 16, N >= 1");
  GPMPC_REQUIRE(A && qdiag && covs && ld_row >= nx, "gpmpc_propagate_cov: null argument or ld_row < nx");
  if (B == 0) return GPMPC_OK;
  propagate_cov_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(nx, N, A, ld_row, stage_stride, qdiag, sigma0, covs, active);
  return gpmpc_check_launch("propagate_cov");
}

extern "C" int gpmpc_pack_traj(int nx, int nu, int B, int N, const double* X, const double* U, double* z, void* stream) {
  const long tot = (long)B * (N * (nx + nu) + nx);
  if (tot == 0) return GPMPC_OK;
  pack_traj_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(nx, nu, B, N, X, U, z, 0, nullptr,
                                                                                    nullptr);
  return gpmpc_check_launch("pack_traj");
}

extern "C" int gpmpc_unpack_traj(int nx, int nu, int B, int N, const double* z, double* X, double* U, void* stream) {
  const long tot = (long)B * (N * (nx + nu) + nx);
  if (tot == 0) return GPMPC_OK;
  pack_traj_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(nx, nu, B, N, nullptr, nullptr,
                                                                                    const_cast<double*>(z), 1, X, U);
  return gpmpc_check_launch("unpack_traj");
}
