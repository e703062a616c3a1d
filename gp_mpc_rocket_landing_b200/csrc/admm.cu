// Batched OSQP-style ADMM for the stage-structured MPC QP of the reference (src/mpc/osqp_rti.py:203-372):
//     min 1/2 z'Pz + q'z   s.t.  x_0 = x_init ;  A_k x_k + B_k u_k - x_{k+1} = b_k ;  lb <= z <= ub
//     z = [x_0,u_0, ..., x_{N-1},u_{N-1}, x_N],  P diagonal,  rows ordered [init | dynamics | identity] exactly as the
//     reference builds them (n = N(nx+nu)+nx variables, m = nx(N+1)+n rows).
// It replaces the osqp.OSQP setup/update/warm_start/solve call sites osqp_rti.py:464-478,496,517,524,527 (the OSQP C
// library itself is third-party and absent; the algorithm is restated from Stellato et al. 2020 -- see oracle/mpc.py,
// which this kernel follows step for step: modified Ruiz equilibration + cost scaling, rho*1e3 on equality rows,
// rho_min on loose rows, over-relaxation alpha, termination / infeasibility checks in unscaled units every
// `check_every` iterations, iteration-count based adaptive rho with refactorisation).
//
// B200 mapping: one CTA per scenario, the whole problem resident in shared memory (~62 KB at N=20, 3 CTAs / SM).
// The reduced KKT system (P + sigma I + A' diag(rho) A) is block tridiagonal in the stage ordering with
// (nx+nu)-square blocks; it is factorised ONCE per rho (block Cholesky, inverse diagonal blocks kept) and every ADMM
// iteration is two block substitutions on warp 0 plus embarrassingly parallel vector work on all 128 threads.  The
// only HBM traffic is the problem in and the solution out (coalesced, ~22 kB per QP) -- the kernel is bound by the
// substitution dependency chain, not by HBM or FP64 throughput (DESIGN.md states this next to the roofline).
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

constexpr int ADMM_THREADS = 128;
constexpr double OSQP_INFTY = 1e30, MIN_SCALING = 1e-4, MAX_SCALING = 1e4;
constexpr double RHO_MIN = 1e-6, RHO_MAX = 1e6, RHO_TOL = 1e-4, RHO_EQ_OVER_INEQ = 1e3;

__device__ __forceinline__ double limit_scaling(double v) {
  v = (v < MIN_SCALING) ? 1.0 : v;
  return (v > MAX_SCALING) ? MAX_SCALING : v;
}

template <int NX, int NU>
struct Prob {
  static constexpr int NS = NX + NU;
  static constexpr int NP = NS * (NS + 1) / 2;  // packed lower triangle
  int N, n, m, neq;
  // shared-memory arrays
  double *AB, *Nd, *Ei, *Bd, *D, *E, *Pb, *qb, *beq, *x, *xt, *z, *y, *zt, *Lp, *Ls, *Mw, *Vw, *red;
  const double *lb, *ub;  // global, shared by all scenarios (n)
  double c, rho, sigma;

  __device__ static size_t smem_doubles_unused(int N) {
    const int n = N * NS + NX, m = NX * (N + 1) + n;
    return (size_t)N * NX * NS * 2 + N * NX + NX + 5 * n + m + NX * (N + 1) + 3 * m + (size_t)(N + 1) * NP + NS * NS +
           2 * NS + 64;
  }
  __device__ void carve(double* s, int N_) {
    N = N_; n = N * NS + NX; neq = NX * (N + 1); m = neq + n;
    AB = s; s += N * NX * NS;
    Ls = s; s += N * NX * NS;
    Nd = s; s += N * NX;
    Ei = s; s += NX;
    Bd = s; s += n;
    D = s; s += n;
    Pb = s; s += n;
    qb = s; s += n;
    x = s; s += n;
    xt = s; s += n;
    E = s; s += m;
    beq = s; s += neq;
    z = s; s += m;
    y = s; s += m;
    zt = s; s += m;
    Lp = s; s += (N + 1) * NP;
    Mw = s; s += NS * NS;
    Vw = s; s += 2 * NS;
    red = s; s += 64;
  }

  // ---- row classes and bounds (scaled) ----------------------------------------------------------------------
  __device__ __forceinline__ double lbar(int r) const { return (r < neq) ? beq[r] : E[r] * lb[r - neq]; }
  __device__ __forceinline__ double ubar(int r) const { return (r < neq) ? beq[r] : E[r] * ub[r - neq]; }
  __device__ __forceinline__ double rho_row(int r) const {
    if (r < neq) return RHO_EQ_OVER_INEQ * rho;
    const double l = E[r] * lb[r - neq], u = E[r] * ub[r - neq];
    if (l < -OSQP_INFTY * MIN_SCALING && u > OSQP_INFTY * MIN_SCALING) return RHO_MIN;
    if (u - l < RHO_TOL) return RHO_EQ_OVER_INEQ * rho;
    return rho;
  }

  // ---- (Abar v)_r and (Abar' w)_j with the scaled blocks ------------------------------------------------------
  __device__ __forceinline__ double A_row(int r, const double* v) const {
    if (r < NX) return Ei[r] * v[r];
    if (r < neq) {
      const int k = (r - NX) / NX, i = (r - NX) % NX;
      const double* a = AB + (k * NX + i) * NS;
      const double* vk = v + k * NS;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int j = 0; j + 1 < NS; j += 2) {
        s0 = fma(a[j], vk[j], s0);
        s1 = fma(a[j + 1], vk[j + 1], s1);
      }
      if (NS & 1) s0 = fma(a[NS - 1], vk[NS - 1], s0);
      return fma(Nd[k * NX + i], v[(k + 1) * NS + i], s0 + s1);
    }
    return Bd[r - neq] * v[r - neq];
  }
  __device__ __forceinline__ double At_col(int j, const double* w) const {
    const int k = j / NS, i = j - k * NS;
    double s = Bd[j] * w[neq + j];
    if (k == 0 && i < NX) s = fma(Ei[i], w[i], s);
    if (k < N) {
      const double* wk = w + NX + k * NX;
      const double* a = AB + k * NX * NS + i;
      double s1 = 0.0;
#pragma unroll
      for (int r = 0; r < NX; ++r) s1 = fma(a[r * NS], wk[r], s1);
      s += s1;
    }
    if (k >= 1 && i < NX) s = fma(Nd[(k - 1) * NX + i], w[NX + (k - 1) * NX + i], s);
    return s;
  }
};

// block-wide reductions of K values at once (max or sum), result broadcast to all threads
template <int K, bool IS_MAX>
__device__ void block_reduce(double (&v)[K], double* red) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int e = 0; e < K; ++e) v[e] = IS_MAX ? warp_max(v[e]) : warp_sum(v[e]);
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int e = 0; e < K; ++e) red[warp * K + e] = v[e];
  __syncthreads();
#pragma unroll
  for (int e = 0; e < K; ++e) {
    double r = red[e];
#pragma unroll
    for (int w = 1; w < ADMM_THREADS / 32; ++w) r = IS_MAX ? fmax(r, red[w * K + e]) : r + red[w * K + e];
    v[e] = r;
  }
}

// ---- factorisation of M = Pb + sigma I + Abar' diag(rho) Abar (block tridiagonal) -------------------------------
template <int NX, int NU>
__device__ void factorize(Prob<NX, NU>& P) {
  constexpr int NS = NX + NU, NP = Prob<NX, NU>::NP;
  const int tid = threadIdx.x, N = P.N;
  const double rho_eq = RHO_EQ_OVER_INEQ * P.rho;
  // phase 1 (all threads): packed lower triangles of the diagonal blocks M_kk
  for (int idx = tid; idx < (N + 1) * NP; idx += ADMM_THREADS) {
    const int k = idx / NP, pidx = idx - k * NP;
    int i = 0;
    while ((i + 1) * (i + 2) / 2 <= pidx) ++i;
    const int j = pidx - i * (i + 1) / 2;
    double v = 0.0;
    if (k == N && i >= NX) {  // last block has only NX variables: keep an identity tail
      P.Lp[idx] = (i == j) ? 1.0 : 0.0;
      continue;
    }
    if (k < N) {
      const double* a = P.AB + k * NX * NS;
      double s = 0.0;
#pragma unroll
      for (int r = 0; r < NX; ++r) s = fma(a[r * NS + i], a[r * NS + j], s);
      v = rho_eq * s;
    }
    if (i == j) {
      const int jj = k * NS + j;
      double d = P.Pb[jj] + P.sigma;
      d = fma(P.rho_row(P.neq + jj), P.Bd[jj] * P.Bd[jj], d);
      if (k == 0 && i < NX) d = fma(rho_eq, P.Ei[i] * P.Ei[i], d);
      if (k >= 1 && i < NX) d = fma(rho_eq, P.Nd[(k - 1) * NX + i] * P.Nd[(k - 1) * NX + i], d);
      v += d;
    }
    P.Lp[idx] = v;
  }
  __syncthreads();
  // phase 2 (warp 0): block Cholesky chain; Lp[k] <- packed inverse of L_kk ; Ls[k-1] = M_{k,k-1} L_{k-1,k-1}^{-T}
  if (tid < 32) {
    const int lane = tid;
    double* Mw = P.Mw;  // dense NS x NS scratch
    for (int k = 0; k <= N; ++k) {
      double* Lk = P.Lp + k * NP;
      if (k > 0) {
        const double* Lip = P.Lp + (k - 1) * NP;  // inverse of previous diagonal factor (packed)
        const double* a = P.AB + (k - 1) * NX * NS;
        double* ls = P.Ls + (k - 1) * NX * NS;
        for (int e = lane; e < NX * NS; e += 32) {
          const int i = e / NS, j = e - i * NS;
          const double sc = rho_eq * P.Nd[(k - 1) * NX + i];
          double s = 0.0;
          for (int t = 0; t <= j; ++t) s = fma(sc * a[i * NS + t], Lip[j * (j + 1) / 2 + t], s);
          ls[e] = s;
        }
        __syncwarp();
        for (int e = lane; e < NX * (NX + 1) / 2; e += 32) {
          int i = 0;
          while ((i + 1) * (i + 2) / 2 <= e) ++i;
          const int j = e - i * (i + 1) / 2;
          double s = 0.0;
#pragma unroll
          for (int t = 0; t < NS; ++t) s = fma(ls[i * NS + t], ls[j * NS + t], s);
          Lk[e] -= s;
        }
        __syncwarp();
      }
      // dense copy
      for (int e = lane; e < NS * NS; e += 32) {
        const int i = e / NS, j = e - i * NS;
        Mw[e] = (j <= i) ? Lk[i * (i + 1) / 2 + j] : 0.0;
      }
      __syncwarp();
      // Cholesky, lane = row
      for (int j = 0; j < NS; ++j) {
        double s = 0.0;
        if (lane >= j && lane < NS) {
          s = Mw[lane * NS + j];
          for (int t = 0; t < j; ++t) s = fma(-Mw[lane * NS + t], Mw[j * NS + t], s);
        }
        const double djj = sqrt(fmax(__shfl_sync(0xffffffffu, s, j), 1e-300));
        if (lane == j) Mw[j * NS + j] = djj;
        else if (lane > j && lane < NS) Mw[lane * NS + j] = s / djj;
        __syncwarp();
      }
      // inverse, lane = column c ; write packed Li into Lk
      if (lane < NS) {
        const int c = lane;
        double col[NS];
#pragma unroll
        for (int i = 0; i < NS; ++i) col[i] = 0.0;
#pragma unroll
        for (int i = 0; i < NS; ++i) {
          if (i >= c) {
            double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
            for (int t = 0; t < NS; ++t)
              if (t >= c && t < i) s = fma(-Mw[i * NS + t], col[t], s);
            col[i] = s / Mw[i * NS + i];
            Lk[i * (i + 1) / 2 + c] = col[i];
          }
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
}

// ---- solve M v = rhs in place (rhs in P.xt), warp 0, lane = row of the stage block ------------------------------------
template <int NX, int NU>
__device__ void block_solve(Prob<NX, NU>& P) {
  constexpr int NS = NX + NU, NP = Prob<NX, NU>::NP;
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x, N = P.N;
  double* v = P.xt;
  double* w = P.Vw;  // NS scratch
  // forward: y_k = Li_k (r_k - Ls_{k-1} y_{k-1})
  for (int k = 0; k <= N; ++k) {
    const int ns = (k < N) ? NS : NX;
    double t = 0.0;
    if (lane < ns) {
      t = v[k * NS + lane];
      if (k > 0 && lane < NX) {
        const double* ls = P.Ls + ((k - 1) * NX + lane) * NS;
        const double* yp = v + (k - 1) * NS;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int j = 0; j + 1 < NS; j += 2) {
          s0 = fma(ls[j], yp[j], s0);
          s1 = fma(ls[j + 1], yp[j + 1], s1);
        }
        if (NS & 1) s0 = fma(ls[NS - 1], yp[NS - 1], s0);
        t -= (s0 + s1);
      }
      w[lane] = t;
    }
    __syncwarp();
    if (lane < ns) {
      const double* li = P.Lp + k * NP + lane * (lane + 1) / 2;
      double s0 = 0.0, s1 = 0.0;
      int j = 0;
      for (; j + 1 <= lane; j += 2) {
        s0 = fma(li[j], w[j], s0);
        s1 = fma(li[j + 1], w[j + 1], s1);
      }
      if (j <= lane) s0 = fma(li[j], w[j], s0);
      v[k * NS + lane] = s0 + s1;
    }
    __syncwarp();
  }
  // backward: x_k = Li_k^T (y_k - Ls_k^T x_{k+1})
  for (int k = N; k >= 0; --k) {
    const int ns = (k < N) ? NS : NX;
    if (lane < ns) {
      double t = v[k * NS + lane];
      if (k < N) {
        const double* ls = P.Ls + k * NX * NS + lane;
        const double* xn = v + (k + 1) * NS;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int i = 0; i + 1 < NX; i += 2) {
          s0 = fma(ls[i * NS], xn[i], s0);
          s1 = fma(ls[(i + 1) * NS], xn[i + 1], s1);
        }
        if (NX & 1) s0 = fma(ls[(NX - 1) * NS], xn[NX - 1], s0);
        t -= (s0 + s1);
      }
      w[lane] = t;
    }
    __syncwarp();
    if (lane < ns) {
      const double* lp = P.Lp + k * NP;
      double s0 = 0.0, s1 = 0.0;
      int i = lane;
      for (; i + 1 < ns; i += 2) {
        s0 = fma(lp[i * (i + 1) / 2 + lane], w[i], s0);
        s1 = fma(lp[(i + 1) * (i + 2) / 2 + lane], w[i + 1], s1);
      }
      if (i < ns) s0 = fma(lp[i * (i + 1) / 2 + lane], w[i], s0);
      v[k * NS + lane] = s0 + s1;
    }
    __syncwarp();
  }
}

struct CheckOut {
  double pri, dua, eps_pri, eps_dua, rho_est;
  int status;
};

// Residuals, termination, infeasibility certificates and the rho estimate (oracle/mpc.py _residuals/_check/_rho_estimate;
// OSQP auxil.c compute_pri_res/compute_dua_res/is_primal_infeasible/is_dual_infeasible/compute_rho_estimate).
// dx is in P.xt, dy in P.zt.  All threads participate.
template <int NX, int NU>
__device__ CheckOut check_termination(Prob<NX, NU>& P, const GpmpcAdmmSettings& S, bool approx) {
  const int tid = threadIdx.x;
  const double INF_T = OSQP_INFTY * MIN_SCALING;
  // --- m-indexed quantities
  double a[6] = {0, 0, 0, 0, 0, 0};  // max: pri_u, |Ax|_u, |z|_u, pri_s, |Ax|_s, |z|_s
  double b[2] = {0, 0};              // max: |E dy_proj| ; (dual infeas) bound violation flag as max
  double lhs[1] = {0};
  double ndx[3] = {0, 0, 0};  // max |D dx|, max |Pb dx / D|, (filled below)
  double qdx[1] = {0};
  for (int j = tid; j < P.n; j += ADMM_THREADS) {
    const double dx = P.xt[j];
    ndx[0] = fmax(ndx[0], fabs(P.D[j] * dx));
    ndx[1] = fmax(ndx[1], fabs(P.Pb[j] * dx / P.D[j]));
    qdx[0] = fma(P.qb[j], dx, qdx[0]);
  }
  block_reduce<3, true>(ndx, P.red);
  block_reduce<1, false>(qdx, P.red);
  const double f = approx ? 10.0 : 1.0;
  const double eps_pinf = f * S.eps_prim_inf, eps_dinf = f * S.eps_dual_inf;
  const double nrm_dx = ndx[0];
  for (int r = tid; r < P.m; r += ADMM_THREADS) {
    const double Ax = P.A_row(r, P.x), zz = P.z[r], Ei = 1.0 / P.E[r];
    a[0] = fmax(a[0], fabs(Ei * (Ax - zz)));
    a[1] = fmax(a[1], fabs(Ei * Ax));
    a[2] = fmax(a[2], fabs(Ei * zz));
    a[3] = fmax(a[3], fabs(Ax - zz));
    a[4] = fmax(a[4], fabs(Ax));
    a[5] = fmax(a[5], fabs(zz));
    // primal infeasibility: project dy on the polar of the recession cone of [l,u]
    const double l = P.lbar(r), u = P.ubar(r);
    double dy = P.zt[r];
    const bool uinf = u > INF_T, linf = l < -INF_T;
    if (uinf && linf) dy = 0.0;
    else if (uinf) dy = fmin(dy, 0.0);
    else if (linf) dy = fmax(dy, 0.0);
    P.zt[r] = dy;  // projected dy (consumed below by At_col)
    b[0] = fmax(b[0], fabs(P.E[r] * dy));
    lhs[0] += u * fmax(dy, 0.0) + l * fmin(dy, 0.0);
    // dual infeasibility: (A dx)_r must stay inside the recession cone
    const double Adx = P.A_row(r, P.xt) * Ei;
    const bool bad = (!uinf && Adx > eps_dinf * nrm_dx) || (!linf && Adx < -eps_dinf * nrm_dx);
    b[1] = fmax(b[1], bad ? 1.0 : 0.0);
  }
  block_reduce<6, true>(a, P.red);
  block_reduce<2, true>(b, P.red);
  block_reduce<1, false>(lhs, P.red);
  // --- n-indexed quantities
  double d[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // dua_u, |Px|_u, |Aty|_u, |q|_u, dua_s, |Px|_s, |Aty|_s, |q|_s
  double e[1] = {0};                       // |D^-1 A' dy_proj|
  for (int j = tid; j < P.n; j += ADMM_THREADS) {
    const double Px = P.Pb[j] * P.x[j], Aty = P.At_col(j, P.y), q = P.qb[j], Di = 1.0 / P.D[j];
    d[0] = fmax(d[0], fabs(Di * (Px + q + Aty)));
    d[1] = fmax(d[1], fabs(Di * Px));
    d[2] = fmax(d[2], fabs(Di * Aty));
    d[3] = fmax(d[3], fabs(Di * q));
    d[4] = fmax(d[4], fabs(Px + q + Aty));
    d[5] = fmax(d[5], fabs(Px));
    d[6] = fmax(d[6], fabs(Aty));
    d[7] = fmax(d[7], fabs(q));
    e[0] = fmax(e[0], fabs(P.At_col(j, P.zt) * Di));
  }
  block_reduce<8, true>(d, P.red);
  block_reduce<1, true>(e, P.red);

  CheckOut o;
  o.pri = a[0];
  o.dua = d[0] / P.c;
  o.eps_pri = f * S.eps_abs + f * S.eps_rel * fmax(a[1], a[2]);
  o.eps_dua = f * S.eps_abs + f * S.eps_rel * fmax(fmax(d[1], d[2]), d[3]) / P.c;
  double pr = a[3] / (fmax(a[5], a[4]) + 1e-10);
  double du = d[4] / (fmax(fmax(d[7], d[6]), d[5]) + 1e-10);
  o.rho_est = fmin(fmax(P.rho * sqrt(pr / (du + 1e-10)), RHO_MIN), RHO_MAX);
  o.status = GPMPC_QP_UNSOLVED;
  bool pinf = false, dinf = false;
  if (b[0] > eps_pinf && lhs[0] < -eps_pinf * b[0]) pinf = e[0] < eps_pinf * b[0];
  if (nrm_dx > eps_dinf && qdx[0] < -P.c * eps_dinf * nrm_dx && ndx[1] < P.c * eps_dinf * nrm_dx) dinf = (b[1] == 0.0);
  if (pinf) o.status = approx ? GPMPC_QP_PRIMAL_INFEASIBLE_INACCURATE : GPMPC_QP_PRIMAL_INFEASIBLE;
  else if (dinf) o.status = approx ? GPMPC_QP_DUAL_INFEASIBLE_INACCURATE : GPMPC_QP_DUAL_INFEASIBLE;
  else if (o.pri <= o.eps_pri && o.dua <= o.eps_dua) o.status = approx ? GPMPC_QP_SOLVED_INACCURATE : GPMPC_QP_SOLVED;
  return o;
}

template <int NX, int NU>
__global__ void __launch_bounds__(ADMM_THREADS) admm_kernel(int N, int nbatch, const double* __restrict__ ABg,
                                                            const double* __restrict__ beqg,
                                                            const double* __restrict__ qg,
                                                            const double* __restrict__ Pdiag,
                                                            const double* __restrict__ lb,
                                                            const double* __restrict__ ub, GpmpcAdmmSettings S,
                                                            const int* __restrict__ active, double* __restrict__ xg,
                                                            double* __restrict__ yg, double* __restrict__ rhog,
                                                            int* __restrict__ status_g, int* __restrict__ iters_g,
                                                            double* __restrict__ obj_g, double* __restrict__ pri_g,
                                                            double* __restrict__ dua_g) {
  constexpr int NS = NX + NU;
  extern __shared__ double smem[];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  Prob<NX, NU> P;
  P.carve(smem, N);
  P.lb = lb; P.ub = ub;
  P.sigma = S.sigma;
  P.rho = (S.rho_from_input && rhog[b] > 0.0) ? rhog[b] : S.rho;
  P.c = 1.0;
  const int n = P.n, m = P.m, neq = P.neq;

  // ---- load (coalesced) -----------------------------------------------------------------------------------------
  for (int i = tid; i < N * NX * NS; i += ADMM_THREADS) P.AB[i] = ABg[(size_t)b * N * NX * NS + i];
  for (int i = tid; i < N * NX; i += ADMM_THREADS) P.Nd[i] = -1.0;
  for (int i = tid; i < NX; i += ADMM_THREADS) P.Ei[i] = 1.0;
  for (int i = tid; i < n; i += ADMM_THREADS) {
    P.Bd[i] = 1.0; P.D[i] = 1.0;
    P.Pb[i] = Pdiag[i];
    P.qb[i] = qg[(size_t)b * n + i];
  }
  for (int i = tid; i < m; i += ADMM_THREADS) P.E[i] = 1.0;
  for (int i = tid; i < neq; i += ADMM_THREADS) P.beq[i] = beqg[(size_t)b * neq + i];
  __syncthreads();

  // ---- modified Ruiz equilibration + cost scaling (OSQP scaling.c; oracle/mpc.py OSQPOracle._scale) --------------------
  for (int it = 0; it < S.scaling; ++it) {
    // column norms -> xt (as Dt), row norms -> zt (as Et); both from the current matrices
    for (int j = tid; j < n; j += ADMM_THREADS) {
      const int k = j / NS, i = j - k * NS;
      double v = fmax(fabs(P.Pb[j]), fabs(P.Bd[j]));
      if (k == 0 && i < NX) v = fmax(v, fabs(P.Ei[i]));
      if (k < N)
#pragma unroll
        for (int r = 0; r < NX; ++r) v = fmax(v, fabs(P.AB[(k * NX + r) * NS + i]));
      if (k >= 1 && i < NX) v = fmax(v, fabs(P.Nd[(k - 1) * NX + i]));
      P.xt[j] = 1.0 / sqrt(limit_scaling(v));
    }
    for (int r = tid; r < m; r += ADMM_THREADS) {
      double v;
      if (r < NX) v = fabs(P.Ei[r]);
      else if (r < neq) {
        const int k = (r - NX) / NX, i = (r - NX) % NX;
        v = fabs(P.Nd[k * NX + i]);
#pragma unroll
        for (int j = 0; j < NS; ++j) v = fmax(v, fabs(P.AB[(k * NX + i) * NS + j]));
      } else v = fabs(P.Bd[r - neq]);
      P.zt[r] = 1.0 / sqrt(limit_scaling(v));
    }
    __syncthreads();
    // P <- Dt P Dt ; A <- Et A Dt ; q <- Dt q ; D *= Dt ; E *= Et
    for (int e = tid; e < N * NX * NS; e += ADMM_THREADS) {
      const int k = e / (NX * NS), rem = e - k * NX * NS, i = rem / NS, j = rem - i * NS;
      P.AB[e] = (P.zt[NX + k * NX + i] * P.AB[e]) * P.xt[k * NS + j];
    }
    for (int e = tid; e < N * NX; e += ADMM_THREADS) {
      const int k = e / NX, i = e - k * NX;
      P.Nd[e] = (P.zt[NX + e] * P.Nd[e]) * P.xt[(k + 1) * NS + i];
    }
    for (int i = tid; i < NX; i += ADMM_THREADS) P.Ei[i] = (P.zt[i] * P.Ei[i]) * P.xt[i];
    double cs[1] = {0.0}, qm[1] = {0.0};
    for (int j = tid; j < n; j += ADMM_THREADS) {
      const double dt = P.xt[j];
      P.Bd[j] = (P.zt[neq + j] * P.Bd[j]) * dt;
      P.Pb[j] = (dt * P.Pb[j]) * dt;
      P.qb[j] = dt * P.qb[j];
      P.D[j] *= dt;
      cs[0] += fabs(P.Pb[j]);
      qm[0] = fmax(qm[0], fabs(P.qb[j]));
    }
    for (int r = tid; r < m; r += ADMM_THREADS) P.E[r] *= P.zt[r];
    block_reduce<1, false>(cs, P.red);
    block_reduce<1, true>(qm, P.red);
    double ct = limit_scaling(fmax(cs[0] / n, limit_scaling(qm[0])));
    ct = 1.0 / ct;
    for (int j = tid; j < n; j += ADMM_THREADS) {
      P.Pb[j] *= ct;
      P.qb[j] *= ct;
    }
    P.c *= ct;
    __syncthreads();
  }
  for (int r = tid; r < neq; r += ADMM_THREADS) P.beq[r] *= P.E[r];
  __syncthreads();

  // ---- iterates (scaled): x = x_in / D ; z = Abar x ; y = c y_in / E -----------------------------------------------------
  for (int j = tid; j < n; j += ADMM_THREADS) P.x[j] = S.warm_x ? xg[(size_t)b * n + j] / P.D[j] : 0.0;
  for (int r = tid; r < m; r += ADMM_THREADS) P.y[r] = S.warm_y ? P.c * yg[(size_t)b * m + r] / P.E[r] : 0.0;
  __syncthreads();
  for (int r = tid; r < m; r += ADMM_THREADS) P.z[r] = S.warm_x ? P.A_row(r, P.x) : 0.0;
  __syncthreads();

  factorize(P);

  int status = GPMPC_QP_UNSOLVED, it = 0;
  CheckOut co;
  co.pri = co.dua = 0.0;
  const double alpha = S.alpha;
  for (it = 1; it <= S.max_iter; ++it) {
    // w = rho*z - y  (into zt)
    for (int r = tid; r < m; r += ADMM_THREADS) P.zt[r] = fma(P.rho_row(r), P.z[r], -P.y[r]);
    __syncthreads();
    // rhs = sigma x - q + Abar' w (into xt)
    for (int j = tid; j < n; j += ADMM_THREADS) P.xt[j] = fma(P.sigma, P.x[j], -P.qb[j]) + P.At_col(j, P.zt);
    __syncthreads();
    block_solve(P);  // xt <- x_tilde
    __syncthreads();
    // z_tilde = Abar x_tilde ; z, y updates ; dy into zt
    for (int r = tid; r < m; r += ADMM_THREADS) {
      const double zt = P.A_row(r, P.xt);
      const double rr = P.rho_row(r), zp = P.z[r], yy = P.y[r];
      const double zr = alpha * zt + (1.0 - alpha) * zp;
      const double zn = fmin(fmax(zr + yy / rr, P.lbar(r)), P.ubar(r));
      const double dy = rr * (zr - zn);
      P.z[r] = zn;
      P.y[r] = yy + dy;
      P.zt[r] = dy;
    }
    __syncthreads();  // all reads of xt (x_tilde) by A_row are done before it is overwritten with dx
    for (int j = tid; j < n; j += ADMM_THREADS) {
      const double xp = P.x[j];
      const double xn = alpha * P.xt[j] + (1.0 - alpha) * xp;
      P.x[j] = xn;
      P.xt[j] = xn - xp;
    }
    __syncthreads();
    const bool can_check = S.check_every > 0 && (it % S.check_every) == 0;
    const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval > 0 && (it % S.adaptive_rho_interval) == 0;
    if (can_check || can_adapt) {
      co = check_termination(P, S, false);
      if (can_check && co.status != GPMPC_QP_UNSOLVED) {
        status = co.status;
        break;
      }
      if (can_adapt && (co.rho_est > P.rho * S.adaptive_rho_tolerance || co.rho_est < P.rho / S.adaptive_rho_tolerance)) {
        P.rho = co.rho_est;
        __syncthreads();
        factorize(P);
      }
    }
  }
  if (it > S.max_iter) it = S.max_iter;
  if (status == GPMPC_QP_UNSOLVED) {
    if (!(S.check_every > 0 && (it % S.check_every) == 0)) {
      co = check_termination(P, S, false);
      status = co.status;
    }
    if (status == GPMPC_QP_UNSOLVED) {
      CheckOut ca = check_termination(P, S, true);
      status = ca.status;
    }
    if (status == GPMPC_QP_UNSOLVED) status = GPMPC_QP_MAX_ITER;
  }
  // ---- objective and solution out (unscaled) ---------------------------------------------------------------------------
  double ob[1] = {0.0};
  for (int j = tid; j < n; j += ADMM_THREADS) ob[0] += (0.5 * P.Pb[j] * P.x[j] + P.qb[j]) * P.x[j];
  block_reduce<1, false>(ob, P.red);
  for (int j = tid; j < n; j += ADMM_THREADS) xg[(size_t)b * n + j] = P.D[j] * P.x[j];
  if (!S.commit_y_solved_only || status == GPMPC_QP_SOLVED)
    for (int r = tid; r < m; r += ADMM_THREADS) yg[(size_t)b * m + r] = P.E[r] * P.y[r] / P.c;
  if (tid == 0) {
    status_g[b] = status;
    iters_g[b] = it;
    obj_g[b] = ob[0] / P.c;
    pri_g[b] = co.pri;
    dua_g[b] = co.dua;
    rhog[b] = P.rho;
  }
}

template <int NX, int NU>
size_t admm_smem_bytes(int N) {
  constexpr int NS = NX + NU, NP = NS * (NS + 1) / 2;
  const size_t n = (size_t)N * NS + NX, m = (size_t)NX * (N + 1) + n;
  size_t d = (size_t)N * NX * NS * 2 + N * NX + NX + 6 * n + m + NX * (N + 1) + 3 * m + (size_t)(N + 1) * NP + NS * NS +
             2 * NS + 64;
  return d * sizeof(double);
}

template <int NX, int NU>
int launch_admm(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag,
                const double* lb, const double* ub, const GpmpcAdmmSettings& S, const int* active, double* x, double* y,
                double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream) {
  const size_t smem = admm_smem_bytes<NX, NU>(N);
  if (smem > 227 * 1024) {
    gpmpc_set_error("gpmpc_admm_solve: horizon too long for the shared-memory resident solver");
    return GPMPC_ERR_ARG;
  }
  auto kern = admm_kernel<NX, NU>;
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // the default carveout heuristic left one CTA per SM resident; ask for the full 228 KB so 3-4 QPs share an SM
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured = smem;
  }
  kern<<<B, ADMM_THREADS, smem, stream>>>(N, B, AB, beq, q, Pdiag, lb, ub, S, active, x, y, rho, status, iters, obj,
                                          pri, dua);
  return gpmpc_check_launch("admm");
}

}  // namespace

extern "C" int gpmpc_admm_solve(int nx, int nu, int N, int B, const double* AB, const double* beq, const double* q,
                                const double* Pdiag, const double* lb, const double* ub,
                                const GpmpcAdmmSettings* settings, const int* active, double* x, double* y, double* rho,
                                int* status, int* iters, double* obj, double* pri_res, double* dua_res, void* stream) {
  GPMPC_REQUIRE(settings != nullptr, "gpmpc_admm_solve: settings is NULL");
  GPMPC_REQUIRE(N >= 1 && B >= 0, "gpmpc_admm_solve: bad N or B");
  if (B == 0) return GPMPC_OK;
  if (nx == 7 && nu == 3)
    return launch_admm<7, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, x, y, rho, status, iters, obj, pri_res,
                             dua_res, (cudaStream_t)stream);
  if (nx == 14 && nu == 3)
    return launch_admm<14, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, x, y, rho, status, iters, obj,
                              pri_res, dua_res, (cudaStream_t)stream);
  gpmpc_set_error("gpmpc_admm_solve: only (nx,nu) = (7,3) and (14,3) are instantiated");
  return GPMPC_ERR_ARG;
}
