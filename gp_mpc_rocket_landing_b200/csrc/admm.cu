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
// B200 mapping: one CTA (4 warps) per scenario, the whole problem resident in shared memory (~75 KB at N=20, so three
// QPs share an SM).  The reduced KKT matrix M = P + sigma I + A' diag(rho) A is block tridiagonal in the stage
// ordering with (nx+nu)-square blocks.  It is factorised once per rho by a block Cholesky chain; the factor is kept
// in the form that makes the per-iteration solve short:
//     forward   y_k = c_k - G_k y_{k-1},   c_k = Li_k r_k        (c for all stages in parallel, then a 1-matvec chain)
//     backward  v_k = d_k - H_k v_{k+1},   d_k = Li_k' y_k       (same)
// with Li_k = L_kk^{-1}, G_k = Li_k Ls_{k-1}, H_k = Li_k' Ls_k'.  The chain is the critical path of an iteration
// (2N dependent small matvecs on one warp); everything else is embarrassingly parallel vector work on 128 threads.
// HBM traffic is the problem in and the solution out (~22 kB per QP, coalesced) plus the scaling vectors D, E which
// live in a global scratch (read only at termination checks): the kernel is bound by the dependency chain of the
// substitutions, not by HBM or FP64 throughput (DESIGN.md states this next to the roofline).
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

constexpr int ADMM_THREADS = 128;
constexpr double OSQP_INFTY = 1e30, MIN_SCALING = 1e-4, MAX_SCALING = 1e4;
constexpr double RHO_MIN = 1e-6, RHO_MAX = 1e6, RHO_TOL = 1e-4, RHO_EQ_OVER_INEQ = 1e3;
constexpr double INF_T = OSQP_INFTY * MIN_SCALING;

__device__ __forceinline__ double limit_scaling(double v) {
  v = (v < MIN_SCALING) ? 1.0 : v;
  return (v > MAX_SCALING) ? MAX_SCALING : v;
}

template <int NX, int NU>
struct Prob {
  static constexpr int NS = NX + NU;
  static constexpr int NP = NS * (NS + 1) / 2;  // packed lower triangle
  static constexpr int GP = (NS + 1) & ~1;      // row pitch of G_k (even: 16-byte vector loads)
  static constexpr int HP = (NX + 1) & ~1;      // row pitch of H_k
  int N, n, m, neq;
  // shared-memory arrays
  double *AB, *G, *H, *Li, *Nd, *Ei, *Bd, *Pb, *qb, *x, *xt, *w, *lbs, *ubs, *z, *y, *zt, *red;
  double *D, *E;          // global scratch (n, m): cumulative Ruiz scalings, needed only at checks / unscaling
  double c, rho, sigma;

  static __host__ __device__ size_t smem_doubles(int N) {
    const size_t n = (size_t)N * NS + NX, m = (size_t)NX * (N + 1) + n;
    return (size_t)N * NX * NS + (size_t)N * NS * GP + (size_t)N * NS * HP + (size_t)(N + 1) * NP + (size_t)N * NX + NX + 2 +
           6 * n + 4 * m + zt_doubles(m) + 64 + 2 * 17;  // + alignment slack of the 17 arrays
  }
  // zt doubles as the factorisation workspace (dense NS x NS block, NS x GP inverse, NX x GP coupling)
  static __host__ __device__ size_t zt_doubles(size_t m) {
    const size_t need = ((NS * NS + 1) & ~1) + (size_t)NS * GP + (size_t)NX * GP;
    return m > need ? m : need;
  }
  __device__ void carve(double* s, int N_) {
    N = N_; n = N * NS + NX; neq = NX * (N + 1); m = neq + n;
    auto take = [&](size_t cnt) { double* p = s; s += (cnt + 1) & ~(size_t)1; return p; };  // keep 16-byte alignment
    AB = take(N * NX * NS);
    G = take(N * NS * GP);
    H = take(N * NS * HP);
    Li = take((N + 1) * NP);
    Nd = take(N * NX);
    Ei = take(NX + 2);
    Bd = take(n); Pb = take(n); qb = take(n); x = take(n); xt = take(n); w = take(n);
    lbs = take(m); ubs = take(m); z = take(m); y = take(m); zt = take(zt_doubles(m));
    red = take(64);
  }

  // ---- row classes (scaled bounds are precomputed: lbs/ubs; equality rows have lbs == ubs == E*beq) ----------------
  __device__ __forceinline__ double rho_row(int r) const {
    if (r < neq) return RHO_EQ_OVER_INEQ * rho;
    const double l = lbs[r], u = ubs[r];
    if (l < -INF_T && u > INF_T) return RHO_MIN;
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
  __device__ __forceinline__ double At_col(int j, const double* wv) const {
    const int k = j / NS, i = j - k * NS;
    double s = Bd[j] * wv[neq + j];
    if (k == 0 && i < NX) s = fma(Ei[i], wv[i], s);
    if (k < N) {
      const double* wk = wv + NX + k * NX;
      const double* a = AB + k * NX * NS + i;
      double s1 = 0.0;
#pragma unroll
      for (int r = 0; r < NX; ++r) s1 = fma(a[r * NS], wk[r], s1);
      s += s1;
    }
    if (k >= 1 && i < NX) s = fma(Nd[(k - 1) * NX + i], wv[NX + (k - 1) * NX + i], s);
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
// Phase 1 (all threads): packed lower triangles of the diagonal blocks M_kk into Li.
// Phase 2 (warp 0): block Cholesky chain.  Per stage k: Ls = C_{k-1} Li_{k-1}' (C = sub-diagonal block), H_{k-1},
// Schur update of M_kk, dense Cholesky (right-looking, lane = row) and triangular inverse (lane = column) -> Li_k,
// then G_k.  Scratch (dense 10x10 workspace, previous Li, Ls) is overlaid on zt, which is free here.
template <int NX, int NU>
__device__ void factorize(Prob<NX, NU>& P) {
  constexpr int NS = NX + NU, NP = Prob<NX, NU>::NP, GP = Prob<NX, NU>::GP, HP = Prob<NX, NU>::HP;
  const int tid = threadIdx.x, N = P.N;
  const double rho_eq = RHO_EQ_OVER_INEQ * P.rho;
  for (int idx = tid; idx < (N + 1) * NP; idx += ADMM_THREADS) {
    const int k = idx / NP, pidx = idx - k * NP;
    int i = 0;
    while ((i + 1) * (i + 2) / 2 <= pidx) ++i;
    const int j = pidx - i * (i + 1) / 2;
    double v = 0.0;
    if (k == N && i >= NX) {  // last block has only NX variables: keep an identity tail
      P.Li[idx] = (i == j) ? 1.0 : 0.0;
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
    P.Li[idx] = v;
  }
  __syncthreads();
  if (tid < 32) {
    const int lane = tid;
    double* Mw = P.zt;                              // NS x NS dense workspace
    double* Ld = Mw + ((NS * NS + 1) & ~1);         // NS x GP dense Li of the current / previous stage
    double* Lsv = Ld + NS * GP;                     // NX x GP
    for (int e = lane; e < NS * GP; e += 32) Ld[e] = 0.0;
    for (int e = lane; e < NX * GP; e += 32) Lsv[e] = 0.0;
    __syncwarp();
    for (int k = 0; k <= N; ++k) {
      double* Lk = P.Li + k * NP;
      if (k > 0) {
        const double* a0 = P.AB + (k - 1) * NX * NS;
        for (int e = lane; e < NX * NS; e += 32) {  // Ls[i][j] = rho_eq Nd_i sum_t AB[i][t] Li_{k-1}[j][t]
          const int i = e / NS, j = e - i * NS;
          const double* a = a0 + i * NS;
          const double* ld = Ld + j * GP;
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int t = 0; t + 1 < NS; t += 2) {
            s0 = fma(a[t], ld[t], s0);
            s1 = fma(a[t + 1], ld[t + 1], s1);
          }
          if (NS & 1) s0 = fma(a[NS - 1], ld[NS - 1], s0);
          const double v = (rho_eq * P.Nd[(k - 1) * NX + i]) * (s0 + s1);
          Lsv[i * GP + j] = v;
          P.H[(size_t)(k - 1) * NS * HP + e] = v;   // kept (row-major NX x NS) for the parallel G / H phase below
        }
        __syncwarp();
        for (int e = lane; e < NX * (NX + 1) / 2; e += 32) {  // Schur: M_kk[i][j] -= sum_t Ls[i][t] Ls[j][t]
          int i = 0;
          while ((i + 1) * (i + 2) / 2 <= e) ++i;
          const int j = e - i * (i + 1) / 2;
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int t = 0; t + 1 < NS; t += 2) {
            s0 = fma(Lsv[i * GP + t], Lsv[j * GP + t], s0);
            s1 = fma(Lsv[i * GP + t + 1], Lsv[j * GP + t + 1], s1);
          }
          if (NS & 1) s0 = fma(Lsv[i * GP + NS - 1], Lsv[j * GP + NS - 1], s0);
          Lk[e] -= (s0 + s1);
        }
        __syncwarp();
      }
      for (int e = lane; e < NS * NS; e += 32) {  // dense copy (lower)
        const int i = e / NS, j = e - i * NS;
        Mw[e] = (j <= i) ? Lk[i * (i + 1) / 2 + j] : 0.0;
      }
      __syncwarp();
      double rd_mine = 1.0;  // 1 / L[lane][lane]
#pragma unroll
      for (int j = 0; j < NS; ++j) {  // right-looking Cholesky, lane = row
        const double v = fmax(Mw[j * NS + j], 1e-300);
        const double rdj = rsqrt(v);
        double lij = 0.0;
        const bool below = lane > j && lane < NS;
        if (lane == j) {
          Mw[j * NS + j] = v * rdj;
          rd_mine = rdj;
        }
        if (below) {
          lij = Mw[lane * NS + j] * rdj;
          Mw[lane * NS + j] = lij;
        }
        __syncwarp();
        if (below) {
#pragma unroll
          for (int c = j + 1; c < NS; ++c)
            if (c <= lane) Mw[lane * NS + c] = fma(-lij, Mw[c * NS + j], Mw[lane * NS + c]);
        }
        __syncwarp();
      }
      double rdv[NS];
#pragma unroll
      for (int i = 0; i < NS; ++i) rdv[i] = __shfl_sync(0xffffffffu, rd_mine, i);
      if (lane < NS) {  // triangular inverse, lane = column c: Li[i][c], i >= c
        double col[NS];
#pragma unroll
        for (int i = 0; i < NS; ++i) {
          double s = (i == lane) ? 1.0 : 0.0;
#pragma unroll
          for (int t = 0; t < i; ++t) s = fma(-Mw[i * NS + t], col[t], s);
          col[i] = (i >= lane) ? s * rdv[i] : 0.0;
          if (i >= lane) Lk[i * (i + 1) / 2 + lane] = col[i];
          Ld[i * GP + lane] = col[i];
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  // Phase 3 (all threads): G_k = Li_k[:, :NX] Ls_{k-1} for k = 1..N  (Ls_{k-1} sits in the H region, Li packed)
  for (int e = tid; e < N * NS * GP; e += ADMM_THREADS) {
    const int k1 = e / (NS * GP), rem = e - k1 * NS * GP, r = rem / GP, c = rem - r * GP;  // k1 = k - 1
    const double* li = P.Li + (size_t)(k1 + 1) * NP + r * (r + 1) / 2;
    const double* ls = P.H + (size_t)k1 * NS * HP + c;
    double sacc = 0.0;
    if (c < NS) {
#pragma unroll
      for (int i = 0; i < NX; ++i)
        if (i <= r) sacc = fma(li[i], ls[i * NS], sacc);
    }
    P.G[e] = sacc;
  }
  __syncthreads();
  // Phase 4 (one warp per stage, in place): H_k[r][i] = sum_{j>=r} Li_k[j][r] Ls_k[i][j]
  {
    const int warp = tid >> 5, lane = tid & 31;
    constexpr int PER = (NS * NX + 31) / 32;
    for (int k = warp; k < N; k += ADMM_THREADS / 32) {
      double* Hk = P.H + (size_t)k * NS * HP;
      const double* lk = P.Li + (size_t)k * NP;
      double out[PER];
#pragma unroll
      for (int t = 0; t < PER; ++t) {
        const int e = lane + 32 * t;
        double sacc = 0.0;
        if (e < NS * NX) {
          const int r = e / NX, i = e - r * NX;
#pragma unroll
          for (int j = 0; j < NS; ++j)
            if (j >= r) sacc = fma(lk[j * (j + 1) / 2 + r], Hk[i * NS + j], sacc);
        }
        out[t] = sacc;
      }
      __syncwarp();
#pragma unroll
      for (int t = 0; t < PER; ++t) {
        const int e = lane + 32 * t;
        if (e < NS * NX) {
          const int r = e / NX, i = e - r * NX;
          Hk[r * HP + i] = out[t];
        }
      }
      __syncwarp();
      for (int e = lane; e < NS; e += 32)
        for (int i = NX; i < HP; ++i) Hk[e * HP + i] = 0.0;   // pad columns
    }
  }
  __syncthreads();
}

// ---- one KKT solve per ADMM iteration: x_tilde = M^{-1} (sigma x - q + Abar' (rho z - y)) into P.xt ---------------------
// Phase A fuses the right-hand side with c_k = Li_k r_k: the columns of a stage are owned by consecutive lanes of ONE
// warp (SPW stages per warp and round), so only a __syncwarp separates the two.
template <int NX, int NU>
__device__ void block_solve(Prob<NX, NU>& P) {
  constexpr int NS = NX + NU, NP = Prob<NX, NU>::NP, GP = Prob<NX, NU>::GP, HP = Prob<NX, NU>::HP;
  constexpr int SPW = 32 / NS;                       // stages per warp per round
  constexpr int SPR = SPW * (ADMM_THREADS / 32);     // stages per round
  const int tid = threadIdx.x, N = P.N, n = P.n;
  {
    const int warp = tid >> 5, lane = tid & 31;
    const int slot = lane / NS, i = lane - slot * NS;
    for (int k0 = 0; k0 <= N; k0 += SPR) {
      const int k = k0 + warp * SPW + slot;
      const bool on = slot < SPW && k <= N && (k < N || i < NX);
      const int row = k * NS + i;
      if (on) P.xt[row] = fma(P.sigma, P.x[row], -P.qb[row]) + P.At_col(row, P.zt);
      __syncwarp();
      if (on) {
        const double* li = P.Li + k * NP + i * (i + 1) / 2;
        const double* r = P.xt + k * NS;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int j = 0; j < NS; j += 2) {
          if (j <= i) s0 = fma(li[j], r[j], s0);
          if (j + 1 <= i && j + 1 < NS) s1 = fma(li[j + 1], r[j + 1], s1);
        }
        P.w[row] = s0 + s1;
      }
    }
  }
  __syncthreads();
  // forward chain y_k = c_k - G_k y_{k-1} (warp 0, lane = row)
  if (tid < NS) {
    constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
    const double* g = P.G + tid * GP;
    for (int k = 1; k <= N; ++k) {
      const double* yp = P.w + (k - 1) * NS;
      double s0 = 0.0, s1 = 0.0;
      if (NS % 2 == 0) {
        const double2* g2 = reinterpret_cast<const double2*>(g);
        const double2* y2 = reinterpret_cast<const double2*>(yp);
#pragma unroll
        for (int c = 0; c < NS / 2; ++c) {
          const double2 a = g2[c], b = y2[c];
          s0 = fma(a.x, b.x, s0);
          s1 = fma(a.y, b.y, s1);
        }
      } else {
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          if (c & 1) s1 = fma(g[c], yp[c], s1);
          else s0 = fma(g[c], yp[c], s0);
        }
      }
      if (k < N || tid < NX) P.w[k * NS + tid] -= (s0 + s1);
      g += NS * GP;
      __syncwarp(mask);
    }
  }
  __syncthreads();
  // d_k = Li_k' y_k for every stage (parallel)
  for (int row = tid; row < n; row += ADMM_THREADS) {
    const int k = row / NS, i = row - k * NS;
    const int ns = (k < N) ? NS : NX;
    const double* lk = P.Li + k * NP + i;
    const double* yv = P.w + k * NS;
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int j = 0; j < NS; j += 2) {
      if (j >= i && j < ns) s0 = fma(lk[j * (j + 1) / 2], yv[j], s0);
      if (j + 1 >= i && j + 1 < ns) s1 = fma(lk[(j + 1) * (j + 2) / 2], yv[j + 1], s1);
    }
    P.xt[row] = s0 + s1;
  }
  __syncthreads();
  // backward chain v_k = d_k - H_k v_{k+1}[0:NX] (warp 0, lane = row)
  if (tid < NS) {
    constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
    const double* h = P.H + (size_t)(N - 1) * NS * HP + tid * HP;
    for (int k = N - 1; k >= 0; --k) {
      const double* vn = P.xt + (k + 1) * NS;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int i = 0; i + 1 < NX; i += 2) {
        s0 = fma(h[i], vn[i], s0);
        s1 = fma(h[i + 1], vn[i + 1], s1);
      }
      if (NX & 1) s0 = fma(h[NX - 1], vn[NX - 1], s0);
      P.xt[k * NS + tid] -= (s0 + s1);
      h -= NS * HP;
      __syncwarp(mask);
    }
  }
}

struct CheckOut {
  double pri, dua, eps_pri, eps_dua, rho_est;
  int status;
};

// Residuals, termination, infeasibility certificates and the rho estimate (oracle/mpc.py _residuals/_check/_rho_estimate;
// OSQP auxil.c compute_pri_res/compute_dua_res/is_primal_infeasible/is_dual_infeasible/compute_rho_estimate).
// dx is in P.w, dy in P.zt.  All threads participate.
template <int NX, int NU>
__device__ CheckOut check_termination(Prob<NX, NU>& P, const GpmpcAdmmSettings& S, bool approx) {
  const int tid = threadIdx.x;
  double a[6] = {0, 0, 0, 0, 0, 0};  // max: pri_u, |Ax|_u, |z|_u, pri_s, |Ax|_s, |z|_s
  double b[2] = {0, 0};              // max: |E dy_proj| ; (dual infeas) bound violation flag as max
  double lhs[1] = {0};
  double ndx[3] = {0, 0, 0};  // max |D dx|, max |Pb dx / D|
  double qdx[1] = {0};
  for (int j = tid; j < P.n; j += ADMM_THREADS) {
    const double dx = P.w[j], Dj = P.D[j];
    ndx[0] = fmax(ndx[0], fabs(Dj * dx));
    ndx[1] = fmax(ndx[1], fabs(P.Pb[j] * dx / Dj));
    qdx[0] = fma(P.qb[j], dx, qdx[0]);
  }
  block_reduce<3, true>(ndx, P.red);
  block_reduce<1, false>(qdx, P.red);
  const double f = approx ? 10.0 : 1.0;
  const double eps_pinf = f * S.eps_prim_inf, eps_dinf = f * S.eps_dual_inf;
  const double nrm_dx = ndx[0];
  for (int r = tid; r < P.m; r += ADMM_THREADS) {
    const double Er = P.E[r];
    const double Ax = P.A_row(r, P.x), zz = P.z[r], Ei = 1.0 / Er;
    a[0] = fmax(a[0], fabs(Ei * (Ax - zz)));
    a[1] = fmax(a[1], fabs(Ei * Ax));
    a[2] = fmax(a[2], fabs(Ei * zz));
    a[3] = fmax(a[3], fabs(Ax - zz));
    a[4] = fmax(a[4], fabs(Ax));
    a[5] = fmax(a[5], fabs(zz));
    // primal infeasibility: project dy on the polar of the recession cone of [l,u]
    const double l = P.lbs[r], u = P.ubs[r];
    double dy = P.zt[r];
    const bool uinf = u > INF_T, linf = l < -INF_T;
    if (uinf && linf) dy = 0.0;
    else if (uinf) dy = fmin(dy, 0.0);
    else if (linf) dy = fmax(dy, 0.0);
    P.zt[r] = dy;  // projected dy (consumed below by At_col)
    b[0] = fmax(b[0], fabs(Er * dy));
    lhs[0] += u * fmax(dy, 0.0) + l * fmin(dy, 0.0);
    // dual infeasibility: (A dx)_r must stay inside the recession cone
    const double Adx = P.A_row(r, P.w) * Ei;
    const bool bad = (!uinf && Adx > eps_dinf * nrm_dx) || (!linf && Adx < -eps_dinf * nrm_dx);
    b[1] = fmax(b[1], bad ? 1.0 : 0.0);
  }
  block_reduce<6, true>(a, P.red);
  block_reduce<2, true>(b, P.red);
  block_reduce<1, false>(lhs, P.red);
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
__global__ void __launch_bounds__(ADMM_THREADS, 3) admm_kernel(int N, int nbatch, const double* __restrict__ ABg,
                                                               const double* __restrict__ beqg,
                                                               const double* __restrict__ qg,
                                                               const double* __restrict__ Pdiag,
                                                               const double* __restrict__ lb,
                                                               const double* __restrict__ ub, GpmpcAdmmSettings S,
                                                               const int* __restrict__ active,
                                                               double* __restrict__ scratch, double* __restrict__ xg,
                                                               double* __restrict__ yg, double* __restrict__ rhog,
                                                               int* __restrict__ status_g, int* __restrict__ iters_g,
                                                               double* __restrict__ obj_g, double* __restrict__ pri_g,
                                                               double* __restrict__ dua_g) {
  constexpr int NS = NX + NU;
  extern __shared__ __align__(16) double smem[];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  Prob<NX, NU> P;
  P.carve(smem, N);
  P.sigma = S.sigma;
  P.rho = (S.rho_from_input && rhog[b] > 0.0) ? rhog[b] : S.rho;
  P.c = 1.0;
  const int n = P.n, m = P.m, neq = P.neq;
  P.D = scratch + (size_t)b * (n + m);
  P.E = P.D + n;

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
  for (int i = tid; i < N * NS * Prob<NX, NU>::HP; i += ADMM_THREADS) P.H[i] = 0.0;  // pad columns stay zero
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
  // scaled bounds: equality rows l = u = E beq ; box rows E lb, E ub
  for (int r = tid; r < m; r += ADMM_THREADS) {
    const double Er = P.E[r];
    if (r < neq) {
      const double v = Er * beqg[(size_t)b * neq + r];
      P.lbs[r] = v;
      P.ubs[r] = v;
    } else {
      P.lbs[r] = Er * fmax(lb[(size_t)b * S.bound_stride + r - neq], -OSQP_INFTY);  // bounds clipped as OSQP's setup does (osqp.h OSQP_INFTY)
      P.ubs[r] = Er * fmin(ub[(size_t)b * S.bound_stride + r - neq], OSQP_INFTY);
    }
  }
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
  // w = rho*z - y (into zt) for the first iteration
  for (int r = tid; r < m; r += ADMM_THREADS) P.zt[r] = (S.max_iter > 0) ? fma(P.rho_row(r), P.z[r], -P.y[r]) : 0.0;
  for (int j = tid; j < n; j += ADMM_THREADS) P.w[j] = 0.0;
  __syncthreads();
  for (it = 1; it <= S.max_iter; ++it) {
    const bool can_check = S.check_every > 0 && (it % S.check_every) == 0;
    const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval > 0 && (it % S.adaptive_rho_interval) == 0;
    const bool checking = can_check || can_adapt || it == S.max_iter;
    block_solve(P);  // xt <- x_tilde  (reads zt = rho z - y)
    __syncthreads();
    // z_tilde = Abar x_tilde ; z, y, x updates.  Normal iterations leave the next w = rho z - y in zt; iterations that
    // run the termination check leave dy in zt and dx in w instead (consumed by check_termination).
    const double rho_eq = RHO_EQ_OVER_INEQ * P.rho;
    const double inv_eq = 1.0 / rho_eq, inv_rho = 1.0 / P.rho, inv_min = 1.0 / RHO_MIN;
    for (int r = tid; r < m; r += ADMM_THREADS) {
      const double zt = P.A_row(r, P.xt);
      const double rr = P.rho_row(r), zp = P.z[r], yy = P.y[r];
      const double ri = (rr == rho_eq) ? inv_eq : ((rr == RHO_MIN) ? inv_min : inv_rho);
      const double zr = alpha * zt + (1.0 - alpha) * zp;
      const double zn = fmin(fmax(fma(yy, ri, zr), P.lbs[r]), P.ubs[r]);
      const double dy = rr * (zr - zn);
      const double yn = yy + dy;
      P.z[r] = zn;
      P.y[r] = yn;
      P.zt[r] = checking ? dy : fma(rr, zn, -yn);
    }
    for (int j = tid; j < n; j += ADMM_THREADS) {
      const double xp = P.x[j];
      const double xn = alpha * P.xt[j] + (1.0 - alpha) * xp;
      P.x[j] = xn;
      if (checking) P.w[j] = xn - xp;
    }
    __syncthreads();
    if (checking) {
      if (can_check || can_adapt) {
        co = check_termination(P, S, false);
        if (can_check && co.status != GPMPC_QP_UNSOLVED) {
          status = co.status;
          break;
        }
        if (can_adapt &&
            (co.rho_est > P.rho * S.adaptive_rho_tolerance || co.rho_est < P.rho / S.adaptive_rho_tolerance)) {
          P.rho = co.rho_est;
          __syncthreads();
          factorize(P);
        }
      }
      if (it < S.max_iter) {  // restore zt = rho z - y for the next iteration
        __syncthreads();
        for (int r = tid; r < m; r += ADMM_THREADS) P.zt[r] = fma(P.rho_row(r), P.z[r], -P.y[r]);
        __syncthreads();
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

// stream-ordered scratch for the cumulative scaling vectors D (n) and E (m) of every scenario
double* alloc_scaling_scratch(size_t doubles, cudaStream_t stream) {
  static GpmpcPerDevice pool_configured;
  if (!pool_configured.here()) {  // keep freed blocks in the pool: without this every call re-allocates from the driver
    int dev = 0;
    cudaMemPool_t pool;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      unsigned long long thr = ~0ULL;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    cudaGetLastError();
    pool_configured.here() = 1;
  }
  double* scratch = nullptr;
  if (cudaMallocAsync(&scratch, sizeof(double) * doubles, stream) != cudaSuccess) {
    cudaGetLastError();
    gpmpc_set_error("gpmpc_admm_solve: could not allocate the scaling scratch");
    return nullptr;
  }
  return scratch;
}

template <int NX, int NU>
int launch_admm(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag,
                const double* lb, const double* ub, const GpmpcAdmmSettings& S, const int* active, double* x, double* y,
                double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream) {
  constexpr int NS = NX + NU;
  const size_t smem = Prob<NX, NU>::smem_doubles(N) * sizeof(double);
  if (smem > 227 * 1024) {
    gpmpc_set_error("gpmpc_admm_solve: horizon too long for the shared-memory resident solver");
    return GPMPC_ERR_ARG;
  }
  auto kern = admm_kernel<NX, NU>;
  static GpmpcPerDevice configured;
  if ((int)smem > configured.here()) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured.here() = (int)smem;
  }
  const size_t n = (size_t)N * NS + NX, m = (size_t)NX * (N + 1) + n;
  double* scratch = alloc_scaling_scratch((n + m) * (size_t)B, stream);
  if (!scratch) return GPMPC_ERR_CUDA;
  kern<<<B, ADMM_THREADS, smem, stream>>>(N, B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters,
                                          obj, pri, dua);
  int rc = gpmpc_check_launch("admm");
  cudaFreeAsync(scratch, stream);
  return rc;
}

}  // namespace

// second-generation kernel (csrc/admm_stage.cu); returns GPMPC_ERR_ARG without launching when the problem does not fit
template <int NX, int NU>
int launch_admm_stage(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag,
                      const double* lb, const double* ub, const GpmpcAdmmSettings& S, const int* active,
                      double* scratch, double* x, double* y, double* rho, int* status, int* iters, double* obj,
                      double* pri, double* dua, cudaStream_t stream);

// third-generation kernel (csrc/admm3.cu); returns GPMPC_ERR_ARG without launching when N is not the compiled horizon
template <int NX, int NU>
size_t admm3_scratch_doubles(int N);
template <int NX, int NU>
int launch_admm3(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag, const double* lb,
                 const double* ub, const GpmpcAdmmSettings& S, const int* active, double* scratch, double* x, double* y,
                 double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream);

extern "C" int gpmpc_admm_solve(int nx, int nu, int N, int B, const double* AB, const double* beq, const double* q,
                                const double* Pdiag, const double* lb, const double* ub,
                                const GpmpcAdmmSettings* settings, const int* active, double* x, double* y, double* rho,
                                int* status, int* iters, double* obj, double* pri_res, double* dua_res, void* stream) {
  GPMPC_REQUIRE(settings != nullptr, "gpmpc_admm_solve: settings is NULL");
  GPMPC_REQUIRE(N >= 1 && B >= 0, "gpmpc_admm_solve: bad N or B");
  if (B == 0) return GPMPC_OK;
  // kernel_variant 0: the fastest kernel measured for the shape -- third generation (csrc/admm3.cu: light iterations +
  // cyclic reduction) for nx = 14 at its compiled horizon N = 30 (ADMM phase of the closed loop 19.5 -> 14.5 ms), second
  // generation (csrc/admm_stage.cu) otherwise: at nx = 7, N = 20 the third generation wins per light iteration (32.3 vs 39.3
  // us per iteration of 4096 QPs) but loses it again in its heavier checking iterations and refactorisations (closed loop:
  // 3.06 vs 2.83 ms).  3 / 2 / 1: force the third / second / first generation.
  const bool gen3 = settings->kernel_variant == 3 || (settings->kernel_variant == 0 && nx == 14);
  if ((nx == 7 || nx == 14) && nu == 3 && gen3) {
    const size_t per = (nx == 7) ? admm3_scratch_doubles<7, 3>(N) : admm3_scratch_doubles<14, 3>(N);
    double* scratch = alloc_scaling_scratch(per * (size_t)B, (cudaStream_t)stream);
    if (!scratch) return GPMPC_ERR_CUDA;
    const int rc = (nx == 7) ? launch_admm3<7, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, scratch, x, y, rho,
                                                  status, iters, obj, pri_res, dua_res, (cudaStream_t)stream)
                             : launch_admm3<14, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, scratch, x, y, rho,
                                                   status, iters, obj, pri_res, dua_res, (cudaStream_t)stream);
    cudaFreeAsync(scratch, (cudaStream_t)stream);
    if (rc != GPMPC_ERR_ARG) return rc;  // GPMPC_ERR_ARG: not the compiled horizon
  }
  if ((nx == 7 || nx == 14) && nu == 3 && settings->kernel_variant != 1) {
    const size_t nv = (size_t)N * (nx + nu) + nx, nm = 2 * nv + (size_t)nx * (N + 1);  // n + m
    double* scratch = alloc_scaling_scratch(nm * (size_t)B, (cudaStream_t)stream);
    if (!scratch) return GPMPC_ERR_CUDA;
    const int rc = (nx == 7) ? launch_admm_stage<7, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, scratch, x, y,
                                                       rho, status, iters, obj, pri_res, dua_res, (cudaStream_t)stream)
                             : launch_admm_stage<14, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, scratch, x, y,
                                                        rho, status, iters, obj, pri_res, dua_res, (cudaStream_t)stream);
    cudaFreeAsync(scratch, (cudaStream_t)stream);
    if (rc != GPMPC_ERR_ARG) return rc;  // GPMPC_ERR_ARG: horizon too long for the stage-parallel kernel
  }
  if (nx == 7 && nu == 3)
    return launch_admm<7, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, x, y, rho, status, iters, obj, pri_res,
                             dua_res, (cudaStream_t)stream);
  if (nx == 14 && nu == 3)
    return launch_admm<14, 3>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, x, y, rho, status, iters, obj,
                              pri_res, dua_res, (cudaStream_t)stream);
  gpmpc_set_error("gpmpc_admm_solve: only (nx,nu) = (7,3) and (14,3) are instantiated");
  return GPMPC_ERR_ARG;
}

// ---- allocation-free variant ---------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  The Ruiz-scaling scratch of the stage kernel (n + m doubles
// per QP) lives in a file-scope device array instead of a stream-ordered allocation; every GPU that loads the library holds
// its GPMPC_NOALLOC_MAX_DOUBLES x 8 bytes (4.6 MB) until exit.  All calls share the array, so calls that overlap on the GPU
// (different streams, or two host threads) race on it.  It keeps the previous call's contents; nothing has to start at zero:
// the kernel writes every scaling entry of a QP (admm_stage.cu, D then E) before it reads it.
static __device__ __align__(256) double g_noalloc_scratch[GPMPC_NOALLOC_MAX_DOUBLES];

extern "C" int gpmpc_admm_solve_noalloc(int N, int B, const double* AB, const double* beq, const double* q,
                                        const double* Pdiag, const double* lb, const double* ub,
                                        const GpmpcAdmmSettings* settings, const int* active, double* x, double* y,
                                        double* rho, int* status, int* iters, double* obj, double* pri_res, double* dua_res,
                                        void* stream) {
  GPMPC_REQUIRE(settings != nullptr, "gpmpc_admm_solve_noalloc: settings is NULL");
  GPMPC_REQUIRE(N >= 1 && B >= 0, "gpmpc_admm_solve_noalloc: bad N or B");
  if (B == 0) return GPMPC_OK;
  constexpr int NX = 7, NU = 3;
  const size_t nv = (size_t)N * (NX + NU) + NX, nm = 2 * nv + (size_t)NX * (N + 1);  // n + m doubles per QP
  if (nm * (size_t)B > (size_t)GPMPC_NOALLOC_MAX_DOUBLES) {
    gpmpc_set_error("gpmpc_admm_solve_noalloc: B x (n + m) exceeds GPMPC_NOALLOC_MAX_DOUBLES (574464 doubles = 1024 QPs at "
                    "N = 20); split the batch or call gpmpc_admm_solve");
    return GPMPC_ERR_ARG;
  }
  // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that
  // device current.
  static double* dev_ptr[GPMPC_MAX_DEVICES] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= GPMPC_MAX_DEVICES) {
    cudaGetLastError();
    gpmpc_set_error("gpmpc_admm_solve_noalloc: no current CUDA device");
    return GPMPC_ERR_CUDA;
  }
  if (!dev_ptr[dev]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      gpmpc_set_error("gpmpc_admm_solve_noalloc: cudaGetSymbolAddress failed for the static scratch");
      return GPMPC_ERR_CUDA;
    }
    dev_ptr[dev] = ptr;
  }
  const int rc = launch_admm_stage<NX, NU>(N, B, AB, beq, q, Pdiag, lb, ub, *settings, active, dev_ptr[dev], x, y, rho, status,
                                           iters, obj, pri_res, dua_res, (cudaStream_t)stream);
  if (rc == GPMPC_ERR_ARG) gpmpc_set_error("gpmpc_admm_solve_noalloc: horizon too long for the stage kernel");
  return rc;
}
