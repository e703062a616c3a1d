// Stage-parallel OSQP-style ADMM for the stage-structured MPC QP (same problem, same algorithm and the same C entry
// point as csrc/admm.cu -- see the header there for the reference call sites osqp_rti.py:464-478,496,517,524,527 and
// for what is restated from Stellato et al. 2020).  This file is the second generation of the kernel:
//
//  * one thread per (stage k, variable i): thread (k,i) owns x_{k,i}, the box row on that variable and -- for the state
//    part i < NX -- the equality row that *defines* x_{k,i} (the initial-state row for k = 0, the dynamics row
//    A_{k-1} x_{k-1} + B_{k-1} u_{k-1} - x_k = b_{k-1} otherwise).  All per-row / per-variable iterates
//    (x, z, y, bounds, scalings, rho class) live in REGISTERS; shared memory holds only the stage blocks, the factor
//    and four exchange vectors, so a QP needs ~50 KB instead of 75 KB.
//  * the reduced KKT matrix M = P + sigma I + A' diag(rho) A (block tridiagonal, NS x NS blocks) is factorised by a
//    TWISTED block elimination ("burn at both ends"): stages 0..p-1 are eliminated top-down by warp 0 while stages
//    N..p+1 are eliminated bottom-up by warp 1; the middle stage p receives both Schur updates.  The factor is stored
//    as explicit inverses S_k^{-1} and coupling products F_k = C_{k-1} S_{k-1}^{-1} (top), Fb_k = C_k' S_{k+1}^{-1}
//    (bottom), so that one KKT solve is
//        forward   yt_k = r_k - F_k yt_{k-1}   (k = 1..p)      ||   yt_k = r_k - Fb_k yt_{k+1}   (k = N-1..p)
//        parallel  d_k  = S_k^{-1} yt_k        (all stages at once)
//        backward  v_k  = d_k - F_{k+1}' v_{k+1} (k = p-1..0)   ||   v_k = d_k - Fb_{k-1}' v_{k-1} (k = p+1..N)
//    i.e. two dependent chains of N/2 one-matvec steps each (the first generation had 2N steps on one warp).
//  * termination checks use ONE block reduction (20 maxima + 2 sums) instead of eight.
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

namespace {

constexpr double OSQP_INFTY = 1e30, MIN_SCALING = 1e-4, MAX_SCALING = 1e4;
constexpr double RHO_MIN = 1e-6, RHO_MAX = 1e6, RHO_TOL = 1e-4, RHO_EQ_OVER_INEQ = 1e3;
constexpr double INF_T = OSQP_INFTY * MIN_SCALING;
constexpr int MAX_WARPS = 8;
#ifndef ADMM_STAGE_MIN_CTAS
#define ADMM_STAGE_MIN_CTAS 3  // resident QPs per SM the register allocation is sized for (128 threads x 3 -> 168 registers;
                               // measured on B200: 4 QPs per SM at 128 registers spills and is 4-10 % slower)
#endif
constexpr int NRED_MAX = 20, NRED_SUM = 2, NRED = NRED_MAX + NRED_SUM;

__device__ __forceinline__ double limit_scaling(double v) {
  v = (v < MIN_SCALING) ? 1.0 : v;
  return (v > MAX_SCALING) ? MAX_SCALING : v;
}

template <int NX, int NU>
struct Stage {
  static constexpr int NS = NX + NU;
  static __host__ __device__ int mid(int N) { return N / 2; }
  static __host__ __device__ size_t smem_doubles(int N) {
    const int p = mid(N);
    size_t s = 0;
    auto add = [&](size_t c) { s += (c + 1) & ~(size_t)1; };
    add((size_t)N * NX * NS);            // AB
    add((size_t)(N + 1) * NS * NS);      // Sinv
    add((size_t)(p > 0 ? p : 1) * NS * NX);  // Ft
    add((size_t)(N - p > 0 ? N - p : 1) * NS * NS);  // Fbt
    for (int t = 0; t < 4; ++t) add((size_t)(N + 1) * NS);  // yt xt sPb sdy
    for (int t = 0; t < 4; ++t) add((size_t)(N + 2) * NX);  // weq tmp ec sbe
    add((size_t)NS);                     // mrg
    add((size_t)MAX_WARPS * NRED);       // red
    return s;
  }
};

// In-place inverse of a symmetric positive definite NS x NS matrix held one row per lane (lanes >= NS hold zeros):
// Gauss-Jordan without pivoting (the pivots are the positive Schur complements of an SPD matrix).
template <int NS>
__device__ __forceinline__ void gj_inverse(double (&a)[NS], int lane) {
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    double rowj[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) rowj[c] = __shfl_sync(0xffffffffu, a[c], j);
    const double rp = 1.0 / rowj[j];
    const double f = a[j];
    if (lane == j) {
#pragma unroll
      for (int c = 0; c < NS; ++c) a[c] = rowj[c] * rp;
      a[j] = rp;
    } else {
      const double g = f * rp;
#pragma unroll
      for (int c = 0; c < NS; ++c) a[c] = fma(-g, rowj[c], a[c]);
      a[j] = -g;
    }
  }
}

// block reduction: v[0..NM) maxima, v[NRED_MAX..NRED_MAX+NSUM) sums; result broadcast to every thread
template <int NM, int NSUM>
__device__ __forceinline__ void block_reduce_mixed(double (&v)[NRED], double* red, int warp, int lane, int nwarps) {
#pragma unroll
  for (int e = 0; e < NM; ++e) v[e] = warp_max(v[e]);
#pragma unroll
  for (int e = 0; e < NSUM; ++e) v[NRED_MAX + e] = warp_sum(v[NRED_MAX + e]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int e = 0; e < NM; ++e) red[warp * NRED + e] = v[e];
#pragma unroll
    for (int e = 0; e < NSUM; ++e) red[warp * NRED + NRED_MAX + e] = v[NRED_MAX + e];
  }
  __syncthreads();
#pragma unroll
  for (int e = 0; e < NM; ++e) {
    double r = red[e];
    for (int w = 1; w < nwarps; ++w) r = fmax(r, red[w * NRED + e]);
    v[e] = r;
  }
#pragma unroll
  for (int e = 0; e < NSUM; ++e) {
    double r = red[NRED_MAX + e];
    for (int w = 1; w < nwarps; ++w) r += red[w * NRED + NRED_MAX + e];
    v[NRED_MAX + e] = r;
  }
}

template <int I>
struct IC { static constexpr int value = I; };

// NT > 0: horizon known at compile time (all shared-memory offsets fold); 0: run time.  NSLOT: (stage, variable) slots
// per thread -- NTHR threads cover (N+1)*NS <= NTHR*NSLOT slots; slot s of thread t is index t + NTHR s.  MINB: resident
// QPs per SM the register allocation is sized for.
template <int NX, int NU, int NT, int NSLOT, int NTHR, int MINB>
__global__ void __launch_bounds__(NTHR, MINB)
admm_stage_kernel(int Nrt, int nbatch, const double* __restrict__ ABg, const double* __restrict__ beqg,
                  const double* __restrict__ qg, const double* __restrict__ Pdiag, const double* __restrict__ lb,
                  const double* __restrict__ ub, GpmpcAdmmSettings S, const int* __restrict__ active,
                  double* __restrict__ scratch, double* __restrict__ xg, double* __restrict__ yg,
                  double* __restrict__ rhog, int* __restrict__ status_g, int* __restrict__ iters_g,
                  double* __restrict__ obj_g, double* __restrict__ pri_g, double* __restrict__ dua_g) {
  using L = Stage<NX, NU>;
  constexpr int NS = L::NS, NWARPS = NTHR / 32;
  extern __shared__ __align__(16) double smem[];
  const int N = (NT > 0) ? NT : Nrt;
  const int b = blockIdx.x, tid = threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  const int warp = tid >> 5, lane = tid & 31;
  const int p = L::mid(N);
  const int n = N * NS + NX, neq = NX * (N + 1), m = neq + n;
  const int npad = (N + 1) * NS, epad = (N + 2) * NX;

  // ---- shared memory carve (must match Stage::smem_doubles) ---------------------------------------------------------
  double* sp = smem;
  auto take = [&](size_t cnt) { double* q_ = sp; sp += (cnt + 1) & ~(size_t)1; return q_; };
  double* AB = take((size_t)N * NX * NS);
  double* Sinv = take((size_t)(N + 1) * NS * NS);
  double* Ft = take((size_t)(p > 0 ? p : 1) * NS * NX);           // Ft[k-1][j][i] = F_k[i][j],   k = 1..p
  double* Fbt = take((size_t)(N - p > 0 ? N - p : 1) * NS * NS);  // Fbt[k-p][c][j] = Fb_k[j][c], k = p..N-1
  double* yt = take(npad);
  double* xt = take(npad);
  double* sPb = take(npad);   // scaled cost diagonal
  double* sdy = take(npad);   // dy of the box rows at termination checks
  double* weq = take(epad);   // equality-row exchange vector, indexed by owner (k,i): k*NX+i ; stage N+1 stays zero
  double* tmp = take(epad);
  double* ec = take(epad);    // own-variable coefficient of every equality row (scaled Ei / Nd)
  double* sbe = take(epad);   // scaled equality right-hand sides
  double* mrg = take((size_t)NS);  // bottom chain's contribution to the middle stage
  double* red = take((size_t)MAX_WARPS * NRED);
  // cumulative Ruiz scalings (needed only at termination checks and for unscaling): global scratch, D (n) then E (m)
  double* gD = scratch + (size_t)b * (n + m);
  double* gE = gD + n;

  // ---- thread <-> (stage, variable) slots ------------------------------------------------------------------------
  int j[NSLOT], er[NSLOT], oABk[NSLOT], oABp[NSLOT], oWn[NSLOT], oVp[NSLOT];
  bool on[NSLOT], live[NSLOT], eqo[NSLOT], hasAB[NSLOT], hasPrev[NSLOT];
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    const int idx = tid + s * NTHR;
    const int k = idx / NS, i = idx - k * NS;
    on[s] = idx < npad;                              // has a slot in the padded (N+1) x NS layout
    live[s] = on[s] && (k < N || i < NX);            // a real variable
    eqo[s] = live[s] && i < NX;                      // owns an equality row
    hasAB[s] = on[s] && k < N;                       // stage has a dynamics block (columns of AB_k)
    hasPrev[s] = eqo[s] && k >= 1;                   // own equality row is a dynamics row (rows of AB_{k-1})
    j[s] = on[s] ? idx : 0;                          // index into the padded per-variable arrays and the global x
    er[s] = eqo[s] ? k * NX + i : 0;                 // index into the equality-row arrays and the global rows
    oABk[s] = (hasAB[s] ? k : 0) * NX * NS + i;                // column i of AB_k, stride NS
    oABp[s] = (hasPrev[s] ? (k - 1) * NX + i : 0) * NS;        // row i of AB_{k-1}
    oWn[s] = ((on[s] ? k : 0) + 1) * NX;                        // equality-row values owned by stage k+1
    oVp[s] = (hasPrev[s] ? k - 1 : 0) * NS;                     // variables of stage k-1
  }

  auto reduce = [&](double (&v)[NRED], auto nmax_c, auto nsum_c) {
    block_reduce_mixed<decltype(nmax_c)::value, decltype(nsum_c)::value>(v, red, warp, lane, NWARPS);
  };

  // ---- load ------------------------------------------------------------------------------------------------------
  for (int e = tid; e < N * NX * NS; e += NTHR) AB[e] = ABg[(size_t)b * N * NX * NS + e];
  for (int e = tid; e < epad; e += NTHR) { weq[e] = 0.0; tmp[e] = 0.0; ec[e] = 0.0; sbe[e] = 0.0; }
  const double sigma = S.sigma;
  double rho = (S.rho_from_input && rhog[b] > 0.0) ? rhog[b] : S.rho, cscale = 1.0;
  // per-slot iterates (registers) and per-slot constants of the scaled problem
  double x[NSLOT], zb[NSLOT], yb[NSLOT], ze[NSLOT], ye[NSLOT], q[NSLOT], Bd[NSLOT], lbs[NSLOT], ubs[NSLOT];
  int cls[NSLOT];  // box row class: 0 loose (rho_min), 1 equality-like (rho_eq), 2 inequality (rho)
  __syncthreads();
  {
    // ---- modified Ruiz equilibration + cost scaling (OSQP scaling.c; oracle/mpc.py OSQPOracle._scale) ---------------
    double Pb[NSLOT], e_[NSLOT], Dj[NSLOT], Eb[NSLOT], Ee[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      x[s] = zb[s] = yb[s] = ze[s] = ye[s] = 0.0;
      q[s] = Pb[s] = Bd[s] = e_[s] = 0.0;
      Dj[s] = Eb[s] = Ee[s] = 1.0;
      cls[s] = 2;
      if (live[s]) {
        Pb[s] = Pdiag[j[s]];
        q[s] = qg[(size_t)b * n + j[s]];
        Bd[s] = 1.0;
        if (eqo[s]) e_[s] = (er[s] < NX) ? 1.0 : -1.0;
      }
    }
    for (int it = 0; it < S.scaling; ++it) {
      double Dt[NSLOT], Etb[NSLOT], Ete[NSLOT];
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        double v = fmax(fabs(Pb[s]), fabs(Bd[s]));
        v = fmax(v, fabs(e_[s]));
        if (hasAB[s])
#pragma unroll
          for (int r = 0; r < NX; ++r) v = fmax(v, fabs(AB[oABk[s] + r * NS]));
        Dt[s] = 1.0 / sqrt(limit_scaling(v));
        Etb[s] = 1.0 / sqrt(limit_scaling(fabs(Bd[s])));
        double ve = fabs(e_[s]);
        if (hasPrev[s])
#pragma unroll
          for (int c = 0; c < NS; ++c) ve = fmax(ve, fabs(AB[oABp[s] + c]));
        Ete[s] = 1.0 / sqrt(limit_scaling(ve));
        if (on[s]) yt[j[s]] = Dt[s];
        if (eqo[s]) weq[er[s]] = Ete[s];
      }
      __syncthreads();
      double rv[NRED];
      rv[0] = 0.0;
      rv[NRED_MAX] = 0.0;
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        if (hasAB[s]) {
#pragma unroll
          for (int r = 0; r < NX; ++r) {
            double* a = AB + oABk[s] + r * NS;
            *a = (weq[oWn[s] + r] * *a) * Dt[s];
          }
        }
        e_[s] = (Ete[s] * e_[s]) * Dt[s];
        Bd[s] = (Etb[s] * Bd[s]) * Dt[s];
        Pb[s] = (Dt[s] * Pb[s]) * Dt[s];
        q[s] = Dt[s] * q[s];
        Dj[s] *= Dt[s]; Eb[s] *= Etb[s]; Ee[s] *= Ete[s];
        if (live[s]) {
          rv[0] = fmax(rv[0], fabs(q[s]));
          rv[NRED_MAX] += fabs(Pb[s]);
        }
      }
      reduce(rv, IC<1>{}, IC<1>{});
      double ct = limit_scaling(fmax(rv[NRED_MAX] / n, limit_scaling(rv[0])));
      ct = 1.0 / ct;
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        Pb[s] *= ct;
        q[s] *= ct;
      }
      cscale *= ct;
      __syncthreads();
    }
    // scaled bounds, row classes, warm start (x = x_in / D ; y = c y_in / E)
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      lbs[s] = ubs[s] = 0.0;
      if (live[s]) {
        lbs[s] = Eb[s] * fmax(lb[(size_t)b * S.bound_stride + j[s]], -OSQP_INFTY);  // bounds clipped as OSQP's setup does (osqp.h OSQP_INFTY)
        ubs[s] = Eb[s] * fmin(ub[(size_t)b * S.bound_stride + j[s]], OSQP_INFTY);
        if (lbs[s] < -INF_T && ubs[s] > INF_T) cls[s] = 0;
        else if (ubs[s] - lbs[s] < RHO_TOL) cls[s] = 1;
        if (S.warm_x) x[s] = xg[(size_t)b * n + j[s]] / Dj[s];
        if (S.warm_y) yb[s] = cscale * yg[(size_t)b * m + neq + j[s]] / Eb[s];
        gD[j[s]] = Dj[s];
        gE[neq + j[s]] = Eb[s];
      }
      if (on[s]) {
        sPb[j[s]] = Pb[s];
        sdy[j[s]] = 0.0; yt[j[s]] = 0.0; xt[j[s]] = x[s];
      }
      if (eqo[s]) {
        ec[er[s]] = e_[s];
        gE[er[s]] = Ee[s];
        sbe[er[s]] = Ee[s] * beqg[(size_t)b * neq + er[s]];
        if (S.warm_y) ye[s] = cscale * yg[(size_t)b * m + er[s]] / Ee[s];
        weq[er[s]] = 0.0;
      }
    }
  }
  __syncthreads();

  // (A v)_own-equality-row given v in shared memory (padded stage layout) and the own value vj  [only for eqo slots]
  auto eq_row = [&](int s, const double* v, double vj) -> double {
    double s0 = 0.0, s1 = 0.0;
    if (hasPrev[s]) {
      const double* a = AB + oABp[s];
      const double* vp = v + oVp[s];
#pragma unroll
      for (int c = 0; c + 1 < NS; c += 2) {
        s0 = fma(a[c], vp[c], s0);
        s1 = fma(a[c + 1], vp[c + 1], s1);
      }
      if (NS & 1) s0 = fma(a[NS - 1], vp[NS - 1], s0);
    }
    return fma(ec[er[s]], vj, s0 + s1);
  };
  // (A' w)_own-column given the own box value wb, the own equality value we and the shared equality vector
  auto at_col = [&](int s, double wb, double we, const double* wv) -> double {
    double r0 = Bd[s] * wb;
    if (eqo[s]) r0 = fma(ec[er[s]], we, r0);
    if (hasAB[s]) {
      const double* a = AB + oABk[s];
      const double* wn = wv + oWn[s];
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int r = 0; r + 1 < NX; r += 2) {
        s0 = fma(a[r * NS], wn[r], s0);
        s1 = fma(a[(r + 1) * NS], wn[r + 1], s1);
      }
      if (NX & 1) s0 = fma(a[(NX - 1) * NS], wn[NX - 1], s0);
      r0 += s0 + s1;
    }
    return r0;
  };

  // z = A x for the warm start
  if (S.warm_x) {
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      zb[s] = Bd[s] * x[s];
      if (eqo[s]) ze[s] = eq_row(s, xt, x[s]);
    }
  }
  __syncthreads();

  // ---- twisted block factorisation ----------------------------------------------------------------------------------
  auto factorize = [&]() {
    const double rho_eq = RHO_EQ_OVER_INEQ * rho;
    // phase 1: row i of M_k into Sinv[k][i][:]
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      if (on[s]) {
        const int k = j[s] / NS, i = j[s] - k * NS;
        double col[NX];
#pragma unroll
        for (int r = 0; r < NX; ++r) col[r] = hasAB[s] ? AB[oABk[s] + r * NS] : 0.0;
        double* out = Sinv + (size_t)j[s] * NS;
        const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
        const double e_ = eqo[s] ? ec[er[s]] : 0.0;
        double d = sPb[j[s]] + sigma;
        d = fma(rb, Bd[s] * Bd[s], d);
        d = fma(rho_eq, e_ * e_, d);
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          double acc = 0.0;
          if (hasAB[s]) {
            const double* a = AB + (size_t)k * NX * NS + c;
#pragma unroll
            for (int r = 0; r < NX; ++r) acc = fma(col[r], a[r * NS], acc);
          }
          double v = rho_eq * acc;
          if (c == i) v = live[s] ? v + d : 1.0;
          out[c] = v;
        }
      }
    }
    __syncthreads();
    const bool row = lane < NS;
    // Schur update of the row held in a[] with the top coupling F_kk (stored at Ft[kk-1]) of stage kk
    auto schur_top = [&](double (&a)[NS], int kk) {
      if (lane < NX) {
        const double* f = Ft + (size_t)(kk - 1) * NS * NX + lane;        // F_kk[lane][jj] at f[jj*NX]
        const double* ab = AB + (size_t)(kk - 1) * NX * NS;              // AB_{kk-1}
        double fr[NS];
#pragma unroll
        for (int jj = 0; jj < NS; ++jj) fr[jj] = f[jj * NX];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          double acc = 0.0;
#pragma unroll
          for (int jj = 0; jj < NS; ++jj) acc = fma(fr[jj], ab[i2 * NS + jj], acc);
          a[i2] = fma(-(rho_eq * ec[kk * NX + i2]), acc, a[i2]);
        }
      }
    };
    // Schur update with the bottom coupling Fb_kk (stored at Fbt[kk-p]) of stage kk
    auto schur_bot = [&](double (&a)[NS], int kk) {
      if (row) {
        const double* fb = Fbt + (size_t)(kk - p) * NS * NS + lane;      // Fb_kk[lane][i2] at fb[i2*NS]
        const double* ab = AB + (size_t)kk * NX * NS;                    // AB_kk
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          const double g = fb[i2 * NS] * (rho_eq * ec[(kk + 1) * NX + i2]);
#pragma unroll
          for (int c = 0; c < NS; ++c) a[c] = fma(-g, ab[i2 * NS + c], a[c]);
        }
      }
    };
    auto load_row = [&](double (&a)[NS], int kk) {
#pragma unroll
      for (int c = 0; c < NS; ++c) a[c] = row ? Sinv[(size_t)kk * NS * NS + lane * NS + c] : 0.0;
    };
    auto store_row = [&](const double (&a)[NS], int kk) {
      if (row)
#pragma unroll
        for (int c = 0; c < NS; ++c) Sinv[(size_t)kk * NS * NS + lane * NS + c] = a[c];
    };
    if (warp == 0) {  // top chain
      for (int kk = 0; kk < p; ++kk) {
        double a[NS];
        load_row(a, kk);
        if (kk > 0) schur_top(a, kk);
        gj_inverse<NS>(a, lane);
        store_row(a, kk);
        // F_{kk+1}[i2][lane] = rho_eq ec[(kk+1),i2] sum_c AB_kk[i2][c] Sinv_kk[c][lane]   (a[] = column `lane` by symmetry)
        if (row) {
          const double* ab = AB + (size_t)kk * NX * NS;
          double* f = Ft + (size_t)kk * NS * NX + lane * NX;
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < NS; ++c) acc = fma(ab[i2 * NS + c], a[c], acc);
            f[i2] = (rho_eq * ec[(kk + 1) * NX + i2]) * acc;
          }
        }
        __syncwarp();
      }
    }
    if (warp == 1) {  // bottom chain
      for (int kk = N; kk > p; --kk) {
        double a[NS];
        load_row(a, kk);
        if (kk < N) schur_bot(a, kk);
        gj_inverse<NS>(a, lane);
        store_row(a, kk);
        // Fb_{kk-1}[jj][lane] = sum_{i2<NX} rho_eq ec[kk,i2] AB_{kk-1}[i2][jj] Sinv_kk[i2][lane]
        if (row) {
          const double* ab = AB + (size_t)(kk - 1) * NX * NS;
          double* f = Fbt + (size_t)(kk - 1 - p) * NS * NS + lane * NS;
          double g[NX];
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) g[i2] = (rho_eq * ec[kk * NX + i2]) * a[i2];
#pragma unroll
          for (int jj = 0; jj < NS; ++jj) {
            double acc = 0.0;
#pragma unroll
            for (int i2 = 0; i2 < NX; ++i2) acc = fma(ab[i2 * NS + jj], g[i2], acc);
            f[jj] = acc;
          }
        }
        __syncwarp();
      }
    }
    __syncthreads();
    if (warp == 0) {
      double a[NS];
      load_row(a, p);
      if (p > 0) schur_top(a, p);
      if (p < N) schur_bot(a, p);
      gj_inverse<NS>(a, lane);
      store_row(a, p);
    }
    __syncthreads();
  };
  bool need_factor = true;   // ONE inlined copy of the factorisation (it is 125 KB of code): run at the top of the loop

  // own - sum_c f[c] v[c] with a short dependency chain (4 partial sums, `own` folded into the first)
  auto chain_dot = [&](const double* f, const double* v, int len, double own) -> double {
    double a0 = own, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
    for (int c = 0; c < NS; c += 4) {
      if (c < len) a0 = fma(-f[c], v[c], a0);
      if (c + 1 < len) a1 = fma(f[c + 1], v[c + 1], a1);
      if (c + 2 < len) a2 = fma(f[c + 2], v[c + 2], a2);
      if (c + 3 < len) a3 = fma(f[c + 3], v[c + 3], a3);
    }
    return (a0 - a1) - (a2 + a3);
  };

  // ---- one KKT solve: xt <- M^{-1} r with r already in yt -------------------------------------------------------------
  auto kkt_solve = [&]() {
    // forward chains
    if (warp == 0) {
      if (lane < NX && p >= 1) {
        double fr[NS];
#pragma unroll
        for (int c = 0; c < NS; ++c) fr[c] = Ft[c * NX + lane];
        double own = yt[NS + lane];
        for (int kk = 1; kk <= p; ++kk) {
          const double* yp = yt + (size_t)(kk - 1) * NS;
          double yv[NS];
#pragma unroll
          for (int c = 0; c < NS; ++c) yv[c] = yp[c];
          yt[kk * NS + lane] = chain_dot(fr, yv, NS, own);
          if (kk < p) {
            const double* f = Ft + (size_t)kk * NS * NX + lane;
#pragma unroll
            for (int c = 0; c < NS; ++c) fr[c] = f[c * NX];
            own = yt[(kk + 1) * NS + lane];
          }
          __syncwarp((1u << NX) - 1u);
        }
      }
    }
    if (warp == 1) {
      if (lane < NS && N - 1 >= p) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        double fr[NS];
#pragma unroll
        for (int c = 0; c < NS; ++c) fr[c] = Fbt[(size_t)(N - 1 - p) * NS * NS + c * NS + lane];
        double own = (N - 1 > p) ? yt[(N - 1) * NS + lane] : 0.0;
        for (int kk = N - 1; kk >= p; --kk) {
          const double* yn = yt + (size_t)(kk + 1) * NS;
          double yv[NS];
#pragma unroll
          for (int c = 0; c < NS; ++c) yv[c] = yn[c];
          const double r = chain_dot(fr, yv, NS, own);
          if (kk > p) yt[kk * NS + lane] = r;
          else mrg[lane] = -r;            // merged into the middle stage by the parallel phase below
          if (kk > p) {
            const double* f = Fbt + (size_t)(kk - 1 - p) * NS * NS + lane;
#pragma unroll
            for (int c = 0; c < NS; ++c) fr[c] = f[c * NS];
            own = (kk - 1 > p) ? yt[(kk - 1) * NS + lane] : 0.0;
          }
          __syncwarp(mask);
        }
      }
    }
    __syncthreads();
    // d_k = S_k^{-1} yt_k for every stage
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      if (on[s]) {
        const int k = j[s] / NS, i = j[s] - k * NS;
        const double* si = Sinv + (size_t)k * NS * NS + i;
        const double* yk = yt + (size_t)k * NS;
        double sv[NS], yv[NS];
#pragma unroll
        for (int c = 0; c < NS; ++c) { sv[c] = si[c * NS]; yv[c] = yk[c]; }
        if (k == p && p < N) {
#pragma unroll
          for (int c = 0; c < NS; ++c) yv[c] -= mrg[c];
        }
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int c = 0; c + 1 < NS; c += 2) {
          s0 = fma(sv[c], yv[c], s0);
          s1 = fma(sv[c + 1], yv[c + 1], s1);
        }
        if (NS & 1) s0 = fma(sv[NS - 1], yv[NS - 1], s0);
        xt[j[s]] = s0 + s1;
      }
    }
    __syncthreads();
    // backward chains (outwards from the middle stage)
    if (warp == 0) {
      if (lane < NS && p >= 1) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        double fr[NS];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) fr[i2] = Ft[(size_t)(p - 1) * NS * NX + lane * NX + i2];
        double own = xt[(p - 1) * NS + lane];
        for (int kk = p - 1; kk >= 0; --kk) {
          const double* vn = xt + (size_t)(kk + 1) * NS;
          double vv[NS];
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) vv[i2] = vn[i2];
          xt[kk * NS + lane] = chain_dot(fr, vv, NX, own);
          if (kk > 0) {
            const double* f = Ft + (size_t)(kk - 1) * NS * NX + lane * NX;   // F_kk[i2][lane]
#pragma unroll
            for (int i2 = 0; i2 < NX; ++i2) fr[i2] = f[i2];
            own = xt[(kk - 1) * NS + lane];
          }
          __syncwarp(mask);
        }
      }
    }
    if (warp == 1) {
      if (lane < NS && p + 1 <= N) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        double fr[NS];
#pragma unroll
        for (int c = 0; c < NS; ++c) fr[c] = Fbt[lane * NS + c];          // Fb_p[jj][lane]
        double own = xt[(p + 1) * NS + lane];
        for (int kk = p + 1; kk <= N; ++kk) {
          const double* vp = xt + (size_t)(kk - 1) * NS;
          double vv[NS];
#pragma unroll
          for (int c = 0; c < NS; ++c) vv[c] = vp[c];
          xt[kk * NS + lane] = chain_dot(fr, vv, NS, own);
          if (kk < N) {
            const double* f = Fbt + (size_t)(kk - p) * NS * NS + lane * NS;  // Fb_kk[jj][lane]
#pragma unroll
            for (int c = 0; c < NS; ++c) fr[c] = f[c];
            own = xt[(kk + 1) * NS + lane];
          }
          __syncwarp(mask);
        }
      }
    }
    __syncthreads();
  };

  // ---- residuals / termination / infeasibility / rho estimate (oracle/mpc.py _residuals, _check, _rho_estimate) --------
  // Reads the exchange vectors written at a checking iteration: x in xt, dx in yt, y_eq in tmp, dy_eq in weq, dy_box in sdy.
  struct CheckOut { double pri, dua, rho_est; int status; };
  auto check = [&](bool approx) -> CheckOut {
    const double f = approx ? 10.0 : 1.0;
    const double eps_pinf = f * S.eps_prim_inf, eps_dinf = f * S.eps_dual_inf;
    double v[NRED];
#pragma unroll
    for (int e = 0; e < NRED; ++e) v[e] = 0.0;
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      if (live[s]) {
        const double Pb = sPb[j[s]], Dj = gD[j[s]], Eb = gE[neq + j[s]];
        const double dx = yt[j[s]], dyb = sdy[j[s]];
        const bool uinf = ubs[s] > INF_T, linf = lbs[s] < -INF_T;
        // box row: projection of dy on the polar of the recession cone
        double dyp = dyb;
        if (uinf && linf) dyp = 0.0;
        else if (uinf) dyp = fmin(dyp, 0.0);
        else if (linf) dyp = fmax(dyp, 0.0);
        v[0] = fmax(v[0], fabs(Dj * dx));
        v[1] = fmax(v[1], fabs(Pb * dx / Dj));
        v[NRED_MAX + 0] += q[s] * dx;
        {
          const double Ax = Bd[s] * x[s], Ei = 1.0 / Eb;
          v[2] = fmax(v[2], fabs(Ei * (Ax - zb[s]))); v[3] = fmax(v[3], fabs(Ei * Ax)); v[4] = fmax(v[4], fabs(Ei * zb[s]));
          v[5] = fmax(v[5], fabs(Ax - zb[s])); v[6] = fmax(v[6], fabs(Ax)); v[7] = fmax(v[7], fabs(zb[s]));
          v[8] = fmax(v[8], fabs(Eb * dyp));
          v[NRED_MAX + 1] += ubs[s] * fmax(dyp, 0.0) + lbs[s] * fmin(dyp, 0.0);
          const double Adx = (Bd[s] * dx) * Ei;
          if (!uinf) v[9] = fmax(v[9], Adx);
          if (!linf) v[10] = fmax(v[10], -Adx);
        }
        double dye = 0.0;
        if (eqo[s]) {
          const double Ee = gE[er[s]], bes = sbe[er[s]];
          dye = weq[er[s]];
          const double Ax = eq_row(s, xt, x[s]), Ei = 1.0 / Ee;
          v[2] = fmax(v[2], fabs(Ei * (Ax - ze[s]))); v[3] = fmax(v[3], fabs(Ei * Ax)); v[4] = fmax(v[4], fabs(Ei * ze[s]));
          v[5] = fmax(v[5], fabs(Ax - ze[s])); v[6] = fmax(v[6], fabs(Ax)); v[7] = fmax(v[7], fabs(ze[s]));
          v[8] = fmax(v[8], fabs(Ee * dye));
          v[NRED_MAX + 1] += bes * fmax(dye, 0.0) + bes * fmin(dye, 0.0);
          const double Adx = eq_row(s, yt, dx) * Ei;
          v[9] = fmax(v[9], Adx);
          v[10] = fmax(v[10], -Adx);
        }
        {
          const double Px = Pb * x[s], Aty = at_col(s, yb[s], ye[s], tmp), Di = 1.0 / Dj;
          v[11] = fmax(v[11], fabs(Di * (Px + q[s] + Aty))); v[12] = fmax(v[12], fabs(Di * Px));
          v[13] = fmax(v[13], fabs(Di * Aty)); v[14] = fmax(v[14], fabs(Di * q[s]));
          v[15] = fmax(v[15], fabs(Px + q[s] + Aty)); v[16] = fmax(v[16], fabs(Px));
          v[17] = fmax(v[17], fabs(Aty)); v[18] = fmax(v[18], fabs(q[s]));
          v[19] = fmax(v[19], fabs(at_col(s, dyp, dye, weq) * Di));
        }
      }
    }
    reduce(v, IC<NRED_MAX>{}, IC<NRED_SUM>{});
    CheckOut o;
    o.pri = v[2];
    o.dua = v[11] / cscale;
    const double eps_pri = f * S.eps_abs + f * S.eps_rel * fmax(v[3], v[4]);
    const double eps_dua = f * S.eps_abs + f * S.eps_rel * fmax(fmax(v[12], v[13]), v[14]) / cscale;
    const double pr = v[5] / (fmax(v[7], v[6]) + 1e-10);
    const double du = v[15] / (fmax(fmax(v[18], v[17]), v[16]) + 1e-10);
    o.rho_est = fmin(fmax(rho * sqrt(pr / (du + 1e-10)), RHO_MIN), RHO_MAX);
    o.status = GPMPC_QP_UNSOLVED;
    const double nrm_dx = v[0], nrm_dy = v[8], lhs = v[NRED_MAX + 1], qdx = v[NRED_MAX + 0];
    bool pinf = false, dinf = false;
    if (nrm_dy > eps_pinf && lhs < -eps_pinf * nrm_dy) pinf = v[19] < eps_pinf * nrm_dy;
    if (nrm_dx > eps_dinf && qdx < -cscale * eps_dinf * nrm_dx && v[1] < cscale * eps_dinf * nrm_dx)
      dinf = !(v[9] > eps_dinf * nrm_dx || v[10] > eps_dinf * nrm_dx);
    if (pinf) o.status = approx ? GPMPC_QP_PRIMAL_INFEASIBLE_INACCURATE : GPMPC_QP_PRIMAL_INFEASIBLE;
    else if (dinf) o.status = approx ? GPMPC_QP_DUAL_INFEASIBLE_INACCURATE : GPMPC_QP_DUAL_INFEASIBLE;
    else if (o.pri <= eps_pri && o.dua <= eps_dua) o.status = approx ? GPMPC_QP_SOLVED_INACCURATE : GPMPC_QP_SOLVED;
    return o;
  };

  // ---- ADMM iterations -----------------------------------------------------------------------------------------------------
  int status = GPMPC_QP_UNSOLVED, it = 0;
  CheckOut co;
  co.pri = co.dua = 0.0; co.rho_est = rho; co.status = GPMPC_QP_UNSOLVED;
  const double alpha = S.alpha;
  double rho_eq = RHO_EQ_OVER_INEQ * rho, rei = 1.0 / rho_eq, rhoi = 1.0 / rho;  // (divisions only when rho changes)
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    if (eqo[s]) {
      if (S.max_iter > 0) weq[er[s]] = fma(RHO_EQ_OVER_INEQ * rho, ze[s], -ye[s]);
      else tmp[er[s]] = ye[s];  // check buffers for a solve without iterations: dx = dy = 0, x in xt
    }
  }
  __syncthreads();
  for (it = 1; it <= S.max_iter; ++it) {
    if (need_factor) {
      factorize();
      need_factor = false;
    }
    const bool can_check = S.check_every > 0 && (it % S.check_every) == 0;
    const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval > 0 && (it % S.adaptive_rho_interval) == 0;
    const bool checking = can_check || can_adapt || it == S.max_iter;
    // right-hand side sigma x - q + A'(rho z - y)
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
      if (on[s])
        yt[j[s]] = fma(sigma, x[s], -q[s]) + at_col(s, fma(rb, zb[s], -yb[s]), eqo[s] ? weq[er[s]] : 0.0, weq);
    }
    __syncthreads();
    kkt_solve();
    // z_tilde = A x_tilde ; z, y, x updates
    double dx[NSLOT], dyb[NSLOT], dye[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
      const double rbi = (cls[s] == 0) ? (1.0 / RHO_MIN) : ((cls[s] == 1) ? rei : rhoi);
      const double xtj = on[s] ? xt[j[s]] : 0.0;
      {
        const double ztb = Bd[s] * xtj;
        const double zr = alpha * ztb + (1.0 - alpha) * zb[s];
        const double zn = fmin(fmax(fma(yb[s], rbi, zr), lbs[s]), ubs[s]);
        dyb[s] = rb * (zr - zn);
        yb[s] += dyb[s];
        zb[s] = zn;
      }
      dye[s] = 0.0;
      if (eqo[s]) {
        const double zte = eq_row(s, xt, xtj), bes = sbe[er[s]];
        const double zr = alpha * zte + (1.0 - alpha) * ze[s];
        const double zn = bes;  // projection of zr + y / rho on the degenerate interval [b, b]
        dye[s] = rho_eq * (zr - zn);
        ye[s] += dye[s];
        ze[s] = zn;
      }
      {
        const double xn = alpha * xtj + (1.0 - alpha) * x[s];
        dx[s] = xn - x[s];
        x[s] = xn;
      }
    }
    if (!checking) {  // (every read of weq of this iteration happened before the barriers inside kkt_solve)
#pragma unroll
      for (int s = 0; s < NSLOT; ++s)
        if (eqo[s]) weq[er[s]] = fma(rho_eq, ze[s], -ye[s]);
      __syncthreads();
      continue;
    }
    __syncthreads();  // every read of xt (eq_row above) is done
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      if (on[s]) { xt[j[s]] = x[s]; yt[j[s]] = dx[s]; sdy[j[s]] = dyb[s]; }
      if (eqo[s]) { tmp[er[s]] = ye[s]; weq[er[s]] = dye[s]; }
    }
    __syncthreads();
    if (can_check || can_adapt) {
      co = check(false);
      if (can_check && co.status != GPMPC_QP_UNSOLVED) {
        status = co.status;
        break;
      }
      if (can_adapt && (co.rho_est > rho * S.adaptive_rho_tolerance || co.rho_est < rho / S.adaptive_rho_tolerance)) {
        rho = co.rho_est;
        rho_eq = RHO_EQ_OVER_INEQ * rho; rei = 1.0 / rho_eq; rhoi = 1.0 / rho;
        need_factor = true;
      }
      if (it < S.max_iter) {  // (the block reduction inside check() separates its reads of weq from this write)
#pragma unroll
        for (int s = 0; s < NSLOT; ++s)
          if (eqo[s]) weq[er[s]] = fma(RHO_EQ_OVER_INEQ * rho, ze[s], -ye[s]);
        __syncthreads();
      }
    }
  }
  if (it > S.max_iter) it = S.max_iter;
  if (status == GPMPC_QP_UNSOLVED) {
    if (!(S.check_every > 0 && (it % S.check_every) == 0)) {
      co = check(false);
      status = co.status;
    }
    if (status == GPMPC_QP_UNSOLVED) {
      CheckOut ca = check(true);
      status = ca.status;
    }
    if (status == GPMPC_QP_UNSOLVED) status = GPMPC_QP_MAX_ITER;
  }
  // ---- objective and solution out (unscaled) ---------------------------------------------------------------------------
  double ob[NRED];
  ob[NRED_MAX] = 0.0;
#pragma unroll
  for (int s = 0; s < NSLOT; ++s)
    if (live[s]) ob[NRED_MAX] += (0.5 * sPb[j[s]] * x[s] + q[s]) * x[s];
  reduce(ob, IC<0>{}, IC<1>{});
  const bool commit_y = !S.commit_y_solved_only || status == GPMPC_QP_SOLVED;
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    if (live[s]) {
      xg[(size_t)b * n + j[s]] = gD[j[s]] * x[s];
      if (commit_y) yg[(size_t)b * m + neq + j[s]] = gE[neq + j[s]] * yb[s] / cscale;
    }
    if (eqo[s] && commit_y) yg[(size_t)b * m + er[s]] = gE[er[s]] * ye[s] / cscale;
  }
  if (tid == 0) {
    status_g[b] = status;
    iters_g[b] = it;
    obj_g[b] = ob[NRED_MAX] / cscale;
    pri_g[b] = co.pri;
    dua_g[b] = co.dua;
    rhog[b] = rho;
  }
}

}  // namespace

// Launch shapes: 3-DoF (nx 7): 128 threads x 2 slots, 3 QPs per SM, horizon 20 compiled in;  6-DoF (nx 14, N = 30 needs
// 220 KB of shared memory): 256 threads x 3 slots, one QP per SM, horizon 30 compiled in.
template <int NX> struct StageShape { static constexpr int NSLOT = 2, NTHR = 128, MINB = ADMM_STAGE_MIN_CTAS, NFIX = 20; };
// 6-DoF shape.  Measured on B200 (tools/prof_admm6.py, 1024 QPs): 3 slots x 256 threads 55.2 us per iteration of the batch,
// 2 slots x 384 threads (168 registers) 56.6 us -- the iteration is bound by the two chain warps, not by the thread count.
// (More than 8 warps also needs MAX_WARPS raised.)
#ifndef ADMM_STAGE14_NSLOT
#define ADMM_STAGE14_NSLOT 3
#define ADMM_STAGE14_NTHR 256
#endif
static_assert(ADMM_STAGE14_NTHR / 32 <= MAX_WARPS, "raise MAX_WARPS for this block size");
template <> struct StageShape<14> { static constexpr int NSLOT = ADMM_STAGE14_NSLOT, NTHR = ADMM_STAGE14_NTHR, MINB = 1, NFIX = 30; };

// Launcher used by gpmpc_admm_solve (csrc/admm.cu).  Returns GPMPC_ERR_ARG without launching when the problem does not
// fit this kernel (the caller then falls back to the first-generation kernel).  `scratch`: B x (n + m) doubles.
template <int NX, int NU>
int launch_admm_stage(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag,
                      const double* lb, const double* ub, const GpmpcAdmmSettings& S, const int* active,
                      double* scratch, double* x, double* y, double* rho, int* status, int* iters, double* obj,
                      double* pri, double* dua, cudaStream_t stream) {
  using L = Stage<NX, NU>;
  using H = StageShape<NX>;
  const size_t smem = L::smem_doubles(N) * sizeof(double);
  if (smem > 227 * 1024 || (N + 1) * L::NS > H::NSLOT * H::NTHR) return GPMPC_ERR_ARG;
  const bool fixed = (N == H::NFIX);
  auto kern = fixed ? admm_stage_kernel<NX, NU, H::NFIX, H::NSLOT, H::NTHR, H::MINB>
                    : admm_stage_kernel<NX, NU, 0, H::NSLOT, H::NTHR, H::MINB>;
  static GpmpcPerDevice configured[2];
  if ((int)smem > configured[fixed].here()) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured[fixed].here() = (int)smem;
  }
  kern<<<B, H::NTHR, smem, stream>>>(N, B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj,
                                     pri, dua);
  return gpmpc_check_launch("admm_stage");
}

template int launch_admm_stage<7, 3>(int, int, const double*, const double*, const double*, const double*,
                                     const double*, const double*, const GpmpcAdmmSettings&, const int*, double*,
                                     double*, double*, double*, int*, int*, double*, double*, double*, cudaStream_t);
template int launch_admm_stage<14, 3>(int, int, const double*, const double*, const double*, const double*,
                                      const double*, const double*, const GpmpcAdmmSettings&, const int*, double*,
                                      double*, double*, double*, int*, int*, double*, double*, double*, cudaStream_t);
