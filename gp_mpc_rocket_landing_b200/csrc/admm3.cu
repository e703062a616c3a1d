// Third generation of the batched OSQP-style ADMM kernel for the stage-structured MPC QP (same problem, same algorithm,
// same C entry point as csrc/admm.cu / csrc/admm_stage.cu -- reference call sites osqp_rti.py:464-478,496,517,524,527;
// algorithm restated from Stellato et al. 2020, see the header of csrc/admm.cu).  What is new over the second generation:
//
//  1. LIGHT ITERATIONS.  The equality rows of the QP project onto a point (z_e = b_e after the first iteration), so their
//     multipliers move by dy_e = rho_e alpha (A_e x~ - b_e) and enter the next right-hand side only through
//     g = A_e'(rho_e b_e - y_e).  Because x~ solves (P + sigma I + A' diag(rho) A) x~ = rhs,
//         rho_e A_e' A_e x~ = rhs - (P + sigma + rho_b B_d^2) .* x~          (element-wise!)
//     so g is advanced per variable without either sparse product:  g <- g - alpha (rhs - Dv x~) + alpha c0,
//     c0 = rho_e A_e' b_e.  Between termination checks an iteration is: element-wise right-hand side, one KKT solve,
//     element-wise updates.  The products A x~ and A' w (two of the three matrix phases of an iteration) are evaluated only
//     at "full" iterations -- the first one and every checking iteration -- where y_e is recovered exactly from the
//     running sum of x~ (y_e += rho_e alpha (A sum x~ - n b_e)) and g is re-synchronised from it, so the recurrence never
//     drifts for more than check_every - 1 iterations.  Iterates equal the second generation's up to rounding.
//  2. ONE LEVEL OF CYCLIC REDUCTION on the four substitution chains of the twisted block factorisation: with the
//     two-step operators G_k = F_k[:,x] F_{k-1} (top) and Gb_k = Fb_k Fb_{k+1} (bottom) formed at factorisation time, every
//     second stage is computed in a parallel phase and the dependent chain visits the others only: 4 dependent steps per
//     chain at N = 20 instead of 10 (7 / 6 instead of 15 at N = 30).
//  3. Bank-conflict-free factor storage (odd row pitch, so that row- and column-wise walks of a 10x10 / 17x17 block both hit
//     distinct banks), stage vectors on an even pitch read with 16-byte loads, and -- for the 3-DoF shape -- the rows of
//     S_k^{-1} held in registers between factorisations.
//  4. N = 30, n_x = 14: the scaled stage blocks [A_k|B_k] live in a global scratch (they are touched only by the
//     factorisation and the full iterations), which makes room for the two-step operators in shared memory.
#include <math.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

#ifdef ADMM3_TRACE
__device__ long long g_admm3_trace[4096];
__device__ int g_admm3_trace_n;
#define TR(id)                                                                    \
  do {                                                                            \
    if (tid == 0 && b == 7 && g_admm3_trace_n < 4090) {                           \
      g_admm3_trace[g_admm3_trace_n++] = (clock64() << 8) | (long long)(id);      \
    }                                                                             \
  } while (0)
extern "C" int gpmpc_admm3_trace_read(long long* out, int cap) {
  int n = 0;
  cudaMemcpyFromSymbol(&n, g_admm3_trace_n, sizeof(int));
  if (n > cap) n = cap;
  cudaMemcpyFromSymbol(out, g_admm3_trace, sizeof(long long) * n);
  int z = 0;
  cudaMemcpyToSymbol(g_admm3_trace_n, &z, sizeof(int));
  return n;
}
#else
#define TR(id)
#endif

namespace {

constexpr double OSQP_INFTY = 1e30, MIN_SCALING = 1e-4, MAX_SCALING = 1e4;
constexpr double RHO_MIN = 1e-6, RHO_MAX = 1e6, RHO_TOL = 1e-4, RHO_EQ_OVER_INEQ = 1e3;
constexpr double INF_T = OSQP_INFTY * MIN_SCALING;
constexpr int MAX_WARPS = 8;
constexpr int NRED_MAX = 20, NRED_SUM = 2, NRED = NRED_MAX + NRED_SUM;

__device__ __forceinline__ double limit_scaling(double v) {
  v = (v < MIN_SCALING) ? 1.0 : v;
  return (v > MAX_SCALING) ? MAX_SCALING : v;
}

template <int NX, int NU, int NT>
struct Lay {
  static constexpr int NS = NX + NU, N = NT, P = NT / 2;
  static constexpr int PS = (NS % 2 == 0) ? NS + 1 : NS;  // odd pitch of the factor blocks
  static constexpr int VS = (NS % 2 == 0) ? NS : NS + 1;  // even pitch of the stage vectors
  static constexpr int NPAD = (N + 1) * VS, EPAD = (N + 2) * NX;
  // top chain (stages 1..P): reduced stages have the parity of P
  static constexpr int KT0 = (P % 2 == 0) ? 2 : 1;          // first reduced stage
  static constexpr int KTG = (P % 2 == 0) ? 2 : 3;          // first reduced stage with a two-step operator (k - 2 >= 0)
  static constexpr int NGT = (P >= KTG) ? (P - KTG) / 2 + 1 : 0;
  // bottom chain (stages N-1..P): reduced stages have the parity of P
  static constexpr int KB0 = ((N - 1 - P) % 2 == 0) ? N - 1 : N - 2;  // first (largest) reduced stage
  static constexpr int KBG = (KB0 == N - 1) ? N - 3 : N - 2;          // largest reduced stage with a two-step operator
  static constexpr int NGB = (KBG >= P) ? (KBG - P) / 2 + 1 : 0;
  static __host__ __device__ constexpr bool red_top(int k) { return k >= 1 && k <= P && ((P - k) & 1) == 0; }
  static __host__ __device__ constexpr bool red_bot(int k) { return k >= P && k <= N - 1 && ((k - P) & 1) == 0; }
  template <bool ABS>
  static __host__ __device__ constexpr size_t smem_doubles() {
    size_t s = 0;
    if (ABS) s += ev((size_t)N * NX * NS);
    s += ev((size_t)(N + 1) * NS * PS);      // Sinv
    s += ev((size_t)P * NX * PS);            // Ft
    s += ev((size_t)(N - P) * NS * PS);      // Fbt
    s += ev((size_t)(NGT > 0 ? NGT : 1) * NX * PS);  // Gt
    s += ev((size_t)(NGB > 0 ? NGB : 1) * NS * PS);  // Gb
    s += 4 * ev((size_t)NPAD);               // yt xt sPb sdy
    s += 5 * ev((size_t)EPAD);               // weq sye ec sbe sze
    s += ev((size_t)VS);                     // cb
    s += ev((size_t)MAX_WARPS * NRED);       // red
    return s;
  }
  static __host__ __device__ constexpr size_t ev(size_t c) { return (c + 1) & ~(size_t)1; }
};

// In-place inverse of a symmetric positive definite NS x NS matrix held one row per lane (lanes >= NS hold zeros):
// Gauss-Jordan without pivoting (the pivots are the positive Schur complements of an SPD matrix).
template <int NS>
__device__ __forceinline__ void gj_inverse3(double (&a)[NS], int lane) {
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

template <int NM, int NSUM>
__device__ __forceinline__ void block_reduce_mixed3(double (&v)[NRED], double* red, int warp, int lane, int nwarps) {
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
struct IC3 { static constexpr int value = I; };

// LEN doubles of a shared-memory vector.  8-byte loads on purpose: measured on B200 (tools/ubench/lds.cu) one warp issues an
// LDS.64 every ~6 cycles but an LDS.128 only every ~16, so two 8-byte loads beat one 16-byte load.
template <int LEN>
__device__ __forceinline__ void load_vec(const double* v, double (&o)[LEN]) {
#pragma unroll
  for (int c = 0; c < LEN; ++c) o[c] = v[c];
}

// own + sgn * sum_c m[c * stride] v[c], four partial sums (dependent FP64 depth LEN/4 + 2)
template <int LEN>
__device__ __forceinline__ double dot4(const double* m, int stride, const double (&v)[LEN], double own,
                                       double sgn) {
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
  for (int c = 0; c < LEN; c += 4) {
    a0 = fma(m[c * stride], v[c], a0);
    if (c + 1 < LEN) a1 = fma(m[(c + 1) * stride], v[c + 1], a1);
    if (c + 2 < LEN) a2 = fma(m[(c + 2) * stride], v[c + 2], a2);
    if (c + 3 < LEN) a3 = fma(m[(c + 3) * stride], v[c + 3], a3);
  }
  return fma(sgn, (a0 + a1) + (a2 + a3), own);
}

// NSLOT (stage, variable) slots per thread: slot s of thread t is logical index t + NTHR s in the (N+1) x NS layout.
// SREG: rows of S_k^{-1} in registers between factorisations.  ABS: scaled stage blocks in shared memory (else global scratch).
template <int NX, int NU, int NT, int NSLOT, int NTHR, int MINB, bool SREG, bool ABS>
__global__ void __launch_bounds__(NTHR, MINB)
admm3_kernel(int nbatch, const double* __restrict__ ABg, const double* __restrict__ beqg, const double* __restrict__ qg,
             const double* __restrict__ Pdiag, const double* __restrict__ lb, const double* __restrict__ ub,
             GpmpcAdmmSettings S, const int* __restrict__ active, double* __restrict__ scratch,
             double* __restrict__ xg, double* __restrict__ yg, double* __restrict__ rhog, int* __restrict__ status_g,
             int* __restrict__ iters_g, double* __restrict__ obj_g, double* __restrict__ pri_g,
             double* __restrict__ dua_g) {
  using L = Lay<NX, NU, NT>;
  constexpr int NS = L::NS, N = L::N, P = L::P, PS = L::PS, VS = L::VS, NWARPS = NTHR / 32;
  constexpr int n = N * NS + NX, neq = NX * (N + 1), m = neq + n;
  constexpr int NPAD = L::NPAD, EPAD = L::EPAD, NLOG = (N + 1) * NS;
  static_assert(NLOG <= NSLOT * NTHR, "not enough (stage, variable) slots");
  static_assert(NWARPS >= 2 && NWARPS <= MAX_WARPS, "two chain warps are needed");
  extern __shared__ __align__(16) double smem3[];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  const int warp = tid >> 5, lane = tid & 31;
  TR(30);

  // ---- shared memory carve (must match Lay::smem_doubles) -------------------------------------------------------------
  double* sp = smem3;
  auto take = [&](size_t cnt) { double* q_ = sp; sp += (cnt + 1) & ~(size_t)1; return q_; };
  double* ABs = ABS ? take((size_t)N * NX * NS) : nullptr;
  double* Sinv = take((size_t)(N + 1) * NS * PS);      // Sinv[k][i][c]                           (row pitch PS)
  double* Ft = take((size_t)P * NX * PS);              // Ft[k-1][i][c]  = F_k[i][c],  k = 1..P    (NX rows)
  double* Fbt = take((size_t)(N - P) * NS * PS);       // Fbt[k-P][j][c] = Fb_k[j][c], k = P..N-1  (NS rows)
  double* Gt = take((size_t)(L::NGT > 0 ? L::NGT : 1) * NX * PS);   // Gt[(k-KTG)/2] = G_k  = F_k[:,x] F_{k-1}
  double* Gb = take((size_t)(L::NGB > 0 ? L::NGB : 1) * NS * PS);   // Gb[(k-P)/2]   = Gb_k = Fb_k Fb_{k+1}
  double* yt = take(NPAD);
  double* xt = take(NPAD);
  double* sPb = take(NPAD);   // scaled cost diagonal
  double* sdy = take(NPAD);   // dy of the box rows at termination checks
  double* weq = take(EPAD);   // equality-row exchange vector, indexed by owner (k,i): k*NX+i ; stage N+1 stays zero
  double* sye = take(EPAD);   // y of the equality rows as of the last full iteration
  double* ec = take(EPAD);    // own-variable coefficient of every equality row (scaled Ei / Nd)
  double* sbe = take(EPAD);   // scaled equality right-hand sides
  double* sze = take(EPAD);   // z of the equality rows (A x of the warm start until the first iteration, then b_e)
  double* cb = take((size_t)VS);   // bottom chain's contribution to the middle stage
  double* red = take((size_t)MAX_WARPS * NRED);
  // global scratch: cumulative Ruiz scalings D (n) then E (m) [+ the scaled stage blocks when they are not in shared memory]
  constexpr size_t SCR = (size_t)(n + m) + (ABS ? 0 : (size_t)N * NX * NS);
  double* gD = scratch + (size_t)b * SCR;
  double* gE = gD + n;
  double* AB = ABS ? ABs : (gE + m);

  // ---- thread <-> (stage, variable) slots ------------------------------------------------------------------------
  int kk_[NSLOT], ii_[NSLOT];
  bool on[NSLOT], live[NSLOT], eqo[NSLOT];
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    const int idx = tid + s * NTHR;
    const int k = idx / NS, i = idx - k * NS;
    on[s] = idx < NLOG;                        // has a slot in the padded (N+1) x NS layout
    kk_[s] = on[s] ? k : 0;
    ii_[s] = on[s] ? i : 0;
    live[s] = on[s] && (k < N || i < NX);      // a real variable
    eqo[s] = live[s] && i < NX;                // owns an equality row
  }
  auto JV = [&](int s) { return kk_[s] * VS + ii_[s]; };      // index into the per-variable shared vectors
  auto JG = [&](int s) { return kk_[s] * NS + ii_[s]; };      // index into the global x and the global box rows
  auto ER = [&](int s) { return kk_[s] * NX + ii_[s]; };      // index into the equality-row arrays / global rows
  auto hasAB = [&](int s) { return on[s] && kk_[s] < N; };    // stage has a dynamics block (columns of AB_k)
  auto hasPrev = [&](int s) { return eqo[s] && kk_[s] >= 1; };  // own equality row is a dynamics row (rows of AB_{k-1})

  auto reduce = [&](double (&v)[NRED], auto nmax_c, auto nsum_c) {
    block_reduce_mixed3<decltype(nmax_c)::value, decltype(nsum_c)::value>(v, red, warp, lane, NWARPS);
  };

  // ---- load ------------------------------------------------------------------------------------------------------
  for (int e = tid; e < N * NX * NS; e += NTHR) AB[e] = ABg[(size_t)b * N * NX * NS + e];
  for (int e = tid; e < EPAD; e += NTHR) { weq[e] = 0.0; sye[e] = 0.0; ec[e] = 0.0; sbe[e] = 0.0; sze[e] = 0.0; }
  for (int e = tid; e < NPAD; e += NTHR) { yt[e] = 0.0; xt[e] = 0.0; sPb[e] = 0.0; sdy[e] = 0.0; }
  if (tid < VS) cb[tid] = 0.0;
  const double sigma = S.sigma;
  double rho = (S.rho_from_input && rhog[b] > 0.0) ? rhog[b] : S.rho, cscale = 1.0;
  double x[NSLOT], zb[NSLOT], yb[NSLOT], q[NSLOT], Bd[NSLOT], lbs[NSLOT], ubs[NSLOT];
  int cls[NSLOT];  // box row class: 0 loose (rho_min), 1 equality-like (rho_eq), 2 inequality (rho)
  __syncthreads();
  TR(31);
  {
    // ---- modified Ruiz equilibration + cost scaling (OSQP scaling.c; oracle/mpc.py OSQPOracle._scale) ---------------
    double Pb[NSLOT], e_[NSLOT], Dj[NSLOT], Eb[NSLOT], Ee[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      x[s] = zb[s] = yb[s] = 0.0;
      q[s] = Pb[s] = Bd[s] = e_[s] = 0.0;
      Dj[s] = Eb[s] = Ee[s] = 1.0;
      cls[s] = 2;
      if (live[s]) {
        Pb[s] = Pdiag[JG(s)];
        q[s] = qg[(size_t)b * n + JG(s)];
        Bd[s] = 1.0;
        if (eqo[s]) e_[s] = (ER(s) < NX) ? 1.0 : -1.0;
      }
    }
    TR(32);
    for (int it = 0; it < S.scaling; ++it) {
      TR(33);
      double Dt[NSLOT], Etb[NSLOT], Ete[NSLOT];
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        double v = fmax(fabs(Pb[s]), fabs(Bd[s]));
        v = fmax(v, fabs(e_[s]));
        if (hasAB(s)) {
          const double* a = AB + (size_t)kk_[s] * NX * NS + ii_[s];
#pragma unroll
          for (int r = 0; r < NX; ++r) v = fmax(v, fabs(a[r * NS]));
        }
        Dt[s] = 1.0 / sqrt(limit_scaling(v));
        Etb[s] = 1.0 / sqrt(limit_scaling(fabs(Bd[s])));
        double ve = fabs(e_[s]);
        if (hasPrev(s)) {
          const double* a = AB + ((size_t)(kk_[s] - 1) * NX + ii_[s]) * NS;
#pragma unroll
          for (int c = 0; c < NS; ++c) ve = fmax(ve, fabs(a[c]));
        }
        Ete[s] = 1.0 / sqrt(limit_scaling(ve));
        if (eqo[s]) weq[ER(s)] = Ete[s];
      }
      __syncthreads();
      double rv[NRED];
      rv[0] = 0.0;
      rv[NRED_MAX] = 0.0;
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        if (hasAB(s)) {
          double* a = AB + (size_t)kk_[s] * NX * NS + ii_[s];
          const double* wn = weq + (kk_[s] + 1) * NX;
#pragma unroll
          for (int r = 0; r < NX; ++r) a[r * NS] = (wn[r] * a[r * NS]) * Dt[s];
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
      reduce(rv, IC3<1>{}, IC3<1>{});
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
    TR(34);
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      lbs[s] = ubs[s] = 0.0;
      if (live[s]) {
        lbs[s] = Eb[s] * fmax(lb[(size_t)b * S.bound_stride + JG(s)], -OSQP_INFTY);  // bounds clipped as OSQP's setup does (osqp.h OSQP_INFTY)
        ubs[s] = Eb[s] * fmin(ub[(size_t)b * S.bound_stride + JG(s)], OSQP_INFTY);
        if (lbs[s] < -INF_T && ubs[s] > INF_T) cls[s] = 0;
        else if (ubs[s] - lbs[s] < RHO_TOL) cls[s] = 1;
        if (S.warm_x) x[s] = xg[(size_t)b * n + JG(s)] / Dj[s];
        if (S.warm_y) yb[s] = cscale * yg[(size_t)b * m + neq + JG(s)] / Eb[s];
        gD[JG(s)] = Dj[s];
        gE[neq + JG(s)] = Eb[s];
      }
      if (on[s]) {
        sPb[JV(s)] = Pb[s];
        xt[JV(s)] = x[s];
      }
      if (eqo[s]) {
        ec[ER(s)] = e_[s];
        gE[ER(s)] = Ee[s];
        sbe[ER(s)] = Ee[s] * beqg[(size_t)b * neq + ER(s)];
        sye[ER(s)] = S.warm_y ? cscale * yg[(size_t)b * m + ER(s)] / Ee[s] : 0.0;
        weq[ER(s)] = 0.0;
      }
    }
  }
  __syncthreads();

  // (A v)_own-equality-row given v in shared memory (stage-vector layout) and the own value vj  [only for eqo slots]
  auto eq_row = [&](int s, const double* v, double vj) -> double {
    double s0 = 0.0, s1 = 0.0;
    if (hasPrev(s)) {
      const double* a = AB + ((size_t)(kk_[s] - 1) * NX + ii_[s]) * NS;
      double vp[NS];
      load_vec<NS>(v + (kk_[s] - 1) * VS, vp);
#pragma unroll
      for (int c = 0; c + 1 < NS; c += 2) {
        s0 = fma(a[c], vp[c], s0);
        s1 = fma(a[c + 1], vp[c + 1], s1);
      }
      if (NS & 1) s0 = fma(a[NS - 1], vp[NS - 1], s0);
    }
    return fma(ec[ER(s)], vj, s0 + s1);
  };
  // (A' w)_own-column given the own box value wb, the own equality value we and the shared equality vector
  auto at_col = [&](int s, double wb, double we, const double* wv) -> double {
    double r0 = Bd[s] * wb;
    if (eqo[s]) r0 = fma(ec[ER(s)], we, r0);
    if (hasAB(s)) {
      const double* a = AB + (size_t)kk_[s] * NX * NS + ii_[s];
      const double* wn = wv + (kk_[s] + 1) * NX;
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
      if (eqo[s]) sze[ER(s)] = eq_row(s, xt, x[s]);
    }
  }
  __syncthreads();

  double srow[SREG ? NSLOT : 1][SREG ? NS : 1];   // rows of S_k^{-1} (SREG)

  // ---- twisted block factorisation + two-step operators ------------------------------------------------------------
  auto factorize = [&]() {
    const double rho_eq = RHO_EQ_OVER_INEQ * rho;
    // phase 1: row i of M_k into Sinv[k][i][:]
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      if (on[s]) {
        const int k = kk_[s], i = ii_[s];
        double col[NX];
#pragma unroll
        for (int r = 0; r < NX; ++r) col[r] = hasAB(s) ? AB[((size_t)k * NX + r) * NS + i] : 0.0;
        double* out = Sinv + ((size_t)k * NS + i) * PS;
        const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
        const double e_ = eqo[s] ? ec[ER(s)] : 0.0;
        double d = sPb[JV(s)] + sigma;
        d = fma(rb, Bd[s] * Bd[s], d);
        d = fma(rho_eq, e_ * e_, d);
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          double acc = 0.0;
          if (hasAB(s)) {
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
    TR(22);
    const bool row = lane < NS;
    // Schur update of the row held in a[] with the top coupling F_kk of stage kk
    auto schur_top = [&](double (&a)[NS], int kk) {
      if (lane < NX) {
        const double* f = Ft + ((size_t)(kk - 1) * NX + lane) * PS;      // F_kk[lane][:]
        const double* ab = AB + (size_t)(kk - 1) * NX * NS;              // AB_{kk-1}
        double fr[NS];
#pragma unroll
        for (int jj = 0; jj < NS; ++jj) fr[jj] = f[jj];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          double acc = 0.0;
#pragma unroll
          for (int jj = 0; jj < NS; ++jj) acc = fma(fr[jj], ab[i2 * NS + jj], acc);
          a[i2] = fma(-(rho_eq * ec[kk * NX + i2]), acc, a[i2]);
        }
      }
    };
    // Schur update with the bottom coupling Fb_kk of stage kk
    auto schur_bot = [&](double (&a)[NS], int kk) {
      if (row) {
        const double* fb = Fbt + ((size_t)(kk - P) * NS + lane) * PS;    // Fb_kk[lane][:]
        const double* ab = AB + (size_t)kk * NX * NS;                    // AB_kk
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          const double g = fb[i2] * (rho_eq * ec[(kk + 1) * NX + i2]);
#pragma unroll
          for (int c = 0; c < NS; ++c) a[c] = fma(-g, ab[i2 * NS + c], a[c]);
        }
      }
    };
    auto load_row = [&](double (&a)[NS], int kk) {
#pragma unroll
      for (int c = 0; c < NS; ++c) a[c] = row ? Sinv[((size_t)kk * NS + lane) * PS + c] : 0.0;
    };
    auto store_row = [&](const double (&a)[NS], int kk) {
      if (row)
#pragma unroll
        for (int c = 0; c < NS; ++c) Sinv[((size_t)kk * NS + lane) * PS + c] = a[c];
    };
    if (warp == 0) {  // top chain
      for (int kk = 0; kk < P; ++kk) {
        double a[NS];
        load_row(a, kk);
        if (kk > 0) schur_top(a, kk);
        gj_inverse3<NS>(a, lane);
        store_row(a, kk);
        // F_{kk+1}[i2][lane] = rho_eq ec[(kk+1),i2] sum_c AB_kk[i2][c] Sinv_kk[c][lane]   (a[] = column `lane` by symmetry)
        if (row) {
          const double* ab = AB + (size_t)kk * NX * NS;
          double* f = Ft + (size_t)kk * NX * PS + lane;
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < NS; ++c) acc = fma(ab[i2 * NS + c], a[c], acc);
            f[i2 * PS] = (rho_eq * ec[(kk + 1) * NX + i2]) * acc;
          }
        }
        __syncwarp();
      }
    }
    if (warp == 1) {  // bottom chain
      for (int kk = N; kk > P; --kk) {
        double a[NS];
        load_row(a, kk);
        if (kk < N) schur_bot(a, kk);
        gj_inverse3<NS>(a, lane);
        store_row(a, kk);
        // Fb_{kk-1}[jj][lane] = sum_{i2<NX} rho_eq ec[kk,i2] AB_{kk-1}[i2][jj] Sinv_kk[i2][lane]
        if (row) {
          const double* ab = AB + (size_t)(kk - 1) * NX * NS;
          double* f = Fbt + (size_t)(kk - 1 - P) * NS * PS + lane;
          double g[NX];
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) g[i2] = (rho_eq * ec[kk * NX + i2]) * a[i2];
#pragma unroll
          for (int jj = 0; jj < NS; ++jj) {
            double acc = 0.0;
#pragma unroll
            for (int i2 = 0; i2 < NX; ++i2) acc = fma(ab[i2 * NS + jj], g[i2], acc);
            f[jj * PS] = acc;
          }
        }
        __syncwarp();
      }
    }
    __syncthreads();
    TR(23);
    if (warp == 0) {
      double a[NS];
      load_row(a, P);
      if (P > 0) schur_top(a, P);
      if (P < N) schur_bot(a, P);
      gj_inverse3<NS>(a, lane);
      store_row(a, P);
    }
    TR(24);
    // two-step operators (independent of the middle stage): G_k = F_k[:, :NX] F_{k-1}, Gb_k = Fb_k Fb_{k+1}
    for (int e = tid; e < L::NGT * NX * NS; e += NTHR) {
      const int g = e / (NX * NS), r = e - g * (NX * NS), i = r / NS, c = r - i * NS;
      const int k = L::KTG + 2 * g;
      const double* fk = Ft + ((size_t)(k - 1) * NX + i) * PS;    // F_k[i][:]
      const double* fp = Ft + (size_t)(k - 2) * NX * PS + c;      // F_{k-1}[:][c]
      double acc = 0.0;
#pragma unroll
      for (int i2 = 0; i2 < NX; ++i2) acc = fma(fk[i2], fp[i2 * PS], acc);
      Gt[((size_t)g * NX + i) * PS + c] = acc;
    }
    for (int e = tid; e < L::NGB * NS * NS; e += NTHR) {
      const int g = e / (NS * NS), r = e - g * (NS * NS), j = r / NS, c = r - j * NS;
      const int k = P + 2 * g;
      const double* fk = Fbt + ((size_t)(k - P) * NS + j) * PS;       // Fb_k[j][:]
      const double* fn = Fbt + (size_t)(k + 1 - P) * NS * PS + c;     // Fb_{k+1}[:][c]
      double acc = 0.0;
#pragma unroll
      for (int c2 = 0; c2 < NS; ++c2) acc = fma(fk[c2], fn[c2 * PS], acc);
      Gb[((size_t)g * NS + j) * PS + c] = acc;
    }
    __syncthreads();
    if (SREG) {
#pragma unroll
      for (int s = 0; s < NSLOT; ++s)
#pragma unroll
        for (int c = 0; c < NS; ++c) srow[SREG ? s : 0][SREG ? c : 0] = on[s] ? Sinv[((size_t)kk_[s] * NS + ii_[s]) * PS + c] : 0.0;
    }
  };
  TR(20);
  factorize();
  TR(21);

  // ---- one KKT solve: xs[s] <- (M^{-1} r)_slot with r already in yt; x~ is also left in xt for every stage --------------
  // Phases (a block barrier after each): P1 reduced right-hand sides | C1 reduced forward chains | P2 skipped stages |
  // P3 d = S^{-1} yt | P4 reduced backward right-hand sides | C2 reduced backward chains | P5 skipped stages | read back.
  // P1 / P2 / P4 / P5 work on a subset of the (stage, row) pairs: the rows are enumerated densely over the threads (row r of
  // the phase -> thread r mod NTHR) so that the warps that work are full, instead of every owner thread branching on its stage.
  constexpr int NRT = (P - L::KT0) / 2 + 1;      // reduced top stages    KT0, KT0+2, .., P
  constexpr int NRB = (L::KB0 - P) / 2 + 1;      // reduced bottom stages P, P+2, .., KB0
  constexpr int NST = P - NRT;                   // skipped top stages    (1..P-1, other parity)
  constexpr int NSB = (N - 1 - P) - (NRB - 1);   // skipped bottom stages (P+1..N-1, other parity)
  constexpr int KST0 = (L::KT0 == 2) ? 1 : 2;    // first skipped top stage
  constexpr int NBT = (P - 2 >= 0) ? (P - 2) / 2 + 1 : 0;   // backward reduced top stages    P-2, P-4, .., (0 or 1)
  constexpr int NBB = (N - (P + 2)) / 2 + 1;                 // backward reduced bottom stages P+2, P+4, .., (N or N-1)
  constexpr int NKT = (P - 1) / 2 + 1;                       // backward skipped top stages    P-1, P-3, .., (1 or 0)
  constexpr int NKB = (N - (P + 1)) / 2 + 1;                 // backward skipped bottom stages P+1, P+3, .., (N-1 or N)
  auto kkt_solve = [&](double (&xs)[NSLOT]) {
    // P1: t_k = r_k - F_k r_{k-1} (+ G_k r_{k-2} for the first reduced stage when k - 2 = 0), top reduced stages, state rows;
    //     t_k = r_k - Fb_k r_{k+1} (+ Gb_k r_{k+2} for the first reduced stage when k + 2 = N), bottom reduced stages;
    //     at the middle stage P the bottom part goes to cb (own term zero).
    //     Reads touch skipped stages and the end stages 0 / N only, writes go to reduced stages: no hazard, no barrier between.
    TR(1);
    for (int r = tid; r < NRT * NX + NRB * NS; r += NTHR) {
      const bool top = r < NRT * NX;
      const int u = top ? r : r - NRT * NX;
      const int gidx = top ? u / NX : u / NS;
      const int i = top ? u - gidx * NX : u - gidx * NS;
      const int k = top ? L::KT0 + 2 * gidx : P + 2 * gidx;
      const double* mrow = top ? Ft + ((size_t)(k - 1) * NX + i) * PS : Fbt + ((size_t)(k - P) * NS + i) * PS;
      double v[NS];
      load_vec<NS>(yt + (top ? k - 1 : k + 1) * VS, v);
      double t = dot4<NS>(mrow, 1, v, (!top && k == P) ? 0.0 : yt[k * VS + i], -1.0);
      if (top ? (L::KT0 == 2 && k == 2) : (L::KB0 == N - 2 && k == N - 2)) {
        load_vec<NS>(top ? yt : yt + N * VS, v);
        t = dot4<NS>(top ? Gt + (size_t)i * PS : Gb + ((size_t)gidx * NS + i) * PS, 1, v, t, 1.0);
      }
      if (!top && k == P) cb[i] = t;
      else yt[k * VS + i] = t;
    }
    __syncthreads();
    TR(2);
    // C1: reduced forward chains
    if (warp == 0) {
      if (lane < NX) {
        constexpr unsigned mask = (1u << NX) - 1u;
        constexpr int K1 = (L::KT0 == 2) ? 4 : 3;   // first dependent stage (stage KT0 was finished in P1)
        for (int k = K1; k <= P; k += 2) {
          double v[NS];
          load_vec<NS>(yt + (k - 2) * VS, v);
          const double r = dot4<NS>(Gt + ((size_t)((k - L::KTG) / 2) * NX + lane) * PS, 1, v, yt[k * VS + lane], 1.0);
          yt[k * VS + lane] = r;
          __syncwarp(mask);
        }
      }
    }
    if (warp == 1) {
      if (lane < NS) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        constexpr int K1 = (L::KB0 == N - 2) ? N - 4 : N - 3;   // first dependent stage
        for (int k = K1; k >= P; k -= 2) {
          double v[NS];
          load_vec<NS>(yt + (k + 2) * VS, v);
          const double own = (k == P) ? cb[lane] : yt[k * VS + lane];
          const double r = dot4<NS>(Gb + ((size_t)((k - P) / 2) * NS + lane) * PS, 1, v, own, 1.0);
          if (k == P) cb[lane] = r;
          else yt[k * VS + lane] = r;
          __syncwarp(mask);
        }
      }
    }
    __syncthreads();
    TR(3);
    // P2: skipped stages  yt_k = r_k - F_k yt_{k-1} (top, state rows) ; yt_k = r_k - Fb_k yt_{k+1} (bottom)
    for (int r = tid; r < NST * NX + NSB * NS; r += NTHR) {
      const bool top = r < NST * NX;
      const int u = top ? r : r - NST * NX;
      const int gidx = top ? u / NX : u / NS;
      const int i = top ? u - gidx * NX : u - gidx * NS;
      const int k = top ? KST0 + 2 * gidx : P + 1 + 2 * gidx;
      const double* mrow = top ? Ft + ((size_t)(k - 1) * NX + i) * PS : Fbt + ((size_t)(k - P) * NS + i) * PS;
      double v[NS];
      load_vec<NS>(yt + (top ? k - 1 : k + 1) * VS, v);
      yt[k * VS + i] = dot4<NS>(mrow, 1, v, yt[k * VS + i], -1.0);
    }
    __syncthreads();
    TR(4);
    // P3: d_k = S_k^{-1} yt_k for every stage (the middle stage adds the bottom chain's contribution)
    double d[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      d[s] = 0.0;
      if (on[s]) {
        const int k = kk_[s], i = ii_[s];
        double v[NS];
        load_vec<NS>(yt + k * VS, v);
        if (k == P && P < N) {
          double c2[NS];
          load_vec<NS>(cb, c2);
#pragma unroll
          for (int c = 0; c < NS; ++c) v[c] += c2[c];
        }
        if (SREG) {
          double s0 = 0.0, s1 = 0.0;
#pragma unroll
          for (int c = 0; c + 1 < NS; c += 2) {
            s0 = fma(srow[SREG ? s : 0][SREG ? c : 0], v[c], s0);
            s1 = fma(srow[SREG ? s : 0][SREG ? c + 1 : 0], v[c + 1], s1);
          }
          if (NS & 1) s0 = fma(srow[SREG ? s : 0][SREG ? NS - 1 : 0], v[NS - 1], s0);
          d[s] = s0 + s1;
        } else {
          d[s] = dot4<NS>(Sinv + ((size_t)k * NS + i) * PS, 1, v, 0.0, 1.0);
        }
        xt[k * VS + i] = d[s];
      }
    }
    __syncthreads();
    TR(5);
    // P4: backward reduced right-hand sides, in place:
    //   top    k in {P-2, P-4, ...}: t'_k = d_k - F_{k+1}' d_{k+1}[x]  (+ G_P' d_P[x] for k = P - 2: v_P = d_P is final)
    //   bottom k in {P+2, P+4, ...}: t'_k = d_k - Fb_{k-1}' d_{k-1}    (+ Gb_P' d_P for k = P + 2)
    //   (reads touch skipped stages and the middle stage only, writes go to reduced stages: no barrier between)
    for (int r = tid; r < (NBT + NBB) * NS; r += NTHR) {
      const bool top = r < NBT * NS;
      const int u = top ? r : r - NBT * NS;
      const int gidx = u / NS, i = u - gidx * NS;
      const int k = top ? P - 2 - 2 * gidx : P + 2 + 2 * gidx;
      double t = xt[k * VS + i];
      if (top) {
        double v[NX];
        load_vec<NX>(xt + (k + 1) * VS, v);
        t = dot4<NX>(Ft + (size_t)k * NX * PS + i, PS, v, t, -1.0);
        if (k == P - 2) {
          load_vec<NX>(xt + P * VS, v);
          t = dot4<NX>(Gt + (size_t)((P - L::KTG) / 2) * NX * PS + i, PS, v, t, 1.0);
        }
      } else {
        double v[NS];
        load_vec<NS>(xt + (k - 1) * VS, v);
        t = dot4<NS>(Fbt + (size_t)(k - 1 - P) * NS * PS + i, PS, v, t, -1.0);
        if (k == P + 2) {
          load_vec<NS>(xt + P * VS, v);
          t = dot4<NS>(Gb + i, PS, v, t, 1.0);
        }
      }
      xt[k * VS + i] = t;
    }
    __syncthreads();
    TR(6);
    // C2: reduced backward chains (outwards from the middle stage)
    if (warp == 0) {
      if (lane < NS) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        for (int k = P - 4; k >= 0; k -= 2) {   // v_k = t'_k + G_{k+2}' v_{k+2}[x]
          double v[NX];
          load_vec<NX>(xt + (k + 2) * VS, v);
          const double r = dot4<NX>(Gt + (size_t)((k + 2 - L::KTG) / 2) * NX * PS + lane, PS, v, xt[k * VS + lane], 1.0);
          xt[k * VS + lane] = r;
          __syncwarp(mask);
        }
      }
    }
    if (warp == 1) {
      if (lane < NS) {
        constexpr unsigned mask = (NS >= 32) ? 0xffffffffu : ((1u << NS) - 1u);
        for (int k = P + 4; k <= N; k += 2) {   // v_k = t'_k + Gb_{k-2}' v_{k-2}
          double v[NS];
          load_vec<NS>(xt + (k - 2) * VS, v);
          const double r = dot4<NS>(Gb + (size_t)((k - 2 - P) / 2) * NS * PS + lane, PS, v, xt[k * VS + lane], 1.0);
          xt[k * VS + lane] = r;
          __syncwarp(mask);
        }
      }
    }
    __syncthreads();
    TR(7);
    // P5: skipped stages  v_k = d_k - F_{k+1}' v_{k+1}[x] (top) ; v_k = d_k - Fb_{k-1}' v_{k-1} (bottom), in place
    for (int r = tid; r < (NKT + NKB) * NS; r += NTHR) {
      const bool top = r < NKT * NS;
      const int u = top ? r : r - NKT * NS;
      const int gidx = u / NS, i = u - gidx * NS;
      const int k = top ? P - 1 - 2 * gidx : P + 1 + 2 * gidx;
      double t = xt[k * VS + i];
      if (top) {
        double v[NX];
        load_vec<NX>(xt + (k + 1) * VS, v);
        t = dot4<NX>(Ft + (size_t)k * NX * PS + i, PS, v, t, -1.0);
      } else {
        double v[NS];
        load_vec<NS>(xt + (k - 1) * VS, v);
        t = dot4<NS>(Fbt + (size_t)(k - 1 - P) * NS * PS + i, PS, v, t, -1.0);
      }
      xt[k * VS + i] = t;
    }
    __syncthreads();
    TR(8);
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) xs[s] = on[s] ? xt[JV(s)] : 0.0;
  };

  // ---- residuals / termination / infeasibility / rho estimate (oracle/mpc.py _residuals, _check, _rho_estimate) --------
  // Reads the exchange vectors written at a checking iteration: x in xt, dx in yt, y_eq in sye, dy_eq in weq, dy_box in sdy.
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
        const double Pb = sPb[JV(s)], Dj = gD[JG(s)], Eb = gE[neq + JG(s)];
        const double dx = yt[JV(s)], dyb = sdy[JV(s)];
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
        double dye = 0.0, yes = 0.0;
        if (eqo[s]) {
          const double Ee = gE[ER(s)], bes = sbe[ER(s)];
          dye = weq[ER(s)];
          yes = sye[ER(s)];
          const double Ax = eq_row(s, xt, x[s]), Ei = 1.0 / Ee;
          const double zes = sze[ER(s)];
          v[2] = fmax(v[2], fabs(Ei * (Ax - zes))); v[3] = fmax(v[3], fabs(Ei * Ax)); v[4] = fmax(v[4], fabs(Ei * zes));
          v[5] = fmax(v[5], fabs(Ax - zes)); v[6] = fmax(v[6], fabs(Ax)); v[7] = fmax(v[7], fabs(zes));
          v[8] = fmax(v[8], fabs(Ee * dye));
          v[NRED_MAX + 1] += bes * fmax(dye, 0.0) + bes * fmin(dye, 0.0);
          const double Adx = eq_row(s, yt, dx) * Ei;
          v[9] = fmax(v[9], Adx);
          v[10] = fmax(v[10], -Adx);
        }
        {
          const double Px = Pb * x[s], Aty = at_col(s, yb[s], yes, sye), Di = 1.0 / Dj;
          v[11] = fmax(v[11], fabs(Di * (Px + q[s] + Aty))); v[12] = fmax(v[12], fabs(Di * Px));
          v[13] = fmax(v[13], fabs(Di * Aty)); v[14] = fmax(v[14], fabs(Di * q[s]));
          v[15] = fmax(v[15], fabs(Px + q[s] + Aty)); v[16] = fmax(v[16], fabs(Px));
          v[17] = fmax(v[17], fabs(Aty)); v[18] = fmax(v[18], fabs(q[s]));
          v[19] = fmax(v[19], fabs(at_col(s, dyp, dye, weq) * Di));
        }
      }
    }
    reduce(v, IC3<NRED_MAX>{}, IC3<NRED_SUM>{});
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
  // light-iteration state: g = A_e'(rho_e z_e - y_e), Dv = P + sigma + rho_b Bd^2, c0 = rho_e A_e' b_e, Xs = sum of x~
  double g[NSLOT], Dv[NSLOT], c0[NSLOT], Xs[NSLOT];
  int n_light = 0;
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    g[s] = Dv[s] = c0[s] = Xs[s] = 0.0;   // (sye also serves a solve without iterations: dx = dy = 0, x in xt)
  }
  // g, c0, Dv from the equality multipliers ye[] and z_e = b_e; stores ye / ze for the next full iteration
  auto sync_light = [&](const double (&ye)[NSLOT]) {
#pragma unroll
    for (int s = 0; s < NSLOT; ++s)
      if (eqo[s]) {
        weq[ER(s)] = fma(rho_eq, sbe[ER(s)], -ye[s]);
        sye[ER(s)] = ye[s];
        sze[ER(s)] = sbe[ER(s)];
      }
    __syncthreads();
    double bown[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
      g[s] = on[s] ? (at_col(s, 0.0, eqo[s] ? weq[ER(s)] : 0.0, weq)) : 0.0;
      Dv[s] = on[s] ? fma(rb, Bd[s] * Bd[s], sPb[JV(s)] + sigma) : 1.0;
      bown[s] = eqo[s] ? rho_eq * sbe[ER(s)] : 0.0;
      Xs[s] = 0.0;
    }
    __syncthreads();   // every read of w is done
#pragma unroll
    for (int s = 0; s < NSLOT; ++s)
      if (eqo[s]) weq[ER(s)] = bown[s];
    __syncthreads();
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) c0[s] = on[s] ? at_col(s, 0.0, bown[s], weq) : 0.0;
    n_light = 0;
    __syncthreads();
  };

  for (it = 1; it <= S.max_iter; ++it) {
    const bool can_check = S.check_every > 0 && (it % S.check_every) == 0;
    const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval > 0 && (it % S.adaptive_rho_interval) == 0;
    const bool checking = can_check || can_adapt || it == S.max_iter;
    const bool full = checking || it == 1;
    double xtil[NSLOT];
    TR(0);
    if (!full) {
      // ---- light iteration: element-wise right-hand side, KKT solve, element-wise updates -----------------------------------
      // g+ = g - alpha (rhs - Dv x~) + alpha c0: the part that does not need x~ is formed before the solve, so that
      // neither rhs nor the old g stays live across it
      double gpre[NSLOT];
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
        const double rhs = fma(sigma, x[s], -q[s]) + (Bd[s] * fma(rb, zb[s], -yb[s]) + g[s]);
        gpre[s] = fma(-alpha, rhs, g[s]);
        if (on[s]) yt[JV(s)] = rhs;
      }
      __syncthreads();
      kkt_solve(xtil);
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
        const double rbi = (cls[s] == 0) ? (1.0 / RHO_MIN) : ((cls[s] == 1) ? rei : rhoi);
        const double xtj = xtil[s];
        g[s] = fma(alpha, fma(Dv[s], xtj, c0[s]), gpre[s]);
        Xs[s] += xtj;
        const double ztb = Bd[s] * xtj;
        const double zr = alpha * ztb + (1.0 - alpha) * zb[s];
        const double zn = fmin(fmax(fma(yb[s], rbi, zr), lbs[s]), ubs[s]);
        yb[s] += rb * (zr - zn);
        zb[s] = zn;
        x[s] = alpha * xtj + (1.0 - alpha) * x[s];
      }
      ++n_light;
      continue;   // (the next write of yt is separated from this solve's reads of yt by the barriers inside kkt_solve)
    }
    // ---- full iteration (the first one and every checking iteration) ------------------------------------------------------------
    double ye[NSLOT], ze[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      ye[s] = eqo[s] ? sye[ER(s)] : 0.0;
      ze[s] = eqo[s] ? sze[ER(s)] : 0.0;
    }
    if (n_light > 0) {
      // equality multipliers as of now: y_e += rho_e alpha (A_e sum x~ - n_light b_e)   (z_e = b_e already)
      __syncthreads();   // the previous (light) iteration's reads of xt are done
#pragma unroll
      for (int s = 0; s < NSLOT; ++s)
        if (on[s]) xt[JV(s)] = Xs[s];
      __syncthreads();
#pragma unroll
      for (int s = 0; s < NSLOT; ++s)
        if (eqo[s]) ye[s] = fma(rho_eq * alpha, eq_row(s, xt, Xs[s]) - (double)n_light * sbe[ER(s)], ye[s]);
      __syncthreads();
    }
#pragma unroll
    for (int s = 0; s < NSLOT; ++s)
      if (eqo[s]) weq[ER(s)] = fma(rho_eq, ze[s], -ye[s]);
    __syncthreads();
    // right-hand side sigma x - q + A'(rho z - y)
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
      if (on[s])
        yt[JV(s)] = fma(sigma, x[s], -q[s]) + at_col(s, fma(rb, zb[s], -yb[s]), eqo[s] ? weq[ER(s)] : 0.0, weq);
    }
    __syncthreads();
    kkt_solve(xtil);   // (x~ is also in xt, behind a barrier)
    // z_tilde = A x_tilde ; z, y, x updates
    double dx[NSLOT], dyb[NSLOT], dye[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
      const double rb = (cls[s] == 0) ? RHO_MIN : ((cls[s] == 1) ? rho_eq : rho);
      const double rbi = (cls[s] == 0) ? (1.0 / RHO_MIN) : ((cls[s] == 1) ? rei : rhoi);
      const double xtj = xtil[s];
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
        const double zte = eq_row(s, xt, xtj), bes = sbe[ER(s)];
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
    __syncthreads();  // every read of xt (eq_row above) is done
    if (checking) {
#pragma unroll
      for (int s = 0; s < NSLOT; ++s) {
        if (on[s]) { xt[JV(s)] = x[s]; yt[JV(s)] = dx[s]; sdy[JV(s)] = dyb[s]; }
        if (eqo[s]) { sye[ER(s)] = ye[s]; sze[ER(s)] = ze[s]; weq[ER(s)] = dye[s]; }
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
          factorize();
        }
      }
      __syncthreads();  // (reads of the check buffers are done before sync_light rewrites weq)
    }
    if (it < S.max_iter) sync_light(ye);
  }
  TR(40);
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
    if (live[s]) ob[NRED_MAX] += (0.5 * sPb[JV(s)] * x[s] + q[s]) * x[s];
  reduce(ob, IC3<0>{}, IC3<1>{});
  const bool commit_y = !S.commit_y_solved_only || status == GPMPC_QP_SOLVED;
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    if (live[s]) {
      xg[(size_t)b * n + JG(s)] = gD[JG(s)] * x[s];
      if (commit_y) yg[(size_t)b * m + neq + JG(s)] = gE[neq + JG(s)] * yb[s] / cscale;
    }
    if (eqo[s] && commit_y) yg[(size_t)b * m + ER(s)] = gE[ER(s)] * sye[ER(s)] / cscale;
  }
  TR(41);
  if (tid == 0) {
    status_g[b] = status;
    iters_g[b] = it;
    obj_g[b] = ob[NRED_MAX] / cscale;
    pri_g[b] = co.pri;
    dua_g[b] = co.dua;
    rhog[b] = rho;
  }
}

#ifndef ADMM3_SREG
#define ADMM3_SREG false
#endif
// Launch shapes, keyed by (n_x, compiled horizon).  n_x = 14 has two: N = 30 (the 6-DoF step, BASELINE configs[3]) and N = 20
// (the 3-DoF QP with the 7 convex-row states, GpmpcCone3: 21 x 17 = 357 logical slots on 256 threads x 2).
template <int NX, int NT> struct Shape3 {
  static constexpr int NSLOT = 2, NTHR = 128, MINB = 3;
  static constexpr bool SREG = ADMM3_SREG, ABS = true;
};
template <> struct Shape3<14, 30> {
  static constexpr int NSLOT = 3, NTHR = 256, MINB = 1;
  static constexpr bool SREG = false, ABS = false;
};
template <> struct Shape3<14, 20> {
  static constexpr int NSLOT = 2, NTHR = 256, MINB = 1;
  static constexpr bool SREG = false, ABS = false;
};

template <int NX, int NU, int NT>
int launch_admm3_fixed(int B, const double* AB, const double* beq, const double* q, const double* Pdiag, const double* lb,
                       const double* ub, const GpmpcAdmmSettings& S, const int* active, double* scratch, double* x, double* y,
                       double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream) {
  using H = Shape3<NX, NT>;
  using L = Lay<NX, NU, NT>;
  constexpr size_t smem = L::template smem_doubles<H::ABS>() * sizeof(double);
  static_assert(smem <= 227 * 1024, "shared memory budget exceeded");
  auto kern = admm3_kernel<NX, NU, NT, H::NSLOT, H::NTHR, H::MINB, H::SREG, H::ABS>;
  static GpmpcPerDevice configured;
  if (!configured.here()) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured.here() = 1;
  }
  kern<<<B, H::NTHR, smem, stream>>>(B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj, pri,
                                     dua);
  return gpmpc_check_launch("admm3");
}

}  // namespace

// Scratch doubles per QP that launch_admm3 needs (cumulative scalings, + the scaled stage blocks for the n_x = 14 shapes).
template <int NX, int NU>
size_t admm3_scratch_doubles(int N) {
  const size_t nv = (size_t)N * (NX + NU) + NX, nm = 2 * nv + (size_t)NX * (N + 1);
  return nm + ((NX == 14) ? (size_t)N * NX * (NX + NU) : 0);
}

// Returns GPMPC_ERR_ARG without launching when the horizon is not a compiled one (the caller falls back to the
// second-generation kernel).
template <int NX, int NU>
int launch_admm3(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag, const double* lb,
                 const double* ub, const GpmpcAdmmSettings& S, const int* active, double* scratch, double* x, double* y,
                 double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream) {
  if (NX == 7 && N == 20)
    return launch_admm3_fixed<7, NU, 20>(B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj, pri,
                                         dua, stream);
  if (NX == 14 && N == 30)
    return launch_admm3_fixed<14, NU, 30>(B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj,
                                          pri, dua, stream);
  if (NX == 14 && N == 20)
    return launch_admm3_fixed<14, NU, 20>(B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj,
                                          pri, dua, stream);
  return GPMPC_ERR_ARG;
}

template size_t admm3_scratch_doubles<7, 3>(int);
template size_t admm3_scratch_doubles<14, 3>(int);
template int launch_admm3<7, 3>(int, int, const double*, const double*, const double*, const double*, const double*,
                                const double*, const GpmpcAdmmSettings&, const int*, double*, double*, double*, double*,
                                int*, int*, double*, double*, double*, cudaStream_t);
template int launch_admm3<14, 3>(int, int, const double*, const double*, const double*, const double*, const double*,
                                 const double*, const GpmpcAdmmSettings&, const int*, double*, double*, double*, double*,
                                 int*, int*, double*, double*, double*, cudaStream_t);
