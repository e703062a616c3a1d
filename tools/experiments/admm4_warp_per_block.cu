// Fourth generation of the batched OSQP-style ADMM kernel, for the 3-DoF shape (n_x = 7, n_u = 3, N = 20): same problem,
// algorithm and C entry point as csrc/admm.cu (reference call sites osqp_rti.py:464-478,496,517,524,527; algorithm restated
// from Stellato et al. 2020, see the header of csrc/admm.cu), same "light iteration" algebra as csrc/admm3.cu.
//
// What changed is the shape of the KKT solve.  The reduced KKT matrix M = P + sigma I + A' diag(rho) A is block tridiagonal
// in the stage ordering; here THREE stages form one 30 x 30 block and ONE WARP owns one block (7 warps, lane = variable):
//
//   * the twisted block elimination runs over 7 blocks instead of 21 stages: 3 dependent steps per side instead of 10;
//     a forward step is 7 (top) or 10 (bottom) rows of a 30-term product, a backward step 30 rows of a 7 / 10-term product;
//   * d_K = S_K^{-1} yt_K is a dense 30 x 30 product INSIDE the owning warp (its row of S_K^{-1} lives in the lane's
//     registers), so it needs no block barrier and overlaps with the chains of the other warps;
//   * the warps are coupled only through their chain neighbours, so a solve uses NAMED barriers between producer and
//     consumer warps (bar.arrive / bar.sync on 64 threads) instead of block-wide barriers: the critical path of one
//     iteration is  rhs -> 3 forward steps -> the middle block's product -> 3 backward steps -> updates,
//     everything else happens off that path in the other warps;
//   * block barriers remain only in the full iterations (first + every checking iteration) and in the set-up.
//
// The per-iteration arithmetic (which products, in which order of accumulation) differs from the stage-wise kernels only
// in the order of floating-point summation inside the KKT solve; tests/test_gpu_mpc.py holds all generations to the same
// oracle (status, iteration count, rho history, x, y).
#include <math.h>
#include <stdio.h>

#include "common.cuh"
#include "gpmpc_b200.h"

using namespace gpmpc;

#ifdef ADMM4_TRACE
// low-overhead phase trace: one thread of one QP appends (clock, id) pairs to a shared-memory buffer (a handful of cycles per
// point), dumped to global memory when the kernel ends
__device__ long long g_admm4_trace[4096];
__device__ int g_admm4_trace_n;
#ifndef TRACE_TID
#define TRACE_TID 96
#endif
#define TR4_DECL __shared__ long long tr4_buf[2048]; int tr4_n = 0;
#define TR4(id)                                                                   \
  do {                                                                            \
    if (tid == TRACE_TID && b == 7 && tr4_n < 2048) tr4_buf[tr4_n++] = (clock64() << 8) | (long long)(id); \
  } while (0)
#define TR4_DUMP                                                                  \
  do {                                                                            \
    if (tid == TRACE_TID && b == 7) {                                             \
      for (int e_ = 0; e_ < tr4_n; ++e_) g_admm4_trace[e_] = tr4_buf[e_];         \
      g_admm4_trace_n = tr4_n;                                                    \
    }                                                                             \
  } while (0)
extern "C" int gpmpc_admm4_trace_read(long long* out, int cap) {
  int n = 0;
  cudaMemcpyFromSymbol(&n, g_admm4_trace_n, sizeof(int));
  if (n > cap) n = cap;
  cudaMemcpyFromSymbol(out, g_admm4_trace, sizeof(long long) * n);
  int z = 0;
  cudaMemcpyToSymbol(g_admm4_trace_n, &z, sizeof(int));
  return n;
}
#else
#define TR4_DECL
#define TR4(id)
#define TR4_DUMP
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

// producer / consumer barriers between two warps (PTX ISA, "bar.arrive ... bar.sync" producer-consumer pattern)
// (the whole warp must be converged at either call)
__device__ __forceinline__ void nb_sync(int id) {
  __syncwarp();
  asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}
__device__ __forceinline__ void nb_arrive(int id) {
  __syncwarp();
  asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory");
}

template <int NX, int NU, int NT, int BS>
struct Lay4 {
  static constexpr int NS = NX + NU, N = NT, NB = (NT + 1) / BS, BV = BS * NS, MID = NB / 2;
  static constexpr int PB = (BV % 2 == 0) ? BV + 1 : BV;   // odd pitch of the factor rows
  static constexpr int NPAD = NB * BV, EPAD = (N + 2) * NX;
  static_assert((NT + 1) % BS == 0, "the stages must split into whole blocks");
  static_assert(BV <= 32 && BV % 2 == 0, "one block per warp, even block size");
  static_assert(4 * MID <= 15 && NB <= MAX_WARPS && NB >= 3 && NB - 1 - MID == MID, "barrier ids / warps");
  static constexpr size_t ev(size_t c) { return (c + 1) & ~(size_t)1; }
  template <bool SREG>
  static constexpr size_t smem_doubles() {
    size_t s = ev((size_t)N * NX * NS);            // AB
    s += ev((size_t)MID * NX * PB);                // Ft
    s += ev((size_t)MID * NS * PB);                // Fbt
    s += ev((size_t)NS * PB) + ev((size_t)NX * PB);  // XT, XB
    if (!SREG) s += ev((size_t)NB * BV * PB);      // Sinv
    s += 4 * ev((size_t)NPAD);                     // yt xt sPb sdy
    s += 5 * ev((size_t)EPAD);                     // weq sye ec sbe sze
    s += ev((size_t)(MID + 1) * ((NX + 1) & ~1)) + ev((size_t)(MID + 1) * ((NS + 1) & ~1));   // ctop, cbot
    s += ev((size_t)NB * 32);                      // sweep column exchange, one row per warp
    s += ev((size_t)MAX_WARPS * NRED);             // red
    return s;
  }
};

template <int NM, int NSUM>
__device__ __forceinline__ void block_reduce_mixed4(double (&v)[NRED], double* red, int warp, int lane, int nwarps) {
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
struct IC4 { static constexpr int value = I; };

template <int LEN>
__device__ __forceinline__ void load_vec4(const double* v, double (&o)[LEN]) {
#pragma unroll
  for (int c = 0; c + 1 < LEN; c += 2) {
    const double2 t = *reinterpret_cast<const double2*>(v + c);
    o[c] = t.x;
    o[c + 1] = t.y;
  }
  if (LEN & 1) o[LEN - 1] = v[LEN - 1];
}

// In-place inverse of the symmetric positive definite BV x BV matrix held one row per lane (lanes >= BV hold zero rows) by
// the symmetric sweep operator (Goodnight 1979): sweeping pivot k maps  a_kk -> -1/d, a_ik -> a_ik/d,
// a_ij -> a_ij - a_ik a_kj / d  (d = a_kk), keeps the matrix symmetric at every step and ends at -A^{-1} (negated on the way
// out).  Symmetry is what makes the row broadcast cheap: row k equals column k, and column k is spread over the lanes
// (lane c holds a_ck), so every lane publishes ONE value per pivot (a conflict-free store) instead of one lane publishing a row.
template <int BV>
__device__ __forceinline__ void sweep_inverse_blk(double (&a)[BV], int lane, double* scr) {
#pragma unroll
  for (int k = 0; k < BV; ++k) {
    scr[lane] = a[k];            // S[lane][k] = S[k][lane]
    __syncwarp();
    const double rp = 1.0 / scr[k];
    const bool piv = lane == k;
    const double g = a[k] * rp;
    const double coef = piv ? rp : -g;
#pragma unroll
    for (int c = 0; c < BV; ++c) a[c] = fma(coef, scr[c], piv ? 0.0 : a[c]);
    a[k] = piv ? -rp : g;
    __syncwarp();
  }
#pragma unroll
  for (int c = 0; c < BV; ++c) a[c] = -a[c];
}

// One warp per block of BS stages; lane l < BV owns variable BV * warp + l of the z ordering.
template <int NX, int NU, int NT, int BS, bool SREG, int MINB>
__global__ void __launch_bounds__(32 * ((NT + 1) / BS), MINB)
admm4_kernel(int nbatch, const double* __restrict__ ABg, const double* __restrict__ beqg, const double* __restrict__ qg,
             const double* __restrict__ Pdiag, const double* __restrict__ lb, const double* __restrict__ ub,
             GpmpcAdmmSettings S, const int* __restrict__ active, double* __restrict__ scratch,
             double* __restrict__ xg, double* __restrict__ yg, double* __restrict__ rhog, int* __restrict__ status_g,
             int* __restrict__ iters_g, double* __restrict__ obj_g, double* __restrict__ pri_g,
             double* __restrict__ dua_g) {
  using L = Lay4<NX, NU, NT, BS>;
  constexpr int NS = L::NS, N = L::N, NB = L::NB, BV = L::BV, PB = L::PB, MID = L::MID, NTHR = 32 * NB;
  constexpr int n = N * NS + NX, neq = NX * (N + 1), m = neq + n;
  constexpr int NPAD = L::NPAD, EPAD = L::EPAD;
  extern __shared__ __align__(16) double smem4[];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (b >= nbatch) return;
  if (active && !active[b]) return;
  const int K = tid >> 5, lane = tid & 31;   // K: block = warp
  TR4_DECL
  TR4(30);

  // ---- shared memory carve (must match Lay4::smem_doubles) ------------------------------------------------------------
  double* sp = smem4;
  auto take = [&](size_t cnt) { double* q_ = sp; sp += (cnt + 1) & ~(size_t)1; return q_; };
  double* AB = take((size_t)N * NX * NS);
  double* Ft = take((size_t)MID * NX * PB);    // Ft[K-1][i][c]    = F_K[i][c],   K = 1..MID        (NX rows of BV)
  double* Fbt = take((size_t)MID * NS * PB);   // Fbt[K-MID][c2][c] = Fb_K[c2][c], K = MID..NB-2    (NS rows of BV)
  double* XT = take((size_t)NS * PB);          // last NS rows of S_{K-1}^{-1} on their way to block K
  double* XB = take((size_t)NX * PB);          // first NX rows of S_{K+1}^{-1} on their way to block K
  double* Sinv = SREG ? nullptr : take((size_t)NB * BV * PB);
  double* yt = take(NPAD);
  double* xt = take(NPAD);
  double* sPb = take(NPAD);   // scaled cost diagonal
  double* sdy = take(NPAD);   // dy of the box rows at termination checks
  double* weq = take(EPAD);   // equality-row exchange vector, indexed by owner (k,i): k*NX+i ; stage N+1 stays zero
  double* sye = take(EPAD);   // y of the equality rows as of the last full iteration
  double* ec = take(EPAD);    // own-variable coefficient of every equality row (scaled Ei / Nd)
  double* sbe = take(EPAD);   // scaled equality right-hand sides
  double* sze = take(EPAD);   // z of the equality rows (A x of the warm start until the first iteration, then b_e)
  constexpr int CT = (NX + 1) & ~1, CB = (NS + 1) & ~1;
  double* ctop = take((size_t)(MID + 1) * CT);   // ctop[K]: correction of the first NX entries of block K's right-hand side
  double* cbot = take((size_t)(MID + 1) * CB) - (size_t)MID * CB;   // cbot[K], K = MID..NB-2: correction of its last NS entries
  double* gjs = take((size_t)NB * 32) + K * 32;
  double* red = take((size_t)MAX_WARPS * NRED);
  double* gD = scratch + (size_t)b * (n + m);   // cumulative Ruiz scalings D (n) then E (m), read at checks / unscaling
  double* gE = gD + n;

  // ---- lane <-> variable ---------------------------------------------------------------------------------------------------
  const bool on = lane < BV;
  const int j = on ? K * BV + lane : 0;        // index into the per-variable vectors and the global x / box rows
  const int k = j / NS, i = j - k * NS;        // stage, variable within the stage
  const bool live = on && (k < N || i < NX);   // a real variable
  const bool eqo = live && i < NX;             // owns an equality row
  const bool hasAB = on && k < N;              // stage has a dynamics block (columns of AB_k)
  const bool hasPrev = eqo && k >= 1;          // own equality row is a dynamics row (rows of AB_{k-1})
  const int er = eqo ? k * NX + i : 0;

  auto reduce = [&](double (&v)[NRED], auto nmax_c, auto nsum_c) {
    block_reduce_mixed4<decltype(nmax_c)::value, decltype(nsum_c)::value>(v, red, K, lane, NB);
  };

  // ---- load ------------------------------------------------------------------------------------------------------
  for (int e = tid; e < N * NX * NS; e += NTHR) AB[e] = ABg[(size_t)b * N * NX * NS + e];
  for (int e = tid; e < EPAD; e += NTHR) { weq[e] = 0.0; sye[e] = 0.0; ec[e] = 0.0; sbe[e] = 0.0; sze[e] = 0.0; }
  for (int e = tid; e < NPAD; e += NTHR) { yt[e] = 0.0; xt[e] = 0.0; sPb[e] = 0.0; sdy[e] = 0.0; }
  const double sigma = S.sigma;
  double rho = (S.rho_from_input && rhog[b] > 0.0) ? rhog[b] : S.rho, cscale = 1.0;
  // own problem data (issued before the first use so that the loads overlap)
  double Pb = 0.0, q = 0.0, lb_in = 0.0, ub_in = 0.0, x_in = 0.0, yb_in = 0.0, ye_in = 0.0, be_in = 0.0;
  if (live) {
    Pb = Pdiag[j];
    q = qg[(size_t)b * n + j];
    lb_in = lb[(size_t)b * S.bound_stride + j];
    ub_in = ub[(size_t)b * S.bound_stride + j];
    if (S.warm_x) x_in = xg[(size_t)b * n + j];
    if (S.warm_y) yb_in = yg[(size_t)b * m + neq + j];
  }
  if (eqo) {
    be_in = beqg[(size_t)b * neq + er];
    if (S.warm_y) ye_in = yg[(size_t)b * m + er];
  }
  double x = 0.0, zb = 0.0, yb = 0.0, Bd = live ? 1.0 : 0.0, lbs = 0.0, ubs = 0.0;
  int cls = 2;  // box row class: 0 loose (rho_min), 1 equality-like (rho_eq), 2 inequality (rho)
  __syncthreads();
  TR4(31);
  {
    // ---- modified Ruiz equilibration + cost scaling (OSQP scaling.c; oracle/mpc.py OSQPOracle._scale) ---------------
    double e_ = eqo ? ((er < NX) ? 1.0 : -1.0) : 0.0, Dj = 1.0, Eb = 1.0, Ee = 1.0;
    for (int it = 0; it < S.scaling; ++it) {
      double v = fmax(fabs(Pb), fabs(Bd));
      v = fmax(v, fabs(e_));
      if (hasAB) {
        const double* a = AB + (size_t)k * NX * NS + i;
#pragma unroll
        for (int r = 0; r < NX; ++r) v = fmax(v, fabs(a[r * NS]));
      }
      const double Dt = 1.0 / sqrt(limit_scaling(v));
      const double Etb = 1.0 / sqrt(limit_scaling(fabs(Bd)));
      double ve = fabs(e_);
      if (hasPrev) {
        const double* a = AB + ((size_t)(k - 1) * NX + i) * NS;
#pragma unroll
        for (int c = 0; c < NS; ++c) ve = fmax(ve, fabs(a[c]));
      }
      const double Ete = 1.0 / sqrt(limit_scaling(ve));
      if (eqo) weq[er] = Ete;
      __syncthreads();
      double rv[NRED];
      rv[0] = 0.0;
      rv[NRED_MAX] = 0.0;
      if (hasAB) {
        double* a = AB + (size_t)k * NX * NS + i;
        const double* wn = weq + (k + 1) * NX;
#pragma unroll
        for (int r = 0; r < NX; ++r) a[r * NS] = (wn[r] * a[r * NS]) * Dt;
      }
      e_ = (Ete * e_) * Dt;
      Bd = (Etb * Bd) * Dt;
      Pb = (Dt * Pb) * Dt;
      q = Dt * q;
      Dj *= Dt; Eb *= Etb; Ee *= Ete;
      if (live) {
        rv[0] = fabs(q);
        rv[NRED_MAX] = fabs(Pb);
      }
      reduce(rv, IC4<1>{}, IC4<1>{});
      double ct = limit_scaling(fmax(rv[NRED_MAX] / n, limit_scaling(rv[0])));
      ct = 1.0 / ct;
      Pb *= ct;
      q *= ct;
      cscale *= ct;
      __syncthreads();
    }
    // scaled bounds, row classes, warm start (x = x_in / D ; y = c y_in / E)
    if (live) {
      lbs = Eb * lb_in;
      ubs = Eb * ub_in;
      if (lbs < -INF_T && ubs > INF_T) cls = 0;
      else if (ubs - lbs < RHO_TOL) cls = 1;
      if (S.warm_x) x = x_in / Dj;
      if (S.warm_y) yb = cscale * yb_in / Eb;
      gD[j] = Dj;
      gE[neq + j] = Eb;
    }
    if (on) {
      sPb[j] = Pb;
      xt[j] = x;
    }
    if (eqo) {
      ec[er] = e_;
      gE[er] = Ee;
      sbe[er] = Ee * be_in;
      sye[er] = S.warm_y ? cscale * ye_in / Ee : 0.0;
      weq[er] = 0.0;
    }
  }
  __syncthreads();
  TR4(32);
#ifdef ADMM4_DEBUG
  if (b == 0 && (tid % 37) == 0) printf("dbg4 tid %d j %d k %d i %d live %d x %g Bd %g lbs %g ubs %g q %g cscale %g gD %g\n", tid, j, k, i, (int)live, x, Bd, lbs, ubs, q, cscale, live ? gD[j] : -1.0);
#endif

  // (A v)_own-equality-row given v in shared memory and the own value vj  [only for eqo lanes]
  auto eq_row = [&](const double* v, double vj) -> double {
    double s0 = 0.0, s1 = 0.0;
    if (hasPrev) {
      const double* a = AB + ((size_t)(k - 1) * NX + i) * NS;
      const double* vp = v + (k - 1) * NS;
#pragma unroll
      for (int c = 0; c + 1 < NS; c += 2) {
        s0 = fma(a[c], vp[c], s0);
        s1 = fma(a[c + 1], vp[c + 1], s1);
      }
      if (NS & 1) s0 = fma(a[NS - 1], vp[NS - 1], s0);
    }
    return fma(ec[er], vj, s0 + s1);
  };
  // (A' w)_own-column given the own box value wb, the own equality value we and the shared equality vector
  auto at_col = [&](double wb, double we, const double* wv) -> double {
    double r0 = Bd * wb;
    if (eqo) r0 = fma(ec[er], we, r0);
    if (hasAB) {
      const double* a = AB + (size_t)k * NX * NS + i;
      const double* wn = wv + (k + 1) * NX;
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
    zb = Bd * x;
    if (eqo) sze[er] = eq_row(xt, x);
  }
  __syncthreads();

  // named barrier ids: forward edge into block Kc from the top (Kc = 1..MID) / from the bottom (Kc = NB-2..MID),
  // backward edge into block Kc above the middle (Kc = MID-1..0) / below it (Kc = MID+1..NB-1)
  auto id_ft = [&](int Kc) { return Kc; };
  auto id_fb = [&](int Kc) { return MID + (NB - 1 - Kc); };
  auto id_bt = [&](int Kc) { return 2 * MID + (MID - Kc); };
  auto id_bb = [&](int Kc) { return 3 * MID + (Kc - MID); };

  double srow[BV];   // row `lane` of S_K^{-1}

  // ---- twisted block factorisation: S_K^{-1} (registers), F_K, Fb_K (shared memory) -----------------------------------
  auto factorize = [&]() {
    const double rho_eq = RHO_EQ_OVER_INEQ * rho;
    TR4(20);
    // row `lane` of M_KK: the three stages of the block couple through the dynamics rows between them
    double a[BV];
#pragma unroll
    for (int c = 0; c < BV; ++c) a[c] = 0.0;
    if (on) {
      const double rb = (cls == 0) ? RHO_MIN : ((cls == 1) ? rho_eq : rho);
      const double e_ = eqo ? ec[er] : 0.0;
      double d = sPb[j] + sigma;
      d = fma(rb, Bd * Bd, d);
      d = fma(rho_eq, e_ * e_, d);
      const int s_own = lane / NS;   // stage of the row within the block
      double col[NX];
#pragma unroll
      for (int r = 0; r < NX; ++r) col[r] = hasAB ? AB[((size_t)k * NX + r) * NS + i] : 0.0;
#pragma unroll
      for (int ic = 0; ic < NS; ++ic) {
        // own stage (diagonal 10 x 10 block): rho_e AB_k' AB_k + own terms
        double acc = 0.0;
        if (hasAB) {
          const double* ac = AB + (size_t)k * NX * NS + ic;
#pragma unroll
          for (int r = 0; r < NX; ++r) acc = fma(col[r], ac[r * NS], acc);
        }
        double vd = rho_eq * acc;
        if (ic == i) vd = live ? vd + d : 1.0;
        // next stage, column (k+1, ic): the equality row (k+1, ic) = ec x_{k+1,ic} + AB_k[ic][:] . (x_k, u_k)
        const double vu = (ic < NX && hasAB) ? (rho_eq * ec[(k + 1) * NX + ic]) * AB[((size_t)k * NX + ic) * NS + i] : 0.0;
        // previous stage, column (k-1, ic): the own equality row (k, i), i < NX
        const double vl = (hasPrev) ? (rho_eq * e_) * AB[((size_t)(k - 1) * NX + i) * NS + ic] : 0.0;
#pragma unroll
        for (int t = 0; t < BS; ++t)
          a[t * NS + ic] = (t == s_own) ? vd : ((t == s_own + 1) ? vu : ((t + 1 == s_own) ? vl : 0.0));
      }
    }
    // Schur updates and couplings
    auto compute_F_top = [&]() {   // F_K[i2][lane] = rho_e ec[kb,i2] sum_c2 AB_{kb-1}[i2][c2] S_{K-1}^{-1}[BV-NS+c2][lane]
      if (on) {
        const int kb = K * BS;
        const double* ab = AB + (size_t)(kb - 1) * NX * NS;
        double xc[NS];
#pragma unroll
        for (int c2 = 0; c2 < NS; ++c2) xc[c2] = XT[c2 * PB + lane];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          double acc = 0.0;
#pragma unroll
          for (int c2 = 0; c2 < NS; ++c2) acc = fma(ab[i2 * NS + c2], xc[c2], acc);
          Ft[((size_t)(K - 1) * NX + i2) * PB + lane] = (rho_eq * ec[kb * NX + i2]) * acc;
        }
      }
    };
    auto schur_top = [&]() {       // S_K[i][i2] -= sum_c2 F_K[i][BV-NS+c2] rho_e ec[kb,i2] AB_{kb-1}[i2][c2], rows i < NX
      if (lane < NX) {
        const int kb = K * BS;
        const double* f = Ft + ((size_t)(K - 1) * NX + lane) * PB + (BV - NS);
        const double* ab = AB + (size_t)(kb - 1) * NX * NS;
        double fr[NS];
#pragma unroll
        for (int c2 = 0; c2 < NS; ++c2) fr[c2] = f[c2];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          double acc = 0.0;
#pragma unroll
          for (int c2 = 0; c2 < NS; ++c2) acc = fma(fr[c2], ab[i2 * NS + c2], acc);
          a[i2] = fma(-(rho_eq * ec[kb * NX + i2]), acc, a[i2]);
        }
      }
    };
    auto compute_F_bot = [&]() {   // Fb_K[c2][lane] = sum_i2 rho_e ec[kn,i2] AB_{kn-1}[i2][c2] S_{K+1}^{-1}[i2][lane]
      if (on) {
        const int kn = (K + 1) * BS;
        const double* ab = AB + (size_t)(kn - 1) * NX * NS;
        double g[NX];
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) g[i2] = (rho_eq * ec[kn * NX + i2]) * XB[i2 * PB + lane];
#pragma unroll
        for (int c2 = 0; c2 < NS; ++c2) {
          double acc = 0.0;
#pragma unroll
          for (int i2 = 0; i2 < NX; ++i2) acc = fma(ab[i2 * NS + c2], g[i2], acc);
          Fbt[((size_t)(K - MID) * NS + c2) * PB + lane] = acc;
        }
      }
    };
    auto schur_bot = [&]() {       // S_K[BV-NS+c2][BV-NS+c] -= sum_i2 Fb_K[c2][i2] rho_e ec[kn,i2] AB_{kn-1}[i2][c]
      if (lane >= BV - NS && lane < BV) {
        const int kn = (K + 1) * BS;
        const double* fb = Fbt + ((size_t)(K - MID) * NS + (lane - (BV - NS))) * PB;
        const double* ab = AB + (size_t)(kn - 1) * NX * NS;
#pragma unroll
        for (int i2 = 0; i2 < NX; ++i2) {
          const double g = fb[i2] * (rho_eq * ec[kn * NX + i2]);
#pragma unroll
          for (int c = 0; c < NS; ++c) a[BV - NS + c] = fma(-g, ab[i2 * NS + c], a[BV - NS + c]);
        }
      }
    };
    auto publish_top = [&]() {     // last NS rows of the inverse -> XT
      if (lane >= BV - NS && lane < BV) {
        double* o = XT + (size_t)(lane - (BV - NS)) * PB;
#pragma unroll
        for (int c = 0; c < BV; ++c) o[c] = a[c];
      }
    };
    auto publish_bot = [&]() {     // first NX rows of the inverse -> XB
      if (lane < NX) {
        double* o = XB + (size_t)lane * PB;
#pragma unroll
        for (int c = 0; c < BV; ++c) o[c] = a[c];
      }
    };
    TR4(21);
#ifdef ADMM4_DEBUG
    if (b == 0 && K == 0 && lane < 12) printf("dbg4-M lane %d: a2 %g a0 %g a1 %g a10 %g a11 %g a29 %g rho_eq %g sPb %g\n", lane, a[2], a[0], a[1], a[10], a[11], a[29], rho_eq, sPb[j]);
#endif
    if (K < MID) {
      if (K >= 1) {
        nb_sync(id_ft(K));
        compute_F_top();
        __syncwarp();
        schur_top();
      }
      sweep_inverse_blk<BV>(a, lane, gjs);
      publish_top();
      nb_arrive(id_ft(K + 1));
    } else if (K > MID) {
      if (K <= NB - 2) {
        nb_sync(id_fb(K));
        compute_F_bot();
        __syncwarp();
        schur_bot();
      }
      sweep_inverse_blk<BV>(a, lane, gjs);
      publish_bot();
      nb_arrive(id_fb(K - 1));
    } else {
      nb_sync(id_ft(MID));
      nb_sync(id_fb(MID));
      compute_F_top();
      compute_F_bot();
      __syncwarp();
      schur_top();
      schur_bot();
      sweep_inverse_blk<BV>(a, lane, gjs);
    }
#pragma unroll
    for (int c = 0; c < BV; ++c) srow[c] = a[c];
    if (!SREG && on) {
      double* o = Sinv + ((size_t)K * BV + lane) * PB;
#pragma unroll
      for (int c = 0; c < BV; ++c) o[c] = a[c];
    }
    __syncthreads();
    TR4(22);
  };
  factorize();
#ifdef ADMM4_DEBUG
  if (b == 0 && (tid % 37) == 0) printf("dbg4-postfact tid %d x %g srow0 %g srow_diag %g\n", tid, x, srow[0], on ? srow[lane % BV] : 0.0);
#endif

  // ---- one KKT solve: returns (M^{-1} r)_own with r (own entry) given; uses yt / xt and the named barriers ---------------------
  // Must be entered by all warps with no pending reads of yt / xt by other warps (block barrier or the previous solve's
  // back edges) and leaves x~ of every block in xt (visible to all after a block barrier).
  //
  // Forward elimination, condensed: block K's right-hand side changes only in its first NX entries (top side, by c_K) or its
  // last NS entries (bottom side, by cb_K):  yt_K = r_K - [c_K; 0]  with  c_{K+1} = F_{K+1} yt_K = F_{K+1} r_K - Phi_{K+1} c_K,
  // Phi_{K+1} = F_{K+1}[:, 0..NX).  The long product F_{K+1} r_K (30 terms) does not depend on the chain: every warp forms it
  // at once, several lanes per row; only the NX x NX (NS x NS) product with the incoming correction is on the critical path.
  constexpr int LT = 4, TT = (BV + LT - 1) / LT;       // top rows: LT lanes per row, TT terms per lane
  constexpr int LB = 3, TB = (BV + LB - 1) / LB;       // bottom rows: LB lanes per row
  static_assert(NX * LT <= 32 && NS * LB <= 32, "lanes per row");
  auto kkt_solve = [&](double r) -> double {
    if (on) yt[j] = r;
    __syncwarp();
    TR4(1);
    if (K < MID) {
      // p = F_{K+1} r_K, row = lane / LT, part = lane % LT
      const int row = lane / LT, part = lane - row * LT;
      double pacc = 0.0;
      if (lane < NX * LT) {
        const double* f = Ft + ((size_t)K * NX + row) * PB + part * TT;
        const double* v = yt + K * BV + part * TT;
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int c = 0; c < TT; c += 2) {
          if (part * TT + c < BV) a0 = fma(f[c], v[c], a0);
          if (c + 1 < TT && part * TT + c + 1 < BV) a1 = fma(f[c + 1], v[c + 1], a1);
        }
        pacc = a0 + a1;
      }
      pacc += __shfl_xor_sync(0xffffffffu, pacc, 1);
      pacc += __shfl_xor_sync(0xffffffffu, pacc, 2);
      if (K >= 1) {
        nb_sync(id_ft(K));
        TR4(2);
        const double* cin = ctop + K * CT;
        if (lane < NX) {           // own correction
          r -= cin[lane];
          yt[j] = r;
        }
        if (lane < NX * LT && part == 0) {   // pacc -= Phi_{K+1} c_K
          const double* f = Ft + ((size_t)K * NX + row) * PB;
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int c = 0; c < NX; c += 2) {
            a0 = fma(f[c], cin[c], a0);
            if (c + 1 < NX) a1 = fma(f[c + 1], cin[c + 1], a1);
          }
          pacc -= a0 + a1;
        }
      }
      if (lane < NX * LT && part == 0) ctop[(K + 1) * CT + row] = pacc;
      nb_arrive(id_ft(K + 1));
    } else if (K > MID) {
      // p = Fb_{K-1} r_K, row = lane / LB (index into the last NS variables of block K-1)
      const int row = lane / LB, part = lane - row * LB;
      double pacc = 0.0;
      if (lane < NS * LB) {
        const double* f = Fbt + ((size_t)(K - 1 - MID) * NS + row) * PB + part * TB;
        const double* v = yt + K * BV + part * TB;
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int c = 0; c < TB; c += 2) {
          if (part * TB + c < BV) a0 = fma(f[c], v[c], a0);
          if (c + 1 < TB && part * TB + c + 1 < BV) a1 = fma(f[c + 1], v[c + 1], a1);
        }
        pacc = a0 + a1;
      }
      {   // sum over the LB = 3 lanes of a row into its first lane
        const double p1 = __shfl_down_sync(0xffffffffu, pacc, 1);
        const double p2 = __shfl_down_sync(0xffffffffu, pacc, 2);
        pacc = pacc + (p1 + p2);
      }
      if (K <= NB - 2) {
        nb_sync(id_fb(K));
        TR4(2);
        const double* cin = cbot + K * CB;
        if (lane >= BV - NS && lane < BV) {   // own correction
          r -= cin[lane - (BV - NS)];
          yt[j] = r;
        }
        if (lane < NS * LB && part == 0) {    // pacc -= Phib_{K-1} cb_K, Phib = Fb_{K-1}[:, BV-NS..BV)
          const double* f = Fbt + ((size_t)(K - 1 - MID) * NS + row) * PB + (BV - NS);
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int c = 0; c < NS; c += 2) {
            a0 = fma(f[c], cin[c], a0);
            if (c + 1 < NS) a1 = fma(f[c + 1], cin[c + 1], a1);
          }
          pacc -= a0 + a1;
        }
      }
      if (lane < NS * LB && part == 0) cbot[(K - 1) * CB + row] = pacc;
      nb_arrive(id_fb(K - 1));
    } else {
      nb_sync(id_ft(MID));
      nb_sync(id_fb(MID));
      TR4(2);
      if (lane < NX) r -= ctop[MID * CT + lane];
      else if (lane >= BV - NS && lane < BV) r -= cbot[MID * CB + lane - (BV - NS)];
      if (on) yt[j] = r;
    }
    __syncwarp();
    TR4(3);
    // d_K = S_K^{-1} yt_K inside the warp
    double d;
    {
      const double* v = yt + K * BV;
      const double* sr = Sinv + ((size_t)K * BV + (on ? lane : 0)) * PB;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
      for (int c = 0; c < BV; c += 4) {
        a0 = fma(SREG ? srow[c] : sr[c], v[c], a0);
        a1 = fma(SREG ? srow[c + 1] : sr[c + 1], v[c + 1], a1);
        if (c + 2 < BV) {
          a2 = fma(SREG ? srow[c + 2] : sr[c + 2], v[c + 2], a2);
          a3 = fma(SREG ? srow[c + 3] : sr[c + 3], v[c + 3], a3);
        }
      }
      d = (a0 + a1) + (a2 + a3);
    }
    // backward chain, outwards from the middle block: v_K = d_K - F_{K+1}' v_{K+1}[0..NX)  (above the middle),
    //                                                 v_K = d_K - Fb_{K-1}' v_{K-1}[BV-NS..BV)  (below it)
    double v = d;
    TR4(4);
    if (K == MID) {
      if (on) xt[j] = v;
      nb_arrive(id_bt(MID - 1));
      nb_arrive(id_bb(MID + 1));
    } else if (K < MID) {
      nb_sync(id_bt(K));
      TR4(5);
      if (on) {
        const double* f = Ft + (size_t)K * NX * PB + lane;   // F_{K+1}[:][lane]
        const double* vn = xt + (K + 1) * BV;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int i2 = 0; i2 + 1 < NX; i2 += 2) {
          s0 = fma(f[i2 * PB], vn[i2], s0);
          s1 = fma(f[(i2 + 1) * PB], vn[i2 + 1], s1);
        }
        if (NX & 1) s0 = fma(f[(NX - 1) * PB], vn[NX - 1], s0);
        v = d - (s0 + s1);
        xt[j] = v;
      }
      if (K >= 1) nb_arrive(id_bt(K - 1));
    } else {
      nb_sync(id_bb(K));
      TR4(5);
      if (on) {
        const double* f = Fbt + (size_t)(K - 1 - MID) * NS * PB + lane;   // Fb_{K-1}[:][lane]
        const double* vn = xt + (K - 1) * BV + (BV - NS);
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int c2 = 0; c2 + 1 < NS; c2 += 2) {
          s0 = fma(f[c2 * PB], vn[c2], s0);
          s1 = fma(f[(c2 + 1) * PB], vn[c2 + 1], s1);
        }
        if (NS & 1) s0 = fma(f[(NS - 1) * PB], vn[NS - 1], s0);
        v = d - (s0 + s1);
        xt[j] = v;
      }
      if (K <= NB - 2) nb_arrive(id_bb(K + 1));
    }
    TR4(6);
    return v;
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
    if (live) {
      const double Pbv = sPb[j], Dj = gD[j], Eb = gE[neq + j];
      const double dx = yt[j], dyb = sdy[j];
      const bool uinf = ubs > INF_T, linf = lbs < -INF_T;
      // box row: projection of dy on the polar of the recession cone
      double dyp = dyb;
      if (uinf && linf) dyp = 0.0;
      else if (uinf) dyp = fmin(dyp, 0.0);
      else if (linf) dyp = fmax(dyp, 0.0);
      v[0] = fabs(Dj * dx);
      v[1] = fabs(Pbv * dx / Dj);
      v[NRED_MAX + 0] = q * dx;
      {
        const double Ax = Bd * x, Ei = 1.0 / Eb;
        v[2] = fabs(Ei * (Ax - zb)); v[3] = fabs(Ei * Ax); v[4] = fabs(Ei * zb);
        v[5] = fabs(Ax - zb); v[6] = fabs(Ax); v[7] = fabs(zb);
        v[8] = fabs(Eb * dyp);
        v[NRED_MAX + 1] = ubs * fmax(dyp, 0.0) + lbs * fmin(dyp, 0.0);
        const double Adx = (Bd * dx) * Ei;
        if (!uinf) v[9] = fmax(v[9], Adx);
        if (!linf) v[10] = fmax(v[10], -Adx);
      }
      double dye = 0.0, yes = 0.0;
      if (eqo) {
        const double Ee = gE[er], bes = sbe[er];
        dye = weq[er];
        yes = sye[er];
        const double Ax = eq_row(xt, x), Ei = 1.0 / Ee;
        const double zes = sze[er];
        v[2] = fmax(v[2], fabs(Ei * (Ax - zes))); v[3] = fmax(v[3], fabs(Ei * Ax)); v[4] = fmax(v[4], fabs(Ei * zes));
        v[5] = fmax(v[5], fabs(Ax - zes)); v[6] = fmax(v[6], fabs(Ax)); v[7] = fmax(v[7], fabs(zes));
        v[8] = fmax(v[8], fabs(Ee * dye));
        v[NRED_MAX + 1] += bes * fmax(dye, 0.0) + bes * fmin(dye, 0.0);
        const double Adx = eq_row(yt, dx) * Ei;
        v[9] = fmax(v[9], Adx);
        v[10] = fmax(v[10], -Adx);
      }
      {
        const double Px = Pbv * x, Aty = at_col(yb, yes, sye), Di = 1.0 / Dj;
        v[11] = fabs(Di * (Px + q + Aty)); v[12] = fabs(Di * Px);
        v[13] = fabs(Di * Aty); v[14] = fabs(Di * q);
        v[15] = fabs(Px + q + Aty); v[16] = fabs(Px);
        v[17] = fabs(Aty); v[18] = fabs(q);
        v[19] = fabs(at_col(dyp, dye, weq) * Di);
      }
    }
    reduce(v, IC4<NRED_MAX>{}, IC4<NRED_SUM>{});
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
  double g = 0.0, Dv = 1.0, c0 = 0.0, Xs = 0.0;
  int n_light = 0;
  // g, c0, Dv from the equality multipliers ye and z_e = b_e; stores ye / ze for the next full iteration
  auto sync_light = [&](double ye) {
    const double rb = (cls == 0) ? RHO_MIN : ((cls == 1) ? rho_eq : rho);
    const double bown = eqo ? rho_eq * sbe[er] : 0.0;
    if (eqo) {
      weq[er] = fma(rho_eq, sbe[er], -ye);
      sye[er] = ye;
      sze[er] = sbe[er];
    }
    __syncthreads();
    g = on ? at_col(0.0, eqo ? weq[er] : 0.0, weq) : 0.0;
    Dv = on ? fma(rb, Bd * Bd, sPb[j] + sigma) : 1.0;
    Xs = 0.0;
    __syncthreads();   // every read of w is done
    if (eqo) weq[er] = bown;
    __syncthreads();
    c0 = on ? at_col(0.0, bown, weq) : 0.0;
    n_light = 0;
    __syncthreads();
  };

  TR4(33);
  for (it = 1; it <= S.max_iter; ++it) {
    const bool can_check = S.check_every > 0 && (it % S.check_every) == 0;
    const bool can_adapt = S.adaptive_rho && S.adaptive_rho_interval > 0 && (it % S.adaptive_rho_interval) == 0;
    const bool checking = can_check || can_adapt || it == S.max_iter;
    const bool full = checking || it == 1;
    const double rb = (cls == 0) ? RHO_MIN : ((cls == 1) ? rho_eq : rho);
    const double rbi = (cls == 0) ? (1.0 / RHO_MIN) : ((cls == 1) ? rei : rhoi);
    TR4(0);
    if (!full) {
      // ---- light iteration: element-wise right-hand side, KKT solve, element-wise updates -----------------------------------
      const double rhs = fma(sigma, x, -q) + (Bd * fma(rb, zb, -yb) + g);
      const double gpre = fma(-alpha, rhs, g);
      const double xtj = kkt_solve(rhs);
      g = fma(alpha, fma(Dv, xtj, c0), gpre);
      Xs += xtj;
      const double zr = alpha * (Bd * xtj) + (1.0 - alpha) * zb;
      const double zn = fmin(fmax(fma(yb, rbi, zr), lbs), ubs);
      yb += rb * (zr - zn);
      zb = zn;
      x = alpha * xtj + (1.0 - alpha) * x;
      ++n_light;
      continue;   // (the next solve's writes of yt / xt are ordered behind this solve's reads by its back edges)
    }
    // ---- full iteration (the first one and every checking iteration) ------------------------------------------------------------
    double ye = eqo ? sye[er] : 0.0, ze = eqo ? sze[er] : 0.0;
    __syncthreads();   // the previous solve's reads of yt / xt are done
    if (n_light > 0) {
      // equality multipliers as of now: y_e += rho_e alpha (A_e sum x~ - n_light b_e)   (z_e = b_e already)
      if (on) xt[j] = Xs;
      __syncthreads();
      if (eqo) ye = fma(rho_eq * alpha, eq_row(xt, Xs) - (double)n_light * sbe[er], ye);
      __syncthreads();
    }
    if (eqo) weq[er] = fma(rho_eq, ze, -ye);
    __syncthreads();
    // right-hand side sigma x - q + A'(rho z - y)
    const double rhs = fma(sigma, x, -q) + at_col(fma(rb, zb, -yb), eqo ? weq[er] : 0.0, weq);
    const double xtj = kkt_solve(rhs);
    __syncthreads();   // x~ of every block is in xt
    // z_tilde = A x_tilde ; z, y, x updates
    double dx, dyb, dye = 0.0;
    {
      const double zr = alpha * (Bd * xtj) + (1.0 - alpha) * zb;
      const double zn = fmin(fmax(fma(yb, rbi, zr), lbs), ubs);
      dyb = rb * (zr - zn);
      yb += dyb;
      zb = zn;
    }
    if (eqo) {
      const double zte = eq_row(xt, xtj), bes = sbe[er];
      const double zr = alpha * zte + (1.0 - alpha) * ze;
      const double zn = bes;  // projection of zr + y / rho on the degenerate interval [b, b]
      dye = rho_eq * (zr - zn);
      ye += dye;
      ze = zn;
    }
    {
      const double xn = alpha * xtj + (1.0 - alpha) * x;
      dx = xn - x;
      x = xn;
    }
    __syncthreads();  // every read of xt (eq_row above) is done
    if (checking) {
      if (on) { xt[j] = x; yt[j] = dx; sdy[j] = dyb; }
      if (eqo) { sye[er] = ye; sze[er] = ze; weq[er] = dye; }
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
  TR4(40);
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
  ob[NRED_MAX] = live ? (0.5 * sPb[j] * x + q) * x : 0.0;
  reduce(ob, IC4<0>{}, IC4<1>{});
  const bool commit_y = !S.commit_y_solved_only || status == GPMPC_QP_SOLVED;
#ifdef ADMM4_DEBUG
  if (b == 0 && (tid % 37) == 0) printf("dbg4-out tid %d x %g gD %g status %d it %d\n", tid, x, live ? gD[j] : -1.0, status, it);
#endif
  if (live) {
    xg[(size_t)b * n + j] = gD[j] * x;
    if (commit_y) yg[(size_t)b * m + neq + j] = gE[neq + j] * yb / cscale;
  }
  if (eqo && commit_y) yg[(size_t)b * m + er] = gE[er] * sye[er] / cscale;
  TR4(41);
  TR4_DUMP;
  if (tid == 0) {
    status_g[b] = status;
    iters_g[b] = it;
    obj_g[b] = ob[NRED_MAX] / cscale;
    pri_g[b] = co.pri;
    dua_g[b] = co.dua;
    rhog[b] = rho;
  }
}

#ifndef ADMM4_SREG
#define ADMM4_SREG false
#endif
constexpr bool kSreg4 = ADMM4_SREG;
constexpr int kMinB4 = 2;

}  // namespace

// 3-DoF shape only: (nx, nu, N) = (7, 3, 20), three stages per block.  Returns GPMPC_ERR_ARG without launching otherwise
// (the caller falls back to the third / second generation).  scratch: B x (n + m) doubles.
int launch_admm4_7_3(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag, const double* lb,
                     const double* ub, const GpmpcAdmmSettings& S, const int* active, double* scratch, double* x, double* y,
                     double* rho, int* status, int* iters, double* obj, double* pri, double* dua, cudaStream_t stream) {
  constexpr int NX = 7, NU = 3, NT = 20, BS = 3;
  if (N != NT) return GPMPC_ERR_ARG;
  using L = Lay4<NX, NU, NT, BS>;
  constexpr size_t smem = L::smem_doubles<kSreg4>() * sizeof(double);
  static_assert(smem <= 227 * 1024, "shared memory budget exceeded");
  auto kern = admm4_kernel<NX, NU, NT, BS, kSreg4, kMinB4>;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured = true;
  }
  kern<<<B, 32 * L::NB, smem, stream>>>(B, AB, beq, q, Pdiag, lb, ub, S, active, scratch, x, y, rho, status, iters, obj, pri,
                                        dua);
  return gpmpc_check_launch("admm4");
}
