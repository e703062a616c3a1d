// FP64 tensor-core GEMM  C(M,N) = alpha * A(M,K) * B(N,K)^T + beta * C   ("NT": both operands k-contiguous, row-major)
// for sm_100a.  FP64 has no tcgen05 kind, so the math is warp-level DMMA.8x8x4 (mma.sync m8n8k4) on register
// fragments, operands staged global->shared by a 3-stage cp.async (LDGSTS) pipeline with 16-byte copies and a
// (BK+4)-double row pitch that makes the 8-byte fragment loads bank-conflict free.
//
// All dimensions are padded by the caller: M % BM == 0, N % BN == 0, K % 16 == 0, leading dimensions even and
// base pointers 16-byte aligned -- no bounds checks in the main loop.
//
// Used by: blocked Cholesky (panel TRSM, trailing SYRK), the blocked triangular solve behind the GP predictive
// variance (reference exact_gp.py:260-261), the sparse-GP M x M products (sparse_gp.py:203,226).
#pragma once
#include "common.cuh"

namespace gpmpc {

constexpr int GEMM_BK = 16;
constexpr int GEMM_PITCH = GEMM_BK + 4;  // doubles; (g*20 + t) mod 16 distinct over a half-warp
constexpr int GEMM_STAGES = 3;
constexpr int GEMM_THREADS = 256;

enum GemmEpilogue : int {
  EPI_STORE = 0,         // C = alpha*acc + beta*C
  EPI_STORE_SUMSQ = 1,   // C = alpha*acc (+beta*C) and sumsq[n] += sum_m (alpha*acc)^2   (one M tile only)
  EPI_SUMSQ_PART = 2,    // no C store: sumsq[(m0/BM)*part_stride + n] = sum_{m in tile} (alpha*acc)^2   (any number of M tiles)
};

struct GemmParams {
  const double* A; long lda; long strideA;   // A[m*lda + k]
  const double* B; long ldb; long strideB;   // B[n*ldb + k]
  double* C; long ldc; long strideC;         // C[m*ldc + n]  (or C[n*ldc + m] when c_trans)
  double* sumsq; long strideS;               // EPI_STORE_SUMSQ: per-column accumulator (length N)
  int M, N, K;
  double alpha, beta;
  int epi;
  int c_trans;      // store transposed
  int k_tri;        // A is lower-triangular in (m,k): limit the k loop to k < m0 + BM
  int lower_only;   // skip tiles strictly above the block diagonal (n0 >= m0 + BM)
  int row_min;      // rows  m < row_min are computed but not stored
  int col_min;      // cols  n < col_min are computed but not stored
  long part_stride; // EPI_SUMSQ_PART: distance between the partial-sum rows of consecutive M tiles
  int m_fast;       // grid = (M tiles, N tiles) with the M tiles of one N tile adjacent and the longest k loops first
  int tri_skip;     // with k_tri: A is EXACTLY zero above its diagonal, so inside the diagonal block the 8-row MMA strips skip
                    // the k chunks that lie entirely above them (44 % of that block's DMMAs)
  int m_valid;      // > 0: rows m >= m_valid of A are padding that contributes nothing (their B columns are zero): the 8-row
                    // strips made of such rows are skipped altogether
};

template <int BM, int BN>
struct GemmSmem {
  double a[GEMM_STAGES][BM * GEMM_PITCH];
  double b[GEMM_STAGES][BN * GEMM_PITCH];
};

// One BM x BN output tile at (m0, n0) of batch entry bz, computed by the whole CTA.
template <int BM, int BN, int WARPS_M, int WARPS_N>
__device__ __forceinline__ void dgemm_nt_tile(const GemmParams& p, int m0, int n0, int bz, GemmSmem<BM, BN>& sm) {
  static_assert(WARPS_M * WARPS_N * 32 == GEMM_THREADS, "8 warps");
  constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
  constexpr int MT = WM / 8, NT = WN / 8;
  static_assert(WM % 8 == 0 && WN % 8 == 0, "warp tile multiple of 8");
  if (p.lower_only && n0 >= m0 + BM) return;
  const double* __restrict__ A = p.A + bz * p.strideA + (long)m0 * p.lda;
  const double* __restrict__ B = p.B + bz * p.strideB + (long)n0 * p.ldb;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int wm = warp / WARPS_N, wn = warp % WARPS_N;

  int K = p.K;
  if (p.k_tri) K = min(K, m0 + BM);
  const int nk = K / GEMM_BK;
  const int rowbase = m0 + wm * WM;                                            // first row of this warp's strips
  const int nk_main = (p.k_tri && p.tri_skip) ? min(nk, m0 / GEMM_BK) : nk;    // chunks entirely left of the diagonal block
  const int i_hi = (p.m_valid > 0) ? max(0, min(MT, (p.m_valid - rowbase + 7) >> 3)) : MT;   // strips with a valid row

  double acc[MT][NT][2];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  auto load_tile = [&](int stage, int kt) {
    const int k0 = kt * GEMM_BK;
    // A: BM rows x 16 doubles = BM*8 chunks of 16 B
#pragma unroll
    for (int c = tid; c < BM * 8; c += GEMM_THREADS) {
      int r = c >> 3, kc = (c & 7) * 2;
      cp_async16(&sm.a[stage][r * GEMM_PITCH + kc], A + (long)r * p.lda + k0 + kc);
    }
#pragma unroll
    for (int c = tid; c < BN * 8; c += GEMM_THREADS) {
      int r = c >> 3, kc = (c & 7) * 2;
      cp_async16(&sm.b[stage][r * GEMM_PITCH + kc], B + (long)r * p.ldb + k0 + kc);
    }
  };

#pragma unroll
  for (int s = 0; s < GEMM_STAGES - 1; ++s) {
    if (s < nk) load_tile(s, s);
    cp_async_commit();
  }

  for (int kt = 0; kt < nk; ++kt) {
    cp_async_wait<GEMM_STAGES - 2>();
    __syncthreads();
    {
      int nxt = kt + GEMM_STAGES - 1;
      if (nxt < nk) load_tile(nxt % GEMM_STAGES, nxt);
      cp_async_commit();
    }
    const double* as = sm.a[kt % GEMM_STAGES] + (wm * WM + g) * GEMM_PITCH + t;
    const double* bs = sm.b[kt % GEMM_STAGES] + (wn * WN + g) * GEMM_PITCH + t;
    if (kt < nk_main && i_hi == MT) {   // warp-uniform: every strip of this warp needs every k of the chunk
#pragma unroll
      for (int kk = 0; kk < GEMM_BK; kk += 4) {
        double af[MT], bf[NT];
#pragma unroll
        for (int i = 0; i < MT; ++i) af[i] = as[i * 8 * GEMM_PITCH + kk];
#pragma unroll
        for (int j = 0; j < NT; ++j) bf[j] = bs[j * 8 * GEMM_PITCH + kk];
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
          for (int j = 0; j < NT; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
      }
    } else {                            // diagonal block of a triangular A and / or padded rows: strip-wise skipping
      // strip i (rows rowbase + 8 i .. + 7) sees the 16 columns of this chunk iff the chunk starts at or left of its last row.
      // One branch per (chunk, strip) around 16 DMMAs: a finer test (per 4 columns) is if-converted by ptxas into predicated
      // DMMAs, and a predicated-off DMMA still occupies the tensor pipe (measured: no gain).
      const int krel = (kt >= nk_main) ? kt * GEMM_BK - rowbase : -(1 << 20);
      double bf[GEMM_BK / 4][NT];
#pragma unroll
      for (int kk = 0; kk < GEMM_BK; kk += 4)
#pragma unroll
        for (int j = 0; j < NT; ++j) bf[kk / 4][j] = bs[j * 8 * GEMM_PITCH + kk];
#pragma unroll
      for (int i = 0; i < MT; ++i) {
        if (i < i_hi && krel <= 8 * i + 7) {
#pragma unroll
          for (int kk = 0; kk < GEMM_BK; kk += 4) {
            const double af = as[i * 8 * GEMM_PITCH + kk];
#pragma unroll
            for (int j = 0; j < NT; ++j) dmma884(acc[i][j][0], acc[i][j][1], af, bf[kk / 4][j]);
          }
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();  // all operand reads (incl. in-place B == C) are complete before any store

  // ---- epilogue -------------------------------------------------------------------------------------------
  double* __restrict__ C = p.C + bz * p.strideC;
  const double alpha = p.alpha, beta = p.beta;
  double colsq[NT][2];
#pragma unroll
  for (int j = 0; j < NT; ++j) colsq[j][0] = colsq[j][1] = 0.0;

#pragma unroll
  for (int i = 0; i < MT; ++i) {
    const int m = m0 + wm * WM + i * 8 + g;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int n = n0 + wn * WN + j * 8 + 2 * t;
      double v0 = alpha * acc[i][j][0], v1 = alpha * acc[i][j][1];
      const bool row_ok = (m >= p.row_min);
      if (p.epi == EPI_SUMSQ_PART) {
      } else if (!p.c_trans) {
        double* cp = C + (long)m * p.ldc + n;
        if (beta != 0.0) {
          double2 old = *reinterpret_cast<const double2*>(cp);
          v0 += beta * old.x;
          v1 += beta * old.y;
        }
        if (row_ok) {
          if (n >= p.col_min && n + 1 >= p.col_min) {
            *reinterpret_cast<double2*>(cp) = make_double2(v0, v1);
          } else if (n + 1 >= p.col_min) {
            cp[1] = v1;
          }
        }
      } else {
        double* cp0 = C + (long)n * p.ldc + m;
        double* cp1 = cp0 + p.ldc;
        if (beta != 0.0) {
          v0 += beta * (*cp0);
          v1 += beta * (*cp1);
        }
        if (row_ok) {
          if (n >= p.col_min) *cp0 = v0;
          if (n + 1 >= p.col_min) *cp1 = v1;
        }
      }
      colsq[j][0] += v0 * v0;
      colsq[j][1] += v1 * v1;
    }
  }

  if (p.epi == EPI_STORE_SUMSQ || p.epi == EPI_SUMSQ_PART) {
    // reduce over the 8 row-groups of the warp (lanes with equal t), then across the WARPS_M warps via shared memory
    __shared__ double red[WARPS_M][BN];
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double v = colsq[j][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (g == 0) red[wm][wn * WN + j * 8 + 2 * t + e] = v;
      }
    __syncthreads();
    double* ss = p.sumsq + bz * p.strideS;
    for (int c = tid; c < BN; c += GEMM_THREADS) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < WARPS_M; ++w) v += red[w][c];
      if (p.epi == EPI_SUMSQ_PART) ss[(long)(m0 / BM) * p.part_stride + n0 + c] = v;
      else ss[n0 + c] += v;
    }
  }
}

template <int BM, int BN, int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(GEMM_THREADS, (BM * BN <= 64 * 128) ? 2 : 1) dgemm_nt_kernel(GemmParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GemmSmem<BM, BN>& sm = *reinterpret_cast<GemmSmem<BM, BN>*>(smem_raw);
  const int m0 = (p.m_fast ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.y) * BM;
  const int n0 = (p.m_fast ? (int)blockIdx.y : (int)blockIdx.x) * BN;
  dgemm_nt_tile<BM, BN, WARPS_M, WARPS_N>(p, m0, n0, blockIdx.z, sm);
}

// Persistent variant: gridDim.x CTAs (one per SM) stride over the tiles in the m_fast order (the M tiles of one N tile
// adjacent, longest k loops first).  Every CTA of the grid is resident from the start, so the blocks of a kernel launched
// on ANOTHER stream are dispatched into the SMs' remaining resources at once -- the work distributor only starts on a
// later kernel after all blocks of an earlier one have been dispatched.
template <int BM, int BN, int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(GEMM_THREADS, 1) dgemm_nt_persistent_kernel(GemmParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GemmSmem<BM, BN>& sm = *reinterpret_cast<GemmSmem<BM, BN>*>(smem_raw);
  const int mt = p.M / BM, nt = p.N / BN;
  for (int tile = blockIdx.x; tile < mt * nt; tile += gridDim.x) {
    const int ni = tile / mt, mi = mt - 1 - (tile - ni * mt);
    dgemm_nt_tile<BM, BN, WARPS_M, WARPS_N>(p, mi * BM, ni * BN, 0, sm);
    __syncthreads();  // the tile buffers and the epilogue's reduction scratch are reused by the next tile
  }
}

template <int BM, int BN, int WARPS_M, int WARPS_N>
int launch_dgemm_nt_persistent(const GemmParams& p, int ctas, int smem_request, cudaStream_t stream) {
  if (p.M % BM || p.N % BN || p.K % GEMM_BK || (p.lda & 1) || (p.ldb & 1)) {
    gpmpc_set_error("dgemm_nt: dimensions must be padded (M%BM, N%BN, K%16, even ld)");
    return GPMPC_ERR_ARG;
  }
  if (p.M == 0 || p.N == 0) return GPMPC_OK;
  auto kern = dgemm_nt_persistent_kernel<BM, BN, WARPS_M, WARPS_N>;
  const int smem = smem_request > (int)sizeof(GemmSmem<BM, BN>) ? smem_request : (int)sizeof(GemmSmem<BM, BN>);
  static GpmpcPerDevice configured;
  if (smem > configured.here()) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured.here() = smem;
  }
  kern<<<ctas, GEMM_THREADS, smem, stream>>>(p);
  return gpmpc_check_launch("dgemm_nt_persistent");
}

template <int BM, int BN, int WARPS_M, int WARPS_N>
int launch_dgemm_nt(const GemmParams& p, int batch, cudaStream_t stream, int smem_request = 0) {
  if (p.M % BM || p.N % BN || p.K % GEMM_BK || (p.lda & 1) || (p.ldb & 1)) {
    gpmpc_set_error("dgemm_nt: dimensions must be padded (M%BM, N%BN, K%16, even ld)");
    return GPMPC_ERR_ARG;
  }
  if (p.M == 0 || p.N == 0) return GPMPC_OK;
  auto kern = dgemm_nt_kernel<BM, BN, WARPS_M, WARPS_N>;
  // smem_request > sizeof(GemmSmem): ask for more shared memory than the tiles need, to cap the resident CTAs per SM
  // (used to leave room for a co-resident kernel on another stream)
  const int smem = smem_request > (int)sizeof(GemmSmem<BM, BN>) ? smem_request : (int)sizeof(GemmSmem<BM, BN>);
  static GpmpcPerDevice configured;
  if (smem > configured.here()) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured.here() = smem;
  }
  dim3 grid(p.N / BN, p.M / BM, batch);
  if (p.m_fast) grid = dim3(p.M / BM, p.N / BN, batch);
  kern<<<grid, GEMM_THREADS, smem, stream>>>(p);
  return gpmpc_check_launch("dgemm_nt");
}

}  // namespace gpmpc
