// Shared device helpers for the sm_100a kernels: FP64 DMMA (mma.sync m8n8k4, the native DMMA.8x8x4 SASS on
// sm_100a -- tcgen05 has no FP64 kind), cp.async staging, warp/block reductions, error plumbing.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define GPMPC_OK 0
#define GPMPC_ERR_ARG -1
#define GPMPC_ERR_CUDA -2

extern "C" void gpmpc_set_error(const char* msg);
int gpmpc_check_launch(const char* what);

#define GPMPC_REQUIRE(cond, msg)          \
  do {                                    \
    if (!(cond)) {                        \
      gpmpc_set_error(msg);               \
      return GPMPC_ERR_ARG;               \
    }                                     \
  } while (0)

// One-time kernel attribute set-up is PER DEVICE (cudaFuncSetAttribute applies to the current device).  The framework runs one
// process per GPU (distributed.py), but a host that switches devices with cudaSetDevice still gets correct launches: the
// "already configured" flags are kept per device ordinal.  (Not thread-safe: concurrent first calls from several host threads
// need external locking, as do the opaque handles -- see include/gpmpc_b200.h.)
constexpr int GPMPC_MAX_DEVICES = 64;
struct GpmpcPerDevice {
  int v[GPMPC_MAX_DEVICES] = {};
  int& here() {
    int d = 0;
    cudaGetDevice(&d);
    return v[(d >= 0 && d < GPMPC_MAX_DEVICES) ? d : 0];
  }
};

// Kernels that may run while a shared-memory heavy kernel (ADMM, DMMA GEMM: carve-out = max shared) is resident on another
// stream ask for the same L1 / shared-memory split: an SM cannot host CTAs that want different carve-outs, so a
// default-carve-out kernel would wait for the SM to drain.
#define GPMPC_PREFER_SHARED(kern)                                                                              \
  do {                                                                                                         \
    static GpmpcPerDevice done_;                                                                               \
    if (!done_.here()) {                                                                                       \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
      done_.here() = 1;                                                                                        \
    }                                                                                                          \
  } while (0)

namespace gpmpc {

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// numpy's pairwise summation order for a contiguous run of n < 128 doubles (np.sum over the last axis of a
// C-contiguous array, numpy/core/src/umath/loops_utils.h pairwise_sum): bit-identical row norms to the reference
// kernels.py:228-229.
__device__ __forceinline__ double numpy_pairwise_sum_small(const double* a, int n) {
  if (n < 8) {
    double r = 0.0;  // numpy starts from the first element; 0.0 + a0 == a0 exactly (except -0.0, irrelevant here)
    for (int i = 0; i < n; ++i) r += a[i];
    return r;
  }
  double r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = a[j];
  int i;
  for (i = 8; i < n - (n % 8); i += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] += a[i + j];
  }
  double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  for (; i < n; ++i) res += a[i];
  return res;
}

// Covariance function value from the scaled squared distance (reference kernels.py:238-262 SE-ARD, :532-545 Matern32,
// :625-638 Matern52), same operation order as the numpy expressions.  kind: 0 SE, 1 Matern-3/2, 2 Matern-5/2, 3 = the distance.
__device__ __forceinline__ double kernel_value(int kind, double d2, double sv) {
  d2 = fmax(d2, 0.0);
  if (kind == 0) return sv * exp(-0.5 * d2);
  if (kind == 3) return d2;  // clamped scaled squared distance itself (isotropic SE, kernels.py:417-421)
  const double r = sqrt(d2);
  if (kind == 1) {
    const double s = 1.7320508075688772 * r;  // np.sqrt(3)
    return (sv * (1.0 + s)) * exp(-s);
  }
  const double s = 2.23606797749979 * r;  // np.sqrt(5)
  return (sv * ((1.0 + s) + (5.0 * (r * r)) / 3.0)) * exp(-s);
}

__device__ __forceinline__ double norm3(double a, double b, double c) {
  return sqrt(__dadd_rn(__dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b)), __dmul_rn(c, c)));
}

// Simple3DoFFeatureExtractor.extract, features.py:403-444 (11 features).
__device__ __forceinline__ void features_3dof_dev(const double* x, const double* u, double* z) {  // features.py:403-444
  const double alt = x[1];
  const double speed = norm3(x[4], x[5], x[6]);
  const double rho = 1.225 * exp(-alt / 8500.0);
  const double q_dyn = 0.5 * rho * (speed * speed);
  const double T = norm3(u[0], u[1], u[2]);
  z[0] = x[4] / 10.0; z[1] = x[5] / 10.0; z[2] = x[6] / 10.0; z[3] = speed / 10.0;
  z[4] = q_dyn / (0.5 * 1.225 * 100.0);
  z[5] = u[0] / 10.0; z[6] = u[1] / 10.0; z[7] = u[2] / 10.0; z[8] = T / 10.0;
  z[9] = alt / 100.0; z[10] = rho / 1.225;
}

__device__ __forceinline__ void quat_to_dcm(const double* q, double C[3][3]) {  // features.py:260-270
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  C[0][0] = 1 - 2 * (y * y + z * z); C[0][1] = 2 * (x * y + w * z);     C[0][2] = 2 * (x * z - w * y);
  C[1][0] = 2 * (x * y - w * z);     C[1][1] = 1 - 2 * (x * x + z * z); C[1][2] = 2 * (y * z + w * x);
  C[2][0] = 2 * (x * z + w * y);     C[2][1] = 2 * (y * z - w * x);     C[2][2] = 1 - 2 * (x * x + y * y);
}

// 6-DoF extractors on the 14-state [m, r(3), v(3), q(4), w(3)]: TranslationalFeatureExtractor.extract (features.py:194-258,
// 13 features, trans != 0) and RotationalFeatureExtractor.extract (features.py:305-354, 12 features).
__device__ __forceinline__ void features_6dof_dev(int trans, const double* x, const double* u, double* z) {
  const double alt = x[1];
  const double speed = norm3(x[4], x[5], x[6]);
  const double rho = 1.225 * exp(-alt / 8500.0);
  const double q_dyn = 0.5 * rho * (speed * speed);
  double C[3][3];
  double q[4] = {x[7], x[8], x[9], x[10]};
  quat_to_dcm(q, C);
  double vB[3];
  for (int r = 0; r < 3; ++r)
    vB[r] = __dadd_rn(__dadd_rn(__dmul_rn(C[r][0], x[4]), __dmul_rn(C[r][1], x[5])), __dmul_rn(C[r][2], x[6]));
  if (trans) {
    double alpha = 0.0, beta = 0.0;
    if (speed > 1e-3) {
      alpha = atan2(-vB[2], vB[0]);
      beta = asin(fmin(fmax(vB[1] / speed, -1.0), 1.0));
    }
    const double T = norm3(u[0], u[1], u[2]);
    z[0] = x[4] / 10.0; z[1] = x[5] / 10.0; z[2] = x[6] / 10.0; z[3] = speed / 10.0;
    z[4] = q_dyn / (0.5 * 1.225 * 100.0); z[5] = alpha; z[6] = beta;
    z[7] = u[0] / 10.0; z[8] = u[1] / 10.0; z[9] = u[2] / 10.0; z[10] = T / 10.0;
    z[11] = alt / 100.0; z[12] = rho / 1.225;
  } else {
    const double om = norm3(x[11], x[12], x[13]);
    z[0] = x[11] / 1.0; z[1] = x[12] / 1.0; z[2] = x[13] / 1.0; z[3] = om / 1.0;
    z[4] = u[0] / 10.0; z[5] = u[1] / 10.0; z[6] = u[2] / 10.0;
    z[7] = vB[0] / 10.0; z[8] = vB[1] / 10.0; z[9] = vB[2] / 10.0; z[10] = speed / 10.0;
    z[11] = q_dyn / (0.5 * 1.225 * 100.0);
  }
}

}  // namespace gpmpc
