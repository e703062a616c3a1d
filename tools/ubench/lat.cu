// Micro-benchmarks of the latencies that bound the ADMM chains (not part of the product): dependent DFMA / DADD chains,
// LDS -> DFMA -> STS -> LDS round trips, SHFL.64, block barrier, on one warp and with several warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_dfma(double* out, long long* t, int n) {
  double a = out[0], b = out[1], c = out[2];
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int j = 0; j < 16; ++j) a = fma(a, b, c);
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) { t[blockIdx.x] = t1 - t0; }
  out[3 + threadIdx.x] = a;
}
__global__ void k_dfma_ilp4(double* out, long long* t, int n) {
  double a0 = out[0], a1 = out[1], a2 = out[2], a3 = out[3], b = out[1], c = out[2];
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c); }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
  out[4 + threadIdx.x] = a0 + a1 + a2 + a3;
}
__global__ void k_ffma(float* out, long long* t, int n) {
  float a = out[0], b = out[1], c = out[2];
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int j = 0; j < 16; ++j) a = fmaf(a, b, c);
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
  out[3 + threadIdx.x] = a;
}
__global__ void k_lds_chain(double* out, long long* t, int n) {
  __shared__ double s[64];
  s[threadIdx.x & 63] = 0.0;
  __syncthreads();
  int idx = threadIdx.x & 31;
  long long t0 = clock64();
  double v = 1.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[idx] = v;            // STS
      __syncwarp();
      v = s[(idx + 1) & 31] + 1.0;   // LDS of another lane's value + 1 dependent DADD
      __syncwarp();
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
  out[threadIdx.x] = v;
}
__global__ void k_shfl(double* out, long long* t, int n) {
  double v = out[threadIdx.x];
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 1) & 31) + 1.0;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
  out[threadIdx.x] = v;
}
__global__ void k_bar(double* out, long long* t, int n) {
  long long t0 = clock64();
  double v = out[threadIdx.x];
  __shared__ double s[512];
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    s[threadIdx.x] = v;
    __syncthreads();
    v = s[(threadIdx.x + 33) % blockDim.x] + 1.0;
    __syncthreads();
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
  out[threadIdx.x] = v;
}
int main() {
  double* d; long long* t; float* f;
  cudaMalloc(&d, 4096 * sizeof(double)); cudaMalloc(&t, 4096 * sizeof(long long)); cudaMalloc(&f, 4096 * sizeof(float));
  cudaMemset(d, 0, 4096 * sizeof(double)); cudaMemset(f, 0, 4096 * sizeof(float));
  long long h[8];
  const int n = 1000;
  auto rd = [&](const char* name, double per) { cudaDeviceSynchronize(); cudaMemcpy(h, t, sizeof(h), cudaMemcpyDeviceToHost); printf("%-44s %8.2f cycles per op\n", name, h[0] / per); };
  for (int warps : {1, 4, 12}) {
    printf("--- %d warp(s) per block, 1 block\n", warps);
    k_dfma<<<1, 32 * warps>>>(d, t, n); rd("dependent DFMA", n * 16.0);
    k_dfma_ilp4<<<1, 32 * warps>>>(d, t, n); rd("DFMA, 4 independent chains (per fma)", n * 16.0);
    k_ffma<<<1, 32 * warps>>>(f, t, n); rd("dependent FFMA", n * 16.0);
    k_lds_chain<<<1, 32 * warps>>>(d, t, n); rd("STS + syncwarp + LDS + DADD + syncwarp", n * 8.0);
    k_shfl<<<1, 32 * warps>>>(d, t, n); rd("SHFL.64 + DADD", n * 8.0);
    k_bar<<<1, 32 * warps>>>(d, t, n); rd("STS + bar + LDS + DADD + bar", n * 1.0);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
