// Micro-benchmark (not part of the product): named-barrier producer/consumer ping-pong between two warps, and a 30-term
// FP64 dot product with row data in shared memory on 7 / 28 lanes.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void nb_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
__global__ void pingpong(long long* t, double* out, int n) {
  __shared__ double s[64];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  s[threadIdx.x & 63] = 1.0;
  __syncthreads();
  double v = 1.0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    if (w == 0) {
      s[lane] = v;
      nb_arrive(1);
      nb_sync(2);
      v = s[32 + lane] + 1.0;
    } else if (w == 1) {
      nb_sync(1);
      v = s[lane] + 1.0;
      s[32 + lane] = v;
      nb_arrive(2);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[0] = t1 - t0;
  out[threadIdx.x] = v;
}
__global__ void dot30(long long* t, double* out, int n, int mode) {
  __shared__ __align__(16) double F[10 * 31];
  __shared__ __align__(16) double v[32];
  const int lane = threadIdx.x & 31;
  for (int e = threadIdx.x; e < 310; e += blockDim.x) F[e] = 1.0 / (1 + e);
  if (threadIdx.x < 32) v[threadIdx.x] = 0.5;
  __syncthreads();
  double acc = 0.0;
  long long t0 = clock64();
  for (int it = 0; it < n; ++it) {
    if (mode == 0) {               // 7 lanes x 30 terms
      if (lane < 7) {
        const double* f = F + lane * 31;
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
        for (int c = 0; c < 30; c += 4) {
          const double2 v0 = *reinterpret_cast<const double2*>(v + c);
          a0 = fma(f[c], v0.x, a0); a1 = fma(f[c + 1], v0.y, a1);
          if (c + 2 < 30) { const double2 v1 = *reinterpret_cast<const double2*>(v + c + 2); a2 = fma(f[c + 2], v1.x, a2); a3 = fma(f[c + 3], v1.y, a3); }
        }
        acc = (a0 + a1) + (a2 + a3);
        v[lane] = acc * 1e-3;
      }
    } else {                        // 28 lanes: 4 lanes per row, 8 terms each (last part 6), then two shuffles
      const int row = lane >> 2, part = lane & 3;
      double a0 = 0, a1 = 0;
      if (lane < 28) {
        const double* f = F + row * 31 + part * 8;
        const double* vv = v + part * 8;
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
          if (part * 8 + c < 30) {
            const double2 v0 = *reinterpret_cast<const double2*>(vv + c);
            a0 = fma(f[c], v0.x, a0); a1 = fma(f[c + 1], v0.y, a1);
          }
        }
      }
      double sacc = a0 + a1;
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
      acc = sacc;
      if (lane < 28 && part == 0) v[row] = acc * 1e-3;
    }
    __syncwarp();
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[0] = t1 - t0;
  out[threadIdx.x] = acc;
}
int main() {
  long long* t; double* d; long long h;
  cudaMalloc(&t, 64); cudaMalloc(&d, 4096 * 8);
  const int n = 2000;
  pingpong<<<1, 64>>>(t, d, n); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
  printf("named-barrier ping-pong (STS, arrive, sync, LDS+DADD, STS, arrive, sync, LDS+DADD): %.1f cycles per round trip\n", (double)h / n);
  pingpong<<<1, 224>>>(t, d, n); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
  printf("  same with 5 idle warps in the block: %.1f\n", (double)h / n);
  for (int mode = 0; mode < 2; ++mode) {
    dot30<<<1, 32>>>(t, d, n, mode); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
    printf("30-term dot, mode %d (%s): %.1f cycles per dependent step\n", mode, mode ? "28 lanes x 8 terms + 2 shuffles" : "7 lanes x 30 terms", (double)h / n);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
