// Micro-benchmark (not part of the product): issue cost of independent shared-memory loads in one warp.
#include <cstdio>
#include <cuda_runtime.h>
template <int NL, int W>   // NL independent loads of W doubles each per iteration
__global__ void lds_rate(long long* t, double* out, int n, int stride, int lanes) {
  __shared__ __align__(16) double s[4096];
  for (int e = threadIdx.x; e < 4096; e += blockDim.x) s[e] = e * 1e-3;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  int base = (lane * stride) & 1023;
  long long t0 = clock64();
  if (lane < lanes) {
    for (int it = 0; it < n; ++it) {
      double r[NL * W];
#pragma unroll
      for (int l = 0; l < NL; ++l) {
        if (W == 1) r[l] = s[base + l * 33];
        else { const double2 q = *reinterpret_cast<const double2*>(s + ((base + l * 34) & ~1)); r[2 * l] = q.x; r[2 * l + 1] = q.y; }
      }
      double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
      for (int l = 0; l < NL * W; l += 4) { a0 += r[l]; a1 += r[l + 1]; a2 += r[l + 2]; a3 += r[l + 3]; }
      acc += (a0 + a1) + (a2 + a3);
      base = (base + (int)(acc * 1e-30)) & 1023;   // dependency between iterations without changing the address
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[0] = t1 - t0;
  out[threadIdx.x] = acc;
}
int main() {
  long long* t; double* d; long long h;
  cudaMalloc(&t, 64); cudaMalloc(&d, 4096 * 8);
  const int n = 2000;
  auto run = [&](const char* name, auto kern, int stride, int lanes, int nl) {
    kern<<<1, 32>>>(t, d, n, stride, lanes); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
    printf("%-60s %7.1f cycles per iteration, %5.2f per load\n", name, (double)h / n, (double)h / n / nl);
  };
  run("8  x LDS.64, 32 lanes, stride 1", lds_rate<8, 1>, 1, 32, 8);
  run("32 x LDS.64, 32 lanes, stride 1", lds_rate<32, 1>, 1, 32, 32);
  run("32 x LDS.64,  7 lanes, stride 31", lds_rate<32, 1>, 31, 7, 32);
  run("32 x LDS.64, 32 lanes, stride 31", lds_rate<32, 1>, 31, 32, 32);
  run("32 x LDS.64, 32 lanes, broadcast (stride 0)", lds_rate<32, 1>, 0, 32, 32);
  run("16 x LDS.128, 32 lanes, stride 2", lds_rate<16, 2>, 2, 32, 16);
  run("16 x LDS.128, 32 lanes, broadcast", lds_rate<16, 2>, 0, 32, 16);
  run("16 x LDS.128,  7 lanes, broadcast", lds_rate<16, 2>, 0, 7, 16);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
