// Micro-benchmark (not part of the product): what a hand-over between the two CTAs of a thread-block cluster costs on B200,
// the number a "one CTA pair per QP" ADMM design would pay at every point where the two stage chains
// meet.  (a) remote DSMEM store of 16 doubles + barrier.cluster arrive.release / wait.acquire + local load + DADD, ping-pong
// between the two CTAs (one round trip = two hand-overs); (b) a dependent remote DSMEM load (ld.shared::cluster) chain;
// (c) for comparison the same ping-pong between two warps of ONE CTA through shared memory and __syncthreads.
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ void cl_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

__global__ void __cluster_dims__(2, 1, 1) pingpong_cluster(long long* t, double* out, int n) {
  __shared__ double s[32];
  cg::cluster_group cl = cg::this_cluster();
  const unsigned rank = cl.block_rank();
  double* remote = cl.map_shared_rank(s, rank ^ 1);
  const int lane = threadIdx.x & 31;
  s[lane] = 0.0;
  cl.sync();
  double v = 1.0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    // phase A: CTA 0 hands v to CTA 1; phase B: CTA 1 hands it back.  Both CTAs pass both barriers (aligned barriers).
    if (rank == 0 && lane < 16) remote[lane] = v;
    cl_arrive(); cl_wait();
    if (rank == 1) { v = s[lane & 15] + 1.0; if (lane < 16) remote[lane] = v; }
    cl_arrive(); cl_wait();
    if (rank == 0) v = s[lane & 15] + 1.0;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && rank == 0) t[0] = t1 - t0;
  out[blockIdx.x * 32 + lane] = v;
}

__global__ void __cluster_dims__(2, 1, 1) remote_load_chain(long long* t, double* out, int n) {
  __shared__ int s[64];
  cg::cluster_group cl = cg::this_cluster();
  const unsigned rank = cl.block_rank();
  for (int i = threadIdx.x; i < 64; i += blockDim.x) s[i] = (i * 7 + 3) & 63;
  cl.sync();
  const int* remote = cl.map_shared_rank(s, rank ^ 1);
  int idx = threadIdx.x & 63;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) idx = remote[idx];
  long long t1 = clock64();
  if (threadIdx.x == 0 && rank == 0) t[1] = t1 - t0;
  out[64 + blockIdx.x * 32 + threadIdx.x] = idx;
  cl.sync();
}

__global__ void pingpong_cta(long long* t, double* out, int n, int nwarps_active) {
  __shared__ double s[64];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 64) s[threadIdx.x] = 0.0;
  __syncthreads();
  double v = 1.0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    if (w == 0 && lane < 16) s[lane] = v;
    __syncthreads();
    if (w == 1) { v = s[lane & 15] + 1.0; if (lane < 16) s[32 + lane] = v; }
    __syncthreads();
    if (w == 0) v = s[32 + (lane & 15)] + 1.0;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[2] = t1 - t0;
  out[128 + threadIdx.x] = v;
}

int main() {
  long long* t; double* out;
  cudaMalloc(&t, 64); cudaMalloc(&out, 8 * 4096);
  const int n = 20000;
  long long h[3];
  for (int threads : {32, 128, 256}) {
    pingpong_cluster<<<2, threads>>>(t, out, n);
    remote_load_chain<<<2, 32>>>(t, out, n);
    pingpong_cta<<<1, threads < 64 ? 64 : threads>>>(t, out, n, 2);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, t, 24, cudaMemcpyDeviceToHost);
    printf("threads/CTA %3d: cluster hand-over (DSMEM store + barrier.cluster + load + DADD) %7.1f cycles | "
           "dependent remote DSMEM load %6.1f cycles | same hand-over inside one CTA (__syncthreads) %6.1f cycles\n",
           threads, h[0] / (2.0 * n), h[1] / (double)n, h[2] / (2.0 * n));
  }
  return 0;
}
