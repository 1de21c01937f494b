// DMMA.8x8x4 (mma.sync.m8n8k4.f64) dependent-issue latency and per-SM-sub-partition throughput on sm_100a.
// One CTA on one SM; W warps (1, 4, 8, 16), each running K independent accumulator chains (1 .. 16): cycles per DMMA per warp.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_latency dmma_latency.cu ; run: ./dmma_latency
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

template <int K>
__global__ void k(long long* out, double a0, double b0, int iters) {
  double acc[K][2];
#pragma unroll
  for (int i = 0; i < K; i++) acc[i][0] = acc[i][1] = threadIdx.x * 1e-9 + i;
  const double a = a0 + threadIdx.x * 1e-12, b = b0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < K; i++) dmma(acc[i], a, b);
  }
  const long long t1 = clock64();
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < K; i++) s += acc[i][0] + acc[i][1];
  if (s == 12345.678) out[1] = 1;
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

template <int K>
void run(int warps, long long* d) {
  const int iters = 2000;
  k<K><<<1, 32 * warps>>>(d, 1.0000001, 0.9999999, iters);
  k<K><<<1, 32 * warps>>>(d, 1.0000001, 0.9999999, iters);
  long long h = 0;
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("warps %2d chains %2d : %.1f cycles per DMMA per warp, %.1f cycles per DMMA per SM sub-partition\n", warps, K,
         (double)h / (iters * K), (double)h / (iters * K) / ((warps + 3) / 4));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  for (int w : {1, 4, 8, 16}) {
    run<1>(w, d); run<2>(w, d); run<4>(w, d); run<8>(w, d); run<16>(w, d);
  }
  return 0;
}
