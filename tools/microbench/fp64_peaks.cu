// FP64 pipe microbenchmarks for B200 (sm_100a): DMMA m8n8k4 rate, DFMA rate, a mixed
// DMMA+DFMA loop (are the two pipes additive?), and an L2-resident read bandwidth probe.
// These set the roofline denominators quoted in DESIGN.md / bench.py (MEASURED_PEAKS.json
// has no FP64 entry). Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peaks fp64_peaks.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int NACC>
__global__ void k_dmma(double* out, const double* in, int iters) {
  double a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)];
  double c[NACC][2];
#pragma unroll
  for (int j = 0; j < NACC; j++) { c[j][0] = 0; c[j][1] = 0; }
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < NACC; j++) dmma(c[j][0], c[j][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < NACC; j++) s += c[j][0] + c[j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void k_dfma(double* out, const double* in, int iters) {
  double a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)];
  double c[NACC];
#pragma unroll
  for (int j = 0; j < NACC; j++) c[j] = j;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < NACC; j++) c[j] = fma(a, c[j], b);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < NACC; j++) s += c[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 8 DMMA + NF DFMA per iteration
template <int NF>
__global__ void k_mixed(double* out, const double* in, int iters) {
  double a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)];
  double c[8][2]; double f[NF > 0 ? NF : 1];
#pragma unroll
  for (int j = 0; j < 8; j++) { c[j][0] = 0; c[j][1] = 0; }
#pragma unroll
  for (int j = 0; j < NF; j++) f[j] = j;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
      dmma(c[j][0], c[j][1], a, b);
#pragma unroll
      for (int q = 0; q < NF / 8; q++) f[j * (NF / 8) + q] = fma(a, f[j * (NF / 8) + q], b);
    }
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
#pragma unroll
  for (int j = 0; j < NF; j++) s += f[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// L2-resident streaming read: every CTA sweeps the same `bytes` window.
__global__ void k_l2read(double* out, const double2* in, size_t n16, int reps) {
  double s = 0;
  for (int r = 0; r < reps; r++) {
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
      double2 v = __ldg(in + ((i + (size_t)r * 977) % n16));
      s += v.x + v.y;
    }
  }
  if (s == 1.2345) out[0] = s;
}

template <typename F>
float timeit(F f, int reps = 3) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, sms, p.clockRate);
  double *in, *out; CK(cudaMalloc(&in, 64 * 8)); CK(cudaMalloc(&out, (size_t)sms * 32 * 1024 * 8));
  double h[64]; for (int i = 0; i < 64; i++) h[i] = 1e-3 * (i + 1);
  CK(cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice));
  const int iters = 20000;
  for (int wps : {4, 8, 16, 32}) {
    int threads = 256, ctas = sms * wps * 32 / threads;
    float ms = timeit([&] { k_dmma<8><<<ctas, threads>>>(out, in, iters); });
    double fl = (double)ctas * (threads / 32) * iters * 8.0 * 512.0;
    printf("{\"bench\": \"dmma_m8n8k4\", \"warps_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", wps, ms, fl / ms * 1e-9);
    ms = timeit([&] { k_dfma<16><<<ctas, threads>>>(out, in, iters); });
    fl = (double)ctas * threads * iters * 16.0 * 2.0;
    printf("{\"bench\": \"dfma\", \"warps_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", wps, ms, fl / ms * 1e-9);
  }
  {
    int wps = 16, threads = 256, ctas = sms * wps * 32 / threads;
    float ms = timeit([&] { k_dmma<2><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"dmma_2acc_latency_bound\", \"ms\": %.3f, \"tflops\": %.2f}\n", ms, (double)ctas * 8 * iters * 2.0 * 512.0 / ms * 1e-9);
    ms = timeit([&] { k_mixed<0><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"mixed_8dmma_0dfma\", \"ms\": %.3f}\n", ms);
    ms = timeit([&] { k_mixed<8><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"mixed_8dmma_8dfma\", \"ms\": %.3f}\n", ms);
    ms = timeit([&] { k_mixed<16><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"mixed_8dmma_16dfma\", \"ms\": %.3f}\n", ms);
    ms = timeit([&] { k_mixed<32><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"mixed_8dmma_32dfma\", \"ms\": %.3f}\n", ms);
    ms = timeit([&] { k_mixed<64><<<ctas, threads>>>(out, in, iters); });
    printf("{\"bench\": \"mixed_8dmma_64dfma\", \"ms\": %.3f}\n", ms);
  }
  for (size_t mb : {16, 32, 64, 512}) {
    size_t bytes = mb << 20; double2* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 0, bytes));
    int reps = mb >= 512 ? 4 : 40;
    float ms = timeit([&] { k_l2read<<<sms * 4, 512>>>(out, buf, bytes / 16, reps); });
    printf("{\"bench\": \"read_window\", \"mb\": %zu, \"ms\": %.3f, \"gbs\": %.1f}\n", mb, ms, (double)bytes * reps / ms * 1e-6);
    CK(cudaFree(buf));
  }
  return 0;
}
