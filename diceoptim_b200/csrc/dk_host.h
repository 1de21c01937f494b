// dk_host.h — host-side helpers shared by the translation units of libdk_b200.so (dk_api*.cu).
// Each kernel family has its own .cu so that the library builds in parallel; the helpers that touch the
// handle (allocation accounting, staging buffers, the forward / gradient launches) live in dk_api.cu.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "dk_common.cuh"

namespace dk {

// cudaMalloc'ed scratch of one call, freed on every exit path
struct DevBuf {
  double* p = nullptr;
  cudaError_t alloc(size_t doubles) { return cudaMalloc(&p, doubles * sizeof(double)); }
  ~DevBuf() { if (p) cudaFree(p); }
};

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

template <class T>
int dev_alloc(dk_model* m, T** ptr, size_t count) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, count * sizeof(T));
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(DK_ERR_NOMEM, "cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e));
  }
  m->device_bytes += count * sizeof(T);
  *ptr = static_cast<T*>(p);
  return DK_OK;
}
template <class T>
void dev_release(dk_model* m, T** ptr, size_t count) {
  if (*ptr) {
    cudaFree(*ptr);
    m->device_bytes -= count * sizeof(T);
    *ptr = nullptr;
  }
}

int ensure_io(dk_model* m, size_t doubles);      // grow-only staging of the host entry points
int ensure_red(dk_model* m, size_t bytes);       // top-k workspace
int ensure_bwd(dk_model* m);                     // backward (flipped T) tiles + W / covariance panels
int check_call(const dk_model* m, const void* x, int64_t N);
// forward pass / forward + backward pass over device-resident candidates
int launch_predict(dk_model* m, const CritParams& P, const double* d_x, int64_t N, double* d_main, double* d_mean,
                   double* d_sd, cudaStream_t st);
int launch_grad(dk_model* m, const CritParams& P, const double* d_x, int64_t N, double* d_grad, double* d_main,
                cudaStream_t st);

// host-pointer entry points: H2D, launch, D2H on the handle's stream, then synchronise
struct HostCall {
  dk_model* m;
  double* d_x;
  double* d_out;  // scratch for outputs after x
  int init(dk_model* mm, const double* x, int64_t N, size_t out_doubles) {
    m = mm;
    const size_t xin = (size_t)N * m->d;
    int rc = ensure_io(m, xin + out_doubles + 16);
    if (rc) return rc;
    d_x = m->d_io;
    d_out = m->d_io + ((xin + 1) & ~(size_t)1);
    if (N > 0) DK_CUDA_TRY(cudaMemcpyAsync(d_x, x, xin * 8, cudaMemcpyHostToDevice, m->stream));
    return DK_OK;
  }
  int fetch(double* host, const double* dev, size_t count) {
    if (host && count) DK_CUDA_TRY(cudaMemcpyAsync(host, dev, count * 8, cudaMemcpyDeviceToHost, m->stream));
    return DK_OK;
  }
  int finish() {
    DK_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return DK_OK;
  }
};

}  // namespace dk
