// dk_host.h — host-side helpers shared by the translation units of libdk_b200.so (dk_api*.cu).
// Each kernel family has its own .cu so that the library builds in parallel; the helpers that touch the
// handle (allocation accounting, staging buffers, the forward / gradient launches) live in dk_api.cu.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
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

// ---- chunked, double-buffered host <-> device pipeline of the host entry points ----
// The caller's matrix x is column-major with leading dimension ldx (R's N x d: ldx = N); rows [row0, row0 + nunits * unit)
// are processed in chunks of `chunk_units` units (unit = 1 candidate, or q rows of a qEI batch).  Each chunk's rows are
// gathered into pinned staging by the calling thread, copied to the device on the handle's copy-in stream, handed to
// `launch` on the handle's compute stream, and the outputs come back on the copy-out stream: the copies of chunk i + 1
// and i - 1 overlap the kernels of chunk i, the caller's memory may be pageable (an R vector is), device memory is two
// chunks whatever N, and a shard of a larger matrix (multi-GPU split) is just another (row0, nunits).
struct HostOut {
  double* host;    // base of the caller's WHOLE output array (NULL: this output is not wanted)
  int cols;        // columns (1, d, ...)
  int64_t ld;      // leading dimension of the caller's array
  bool per_unit;   // one row per unit (qEI value) instead of one per x row
};
// launch(d_x, rows, d_out[nouts], unit0, stream): d_x is rows x d column-major (ld = rows); d_out[j] is cols_j x count
// column-major (ld = count) or NULL when outs[j].host is NULL; unit0 = index of the chunk's first unit within this call.
using ChunkLaunch = std::function<int(const double*, int64_t, double* const*, int64_t, cudaStream_t)>;
int run_chunked(dk_model* m, const double* x, int64_t ldx, int64_t row0, int64_t nunits, int unit, int64_t chunk_units,
                const HostOut* outs, int nouts, const ChunkLaunch& launch);
// chunk size (in units) for tiles of `tile_units` units: whole waves of the persistent grid, about 24 MB of input
int64_t chunk_units_for(const dk_model* m, int unit, int tile_units);

// Host-pointer forms over rows [row0, row0 + n) of the caller's column-major matrix (leading dimension ldx); outputs
// are the caller's whole arrays, written at the same rows (gradients: n x d with leading dimension ldx).
int host_predict(dk_model* m, const CritParams& P, const double* x, int64_t ldx, int64_t row0, int64_t n,
                 double* out_main, double* out_mean, double* out_sd);
int host_grad(dk_model* m, const CritParams& P, const double* x, int64_t ldx, int64_t row0, int64_t n, double* out_grad,
              double* out_main);
struct ScreenSpec {   // the criterion a screen ranks by
  int criterion;      // DK_CRIT_EI / DK_CRIT_QUANTILE / DK_CRIT_AKG
  CritParams P;       // EI / quantile parameters (type for all three)
  double tau2;        // AKG: new.noise.var
};
int make_screen_spec(const dk_model* m, const dk_crit_spec* c, ScreenSpec* S);
int host_screen(dk_model* m, const ScreenSpec& S, const double* x, int x_on_device, const dk_uniform_spec* gen, int64_t ldx,
                int64_t row0, int64_t n, int k, int64_t* out_idx, double* out_val, cudaStream_t user_stream);
void generated_rows(const dk_uniform_spec* gen, int d, const int64_t* idx, int k, double* out_x);
int host_weights_to_vec(dk_model* m, const ScreenSpec& S, const double* x, int64_t N);
int prepare_akg(dk_model* m, bool with_grad, cudaStream_t st);                    // dk_api_akg.cu
int host_akg(dk_model* m, const double* x, int64_t ldx, int64_t row0, int64_t n, double new_noise_var, int type,
             double* out);                                                       // dk_api_akg.cu
int host_qei(dk_model* m, const double* x, int64_t ldx, int64_t b0, int64_t nb, int q, double plugin, int type,
             int minimization, double* out, double* out_grad);                    // dk_api_qei.cu

// small calls (a few launches on one staged copy of x): H2D, launches, D2H on the handle's stream, then synchronise
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
