// dk_api_akg.cu — C ABI entry points of the AKG family (include/dk_b200.h): dk_akg*, dk_akg_grad*.
#include "dk_host.h"
#include "dk_akg.cuh"

using namespace dk;


// ------------------------------------------------------------------------------------------
// AKG (R/AKG.R:93-229)
// ------------------------------------------------------------------------------------------

namespace {
bool g_akg_attr_done[64] = {};

int ensure_akg(dk_model* m) {
  if (m->device < 64 && !g_akg_attr_done[m->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_akg_fused<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    g_akg_attr_done[m->device] = true;
  }
  if (m->d_akg_const) return DK_OK;
  const int n = m->n;
  if (!m->h_y) return fail(DK_ERR_ARG, "AKG needs model@y (dk_model_desc.y was NULL)");
  std::vector<double> alpha(n), cst(2 * (size_t)n), a(n);
  DK_CUDA_TRY(cudaMemcpy(alpha.data(), m->dm.alpha, (size_t)n * 8, cudaMemcpyDeviceToHost));
  // mk.X = y - delta .* alpha (predict.km at the design, hoisted out of R/AKG.R:121)
  double min_mkX = INFINITY;
  for (int i = 0; i < n; i++) {
    const double delta = m->h_nv ? m->h_nv[i] : 0.0;
    const double mkX = m->h_y[i] - delta * alpha[i];
    a[i] = -mkX;
    cst[n + i] = delta;
    min_mkX = std::min(min_mkX, mkX);
  }
  std::vector<int> perm(n);
  for (int i = 0; i < n; i++) perm[i] = i;
  std::stable_sort(perm.begin(), perm.end(), [&](int u, int v) { return a[u] < a[v]; });
  for (int j = 0; j < n; j++) cst[j] = a[perm[j]];
  // locals first, published together: a failed call leaves the handle unchanged and the next call retries
  double *d_const = nullptr, *d_scr = nullptr;
  int* d_perm = nullptr;
  const size_t scr_count = (size_t)m->grid * akg_scratch_doubles(n);
  int rc = dev_alloc(m, &d_const, cst.size());
  if (!rc) rc = dev_alloc(m, &d_perm, (size_t)n);
  if (!rc) rc = dev_alloc(m, &d_scr, scr_count);
  if (!rc) {
    cudaError_t e = cudaMemcpy(d_const, cst.data(), cst.size() * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_perm, perm.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) rc = fail(DK_ERR_CUDA, "upload of the AKG constants failed: %s", cudaGetErrorString(e));
  }
  if (rc) {
    dev_release(m, &d_const, cst.size());
    dev_release(m, &d_perm, (size_t)n);
    dev_release(m, &d_scr, scr_count);
    return rc;
  }
  m->d_akg_perm = d_perm;
  m->d_akg_scratch = d_scr;
  m->d_akg_const = d_const;
  m->dm.akg_a_sorted = m->d_akg_const;
  m->dm.akg_delta = m->d_akg_const + n;
  m->dm.akg_perm = m->d_akg_perm;
  m->dm.akg_min_mkX = min_mkX;
  return DK_OK;
}
}  // namespace

extern "C" int dk_akg_dev(dk_model* m, const double* d_x, int64_t N, double new_noise_var, int type, double* d_out,
                          void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  if (m->n + 1 > AKG_MAX_SORT) return fail(DK_ERR_ARG, "AKG: n + 1 = %d exceeds the %d keys the in-kernel sort holds", m->n + 1, AKG_MAX_SORT);
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  if (N == 0) return DK_OK;
  DeviceGuard guard(m->device);
  if ((rc = ensure_bwd(m))) return rc;
  if ((rc = ensure_akg(m))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t ntiles = (N + BN - 1) / BN;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  k_akg_fused<false><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, new_noise_var, d_x, N, ntiles, m->d_vscratch,
                                                      m->d_wscratch, m->scratch_doubles_per_cta, m->d_akg_scratch, d_out,
                                                      nullptr, nullptr, nullptr, nullptr);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

int dk::host_akg(dk_model* m, const double* x, int64_t ldx, int64_t row0, int64_t n, double new_noise_var, int type,
                 double* out) {
  DeviceGuard guard(m->device);
  const HostOut outs[1] = {{out, 1, ldx, false}};
  int rc = run_chunked(m, x, ldx, row0, n, 1, chunk_units_for(m, 1, BN), outs, 1,
                       [&](const double* d_x, int64_t rows, double* const* o, int64_t, cudaStream_t st) {
                         return dk_akg_dev(m, d_x, rows, new_noise_var, type, o[0], st);
                       });
  if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); cudaStreamSynchronize(m->copy_out); }
  return rc;
}

extern "C" int dk_akg(dk_model* m, const double* x, int64_t N, double new_noise_var, int type, double* out) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  return host_akg(m, x, N, 0, N, new_noise_var, type, out);
}


// ------------------------------------------------------------------------------------------
// AKG.grad (R/AKG.grad.R:65-281)
// ------------------------------------------------------------------------------------------
namespace {
// K^-1 is built on the CALLER's stream (it shares the V / W panels with whatever that stream already holds for this
// handle, and the *_dev contract is "enqueue without synchronising"); the first call still allocates.
int ensure_akg_grad(dk_model* m, cudaStream_t st) {
  static bool attr_done[64] = {};
  if (m->device < 64 && !attr_done[m->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_akg_fused<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    DK_CUDA_TRY(cudaFuncSetAttribute(k_kinv, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    attr_done[m->device] = true;
  }
  int rc;
  if (!m->d_akg_gscratch && (rc = dev_alloc(m, &m->d_akg_gscratch, (size_t)m->grid * akg_grad_scratch_doubles(m->d)))) return rc;
  if (!m->d_Kinv) {
    const size_t n = m->n;
    if ((rc = dev_alloc(m, &m->d_Kinv, n * n))) return rc;
    const int64_t ntiles = ((int64_t)n + BN - 1) / BN;
    const int grid = (int)std::min<int64_t>(ntiles, m->grid);
    k_kinv<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, ntiles, m->d_vscratch, m->d_wscratch,
                                                   m->scratch_doubles_per_cta, m->d_Kinv);
    m->launches++;
    DK_CUDA_TRY(cudaGetLastError());
  }
  return DK_OK;
}
}  // namespace

extern "C" int dk_akg_grad_dev(dk_model* m, const double* d_x, int64_t N, double new_noise_var, int type, double* d_grad,
                               double* d_akg, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_grad) return fail(DK_ERR_ARG, "d_grad is NULL");
  if (m->n + 1 > AKG_MAX_SORT) return fail(DK_ERR_ARG, "AKG: n + 1 = %d exceeds the %d keys the in-kernel sort holds", m->n + 1, AKG_MAX_SORT);
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  if (N == 0) return DK_OK;
  DeviceGuard guard(m->device);
  if ((rc = ensure_bwd(m))) return rc;
  if ((rc = ensure_akg(m))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if ((rc = ensure_akg_grad(m, st))) return rc;
  const int64_t ntiles = (N + BN - 1) / BN;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  k_akg_fused<true><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, new_noise_var, d_x, N, ntiles, m->d_vscratch,
                                                     m->d_wscratch, m->scratch_doubles_per_cta, m->d_akg_scratch, d_akg,
                                                     m->d_cscratch, m->d_Kinv, m->d_akg_gscratch, d_grad);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

extern "C" int dk_akg_grad(dk_model* m, const double* x, int64_t N, double new_noise_var, int type, double* out_grad,
                           double* out_akg) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  DeviceGuard guard(m->device);
  const HostOut outs[2] = {{out_grad, m->d, N, false}, {out_akg, 1, N, false}};
  rc = run_chunked(m, x, N, 0, N, 1, chunk_units_for(m, 1, BN), outs, 2,
                   [&](const double* d_x, int64_t rows, double* const* o, int64_t, cudaStream_t st) {
                     return dk_akg_grad_dev(m, d_x, rows, new_noise_var, type, o[0], o[1], st);
                   });
  if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); cudaStreamSynchronize(m->copy_out); }
  return rc;
}

// the AKG family's lazily built pieces, for dk_model_prepare (dk_api.cu)
int dk::prepare_akg(dk_model* m, bool with_grad, cudaStream_t st) {
  int rc = ensure_bwd(m);
  if (!rc) rc = ensure_akg(m);
  if (!rc && with_grad) rc = ensure_akg_grad(m, st);
  return rc;
}
