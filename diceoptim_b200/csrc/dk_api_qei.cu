// dk_api_qei.cu — C ABI entry points of the qEI family (include/dk_b200.h): dk_qei*, dk_qei_grad*, dk_mvn_cdf.
#include "dk_host.h"
#include "dk_qei_grad_gen.cuh"

using namespace dk;


// ------------------------------------------------------------------------------------------
// qEI (R/qEI.R:119-261): batches of q <= DK_MAX_Q points
// ------------------------------------------------------------------------------------------

namespace {
bool g_qei_attr_done[64] = {};

int ensure_mom(dk_model* m, size_t need);

int ensure_qei(dk_model* m, int64_t B) {
  if (m->device < 64 && !g_qei_attr_done[m->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_qei_moments, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    g_qei_attr_done[m->device] = true;
  }
  if (!m->d_aux) {
    int rc = dev_alloc(m, &m->d_aux, (size_t)m->grid * QEI_AUX_DOUBLES);
    if (rc) return rc;
  }
  return ensure_mom(m, (size_t)std::max<int64_t>(B, 1) * MOM_STRIDE);
}

int ensure_mom(dk_model* m, size_t need) {
  if (m->mom_doubles < need) {
    if (m->d_mom) { cudaFree(m->d_mom); m->device_bytes -= m->mom_doubles * 8; m->d_mom = nullptr; m->mom_doubles = 0; }
    int rc = dev_alloc(m, &m->d_mom, need);
    if (rc) return rc;
    m->mom_doubles = need;
  }
  return DK_OK;
}

}  // namespace

extern "C" int dk_qei_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization,
                          double* d_out, void* stream) {
  int rc = check_call(m, d_x, B);
  if (rc) return rc;
  if (q < 1 || q > DK_MAX_Q_LATTICE) return fail(DK_ERR_ARG, "qEI batch size q=%d outside 1..%d", q, DK_MAX_Q_LATTICE);
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  if (B == 0) return DK_OK;
  DeviceGuard guard(m->device);
  if (q > DK_MAX_Q) {
    // 5 .. 10 points per batch: general-q moments + one CTA per batch for the lattice CDFs (dk_qei_gen.cuh)
    static bool gen_attr_done[64] = {};
    if (m->device < 64 && !gen_attr_done[m->device]) {
      DK_CUDA_TRY(cudaFuncSetAttribute(k_qei_moments_gen, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
      gen_attr_done[m->device] = true;
    }
    if ((rc = ensure_mom(m, (size_t)B * gen_mom_stride(q)))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = BN / q;
    const int64_t ntiles = (B + nb - 1) / nb;
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
    k_qei_moments_gen<<<(int)std::min<int64_t>(ntiles, m->grid), NT, SM_TOTAL + PIPE_SMEM, st>>>(
        m->dm, type, q, d_x, B, ntiles, m->d_vscratch, m->scratch_doubles_per_cta, m->d_mom);
    DK_CUDA_TRY(cudaGetLastError());
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev_mid, st));
    k_qei_cdf_gen<<<(unsigned)std::min<int64_t>(B, 8 * (int64_t)m->num_sms), CDF_THREADS, 0, st>>>(m->d_mom, B, q, plugin,
                                                                                                minimization, d_out);
    if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = true; }
    m->launches += 2;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  if ((rc = ensure_qei(m, B))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t ntiles = (B + QB - 1) / QB;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  k_qei_moments<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, q, d_x, B, ntiles, m->d_vscratch, m->scratch_doubles_per_cta,
                                             m->d_aux, m->d_mom);
  DK_CUDA_TRY(cudaGetLastError());
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev_mid, st));
  static const int64_t warp_below = [] { const char* e = std::getenv("DK_QEI_WARP_BELOW"); return e ? std::atoll(e) : (int64_t)12288; }();  // measured crossover (cfg3): 2048 batches 0.64 vs 2.06 ms, 18944: 5.4 vs 5.0 ms
  if (B < warp_below)  // few batches: a warp per batch (path-integral nodes spread over lanes) hides latency better
    k_qei_cdf_warp<<<(unsigned)((B + 3) / 4), 128, 0, st>>>(m->d_mom, B, q, plugin, minimization, d_out);
  else
    k_qei_cdf<<<(unsigned)((B * 4 + 127) / 128), 128, 0, st>>>(m->d_mom, B, q, plugin, minimization, d_out);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = true; }
  m->launches += 2;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

// Batches [b0, b0 + nb) of the caller's (B q) x d matrix (leading dimension ldx = B q): qEI values into out[b] and / or
// qEI.grad rows into out_grad (leading dimension ldx), through the chunked pipeline (whole tiles of QB batches).
int dk::host_qei(dk_model* m, const double* x, int64_t ldx, int64_t b0, int64_t nb, int q, double plugin, int type,
                 int minimization, double* out, double* out_grad) {
  DeviceGuard guard(m->device);
  const HostOut outs[2] = {{out, 1, ldx, true}, {out_grad, m->d, ldx, false}};
  int rc = run_chunked(m, x, ldx, b0 * q, nb, q, chunk_units_for(m, q, q > DK_MAX_Q ? BN / q : QB), outs, 2,
                       [&](const double* d_x, int64_t rows, double* const* o, int64_t, cudaStream_t st) {
                         int r = DK_OK;
                         if (o[0]) r = dk_qei_dev(m, d_x, rows / q, q, plugin, type, minimization, o[0], st);
                         if (!r && o[1]) r = dk_qei_grad_dev(m, d_x, rows / q, q, plugin, type, minimization, o[1], st);
                         return r;
                       });
  if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); cudaStreamSynchronize(m->copy_out); }
  return rc;
}

extern "C" int dk_qei(dk_model* m, const double* x, int64_t B, int q, double plugin, int type, int minimization,
                      double* out) {
  int rc = check_call(m, x, B);
  if (rc) return rc;
  if (q < 1 || q > DK_MAX_Q_LATTICE) return fail(DK_ERR_ARG, "qEI batch size q=%d outside 1..%d", q, DK_MAX_Q_LATTICE);
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  return host_qei(m, x, B * q, 0, B, q, plugin, type, minimization, out, nullptr);
}

namespace {
// the value CDF kernels of dk_qei_dev on moments that are already in m->d_mom
int launch_qei_value_cdf(dk_model* m, int64_t B, int q, double plugin, int minimization, double* d_out, cudaStream_t st) {
  if (q > DK_MAX_Q) {
    k_qei_cdf_gen<<<(unsigned)std::min<int64_t>(B, 8 * (int64_t)m->num_sms), CDF_THREADS, 0, st>>>(m->d_mom, B, q, plugin,
                                                                                                minimization, d_out);
  } else {
    static const int64_t warp_below = [] { const char* e = std::getenv("DK_QEI_WARP_BELOW"); return e ? std::atoll(e) : (int64_t)12288; }();
    if (B < warp_below) k_qei_cdf_warp<<<(unsigned)((B + 3) / 4), 128, 0, st>>>(m->d_mom, B, q, plugin, minimization, d_out);
    else k_qei_cdf<<<(unsigned)((B * 4 + 127) / 128), 128, 0, st>>>(m->d_mom, B, q, plugin, minimization, d_out);
  }
  m->launches += 1;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

int qei_grad_impl(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization, double* d_val,
                  double* d_grad, void* stream);
}  // namespace

extern "C" int dk_qei_grad_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type,
                               int minimization, double* d_grad, void* stream) {
  return qei_grad_impl(m, d_x, B, q, plugin, type, minimization, nullptr, d_grad, stream);
}

/* qEI and qEI.grad of the same batches from ONE pass over the model: what qEI(x, envir) followed by qEI.grad(x, envir) saves
 * in the reference (R/qEI.R:190-198 stores the kriging mean / covariance and the CDF terms, R/qEI.grad.R:299-316 reads them).
 * The value is taken at x itself, as the gradient is (dk_qei rounds x to 8 digits first, R/qEI.R:126-128). */
extern "C" int dk_qei_value_grad_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type,
                                     int minimization, double* d_val, double* d_grad, void* stream) {
  if (!d_val) return fail(DK_ERR_ARG, "d_val is NULL");
  return qei_grad_impl(m, d_x, B, q, plugin, type, minimization, d_val, d_grad, stream);
}

namespace {
int qei_grad_impl(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization, double* d_val,
                  double* d_grad, void* stream) {
  int rc = check_call(m, d_x, B);
  if (rc) return rc;
  if (q < 2 || q > DK_MAX_Q_LATTICE)
    return fail(DK_ERR_ARG, "qEI.grad batch size q=%d outside 2..%d (use dk_ei_grad for q = 1)", q, DK_MAX_Q_LATTICE);
  if (!d_grad) return fail(DK_ERR_ARG, "d_grad is NULL");
  if (B == 0) return DK_OK;
  DeviceGuard guard(m->device);
  if ((rc = ensure_bwd(m))) return rc;
  if (q > DK_MAX_Q) {
    // 5 .. 10 points per batch: general-q derivatives + one CTA per batch for the gradient of the lattice CDFs (dk_qei_grad_gen.cuh)
    static bool gen_attr_done[64] = {};
    if (m->device < 64 && !gen_attr_done[m->device]) {
      DK_CUDA_TRY(cudaFuncSetAttribute(k_qei_grad_moments_gen, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
      DK_CUDA_TRY(cudaFuncSetAttribute(k_qei_grad_cdf_gen, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GenGradShared)));
      gen_attr_done[m->device] = true;
    }
    if ((rc = ensure_mom(m, (size_t)B * gen_mom_stride(q)))) return rc;
    // derivatives of every batch, then the q summands of every batch's gradient
    const size_t deriv_part = (size_t)B * qei_deriv_stride_gen(q, m->d);
    const size_t need = deriv_part + (size_t)B * q * q * m->d;
    if (m->deriv_doubles < need) {
      if (m->d_deriv) { cudaFree(m->d_deriv); m->device_bytes -= m->deriv_doubles * 8; m->d_deriv = nullptr; m->deriv_doubles = 0; }
      if ((rc = dev_alloc(m, &m->d_deriv, need))) return rc;
      m->deriv_doubles = need;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = BN / q;
    const int64_t ntiles = (B + nb - 1) / nb;
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
    k_qei_grad_moments_gen<<<(int)std::min<int64_t>(ntiles, m->grid), NT, SM_TOTAL + PIPE_SMEM, st>>>(
        m->dm, type, q, d_x, B, ntiles, m->d_vscratch, m->d_wscratch, m->d_cscratch, m->scratch_doubles_per_cta, m->d_mom,
        m->d_deriv);
    DK_CUDA_TRY(cudaGetLastError());
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev_mid, st));
    if (d_val && (rc = launch_qei_value_cdf(m, B, q, plugin, minimization, d_val, st))) return rc;
    double* d_partial = m->d_deriv + deriv_part;
    k_qei_grad_cdf_gen<<<(unsigned)std::min<int64_t>(B * q, 8 * (int64_t)m->num_sms), CDF_THREADS, sizeof(GenGradShared), st>>>(
        m->d_mom, m->d_deriv, B, q, m->d, plugin, minimization, m->dm.sd2, d_partial);
    DK_CUDA_TRY(cudaGetLastError());
    k_qei_grad_sum_gen<<<(unsigned)std::min<int64_t>(B, 8 * (int64_t)m->num_sms), CDF_THREADS, 0, st>>>(d_partial, B, q, m->d, d_grad);
    if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = true; }
    m->launches += 3;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  if ((rc = ensure_qei(m, B))) return rc;
  static bool attr_done[64] = {};
  if (m->device < 64 && !attr_done[m->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_qei_grad_moments, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    attr_done[m->device] = true;
  }
  const size_t need = (size_t)B * qei_deriv_stride(m->d);
  if (m->deriv_doubles < need) {
    if (m->d_deriv) { cudaFree(m->d_deriv); m->device_bytes -= m->deriv_doubles * 8; m->d_deriv = nullptr; m->deriv_doubles = 0; }
    if ((rc = dev_alloc(m, &m->d_deriv, need))) return rc;
    m->deriv_doubles = need;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t ntiles = (B + QB - 1) / QB;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  k_qei_grad_moments<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, q, d_x, B, ntiles, m->d_vscratch, m->d_wscratch,
                                                  m->d_cscratch, m->scratch_doubles_per_cta, m->d_aux, m->d_mom, m->d_deriv);
  DK_CUDA_TRY(cudaGetLastError());
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev_mid, st));
  if (d_val && (rc = launch_qei_value_cdf(m, B, q, plugin, minimization, d_val, st))) return rc;
  const int64_t threads = B * 4;
  k_qei_grad_cdf<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(m->d_mom, m->d_deriv, B, q, m->d, plugin, minimization,
                                                                   m->dm.sd2, d_grad);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = true; }
  m->launches += 2;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}
}  // namespace

extern "C" int dk_qei_grad(dk_model* m, const double* x, int64_t B, int q, double plugin, int type, int minimization,
                           double* out_grad) {
  int rc = check_call(m, x, B);
  if (rc) return rc;
  if (q < 2 || q > DK_MAX_Q_LATTICE)
    return fail(DK_ERR_ARG, "qEI.grad batch size q=%d outside 2..%d (use dk_ei_grad for q = 1)", q, DK_MAX_Q_LATTICE);
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  return host_qei(m, x, B * q, 0, B, q, plugin, type, minimization, nullptr, out_grad);
}

/* Test hook: pmnorm(h_b, varcov = S_b) for B independent q-variate problems (q <= 4), host pointers,
 * device 0.  h is B x q row-major, S is B x q x q row-major. */
extern "C" int dk_mvn_cdf(int q, int64_t B, const double* h, const double* S, double* out) {
  if (q < 1 || q > DK_MAX_Q_LATTICE) return fail(DK_ERR_ARG, "dimension q=%d outside 1..%d", q, DK_MAX_Q_LATTICE);
  if (!h || !S || !out || B < 0) return fail(DK_ERR_ARG, "dk_mvn_cdf: bad argument");
  if (B == 0) return DK_OK;
  DevBuf bh, bS, bo;
  DK_CUDA_TRY(bh.alloc((size_t)B * q));
  DK_CUDA_TRY(bS.alloc((size_t)B * q * q));
  DK_CUDA_TRY(bo.alloc((size_t)B));
  double *d_h = bh.p, *d_S = bS.p, *d_o = bo.p;
  DK_CUDA_TRY(cudaMemcpy(d_h, h, B * q * 8, cudaMemcpyHostToDevice));
  DK_CUDA_TRY(cudaMemcpy(d_S, S, B * q * q * 8, cudaMemcpyHostToDevice));
  if (q > DK_MAX_Q) k_mvn_lattice_test<<<(unsigned)std::min<int64_t>(B, 1184), CDF_THREADS>>>(q, B, d_h, d_S, d_o);
  else k_mvn_cdf_test<<<(unsigned)((B + 63) / 64), 64>>>(q, B, d_h, d_S, d_o);
  DK_CUDA_TRY(cudaGetLastError());
  DK_CUDA_TRY(cudaMemcpy(out, d_o, B * 8, cudaMemcpyDeviceToHost));
  return DK_OK;
}
