// dk_api_cst.cu — C ABI entry points of the constrained criteria (include/dk_b200.h): dk_check_*,
// dk_feasible_improvement, dk_crit_sur_cst.
#include "dk_host.h"
#include "dk_check.cuh"
#include "dk_sur.cuh"

using namespace dk;


// ------------------------------------------------------------------------------------------
// Constrained criteria: Feas (R/crit_EFI.R:231-238), checkPredict (R/checkPredict.R:18-75),
// crit_EFI (R/crit_EFI.R:125-228) and prob.of.feasible.improvement (R/prob.of.feasible.improvement.R:1-81)
// over a batch of candidates and several km models (objective + constraints) on one device.
// ------------------------------------------------------------------------------------------
namespace {
using namespace dk;
bool g_covdist_attr_done[64] = {};

int ensure_delta(dk_model* m) {
  if (m->dm.akg_delta) return DK_OK;
  std::vector<double> delta((size_t)m->n, 0.0);
  if (m->h_nv) std::copy(m->h_nv, m->h_nv + m->n, delta.begin());
  int rc = dev_alloc(m, &m->d_delta, (size_t)m->n);
  if (rc) return rc;
  DK_CUDA_TRY(cudaMemcpy(m->d_delta, delta.data(), (size_t)m->n * 8, cudaMemcpyHostToDevice));
  m->dm.akg_delta = m->d_delta;
  return DK_OK;
}

// kn_yy = sd(X_j)^2 by the predict kernel over the design itself, once per (model, type).  Uses the model's host
// staging buffer: call it before a HostCall on the same model stages the caller's candidates there.
int ensure_knyy(dk_model* m, int type, cudaStream_t st) {
  const int t = type == DK_UK ? 1 : 0;
  if (m->d_knyy[t]) return DK_OK;
  const int n = m->n, d = m->d;
  DevBuf buf;  // handed to the model only when filled (freed on every error path)
  if (buf.alloc((size_t)n) != cudaSuccess) { cudaGetLastError(); return fail(DK_ERR_NOMEM, "cudaMalloc of %zu bytes failed", (size_t)n * 8); }
  int rc = DK_OK;
  // candidates are n x d column-major; d_Xrm is row-major: transpose a host copy (small, once)
  std::vector<double> Xrm((size_t)n * d), Xcm((size_t)n * d);
  DK_CUDA_TRY(cudaMemcpy(Xrm.data(), m->d_Xrm, Xrm.size() * 8, cudaMemcpyDeviceToHost));
  for (int i = 0; i < n; i++)
    for (int k = 0; k < d; k++) Xcm[i + (size_t)k * n] = Xrm[(size_t)i * d + k];
  if ((rc = ensure_io(m, (size_t)n * d + 2 * (size_t)n + 16))) return rc;
  double* d_x = m->d_io;
  double* d_sd = m->d_io + (size_t)n * d;
  DK_CUDA_TRY(cudaMemcpyAsync(d_x, Xcm.data(), Xcm.size() * 8, cudaMemcpyHostToDevice, st));
  CritParams P{MODE_PREDICT, type, 1, 0, 0.0, 0.0};
  if ((rc = launch_predict(m, P, d_x, n, nullptr, nullptr, d_sd, st))) return rc;
  std::vector<double> sd((size_t)n);
  DK_CUDA_TRY(cudaMemcpyAsync(sd.data(), d_sd, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaStreamSynchronize(st));
  for (auto& v : sd) v = v * v;
  DK_CUDA_TRY(cudaMemcpy(buf.p, sd.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
  m->d_knyy[t] = buf.p;
  m->device_bytes += (size_t)n * 8;
  buf.p = nullptr;
  return DK_OK;
}
}  // namespace

extern "C" int dk_feas_dev(dk_model* m, const double* d_x, int64_t N, int type, int equality, double threshold,
                           int multiply, double* d_out, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_FEAS, type, 1, equality ? 1 : 0, threshold, 0.0, 0.0, m->dm.sd2};
  P.combine = multiply ? COMBINE_MUL : COMBINE_STORE;
  return launch_predict(m, P, d_x, N, d_out, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int dk_check_distance_dev(dk_model* m, const double* d_x, int64_t N, int type, int distance, int running_min,
                                     double* d_out, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  if (N == 0) return DK_OK;
  DeviceGuard guard(m->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (distance == DK_DIST_COVRATIO) {
    CritParams P{MODE_COVRATIO, type, 1, 0, 0.0, 0.0, 0.0, m->dm.sd2};
    P.combine = running_min ? COMBINE_MIN : COMBINE_STORE;
    return launch_predict(m, P, d_x, N, d_out, nullptr, nullptr, st);
  }
  if (distance == DK_DIST_EUCLIDEAN) {
    if (running_min) return fail(DK_ERR_ARG, "the euclidean distance uses the first model only (R/checkPredict.R:35)");
    k_euclid_min<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(m->dm, d_x, N, d_out);
    m->launches++;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  if (distance != DK_DIST_COVDIST) return fail(DK_ERR_ARG, "distance = %d is not one of euclidean / covdist / covratio", distance);
  if (m->device < 64 && !g_covdist_attr_done[m->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_covdist, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    g_covdist_attr_done[m->device] = true;
  }
  if ((rc = ensure_bwd(m))) return rc;
  if ((rc = ensure_delta(m))) return rc;
  if ((rc = ensure_knyy(m, type, st))) return rc;
  const int64_t ntiles = (N + BN - 1) / BN;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  k_covdist<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, d_x, N, ntiles, m->d_vscratch, m->d_wscratch,
                                                   m->scratch_doubles_per_cta, m->d_knyy[type == DK_UK ? 1 : 0],
                                                   running_min ? 1 : 0, d_out);
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

namespace {
int check_models(dk_model* const* models, int n_models, const char* what) {
  if (!models || n_models < 1) return fail(DK_ERR_ARG, "%s: no models", what);
  for (int i = 0; i < n_models; i++) {
    if (!models[i]) return fail(DK_ERR_ARG, "%s: model %d is NULL", what, i);
    if (models[i]->device != models[0]->device)
      return fail(DK_ERR_ARG, "%s: model %d lives on device %d, model 0 on device %d", what, i, models[i]->device, models[0]->device);
    if (models[i]->d != models[0]->d)
      return fail(DK_ERR_ARG, "%s: model %d has d = %d, model 0 has d = %d", what, i, models[i]->d, models[0]->d);
  }
  return DK_OK;
}
}  // namespace

// checkPredict over a list of models: out[i] = min over the models of the chosen distance (R/checkPredict.R:40-66);
// the caller compares with its threshold.  x is uploaded once (staging of models[0]) and shared by every model.
extern "C" int dk_check_predict(dk_model* const* models, int n_models, const double* x, int64_t N, int type, int distance,
                                double* out_mindist) {
  int rc = check_models(models, n_models, "dk_check_predict");
  if (rc) return rc;
  dk_model* m0 = models[0];
  if ((rc = check_call(m0, x, N))) return rc;
  if (!out_mindist) return fail(DK_ERR_ARG, "out_mindist is NULL");
  DeviceGuard guard(m0->device);
  if (distance == DK_DIST_COVDIST)
    for (int i = 0; i < n_models; i++) {
      if ((rc = ensure_bwd(models[i])) || (rc = ensure_delta(models[i])) || (rc = ensure_knyy(models[i], type, m0->stream)))
        return rc;
    }
  HostCall h;
  if ((rc = h.init(m0, x, N, (size_t)N))) return rc;
  const int used = distance == DK_DIST_EUCLIDEAN ? 1 : n_models;
  for (int i = 0; i < used; i++)
    if ((rc = dk_check_distance_dev(models[i], h.d_x, N, type, distance, i > 0, h.d_out, m0->stream))) return rc;
  if ((rc = h.fetch(out_mindist, h.d_out, N))) return rc;
  return h.finish();
}

// kind = DK_FI_EXPECTED: crit_EFI's product EI(x; fun, plugin) * prod_i Feas_i(x) (R/crit_EFI.R:158-191; constraints
// predicted with type "UK" whatever `type` says, :185-189; plugin = +Inf -> the EI factor is 1, :160-161).
// kind = DK_FI_PROBABILITY: prob.of.feasible.improvement, pnorm((plugin - mean) / sd) * prod_i Feas_i
// (R/prob.of.feasible.improvement.R:39-79; constraints predicted with `type`).
extern "C" int dk_feasible_improvement(dk_model* fun, dk_model* const* cst, int n_cst, const double* x, int64_t N,
                                       double plugin, int type, int kind, const int* equality, const double* tol,
                                       double* out) {
  if (!fun) return fail(DK_ERR_ARG, "objective model is NULL");
  if (n_cst < 0 || (n_cst > 0 && !cst)) return fail(DK_ERR_ARG, "constraint models missing");
  std::vector<dk_model*> all;
  all.push_back(fun);
  for (int i = 0; i < n_cst; i++) all.push_back(cst[i]);
  int rc = check_models(all.data(), (int)all.size(), "dk_feasible_improvement");
  if (rc) return rc;
  if ((rc = check_call(fun, x, N))) return rc;
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  if (kind != DK_FI_EXPECTED && kind != DK_FI_PROBABILITY) return fail(DK_ERR_ARG, "kind = %d", kind);
  DeviceGuard guard(fun->device);
  HostCall h;
  if ((rc = h.init(fun, x, N, (size_t)N))) return rc;
  cudaStream_t st = fun->stream;
  int started = 0;
  if (plugin == INFINITY) {
    // no feasible observation yet: the improvement factor is 1 and the product starts with the first constraint
  } else if (kind == DK_FI_EXPECTED) {
    if ((rc = dk_ei_dev(fun, h.d_x, N, plugin, type, 1, 0, h.d_out, nullptr, nullptr, st))) return rc;
    started = 1;
  } else {
    if ((rc = dk_feas_dev(fun, h.d_x, N, type, 0, plugin, 0, h.d_out, st))) return rc;
    started = 1;
  }
  for (int i = 0; i < n_cst; i++) {
    const int eq = equality ? equality[i] : 0;
    const double thr = tol ? tol[i] : 0.0;
    if ((rc = dk_feas_dev(cst[i], h.d_x, N, kind == DK_FI_EXPECTED ? DK_UK : type, eq, thr, started, h.d_out, st))) return rc;
    started = 1;
  }
  if (!started) {
    for (int64_t i = 0; i < N; i++) out[i] = 1.0;
    return h.finish();
  }
  if ((rc = h.fetch(out, h.d_out, N))) return rc;
  return h.finish();
}

// ------------------------------------------------------------------------------------------
// crit_SUR_cst (R/crit_SUR_cst.R:120-378) for every row of x, with precomputeUpdateData and computeQuickKrigcov2
// (SURVEY 8(f) rank 4).  Stateless: the integration-point pass (two solves per model over M points) is repeated per call;
// it costs M / N of the candidate pass.
// ------------------------------------------------------------------------------------------
namespace {
bool g_sur_attr_done[64] = {};
struct SurModelBufs {
  DevBuf mn, sn, first, W, mk, sk, U, Cnew, kn;
};
}  // namespace

extern "C" int dk_crit_sur_cst(dk_model* fun, dk_model* const* cst, int n_cst, const double* x, int64_t N,
                               const double* integration_points, int64_t Mi, const double* integration_weights,
                               double obj_min, int type, const int* equality, const double* tol, double* out) {
  if (!fun) return fail(DK_ERR_ARG, "objective model is NULL");
  if (n_cst < 1 || n_cst > 3 || !cst) return fail(DK_ERR_ARG, "crit_SUR_cst takes 1 to 3 constraint models, got %d", n_cst);
  std::vector<dk_model*> all;
  all.push_back(fun);
  for (int i = 0; i < n_cst; i++) all.push_back(cst[i]);
  int rc = check_models(all.data(), (int)all.size(), "dk_crit_sur_cst");
  if (rc) return rc;
  if ((rc = check_call(fun, x, N))) return rc;
  if (!integration_points || Mi < 1) return fail(DK_ERR_ARG, "integration points missing");
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  if (N == 0) return DK_OK;
  DeviceGuard guard(fun->device);
  const int d = fun->d, nm = (int)all.size();
  cudaStream_t st = fun->stream;
  if (fun->device < 64 && !g_sur_attr_done[fun->device]) {
    DK_CUDA_TRY(cudaFuncSetAttribute(k_sur_points, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    DK_CUDA_TRY(cudaFuncSetAttribute(k_sur_cand, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
    g_sur_attr_done[fun->device] = true;
  }
  for (dk_model* m : all)
    if ((rc = ensure_bwd(m))) return rc;
  // candidate chunk: all per-candidate buffers of all models within ~1.5 GB
  size_t per_cand = 0;
  for (dk_model* m : all) per_cand += ((size_t)m->n + (size_t)Mi + m->p + 2) * 8;
  int64_t NC = (int64_t)std::max<size_t>(64, (size_t)1500000000 / per_cand);
  NC = std::min<int64_t>(NC, (int64_t)1 << 21);   // k_sur_kn's grid.y = NC / 64 stays far below 65535
  NC = std::min<int64_t>(N, NC & ~(int64_t)63);
  if (NC < 1) NC = std::min<int64_t>(N, 64);

  DevBuf d_pts, d_w, d_x, d_out;
  std::vector<SurModelBufs> B(nm);
#define SUR_ALLOC(buf, count)                                                                              \
  if ((buf).alloc(count) != cudaSuccess) { cudaGetLastError(); return fail(DK_ERR_NOMEM, "cudaMalloc of %zu doubles failed", (size_t)(count)); }
  SUR_ALLOC(d_pts, (size_t)Mi * d);
  SUR_ALLOC(d_w, (size_t)Mi);
  SUR_ALLOC(d_x, (size_t)NC * d);
  SUR_ALLOC(d_out, (size_t)NC);
  for (int k = 0; k < nm; k++) {
    dk_model* m = all[k];
    SUR_ALLOC(B[k].mn, (size_t)Mi); SUR_ALLOC(B[k].sn, (size_t)Mi); SUR_ALLOC(B[k].first, (size_t)Mi * m->p);
    SUR_ALLOC(B[k].W, (size_t)m->n * Mi);
    SUR_ALLOC(B[k].mk, (size_t)NC); SUR_ALLOC(B[k].sk, (size_t)NC); SUR_ALLOC(B[k].U, (size_t)NC * m->p);
    SUR_ALLOC(B[k].Cnew, (size_t)m->n * NC); SUR_ALLOC(B[k].kn, (size_t)NC * Mi);
  }
#undef SUR_ALLOC
  std::vector<double> hw((size_t)Mi, 1.0 / (double)Mi);   // rep(1 / n.integration.points, n.integration.points) (:150)
  if (integration_weights) std::copy(integration_weights, integration_weights + Mi, hw.begin());
  DK_CUDA_TRY(cudaMemcpyAsync(d_pts.p, integration_points, (size_t)Mi * d * 8, cudaMemcpyHostToDevice, st));
  DK_CUDA_TRY(cudaMemcpyAsync(d_w.p, hw.data(), (size_t)Mi * 8, cudaMemcpyHostToDevice, st));
  const int64_t ptiles = (Mi + BN - 1) / BN;
  for (int k = 0; k < nm; k++) {
    dk_model* m = all[k];
    const int grid = (int)std::min<int64_t>(ptiles, m->grid);
    k_sur_points<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, d_pts.p, Mi, ptiles, m->d_vscratch, m->d_wscratch,
                                                        m->scratch_doubles_per_cta, B[k].mn.p, B[k].sn.p, B[k].first.p, B[k].W.p);
    m->launches++;
  }
  DK_CUDA_TRY(cudaGetLastError());
  std::vector<double> hx((size_t)NC * d);
  for (int64_t c0 = 0; c0 < N; c0 += NC) {
    const int64_t nc = std::min<int64_t>(NC, N - c0);
    for (int k = 0; k < d; k++) std::copy(x + c0 + (int64_t)k * N, x + c0 + (int64_t)k * N + nc, hx.begin() + (size_t)k * nc);
    DK_CUDA_TRY(cudaMemcpyAsync(d_x.p, hx.data(), (size_t)nc * d * 8, cudaMemcpyHostToDevice, st));
    const int64_t ctiles = (nc + BN - 1) / BN;
    SurParams S{};
    S.nc = n_cst; S.any_eq = 0; S.obj_min = obj_min; S.w = d_w.p; S.Mi = Mi;
    for (int i = 0; i < n_cst; i++) S.any_eq |= (equality && equality[i]) ? 1 : 0;
    for (int k = 0; k < nm; k++) {
      dk_model* m = all[k];
      const int grid = (int)std::min<int64_t>(ctiles, m->grid);
      k_sur_cand<<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, type, d_x.p, nc, ctiles, m->d_vscratch, m->d_cscratch,
                                                        m->scratch_doubles_per_cta, B[k].mk.p, B[k].sk.p, B[k].U.p, B[k].Cnew.p);
      dim3 g((unsigned)((Mi + 63) / 64), (unsigned)((nc + 63) / 64));
      k_sur_kn<<<g, 256, 0, st>>>(m->dm, d_pts.p, Mi, d_x.p, nc, B[k].W.p, B[k].Cnew.p, B[k].first.p, B[k].U.p, B[k].kn.p);
      m->launches += 2;
      SurModel sm{B[k].mn.p, B[k].sn.p, B[k].mk.p, B[k].sk.p, B[k].kn.p, std::sqrt(m->dm.sd2) / 1e4,
                  (k > 0 && tol) ? tol[k - 1] : 0.0, (k > 0 && equality) ? equality[k - 1] : 0};
      if (k == 0) S.obj = sm; else S.cst[k - 1] = sm;
    }
    k_sur_reduce<<<(unsigned)nc, 128, 0, st>>>(S, nc, d_out.p);
    fun->launches++;
    DK_CUDA_TRY(cudaGetLastError());
    DK_CUDA_TRY(cudaMemcpyAsync(out + c0, d_out.p, (size_t)nc * 8, cudaMemcpyDeviceToHost, st));
    DK_CUDA_TRY(cudaStreamSynchronize(st));   // hx is reused by the next chunk
  }
  return DK_OK;
}

