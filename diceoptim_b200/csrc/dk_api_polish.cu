// dk_api_polish.cu — the multistart callers' inner loops on the device:
//   dk_polish          batched projected L-BFGS over S starts: the optimiser state (iterates, gradients, the (s, y) history,
//                      step lengths, per-start phase) lives in device memory and the update / line-search kernels consume the
//                      criterion kernels' value + gradient outputs in place.  Replaces the restart-by-restart
//                      optim(method = "L-BFGS-B") loops of /root/reference/R/max_crit.R:114-122, R/max_qEI.R:204-222 and the
//                      BFGS stage of genoud in R/max_EI.R:207-219 (and max_AEI / max_AKG / max_EQI / min_quantile).
//   dk_sample_from_ei  the EI-proportional draw without replacement of R/sampleFromEI.R:119-135 on the device.
//
// One trip of the loop = ONE criterion launch for all starts.  Every start runs its own state machine (new direction ->
// Armijo backtracking along the projection arc -> accept), so starts do not wait for each other's line searches.  Per trip
// the host reads one int (the number of unfinished starts) to know when to stop; nothing else crosses the bus until the end.
// Method (same as diceoptim_b200/optim.py, which remains as the host-side cross-check): variables at a bound whose
// gradient points outward are frozen, two-loop recursion on the free ones, projected step, Armijo c1 = 1e-4, halving;
// a non-descent direction falls back to the projected gradient; stops on the projected-gradient norm (pgtol), on the
// relative decrease (factr x machine epsilon, as optim's L-BFGS-B) or at maxit.
#include "dk_host.h"
#include "dk_reduce.cuh"

using namespace dk;

namespace {

struct PolishState {
  int S, D, q, d, m;        // starts, variables per start (= q d), points per start, input dimension, history length
  int64_t rows;             // S q: rows of the candidate matrices
  double *x, *g, *xt, *gt, *dir, *sh, *yh;   // [rows x d] candidate layout; histories m of them
  double *f, *ft, *t, *rho;                   // [S], rho [m x S]
  int *phase, *iters, *nbt;                   // [S]
  int* remaining;                             // [1]
  double lower[DK_MAX_D], upper[DK_MAX_D];
  double sign;              // f = sign * criterion (sign = -1 maximises)
  double pgtol, factr_eps, c1;
  int maxit, max_backtracks;
};

// component c = (l, k) of start s in the candidate layout (rows x d column-major, rows b q .. b q + q - 1 = start b)
__device__ __forceinline__ int64_t cidx(const PolishState& P, int s, int c) {
  const int l = c % P.q, k = c / P.q;
  return ((int64_t)s * P.q + l) + (int64_t)k * P.rows;
}
__device__ __forceinline__ double clampd(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

// first evaluation (at the clipped starts): f, g <- sign * (value, gradient)
__global__ void k_polish_init(PolishState P) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= P.S) return;
  bool bad = !isfinite(P.ft[s]);
  P.f[s] = P.sign * P.ft[s];
  for (int c = 0; c < P.D; c++) {
    const int64_t i = cidx(P, s, c);
    const double gv = P.sign * P.gt[i];
    if (!isfinite(gv)) bad = true;
    P.g[i] = gv;
    P.x[i] = P.xt[i];
  }
  if (bad)
    for (int c = 0; c < P.D; c++) P.g[cidx(P, s, c)] = 0.0;
  P.phase[s] = bad ? 2 : 0;
  P.iters[s] = 0;
  P.nbt[s] = 0;
  P.t[s] = 1.0;
  if (!bad) atomicAdd(P.remaining, 1);
}

// next trial point of every start
__global__ void k_polish_prepare(PolishState P) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= P.S) return;
  const int D = P.D, m = P.m;
  int phase = P.phase[s];
  if (phase == 0) {
    // ---- new direction: projected gradient, two-loop recursion on the free variables ----
    double pgmax = 0.0, pgnorm2 = 0.0;
    for (int c = 0; c < D; c++) {
      const int64_t i = cidx(P, s, c);
      const int k = c / P.q;
      const double xv = P.x[i], gv = P.g[i];
      const bool frozen = (xv <= P.lower[k] && gv > 0.0) || (xv >= P.upper[k] && gv < 0.0);
      const double pg = frozen ? 0.0 : gv;
      P.dir[i] = pg;   // q of the recursion, for now
      pgmax = fmax(pgmax, fabs(pg));
      pgnorm2 = fma(pg, pg, pgnorm2);
    }
    if (pgmax <= P.pgtol) {
      phase = 2;
    } else {
      const int it = P.iters[s], nh = it < m ? it : m;
      double alpha[16];
      for (int j = 0; j < nh; j++) {           // newest first
        const int slot = (it - 1 - j) % m;
        const double* sj = P.sh + (size_t)slot * P.rows * P.d;
        const double* yj = P.yh + (size_t)slot * P.rows * P.d;
        double a = 0.0;
        for (int c = 0; c < D; c++) { const int64_t i = cidx(P, s, c); a = fma(sj[i], P.dir[i], a); }
        a *= P.rho[(size_t)slot * P.S + s];
        alpha[j] = a;
        for (int c = 0; c < D; c++) { const int64_t i = cidx(P, s, c); P.dir[i] = fma(-a, yj[i], P.dir[i]); }
      }
      double gamma;
      if (nh > 0) {
        const int slot = (it - 1) % m;
        const double* yj = P.yh + (size_t)slot * P.rows * P.d;
        double yy = 0.0;
        for (int c = 0; c < D; c++) { const int64_t i = cidx(P, s, c); yy = fma(yj[i], yj[i], yy); }
        const double r0 = P.rho[(size_t)slot * P.S + s];
        gamma = (r0 > 0.0 && yy > 0.0) ? 1.0 / fmax(r0 * yy, 1e-300) : 1.0;
      } else {
        gamma = 1.0 / fmax(sqrt(pgnorm2), 1.0);   // first step: unit length at most
      }
      for (int c = 0; c < D; c++) P.dir[cidx(P, s, c)] *= gamma;
      for (int j = nh - 1; j >= 0; j--) {        // oldest first
        const int slot = (it - 1 - j) % m;
        const double* sj = P.sh + (size_t)slot * P.rows * P.d;
        const double* yj = P.yh + (size_t)slot * P.rows * P.d;
        double b = 0.0;
        for (int c = 0; c < D; c++) { const int64_t i = cidx(P, s, c); b = fma(yj[i], P.dir[i], b); }
        b *= P.rho[(size_t)slot * P.S + s];
        const double coef = alpha[j] - b;
        for (int c = 0; c < D; c++) { const int64_t i = cidx(P, s, c); P.dir[i] = fma(coef, sj[i], P.dir[i]); }
      }
      // direction = -r on the free variables; not a descent direction -> scaled projected gradient
      double slope = 0.0;
      for (int c = 0; c < D; c++) {
        const int64_t i = cidx(P, s, c);
        const int k = c / P.q;
        const double xv = P.x[i], gv = P.g[i];
        const bool frozen = (xv <= P.lower[k] && gv > 0.0) || (xv >= P.upper[k] && gv < 0.0);
        const double dv = frozen ? 0.0 : -P.dir[i];
        P.dir[i] = dv;
        slope = fma(dv, gv, slope);
      }
      if (!(slope < 0.0)) {
        const double sc = 1.0 / fmax(sqrt(pgnorm2), 1.0);
        for (int c = 0; c < D; c++) {
          const int64_t i = cidx(P, s, c);
          const int k = c / P.q;
          const double xv = P.x[i], gv = P.g[i];
          const bool frozen = (xv <= P.lower[k] && gv > 0.0) || (xv >= P.upper[k] && gv < 0.0);
          P.dir[i] = frozen ? 0.0 : -gv * sc;
        }
      }
      P.t[s] = 1.0;
      P.nbt[s] = 0;
      phase = 1;
    }
    P.phase[s] = phase;
  }
  const double t = P.t[s];
  for (int c = 0; c < D; c++) {
    const int64_t i = cidx(P, s, c);
    const int k = c / P.q;
    P.xt[i] = (phase == 1) ? clampd(fma(t, P.dir[i], P.x[i]), P.lower[k], P.upper[k]) : P.x[i];
  }
}

// Armijo test on the evaluated trial points; accepted starts update their history and go back to phase 0
__global__ void k_polish_accept(PolishState P) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= P.S) return;
  if (P.phase[s] != 1) return;
  const int D = P.D;
  const double ftv = P.sign * P.ft[s], fv = P.f[s];
  bool finite = isfinite(ftv);
  double dec = 0.0, ss = 0.0, yy = 0.0, sy = 0.0;
  for (int c = 0; c < D; c++) {
    const int64_t i = cidx(P, s, c);
    const double gtv = P.sign * P.gt[i];
    if (!isfinite(gtv)) finite = false;
    const double sv = P.xt[i] - P.x[i], yv = gtv - P.g[i];
    dec = fma(P.g[i], sv, dec);
    ss = fma(sv, sv, ss); yy = fma(yv, yv, yy); sy = fma(sv, yv, sy);
  }
  const bool ok = finite && (ftv <= fv + P.c1 * dec);
  if (!ok) {
    const int nb = P.nbt[s] + 1;
    P.nbt[s] = nb;
    if (nb >= P.max_backtracks) { P.phase[s] = 2; atomicSub(P.remaining, 1); }   // no acceptable step: stop there
    else P.t[s] *= 0.5;
    return;
  }
  const int it = P.iters[s], slot = it % P.m;
  const bool keep = (sy > 1e-10 * sqrt(ss) * sqrt(yy)) && (sy > 0.0);
  double* sj = P.sh + (size_t)slot * P.rows * P.d;
  double* yj = P.yh + (size_t)slot * P.rows * P.d;
  for (int c = 0; c < D; c++) {
    const int64_t i = cidx(P, s, c);
    const double gtv = P.sign * P.gt[i];
    sj[i] = keep ? P.xt[i] - P.x[i] : 0.0;
    yj[i] = keep ? gtv - P.g[i] : 0.0;
    P.x[i] = P.xt[i];
    P.g[i] = gtv;
  }
  P.rho[(size_t)slot * P.S + s] = keep ? 1.0 / sy : 0.0;
  P.f[s] = ftv;
  P.iters[s] = it + 1;
  const bool small = (fv - ftv) <= P.factr_eps * fmax(fmax(fabs(fv), fabs(ftv)), 1.0);
  if (small || it + 1 >= P.maxit) { P.phase[s] = 2; atomicSub(P.remaining, 1); }
  else P.phase[s] = 0;
}

// criterion value + gradient at rows x d device candidates
int evaluate(dk_model* m, const dk_crit_spec& c, int q, const double* d_x, int64_t rows, double* d_val, double* d_grad,
             cudaStream_t st) {
  switch (c.criterion) {
    case DK_CRIT_EI: {
      CritParams P{MODE_EI, c.type, c.minimization, c.proxy, c.plugin, 0.0};
      return launch_grad(m, P, d_x, rows, d_grad, d_val, st);
    }
    case DK_CRIT_QUANTILE: {
      CritParams P{MODE_QUANTILE, c.type, 1, 0, 0.0, c.qnorm_beta};
      return launch_grad(m, P, d_x, rows, d_grad, d_val, st);
    }
    case DK_CRIT_AEI: return dk_aei_grad_dev(m, d_x, rows, c.plugin, c.new_noise_var, c.type, d_grad, d_val, st);
    case DK_CRIT_EQI: return dk_eqi_grad_dev(m, d_x, rows, c.plugin, c.new_noise_var, c.qnorm_beta, c.type, d_grad, d_val, st);
    case DK_CRIT_AKG: return dk_akg_grad_dev(m, d_x, rows, c.new_noise_var, c.type, d_grad, d_val, st);
    case DK_CRIT_QEI: {
      // value and gradient from one pass over the model (the envir reuse of R/max_qEI.R:205-222's optim loop)
      return dk_qei_value_grad_dev(m, d_x, rows / q, q, c.plugin, c.type, c.minimization, d_val, d_grad, st);
    }
  }
  return fail(DK_ERR_ARG, "criterion %d has no gradient entry point", c.criterion);
}

// ---- EI-proportional draw without replacement ----
// Successive sampling (R's sample(prob = w, replace = FALSE)) has the same law as taking the n largest keys
// log(u_i) / w_i with independent uniforms u_i (Efraimidis & Spirakis 2006): one pass + the device top-k.
__global__ void k_weight_stats(const double* __restrict__ w, int64_t N, unsigned long long* nonzero, unsigned long long* maxbits) {
  unsigned long long cnt = 0;
  double mx = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = w[i];
    if (v != 0.0) cnt++;
    mx = fmax(mx, v);
  }
  for (int off = 16; off > 0; off >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(nonzero, cnt);
    atomicMax(maxbits, (unsigned long long)__double_as_longlong(mx));   // weights are >= 0: the bit pattern is monotone
  }
}
__global__ void k_weight_keys(double* __restrict__ w, int64_t N, int n_draw, double base_count, unsigned long long seed,
                              const unsigned long long* nonzero, const unsigned long long* maxbits) {
  // R/sampleFromEI.R:126-133: fewer positive weights than draws -> ei / max(ei) + 1 / (1000 d), or all ones
  const bool few = *nonzero < (unsigned long long)n_draw;
  const double mx = __longlong_as_double((long long)*maxbits);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
    double v = w[i];
    if (few) v = (mx > 0.0) ? v / mx + 1.0 / base_count : 1.0;
    unsigned r[4];
    philox4x32_10((unsigned)i, (unsigned)((unsigned long long)i >> 32), 0x5eedu, 1u, (unsigned)seed, (unsigned)(seed >> 32), r);
    const unsigned long long bits = ((unsigned long long)r[0] << 21) | (r[1] >> 11);
    const double u = ((double)bits + 0.5) * 1.1102230246251565e-16;    // (0, 1)
    w[i] = (v > 0.0) ? log(u) / v : -INFINITY;                          // larger key = drawn earlier
  }
}

}  // namespace

extern "C" int dk_polish(dk_model* m, const dk_crit_spec* crit, int q, const double* x0, int S, const double* lower,
                         const double* upper, const dk_lbfgs_options* opt, double* out_x, double* out_val, int* out_iters) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  if (!crit || !x0 || !out_x || !out_val) return fail(DK_ERR_ARG, "dk_polish: NULL argument");
  if (S < 0) return fail(DK_ERR_ARG, "S = %d is negative", S);
  if (S == 0) return DK_OK;
  const bool qei = crit->criterion == DK_CRIT_QEI;
  if (!qei) q = 1;
  if (q < 1 || q > DK_MAX_Q_LATTICE) return fail(DK_ERR_ARG, "dk_polish: q = %d outside 1..%d", q, DK_MAX_Q_LATTICE);
  if (qei && q < 2) return fail(DK_ERR_ARG, "dk_polish: qEI needs q >= 2 (q = 1 is EI)");
  dk_lbfgs_options o{100, 8, 0.0, 1e7, 25};
  if (opt) o = *opt;
  if (o.history < 1 || o.history > 16) return fail(DK_ERR_ARG, "history = %d outside 1..16", o.history);
  if (o.maxit < 0 || o.max_backtracks < 1) return fail(DK_ERR_ARG, "bad maxit / max_backtracks");
  DeviceGuard guard(m->device);
  const int d = m->d;
  const int64_t rows = (int64_t)S * q;
  const size_t mat = (size_t)rows * d;
  PolishState P{};
  P.S = S; P.D = q * d; P.q = q; P.d = d; P.m = o.history; P.rows = rows;
  P.sign = (crit->criterion == DK_CRIT_QUANTILE) ? 1.0 : -1.0;
  P.pgtol = o.pgtol; P.factr_eps = o.factr * 2.220446049250313e-16; P.c1 = 1e-4;
  P.maxit = o.maxit; P.max_backtracks = o.max_backtracks;
  for (int k = 0; k < d; k++) {
    P.lower[k] = lower ? lower[k] : 0.0;
    P.upper[k] = upper ? upper[k] : 1.0;
    if (!(P.upper[k] >= P.lower[k])) return fail(DK_ERR_ARG, "upper[%d] < lower[%d]", k, k);
  }
  // one allocation for the whole state
  const size_t nd = mat * (5 + 2 * (size_t)o.history) + (size_t)S * (3 + o.history);
  const size_t ni = (size_t)S * 3 + 4;
  DevBuf buf;
  if (buf.alloc(nd + (ni * sizeof(int) + 7) / 8) != cudaSuccess) { cudaGetLastError(); return fail(DK_ERR_NOMEM, "device allocation of the optimiser state failed"); }
  double* p = buf.p;
  P.x = p; p += mat; P.g = p; p += mat; P.xt = p; p += mat; P.gt = p; p += mat; P.dir = p; p += mat;
  P.sh = p; p += mat * o.history; P.yh = p; p += mat * o.history;
  P.f = p; p += S; P.ft = p; p += S; P.t = p; p += S; P.rho = p; p += (size_t)S * o.history;
  int* ip = reinterpret_cast<int*>(p);
  P.phase = ip; P.iters = ip + S; P.nbt = ip + 2 * S; P.remaining = ip + 3 * S;
  cudaStream_t st = m->stream;
  DK_CUDA_TRY(cudaMemsetAsync(buf.p, 0, (nd + (ni * sizeof(int) + 7) / 8) * 8, st));
  // starts, clipped into the box (host side: S x D numbers)
  std::vector<double> h((size_t)mat);
  for (int64_t r = 0; r < rows; r++)
    for (int k = 0; k < d; k++) h[r + (size_t)k * rows] = std::min(std::max(x0[r + (size_t)k * rows], P.lower[k]), P.upper[k]);
  DK_CUDA_TRY(cudaMemcpyAsync(P.xt, h.data(), mat * 8, cudaMemcpyHostToDevice, st));
  const int tb = 64, gb = (S + tb - 1) / tb;
  int rc = evaluate(m, *crit, q, P.xt, rows, P.ft, P.gt, st);
  if (rc) { cudaStreamSynchronize(st); return rc; }
  k_polish_init<<<gb, tb, 0, st>>>(P);
  m->launches++;
  int remaining = 0;
  DK_CUDA_TRY(cudaMemcpyAsync(&remaining, P.remaining, sizeof(int), cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaStreamSynchronize(st));
  const long long max_trips = (long long)o.maxit * o.max_backtracks + 2;
  for (long long trip = 0; remaining > 0 && trip < max_trips; trip++) {
    k_polish_prepare<<<gb, tb, 0, st>>>(P);
    if ((rc = evaluate(m, *crit, q, P.xt, rows, P.ft, P.gt, st))) { cudaStreamSynchronize(st); return rc; }
    k_polish_accept<<<gb, tb, 0, st>>>(P);
    m->launches += 2;
    DK_CUDA_TRY(cudaMemcpyAsync(&remaining, P.remaining, sizeof(int), cudaMemcpyDeviceToHost, st));   // 4 bytes per trip
    DK_CUDA_TRY(cudaStreamSynchronize(st));
  }
  DK_CUDA_TRY(cudaGetLastError());
  std::vector<double> hf((size_t)S);
  std::vector<int> hi((size_t)S);
  DK_CUDA_TRY(cudaMemcpyAsync(out_x, P.x, mat * 8, cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaMemcpyAsync(hf.data(), P.f, (size_t)S * 8, cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaMemcpyAsync(hi.data(), P.iters, (size_t)S * sizeof(int), cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaStreamSynchronize(st));
  for (int s = 0; s < S; s++) {
    out_val[s] = P.sign * hf[s];
    if (out_iters) out_iters[s] = hi[s];
  }
  return DK_OK;
}

extern "C" int dk_sample_from_ei(dk_model* m, const double* x, int64_t N, double T, int minimization, int n_draw,
                                 unsigned long long seed, int64_t* out_idx) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_idx) return fail(DK_ERR_ARG, "out_idx is NULL");
  if (n_draw < 1 || n_draw > DK_TOPK_MAX) return fail(DK_ERR_ARG, "n = %d draws outside 1..%d", n_draw, DK_TOPK_MAX);
  if (n_draw > N) return fail(DK_ERR_ARG, "cannot take a sample larger than the population (n = %d, N = %lld)", n_draw, (long long)N);
  // weights through the chunked pipeline into a device vector (the screen's own), keys in place, top-n
  ScreenSpec S{DK_CRIT_EI, CritParams{MODE_EI_WEIGHT, DK_UK, minimization, 0, T, 0.0}, 0.0};
  DeviceGuard guard(m->device);
  std::vector<int64_t> idx(n_draw);
  std::vector<double> val(n_draw);
  // host_screen fills m->d_vec and ranks it; here the ranking must happen on the KEYS, so run its two halves by hand
  if ((rc = host_weights_to_vec(m, S, x, N))) return rc;
  cudaStream_t st = m->stream;
  if ((rc = ensure_red(m, topk_workspace_bytes(n_draw) + 64))) return rc;
  unsigned long long* stats = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(m->d_red) + topk_workspace_bytes(n_draw));
  DK_CUDA_TRY(cudaMemsetAsync(stats, 0, 16, st));
  k_weight_stats<<<296, 256, 0, st>>>(m->d_vec, N, stats, stats + 1);
  k_weight_keys<<<296, 256, 0, st>>>(m->d_vec, N, n_draw, (double)N, seed, stats, stats + 1);
  m->launches += 2;
  DK_CUDA_TRY(cudaGetLastError());
  if ((rc = launch_topk(m->d_vec, N, n_draw, m->d_red, st))) return rc;
  m->launches += n_draw;
  std::vector<TopkEntry> host(n_draw);
  DK_CUDA_TRY(cudaMemcpyAsync(host.data(), topk_result_ptr(m->d_red), sizeof(TopkEntry) * n_draw, cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaStreamSynchronize(st));
  for (int i = 0; i < n_draw; i++) out_idx[i] = host[i].idx;
  return DK_OK;
}
