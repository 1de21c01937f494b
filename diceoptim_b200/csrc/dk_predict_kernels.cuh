// dk_predict_kernels.cuh — the __global__ entry points of the fused forward pass (device code: dk_kernels.cuh).
// Included by dk_api.cu only: the other translation units use the device functions, not these kernels.
#pragma once
#include "dk_kernels.cuh"

namespace dk {

// ---------------------------------------------------------------------------------------
// The fused forward kernel: persistent CTAs, two per SM, each looping over candidate tiles.
// ---------------------------------------------------------------------------------------
// Same kernel with the GEMM phase fed by bulk asynchronous copies + mbarriers (SM_TOTAL + 64 B of shared memory).
template <int KSEL, bool NUGSEL>
__global__ void __launch_bounds__(NT, 2)
k_predict_fused_tma(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
                    double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ out_main,
                    double* __restrict__ out_mean, double* __restrict__ out_sd) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#ifdef DK_X_STAGGER_NS  // what-if build: the second resident CTA of each SM starts late, so the pair is out of phase
  if (blockIdx.x >= gridDim.x / 2) {
    const long long t_end = clock64() + (long long)(DK_X_STAGGER_NS * 1.965);
    while (clock64() < t_end) __nanosleep(1000);
  }
#endif
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true, KSEL, NUGSEL>(M, x, N, c0, panel, smem, tid, &stage_base);
    if (tid < BN && c0 + tid < N)
      crit_epilogue(M, P, panel, xs, s2acc, tid, c0 + tid, out_main, out_mean, out_sd);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(NT, 2)
k_predict_fused(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
                double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ out_main,
                double* __restrict__ out_mean, double* __restrict__ out_sd) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile(M, x, N, c0, panel, smem, tid);
    if (tid < BN && c0 + tid < N)
      crit_epilogue(M, P, panel, xs, s2acc, tid, c0 + tid, out_main, out_mean, out_sd);
    __syncthreads();  // xs / s2acc / panel are reused by the next tile
  }
}

// ---------------------------------------------------------------------------------------
// Plain one-thread-per-candidate forward substitution (no tiles, no tensor pipe).  Debug aid
// only: tests use it to tell a tiling bug from an epilogue bug.  v lives in `vcol` (n x N, column j
// strided by N so that consecutive threads touch consecutive addresses).
// ---------------------------------------------------------------------------------------
__global__ void k_predict_plain(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N,
                                const double* __restrict__ Tcm, const double* __restrict__ zM, double* __restrict__ vcol,
                                double* __restrict__ out_main, double* __restrict__ out_mean,
                                double* __restrict__ out_sd) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= N) return;
  const int n = M.n, d = M.d, p = M.p;
  double s2 = 0.0, zv = 0.0, Mv[DK_MAX_P];
  for (int q = 0; q < p; q++) Mv[q] = 0.0;
  for (int i = 0; i < n; i++) {
    double s = 0.0, pr = 1.0, hs = 0.0;
    for (int k = 0; k < d; k++) {
      const double h = fabs(M.Xrm[(size_t)i * d + k] - x[j + (int64_t)k * N]);
      const double a = h * M.w[k];
      hs += h;
      switch (M.kernel) {
        case DK_GAUSS: s = fma(0.5 * a, a, s); break;
        case DK_EXP: s += a; break;
        case DK_MATERN3_2: pr *= (1.0 + a); s += a; break;
        case DK_MATERN5_2: pr *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0); s += a; break;
        default: s += pow(a, M.shape[k]); break;
      }
    }
    double c = M.sd2 * pr * exp(-s);
    if (M.nugget_flag && hs < 1e-15) c += M.nugget;
    // row i of T' = column i of T (upper): T[r + i n], r < i
    double acc = c;
    for (int r = 0; r < i; r++) acc = fma(-Tcm[r + (size_t)i * n], vcol[j + (int64_t)r * N], acc);
    const double v = acc / Tcm[i + (size_t)i * n];
    vcol[j + (int64_t)i * N] = v;
    s2 = fma(v, v, s2);
    zv = fma(v, zM[i], zv);
    for (int q = 0; q < p; q++) Mv[q] = fma(v, zM[(size_t)(q + 1) * n + i], Mv[q]);
  }
  double f[DK_MAX_P], S[DK_MAX_P];
  f[0] = 1.0;
  if (M.trend == DK_TREND_LINEAR)
    for (int k = 0; k < d; k++) f[1 + k] = x[j + (int64_t)k * N];
  double mean = 0.0;
  for (int q = 0; q < p; q++) mean = fma(f[q], M.beta[q], mean);
  mean += zv;
  double var = M.total_sd2 - s2;
  if (P.type == DK_UK) {
    double s22 = 0.0;
    for (int q = 0; q < p; q++) {
      double u = f[q] - Mv[q];
      for (int i = 0; i < q; i++) u -= M.RM[i + q * p] * S[i];
      S[q] = u / M.RM[q + q * p];
      s22 = fma(S[q], S[q], s22);
    }
    var += s22;
  }
  var = fmax(var, 0.0);
  const double sd = sqrt(var);
  if (out_mean) out_mean[j] = mean;
  if (out_sd) out_sd[j] = sd;
  double res;
  if (P.mode == MODE_PREDICT) res = mean;
  else if (P.mode == MODE_QUANTILE) res = mean + P.qnorm_beta * sd;
  else {
    const double km = P.minimization ? mean : -mean;
    const double diff = P.plugin - km;
    if (P.proxy) res = diff;
    else {
      const double xcr = diff / sd;
      res = diff * normcdf(xcr) + sd * dnorm_std(xcr);
    }
  }
  if (out_main) out_main[j] = res;
}

}  // namespace dk
