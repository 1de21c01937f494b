// dk_check.cuh — checkPredict's distances for a batch of candidates, /root/reference/R/checkPredict.R:18-75.
//
// "covdist" (the default, :49-66), per model and per candidate x:
//     kn_xx = sd(x)^2, kn_yy[j] = sd(X_j)^2 (predict.km with the caller's type),
//     kn_xy = kxy - kxy chol2inv(T) kyy,   kyy = covMatrix(model@X)$C,
//     mindist = sqrt(max(0, kn_xx - 2 max_j(kn_xy[j] - kn_yy[j])) / sd2).
// The reference forms kn_xy with an n x n product per call.  With delta = noise.var (0 for nugget / noise-free
// models, where kyy is the matrix T factors) kyy = T'T - diag(delta), hence
//     kn_xy[j] = delta_j (T^-1 v)_j,     v = T'^-1 kxy  (the forward solve the predictor runs anyway),
// one extra back substitution on the DMMA path; exact zeros where the reference holds rounding noise.
// kn_yy is candidate-independent and computed once per (model, type) by the predict kernel.
//
// "euclidean" (:35-38): min_j ||x - X_j||_2 against the FIRST model's design; "covratio" (:40-48) is an epilogue
// of the predict kernel (MODE_COVRATIO).  Over several models the caller keeps the running pmin in out[].
#pragma once
#include "dk_grad.cuh"

namespace dk {

__global__ void __launch_bounds__(NT, 2)
k_covdist(const DevModel M, int type, const double* __restrict__ x, int64_t N, int64_t ntiles,
          double* __restrict__ vscratch, double* __restrict__ wscratch, size_t panel_doubles,
          const double* __restrict__ knyy, int combine_min, double* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;  // row 0: kn_xx
  const int n = M.n;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true>(M, x, N, c0, vpanel, smem, tid, &stage_base, nullptr);
    if (tid < BN) {
      double mean, sd, u[DK_MAX_P];
      cand_predict(M, type, vpanel, xs, s2acc, tid, mean, sd, u);
      coef[tid] = sd * sd;
    }
    __syncthreads();

    // ---- backward blocked solve w = T^-1 v (flipped row order in wpanel) ----
#pragma unroll 1
    for (int I = 0; I < M.nblk_b; I++) {
      double acc[4][4][2];
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const int ip = I * BM + 32 * wm + 8 * mi + (lane >> 2);
        const int i = n - 1 - ip;
#pragma unroll
        for (int ni = 0; ni < 4; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int c = 32 * wn + 8 * ni + 2 * (lane & 3) + e;
            acc[mi][ni][e] = (ip < n) ? panel_at(vpanel, i, c) : 0.0;
          }
      }
      solve_block_row<true>(acc, M.tiles_bwd, I, wpanel, smem, tid, &stage_base);
      __syncthreads();
    }

    // ---- max_j (delta_j w_j - kn_yy[j]) per candidate: rows j = slice, slice + 4, ... then the four slices ----
    const int c = tid & (BN - 1), slice = tid >> 6;
    double best = -INFINITY;
    for (int j = slice; j < n; j += 4)
      best = fmax(best, fma(__ldg(M.akg_delta + j), panel_at(wpanel, n - 1 - j, c), -__ldg(knyy + j)));
    double* red = reinterpret_cast<double*>(smem);  // [4][64]
    red[slice * BN + c] = best;
    __syncthreads();
    if (tid < BN && c0 + tid < N) {
      const double mx = fmax(fmax(red[tid], red[BN + tid]), fmax(red[2 * BN + tid], red[3 * BN + tid]));
      double v = sqrt(fmax(0.0, __ldcg(coef + tid) - 2.0 * mx) / M.sd2);
      if (combine_min) v = fmin(v, out[c0 + tid]);
      out[c0 + tid] = v;
    }
    __syncthreads();
  }
}

// min_j ||x - X_j|| per candidate; one thread per candidate, design rows broadcast from L1 / L2
__global__ void __launch_bounds__(256)
k_euclid_min(const DevModel M, const double* __restrict__ x, int64_t N, double* __restrict__ out) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= N) return;
  const int n = M.n, d = M.d;
  double xr[DK_MAX_D];
  for (int k = 0; k < d; k++) xr[k] = x[g + (int64_t)k * N];
  double best = INFINITY;
  for (int j = 0; j < n; j++) {
    const double* Xj = M.Xrm + (size_t)j * d;
    double s = 0.0;
    for (int k = 0; k < d; k++) {
      const double t = __ldg(Xj + k) - xr[k];   // tp1 - x, squared, summed over the columns in order (:36-37)
      s += t * t;
    }
    best = fmin(best, sqrt(s));
  }
  out[g] = best;
}

}  // namespace dk
