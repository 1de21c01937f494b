// dk_grad.cuh — batched EI.grad / kriging.quantile.grad in adjoint form.
//
// The reference solves d extra triangular systems per point, W = backsolve(t(T), dc)
// (R/EI.grad.R:161, R/kriging.quantile.grad.R:120), but W only ever appears as W'z, v'W and W'M:
//     W'z           = dc' alpha                  alpha = T^-1 z          (per model)
//     (v + M r)'W   = dc' T^-1 (v + M r)         ONE back substitution   (per candidate)
// with r = tuuinv (F.newdata - v'M)'.  So per candidate tile this kernel runs
//   1. the forward pass of dk_kernels.cuh (V panel, mean, sd),
//   2. a backward blocked solve  T wt = v + M r  with the flipped factor's tiles (same DMMA code),
//   3. one reduction  grad_k = sum_i c_i dlogk_k(i) (cA alpha_i + cW wt_i) + trend terms,
// where (cA, cW) fold the closed-form outer derivative:
//   EI       : cA = -/+ Phi(xcr) (min / max), cW = -phi(xcr)/sd ; all zero when sd/sqrt(sd2) < 1e-6 (R/EI.grad.R:153)
//   EI proxy : cA = -/+ 1, cW = 0
//   quantile : cA = 1, cW = -qnorm(beta)/sd ; cW = 0 when sd < sqrt(sd2)/1e6 (R/kriging.quantile.grad.R:123)
#pragma once
#include "dk_kernels.cuh"

namespace dk {

constexpr int COEF_ROWS = 2 + DK_MAX_P;  // cA, cW, r[p]
constexpr int COEF_DOUBLES = COEF_ROWS * BN;

// 1 / x for x >= 1: hardware seed (about 20 bits) + two Newton steps; a few ulp, no special cases needed here
__device__ __forceinline__ double rcp_pos(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

// d log k / d x_k of one design row.  With t = (x_k - X_ik) w_k and a = |t|:  gauss -t w ; exp -sign(t) w ;
// matern3_2 -w t / (1 + a) ; matern5_2 -w (t / 3)(1 + a) / (1 + a + a^2 / 3) ; powexp -sign(t) p a^(p-1) w.
template <int KERNEL>
__device__ __forceinline__ double dlogk_scaled(const DevModel& M, int k, double t, double wk) {
  const double a = fabs(t);
  if (KERNEL == DK_GAUSS) return -t * wk;
  if (KERNEL == DK_EXP) return (t > 0.0) ? -wk : ((t < 0.0) ? wk : 0.0);
  if (KERNEL == DK_MATERN3_2) return -wk * t * rcp_pos(1.0 + a);
  if (KERNEL == DK_MATERN5_2) return -(wk * (1.0 / 3.0)) * t * (1.0 + a) * rcp_pos(fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0));
  return (a > 0.0) ? -copysign(M.shape[k] * pow(a, M.shape[k] - 1.0) * wk, t) : 0.0;
}

// g_k += cwv * dlogk_k(x - X_i) for all k (covVector.dx of one design row, already weighted).
template <int KERNEL>
__device__ __forceinline__ void accum_dlogk(const DevModel& M, const double* __restrict__ xs,
                                            const double* __restrict__ Xi, int c, double cwv, double (&g)[DK_MAX_D]) {
  const int d = M.d;
#pragma unroll
  for (int k = 0; k < DK_MAX_D; k++) {
    if (k < d) {
      const double wk = M.w[k];
      const double t = (xs[k * BN + c] - __ldg(Xi + k)) * wk;
      g[k] = fma(cwv, dlogk_scaled<KERNEL>(M, k, t, wk), g[k]);
    }
  }
}

// g_k += sum_i c_i omega_i dlogk_k(x - X_i),  omega_i = cA alpha_i + cW wt_i[cw]  (covVector.dx contracted with the
// adjoint weights).  c_i comes from the covariance panel the forward pass kept; rows i = slice, slice + 4, ...
template <int KERNEL>
__device__ __forceinline__ void grad_accumulate(const DevModel& M, const double* __restrict__ xs,
                                                const double* __restrict__ cpanel, const double* __restrict__ wpanel,
                                                int c, int slice, double cA, double cW, double (&g)[DK_MAX_D],
                                                int cw = -1) {
  if (cw < 0) cw = c;  // column of the W panel that supplies the weights (qEI.grad pairs point c with point cw)
  const int n = M.n, d = M.d;
  // two design rows per step: their loads are issued together (the pass is latency-bound otherwise)
  int i = slice;
#pragma unroll 1
  for (; i + 4 < n; i += 8) {
    const double om0 = cA * __ldg(M.alpha + i) + cW * panel_at(wpanel, n - 1 - i, cw);
    const double om1 = cA * __ldg(M.alpha + i + 4) + cW * panel_at(wpanel, n - 1 - (i + 4), cw);
    const double c0 = __ldcg(cpanel + (size_t)i * BN + c), c1 = __ldcg(cpanel + (size_t)(i + 4) * BN + c);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)i * d, c, c0 * om0, g);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)(i + 4) * d, c, c1 * om1, g);
  }
  if (i < n) {
    const double omega = cA * __ldg(M.alpha + i) + cW * panel_at(wpanel, n - 1 - i, cw);
    const double cwv = __ldcg(cpanel + (size_t)i * BN + c) * omega;
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)i * d, c, cwv, g);
  }
}

// Same contraction with the weights read from a plain vector (stride 1): rows of K^-1, columns of T^-1 M, alpha.
template <int KERNEL>
__device__ __forceinline__ void grad_accumulate_vec(const DevModel& M, const double* __restrict__ xs,
                                                    const double* __restrict__ cpanel, const double* __restrict__ wvec,
                                                    int c, int slice, double (&g)[DK_MAX_D]) {
  const int n = M.n, d = M.d;
#pragma unroll 1
  for (int i = slice; i < n; i += 4) {
    const double cwv = __ldcg(cpanel + (size_t)i * BN + c) * __ldg(wvec + i);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)i * d, c, cwv, g);
  }
}

// KSEL = -1: any covariance kernel (runtime switch); KSEL = DK_MATERN5_2: the km default, without nugget, as its own
// instantiation so that it carries only the code it runs.
template <int KSEL>
__global__ void __launch_bounds__(NT, 2)
k_grad_fused(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
             double* __restrict__ vscratch, double* __restrict__ wscratch, double* __restrict__ cscratch,
             size_t panel_doubles, double* __restrict__ out_grad, double* __restrict__ out_main) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;  // [COEF_ROWS][64]
  double* cpanel = cscratch + (size_t)blockIdx.x * panel_doubles;  // covariance columns, row-major [i][64]
  const int n = M.n, d = M.d, p = M.p;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true, KSEL, false>(M, x, N, c0, vpanel, smem, tid, &stage_base, cpanel);

    // ---- epilogue 1: value + outer-derivative coefficients per candidate ----
    if (tid < BN) {
      const int c = tid;
      double mean, sd, u[DK_MAX_P];
      cand_predict(M, P.type, vpanel, xs, s2acc, c, mean, sd, u);
      if (out_main && c0 + c < N) out_main[c0 + c] = crit_value(P, mean, sd);
      double cA, cW;
      bool drop_r = false;
      if (P.mode == MODE_QUANTILE) {
        cA = 1.0;
        if (sd < sqrt(M.sd2) / 1e6) { cW = 0.0; drop_r = true; }
        else cW = -P.qnorm_beta / sd;
      } else if (P.mode == MODE_AEI) {  // R/AEI.grad.R:117-147: ei.grad * pen + ei.val * pen.grad
        if (sd < sqrt(P.sd2) / 1e6) { cA = 0.0; cW = 0.0; drop_r = true; }
        else {
          const double diff = P.plugin - mean, xcr = diff / sd, Phi = normcdf(xcr), phi = dnorm_std(xcr);
          const double v2 = P.tau2 + sd * sd, pen = 1.0 - sqrt(P.tau2) / sqrt(v2);
          const double eival = diff * Phi + sd * phi;
          const double kappa = phi * pen / (2.0 * sd) + eival * sqrt(P.tau2) / (2.0 * v2 * sqrt(v2));  // on sk2.grad
          cA = -Phi * pen;
          cW = -2.0 * kappa;
        }
      } else if (P.mode == MODE_EQI) {  // R/EQI.grad.R:112-141: -Phi mq.grad + phi sq.grad
        if (sd < sqrt(P.sd2) / 1e6) { cA = 0.0; cW = 0.0; drop_r = true; }
        else {
          const double t2 = P.tau2, v2 = t2 + sd * sd;
          const double mq = mean + P.qnorm_beta * sqrt(t2 * sd * sd / v2), sq = sd * sd / sqrt(v2);
          const double xcr = (P.plugin - mq) / sq, Phi = normcdf(xcr), phi = dnorm_std(xcr);
          const double g1 = P.qnorm_beta * t2 * sqrt(t2) / (2.0 * sd * v2 * sqrt(v2));
          const double kappa = -Phi * g1 + phi * (sd * sd * (2.0 * t2 + sd * sd) / (v2 * v2)) / (2.0 * sq);
          cA = -Phi;
          cW = -2.0 * kappa;
        }
      } else {  // EI
        const double sgn = P.minimization ? 1.0 : -1.0;
        if (sd / sqrt(M.sd2) < 1e-6) { cA = 0.0; cW = 0.0; drop_r = true; }
        else if (P.proxy) { cA = -sgn; cW = 0.0; drop_r = true; }
        else {
          const double xcr = (P.plugin - sgn * mean) / sd;
          cA = -sgn * normcdf(xcr);
          cW = -dnorm_std(xcr) / sd;
        }
      }
      coef[c] = cA;
      coef[BN + c] = cW;
      for (int j = 0; j < p; j++) {
        double r = 0.0;
        if (P.type == DK_UK && !drop_r)
          for (int i = 0; i < p; i++) r = fma(__ldg(M.tuuinv + j + i * p), u[i], r);
        coef[(2 + j) * BN + c] = r;
      }
    }
    __syncthreads();

    // ---- backward blocked solve with the flipped factor: wt' = flip(T)^-1 flip(v + M r) ----
#ifdef DK_X_NOBWD  // what-if build: skip the backward solve (timing only)
    const int nbwd = 0;
#else
    const int nbwd = M.nblk_b;
#endif
#pragma unroll 1
    for (int I = 0; I < nbwd; I++) {
      double acc[4][4][2];
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const int ip = I * BM + 32 * wm + 8 * mi + (lane >> 2);  // flipped row
        const int i = n - 1 - ip;
#pragma unroll
        for (int ni = 0; ni < 4; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int c = 32 * wn + 8 * ni + 2 * (lane & 3) + e;
            double v = 0.0;
            if (ip < n) {
              v = panel_at(vpanel, i, c);
              if (P.type == DK_UK)
                for (int j = 0; j < p; j++) v = fma(__ldg(M.Mcm + i + (size_t)j * n), __ldcg(coef + (2 + j) * BN + c), v);
            }
            acc[mi][ni][e] = v;
          }
      }
      solve_block_row<true>(acc, M.tiles_bwd, I, wpanel, smem, tid, &stage_base, min(BM, n - I * BM));
      __syncthreads();
    }

    // ---- gradient reduction: 4 row slices x 64 candidates ----
    {
      const int c = tid & (BN - 1), slice = tid >> 6;
      const double cA = __ldcg(coef + c), cW = __ldcg(coef + BN + c);
      double g[DK_MAX_D];
#pragma unroll
      for (int k = 0; k < DK_MAX_D; k++) g[k] = 0.0;
#ifdef DK_X_NOGRADACC  // what-if build: skip the gradient reduction (timing only)
      if (M.n < 0)
#endif
      if constexpr (KSEL >= 0) grad_accumulate<KSEL>(M, xs, cpanel, wpanel, c, slice, cA, cW, g);
      else switch (M.kernel) {
        case DK_GAUSS: grad_accumulate<DK_GAUSS>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_EXP: grad_accumulate<DK_EXP>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_MATERN3_2: grad_accumulate<DK_MATERN3_2>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_MATERN5_2: grad_accumulate<DK_MATERN5_2>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        default: grad_accumulate<DK_POWEXP>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
      }
      double* gbuf = reinterpret_cast<double*>(smem);  // [4][DK_MAX_D][64] in region 0
#pragma unroll
      for (int k = 0; k < DK_MAX_D; k++)
        if (k < d) gbuf[(slice * DK_MAX_D + k) * BN + c] = g[k];
      __syncthreads();
      if (c0 + c < N) {
        for (int k = slice; k < d; k += 4) {
          double s = ((gbuf[(0 * DK_MAX_D + k) * BN + c] + gbuf[(1 * DK_MAX_D + k) * BN + c]) +
                      gbuf[(2 * DK_MAX_D + k) * BN + c]) + gbuf[(3 * DK_MAX_D + k) * BN + c];
          if (M.trend == DK_TREND_LINEAR)  // trend.deltax = rbind(0, I_d): beta_{1+k} and r_{1+k}
            s += cA * __ldg(M.beta + 1 + k) - cW * __ldcg(coef + (2 + 1 + k) * BN + c);
          out_grad[(c0 + c) + (int64_t)k * N] = s;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace dk
