// dk_sur.cuh — crit_SUR_cst for a batch of candidates, /root/reference/R/crit_SUR_cst.R:120-378, with its helpers
// precomputeUpdateData (R/precomputeUpdateData.R:1-22) and computeQuickKrigcov2 (R/computeQuickKrigcov2.R:22-33).
//
// Per km model (objective and up to three constraints), over M integration points P and N candidates x:
//   k_sur_points  : predict at P (mn, sn) and precomputeUpdateData - "a big backsolve", twice: the forward and the
//                   backward blocked DMMA solves give  W = Kinv.c.olddata = T^-1 T'^-1 c(P)  (n x M, row-major) and
//                   first.member = (f(P) - c'Kinv.F)(M'M)^-1 = u tuuinv  (u is the predictor's own F - v'M).
//   k_sur_cand    : predict at x (mk, sk), the covariance columns c.newdata (n x N, row-major) and
//                   u = F.newdata - c.newdata' Kinv.F.
//   k_sur_kn      : kn[j][i] = k(P_i, x_j) - W[:, i]'c.newdata[:, j] + first.member[i, :] u[j, :]'  (the M x N kriging
//                   covariances of computeQuickKrigcov2: one FP64 GEMM over n with the cross-covariance and the trend
//                   term fused into its epilogue).
// k_sur_reduce then walks the integration points of each candidate: the bivariate-normal terms (pbivnorm -> the BVND
// of dk_mvn.cuh), oldcrit - newcrit, and mean(a w, na.rm = TRUE) (:370-374) by a fixed-order tree.
#pragma once
#include "dk_grad.cuh"
#include "dk_mvn.cuh"

namespace dk {

// Integration points as the "candidates" of a forward + backward pass; writes mn, sn, first.member and W.
__global__ void __launch_bounds__(NT, 2)
k_sur_points(const DevModel M, int type, const double* __restrict__ pts, int64_t Mi, int64_t ntiles,
             double* __restrict__ vscratch, double* __restrict__ wscratch, size_t panel_doubles,
             double* __restrict__ mn, double* __restrict__ sn, double* __restrict__ first /* Mi x p, row-major */,
             double* __restrict__ W /* n x Mi, row-major */) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  const int n = M.n, p = M.p;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true>(M, pts, Mi, c0, vpanel, smem, tid, &stage_base, nullptr);
    if (tid < BN && c0 + tid < Mi) {
      double mean, sd, u[DK_MAX_P];
      cand_predict(M, type, vpanel, xs, s2acc, tid, mean, sd, u);
      mn[c0 + tid] = mean;
      sn[c0 + tid] = sd;
      for (int j = 0; j < p; j++) {
        double r = 0.0;
        for (int i = 0; i < p; i++) r = fma(u[i], __ldg(M.tuuinv + i + j * p), r);
        first[(c0 + tid) * p + j] = r;
      }
    }
    __syncthreads();
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
    // W[r][c0 + c] = (T^-1 v)_r: 64 consecutive columns per row, rows strided over the four thread groups
    const int c = tid & (BN - 1), slice = tid >> 6;
    if (c0 + c < Mi)
      for (int r = slice; r < n; r += 4) W[(size_t)r * Mi + c0 + c] = panel_at(wpanel, n - 1 - r, c);
    __syncthreads();
  }
}

// Candidates: forward pass only; keeps mk, sk, u = F.newdata - c'Kinv.F and the covariance columns.
__global__ void __launch_bounds__(NT, 2)
k_sur_cand(const DevModel M, int type, const double* __restrict__ x, int64_t N, int64_t ntiles,
           double* __restrict__ vscratch, double* __restrict__ cscratch, size_t panel_doubles,
           double* __restrict__ mk, double* __restrict__ sk, double* __restrict__ U /* N x p, row-major */,
           double* __restrict__ Cnew /* n x N, row-major */) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* cpanel = cscratch + (size_t)blockIdx.x * panel_doubles;
  const int n = M.n, p = M.p;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true>(M, x, N, c0, vpanel, smem, tid, &stage_base, cpanel);
    if (tid < BN && c0 + tid < N) {
      double mean, sd, u[DK_MAX_P];
      cand_predict(M, type, vpanel, xs, s2acc, tid, mean, sd, u);
      mk[c0 + tid] = mean;
      sk[c0 + tid] = sd;
      for (int j = 0; j < p; j++) U[(c0 + tid) * p + j] = u[j];
    }
    const int c = tid & (BN - 1), slice = tid >> 6;
    if (c0 + c < N)
      for (int r = slice; r < n; r += 4) Cnew[(size_t)r * N + c0 + c] = __ldcg(cpanel + (size_t)r * BN + c);
    __syncthreads();
  }
}

// covMat1Mat2 entry between integration point i and candidate j (nugget on coincident points when nugget.flag)
__device__ __forceinline__ double cov_point_pair(const DevModel& M, const double* __restrict__ pts, int64_t Mi, int64_t i,
                                                 const double* __restrict__ x, int64_t N, int64_t j) {
  double s = 0.0, pr = 1.0, hs = 0.0;
  for (int k = 0; k < M.d; k++) {
    const double h = fabs(pts[i + (int64_t)k * Mi] - x[j + (int64_t)k * N]);
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
  return c;
}

// kn[j * Mi + i] = k(P_i, x_j) - sum_r W[r][i] Cnew[r][j] + first[i,:] . U[j,:]   (computeQuickKrigcov2)
// 64 x 64 output tile per CTA of 16 x 16 threads (4 x 4 each), the n-loop staged through shared memory 16 rows at a time.
__global__ void __launch_bounds__(256)
k_sur_kn(const DevModel M, const double* __restrict__ pts, int64_t Mi, const double* __restrict__ x, int64_t N,
         const double* __restrict__ W, const double* __restrict__ Cnew, const double* __restrict__ first,
         const double* __restrict__ U, double* __restrict__ kn) {
  __shared__ double Ws[16][64], Cs[16][64];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t i0 = (int64_t)blockIdx.x * 64, j0 = (int64_t)blockIdx.y * 64;
  const int n = M.n, p = M.p;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
  for (int r0 = 0; r0 < n; r0 += 16) {
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const int idx = threadIdx.x + 256 * e, rr = idx >> 6, cc = idx & 63;
      const int r = r0 + rr;
      Ws[rr][cc] = (r < n && i0 + cc < Mi) ? W[(size_t)r * Mi + i0 + cc] : 0.0;
      Cs[rr][cc] = (r < n && j0 + cc < N) ? Cnew[(size_t)r * N + j0 + cc] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; k++) {
      double wv[4], cv[4];
#pragma unroll
      for (int a = 0; a < 4; a++) { wv[a] = Ws[k][tx + 16 * a]; cv[a] = Cs[k][ty + 16 * a]; }
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = fma(wv[a], cv[b], acc[a][b]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int b = 0; b < 4; b++) {
    const int64_t j = j0 + ty + 16 * b;
    if (j >= N) continue;
#pragma unroll
    for (int a = 0; a < 4; a++) {
      const int64_t i = i0 + tx + 16 * a;
      if (i >= Mi) continue;
      double tr = 0.0;
      for (int q = 0; q < p; q++) tr = fma(first[i * p + q], U[j * p + q], tr);
      kn[j * Mi + i] = tr + (cov_point_pair(M, pts, Mi, i, x, N, j) - acc[a][b]);
    }
  }
}

struct SurModel {
  const double *mn, *sn;   // predict at the integration points
  const double *mk, *sk;   // predict at the candidates
  const double* kn;        // N x Mi kriging covariances
  double small_sd;         // sqrt(sd2) / 1e4 (:192)
  double tol;
  int equality;
};
struct SurParams {
  SurModel obj, cst[3];
  int nc, any_eq;
  double obj_min;
  const double* w;         // integration weights
  int64_t Mi;
};

// pbivnorm(x, y, rho) with rho clamped to [-1, 1]; NaN in -> NaN out
__device__ __forceinline__ double pbiv(double x, double y, double rho) {
  if (x != x || y != y || rho != rho) return NAN;
  return bvn_cdf_dev(x, y, fmin(1.0, fmax(-1.0, rho)));
}

__global__ void __launch_bounds__(128)
k_sur_reduce(const SurParams S, int64_t N, double* __restrict__ out) {
  __shared__ double ssum[128];
  __shared__ int scnt[128];
  const int64_t j = blockIdx.x;
  const int64_t Mi = S.Mi;
  double sum = 0.0;
  int cnt = 0;
  for (int64_t i = threadIdx.x; i < Mi; i += 128) {
    bool all_small = true;                                      // :192-193 drop points every constraint is sure about
    for (int c = 0; c < S.nc; c++) all_small = all_small && (S.cst[c].sn[i] < S.cst[c].small_sd);
    if (all_small) continue;
    // ---- objective (:229-262) ----
    const double mk = S.obj.mk[j], sk = S.obj.sk[j], mn = S.obj.mn[i], sn = S.obj.sn[i], kn = S.obj.kn[j * Mi + i];
    const double rho = kn / (sk * sn), den = sqrt(sk * sk + sn * sn - 2.0 * kn);
    const double eta = (mk - mn) / den, nu = (kn - sk * sk) / sk / den;
    const double omt = (S.obj_min - mn) / sn, p_omt = phi_cdf(omt), omb = (S.obj_min - mk) / sk;
    double w1 = 0.0, w2 = 0.0;
    if (omb > -10.0) {
      if (omb == INFINITY) w1 = phi_cdf(eta);
      else if (eta > -10.0) w1 = pbiv(omb, eta, nu);
    }
    if (-omb > -10.0) {
      if (-omb == INFINITY) w2 = p_omt;
      else if (omt > -10.0) w2 = pbiv(-omb, omt, -rho);
    }
    const double pf_minus = w1 + w2, pf_plus = p_omt;
    // ---- constraints (:274-348) ----
    double pgm[3], pgp[3], PT[3], pf = 1.0;
    for (int c = 0; c < S.nc; c++) {
      const SurModel& Cm = S.cst[c];
      const double mkc = Cm.mk[j], skc = Cm.sk[j], mnc = Cm.mn[i], snc = Cm.sn[i];
      const double rc = Cm.kn[j * Mi + i] / (skc * snc);
      if (S.any_eq && Cm.equality) {
        const double Tbp = (Cm.tol - mkc) / skc, Ttp = (Cm.tol - mnc) / snc;
        const double Tbm = (-Cm.tol - mkc) / skc, Ttm = (-Cm.tol - mnc) / snc;
        const double band = phi_cdf(Ttp) - phi_cdf(Ttm);
        pf *= band;
        pgm[c] = pbiv(Tbp, Ttp, rc) - pbiv(Tbp, Ttm, rc) - pbiv(Tbm, Ttp, rc) + pbiv(Tbm, Ttm, rc);
        pgp[c] = band - pgm[c];
        PT[c] = 0.5;                                            // pnorm(Ttilde) with Ttilde left at 0 (:272, :353-367)
      } else {
        const double Tb = (-Cm.tol - mkc) / skc, Tt = (-Cm.tol - mnc) / snc;
        PT[c] = phi_cdf(Tt);
        pf *= PT[c];
        pgm[c] = (Tb > -10.0 && Tt > -10.0) ? pbiv(Tb, Tt, rc) : 0.0;
        pgp[c] = PT[c] - pgm[c];
      }
    }
    const double oldcrit = p_omt * pf;
    double newcrit;
    if (S.nc == 1) newcrit = pf_minus * pgm[0] + pf_plus * pgp[0];
    else if (S.nc == 2)
      newcrit = pf_minus * pgm[0] * pgm[1] + pf_plus * (PT[0] * pgp[1] + PT[1] * pgp[0] - pgp[0] * pgp[1]);
    else
      newcrit = pf_minus * pgm[0] * pgm[1] * pgm[2] +
                pf_plus * (PT[1] * PT[2] * pgp[0] + PT[0] * PT[2] * pgp[1] + PT[0] * PT[1] * pgp[2] -
                           PT[2] * pgp[0] * pgp[1] - PT[1] * pgp[0] * pgp[2] - PT[0] * pgp[1] * pgp[2] +
                           pgp[0] * pgp[1] * pgp[2]);
    const double a = (oldcrit - newcrit) * S.w[i];
    if (a == a) { sum += a; cnt++; }                            // mean(., na.rm = TRUE)
  }
  ssum[threadIdx.x] = sum;
  scnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if (threadIdx.x < o) { ssum[threadIdx.x] += ssum[threadIdx.x + o]; scnt[threadIdx.x] += scnt[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) out[j] = scnt[0] > 0 ? ssum[0] / scnt[0] : NAN;
}

}  // namespace dk
