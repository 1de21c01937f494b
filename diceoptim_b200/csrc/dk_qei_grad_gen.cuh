// dk_qei_grad_gen.cuh — qEI.grad for batches of 5 .. DK_MAX_Q_LATTICE points (/root/reference/R/qEI.grad.R:267-424 takes
// any q; dk_qei.cuh serves 2 .. 4 with four point slots per batch and the deterministic CDF rules).
//
//   k_qei_grad_moments_gen : tiles of floor(64 / q) batches.  Forward solve (covariance columns kept), mu / Sigma / duplicate
//                            mask as k_qei_moments_gen but at x itself (only the duplicate test rounds, R/qEI.grad.R:292-295),
//                            ONE backward DMMA solve wt_l = T^-1 (v_l + M r_l) for all columns, then krigingDeriv
//                            (R/qEI.grad.R:62-142) in adjoint form, the q + 1 weight vectors of a point (alpha and the wt of
//                            its batch) swept five at a time:
//                              d m_k / d x_k      = dc_k' alpha + beta f.deltax
//                              d Sigma_kl / d x_k = ker.grad(k, l) - dc_k' wt_l + f.deltax' r_l
//   k_qei_grad_cdf_gen     : one CTA per batch, R/qEI.grad.R:337-421 as written there: for every k the q-variate CDF p_k, its
//                            gradient and Hessian (GPhi / HPhi by decompo, R/qEI.grad.R:1-60) and, for every i, the same three
//                            of the (q - 1)-variate conditional law, then A1 + A2 + B1 + B2 + B3 per (point, dimension), the
//                            entries dealt over the threads.  Normal CDFs of dimension >= 5 by the fixed lattice of
//                            dk_qei_gen.cuh (all threads share the points), <= 4 by the deterministic rules (one CDF per thread).
// Cost at q = 10: 5160 lattice CDFs per batch (about 0.15 s); the formula is the reference's, which spends minutes there.
#pragma once
#include "dk_grad.cuh"
#include "dk_qei_gen.cuh"

namespace dk {

// per batch: mj[a][j] = d m_a / d x_aj (q x d), then cj[a][l][j] = one-sided d Sigma_al / d x_aj (q x q x d)
__host__ __device__ inline size_t qei_deriv_stride_gen(int q, int d) { return (size_t)(q + q * q) * d; }

// Five of the q + 1 derivative contractions of batch point c, dims k0 .. k0 + 3, in one sweep over the design rows
// i = slice, slice + 4, ...: weight widx0 + ps of the list [alpha, wt(colbase), wt(colbase + 1), ...].
template <int KERNEL>
__device__ __forceinline__ void qei_deriv_accumulate_gen(const DevModel& M, const double* __restrict__ xs,
                                                         const double* __restrict__ cpanel,
                                                         const double* __restrict__ wpanel, int c, int slice, int q,
                                                         int colbase, int widx0, int k0, double (&acc)[5][4]) {
  const int n = M.n, d = M.d;
  double xk[4], wk[4];
#pragma unroll
  for (int kk = 0; kk < 4; kk++) {
    const int k = k0 + kk;
    xk[kk] = (k < d) ? xs[k * BN + c] : 0.0;
    wk[kk] = (k < d) ? M.w[k] : 0.0;
  }
#pragma unroll 2
  for (int i = slice; i < n; i += 4) {
    const double ci = __ldcg(cpanel + (size_t)i * BN + c);
    double om[5];
#pragma unroll
    for (int ps = 0; ps < 5; ps++) {
      const int w = widx0 + ps;
      om[ps] = (w == 0) ? ci * __ldg(M.alpha + i) : ((w <= q) ? ci * panel_at(wpanel, n - 1 - i, colbase + w - 1) : 0.0);
    }
    const double* Xi = M.Xrm + (size_t)i * d;
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
      const int k = k0 + kk;
      if (k < d) {
        const double dl = dlogk_scaled<KERNEL>(M, k, (xk[kk] - __ldg(Xi + k)) * wk[kk], wk[kk]);
#pragma unroll
        for (int ps = 0; ps < 5; ps++) acc[ps][kk] = fma(om[ps], dl, acc[ps][kk]);
      }
    }
  }
}

__global__ void __launch_bounds__(NT, 2)
k_qei_grad_moments_gen(const DevModel M, int type, int q, const double* __restrict__ x, int64_t B, int64_t ntiles,
                       double* __restrict__ vscratch, double* __restrict__ wscratch, double* __restrict__ cscratch,
                       size_t panel_doubles, double* __restrict__ moments, double* __restrict__ deriv) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;
  double* cpanel = cscratch + (size_t)blockIdx.x * panel_doubles;
  const int n = M.n, d = M.d, p = M.p;
  const int nb = BN / q;            // batches per tile
  const int64_t N = B * q;
  const int mstride = gen_mom_stride(q);
  const size_t dstride = qei_deriv_stride_gen(q, d);
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b0 = tile * nb;
    // column c = q g + a is point a of batch b0 + g, at x itself; columns beyond the tile's batches repeat the last row
    for (int i = tid; i < d * BN; i += NT) {
      const int k = i / BN, c = i % BN;
      int64_t row = b0 * q + c;
      if (c >= nb * q || row >= N) row = N - 1;
      xs[i] = x[row + (int64_t)k * N];
    }
    forward_tile<true, -1, false, true>(M, x, N, 0, vpanel, smem, tid, &stage_base, cpanel);

    // ---- per column: duplicate flag (on the rounded rows), mean, S (UK), r = tuuinv (F.newdata - v'M)' ----
    int* dupflag = reinterpret_cast<int*>(smem);            // [64]
    double* mu_s = reinterpret_cast<double*>(smem + 1024);  // [64]
    double* S_s = reinterpret_cast<double*>(smem + 2048);   // [64][p]
    if (tid < BN) dupflag[tid] = 0;
    __syncthreads();
    {
      const int c = tid & (BN - 1), slice = tid >> 6;
      if (design_duplicate(M, xs, c, slice)) atomicOr(&dupflag[c], 1);
    }
    if (tid < BN) {
      const int c = tid;
      double mean = 0.0, u[DK_MAX_P];
      for (int j = 0; j < p; j++) {
        const double fj = (j == 0) ? 1.0 : xs[(j - 1) * BN + c];
        mean = fma(fj, __ldg(M.beta + j), mean);
        u[j] = fj + panel_at(vpanel, n + 1 + j, c);
      }
      mu_s[c] = mean - panel_at(vpanel, n, c);
      if (type == DK_UK)
        for (int j = 0; j < p; j++) {
          double t = u[j];
          for (int i = 0; i < j; i++) t -= __ldg(M.RM + i + j * p) * S_s[c * p + i];
          S_s[c * p + j] = t / __ldg(M.RM + j + j * p);
        }
      for (int j = 0; j < p; j++) {
        double r = 0.0;
        if (type == DK_UK)
          for (int i = 0; i < p; i++) r = fma(__ldg(M.tuuinv + j + i * p), u[i], r);
        coef[(2 + j) * BN + c] = r;
      }
    }
    __syncthreads();
    // ---- Sigma: one warp per (batch, a <= b) pair, lanes over the design rows, fixed-order reduction ----
    const int npair = q * (q + 1) / 2;
    for (int job = warp; job < nb * npair; job += NT / 32) {
      const int g = job / npair;
      int t = job % npair, a = 0;
      while (t >= q - a) { t -= q - a; a++; }
      const int bb = a + t;
      const int64_t bidx = b0 + g;
      if (bidx >= B) continue;
      const int ca = q * g + a, cb = q * g + bb;
      double gsum;
      if (a == bb) {
        gsum = ((s2acc[ca] + s2acc[BN + ca]) + s2acc[2 * BN + ca]) + s2acc[3 * BN + ca];   // as cand_predict
      } else {
        double part = 0.0;
        for (int i = lane; i < n; i += 32) part = fma(panel_at(vpanel, i, ca), panel_at(vpanel, i, cb), part);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        gsum = part;
      }
      if (lane == 0) {
        const double kab = (a == bb) ? M.total_sd2 : cov_pair(M, xs, ca, cb);
        double sab = 0.0;
        if (type == DK_UK)
          for (int j = 0; j < p; j++) sab = fma(S_s[ca * p + j], S_s[cb * p + j], sab);
        const double v = kab - gsum + sab;
        double* mom = moments + bidx * mstride;
        mom[q + a * q + bb] = v;
        mom[q + bb * q + a] = v;
      }
    }
    if (tid < nb) {
      const int g = tid;
      const int64_t bidx = b0 + g;
      if (bidx < B) {
        double* mom = moments + bidx * mstride;
        int mask = 0;
        for (int aa = 0; aa < q; aa++) {
          mom[aa] = mu_s[q * g + aa];
          bool drop = dupflag[q * g + aa] != 0;
          for (int e = 0; e < aa && !drop; e++) {
            bool same = true;
            for (int k = 0; k < d && same; k++) same = (round8(xs[k * BN + q * g + aa]) == round8(xs[k * BN + q * g + e]));
            drop = same;
          }
          if (!drop) mask |= 1 << aa;
        }
        mom[q + q * q] = (double)mask;
      }
    }
    __syncthreads();

    // ---- backward blocked solve: wt_l = T^-1 (v_l + M r_l) for all 64 columns ----
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
            double v = 0.0;
            if (ip < n) {
              v = panel_at(vpanel, i, c);
              if (type == DK_UK)
                for (int j = 0; j < p; j++) v = fma(__ldg(M.Mcm + i + (size_t)j * n), __ldcg(coef + (2 + j) * BN + c), v);
            }
            acc[mi][ni][e] = v;
          }
      }
      solve_block_row<true>(acc, M.tiles_bwd, I, wpanel, smem, tid, &stage_base, min(BM, n - I * BM));
      __syncthreads();
    }

    // ---- derivative contractions: four dims and five weights per sweep ----
    const int c = tid & (BN - 1), slice = tid >> 6;
    const bool in_tile = c < nb * q;
    const int g = in_tile ? c / q : 0, apt = in_tile ? c - g * q : 0;   // the columns past the last batch compute nothing kept
    double* gbuf = reinterpret_cast<double*>(smem);  // [4 slices][5 weights][4 dims][64]
    const int64_t b = b0 + g;
    const bool owns = in_tile && b < B;
#pragma unroll 1
    for (int k0 = 0; k0 < d; k0 += 4)
#pragma unroll 1
      for (int widx0 = 0; widx0 <= q; widx0 += 5) {
        double acc[5][4];
#pragma unroll
        for (int ps = 0; ps < 5; ps++)
#pragma unroll
          for (int kk = 0; kk < 4; kk++) acc[ps][kk] = 0.0;
        switch (M.kernel) {
          case DK_GAUSS: qei_deriv_accumulate_gen<DK_GAUSS>(M, xs, cpanel, wpanel, c, slice, q, q * g, widx0, k0, acc); break;
          case DK_EXP: qei_deriv_accumulate_gen<DK_EXP>(M, xs, cpanel, wpanel, c, slice, q, q * g, widx0, k0, acc); break;
          case DK_MATERN3_2: qei_deriv_accumulate_gen<DK_MATERN3_2>(M, xs, cpanel, wpanel, c, slice, q, q * g, widx0, k0, acc); break;
          case DK_MATERN5_2: qei_deriv_accumulate_gen<DK_MATERN5_2>(M, xs, cpanel, wpanel, c, slice, q, q * g, widx0, k0, acc); break;
          default: qei_deriv_accumulate_gen<DK_POWEXP>(M, xs, cpanel, wpanel, c, slice, q, q * g, widx0, k0, acc); break;
        }
#pragma unroll
        for (int ps = 0; ps < 5; ps++)
#pragma unroll
          for (int kk = 0; kk < 4; kk++) gbuf[((slice * 5 + ps) * 4 + kk) * BN + c] = acc[ps][kk];
        __syncthreads();
        // thread (c, slice) finishes dim k0 + slice for the five weights of this sweep
        const int k = k0 + slice;
        if (owns && k < d) {
          double* dv = deriv + b * dstride;
          for (int ps = 0; ps < 5 && widx0 + ps <= q; ps++) {
            const double s = ((gbuf[((0 * 5 + ps) * 4 + slice) * BN + c] + gbuf[((1 * 5 + ps) * 4 + slice) * BN + c]) +
                              gbuf[((2 * 5 + ps) * 4 + slice) * BN + c]) + gbuf[((3 * 5 + ps) * 4 + slice) * BN + c];
            if (widx0 + ps == 0) {
              double v = s;  // dc_k' alpha
              if (M.trend == DK_TREND_LINEAR) v += __ldg(M.beta + 1 + k);
              dv[apt * d + k] = v;
            } else {
              // ker.grad(k, l) - dc_k' wt_l + f.deltax' r_l
              const int l = widx0 + ps - 1;
              double v = -s;
              if (l != apt) v += cov_pair(M, xs, c, q * g + l) * dlogk_dev(M, k, xs[k * BN + c] - xs[k * BN + q * g + l]);
              if (M.trend == DK_TREND_LINEAR) v += __ldcg(coef + (2 + 1 + k) * BN + q * g + l);
              dv[(size_t)(q + apt * q + l) * d + k] = v;
            }
          }
        }
        __syncthreads();
      }
  }
}

// ---------------------------------------------------------------------------------------------------------
// decompo / GPhi / HPhi (R/qEI.grad.R:1-60) for any dimension, matrices with a leading dimension
// ---------------------------------------------------------------------------------------------------------
// decompo(x, mu, Sigma, k = {i} or {i, j}) without its CDF factor: returns the density of the components k and writes the
// conditional law of the rest (x2 - moy, varcov with the reference's repair of non-positive variances).  j < 0: k = {i}.
__device__ inline double decompo_build(int m, const double* x, const double* mu, const double* S, int ld, int i, int j,
                                       double* x2, double* V, int ldv) {
  double dens;
  int r = 0;
  if (j < 0) {
    const double sii = S[i * ld + i];
    dens = dmnorm1(x[i], mu[i], sii);
    for (int u = 0; u < m; u++) {
      if (u == i) continue;
      x2[r] = x[u] - (mu[u] + S[u * ld + i] / sii * (x[i] - mu[i]));
      int cc = 0;
      for (int v = 0; v < m; v++) {
        if (v == i) continue;
        const double a = S[u * ld + v] - S[u * ld + i] * S[v * ld + i] / sii;
        const double b = S[v * ld + u] - S[v * ld + i] * S[u * ld + i] / sii;
        V[r * ldv + cc] = 0.5 * (a + b);
        cc++;
      }
      r++;
    }
  } else {
    const double sii = S[i * ld + i], sij = S[i * ld + j], sjj = S[j * ld + j];
    const double di = x[i] - mu[i], dj = x[j] - mu[j];
    dens = dmnorm2(di, dj, sii, sij, sjj);
    const double det = sii * sjj - sij * sij;
    for (int u = 0; u < m; u++) {
      if (u == i || u == j) continue;
      const double a_ui = S[u * ld + i], a_uj = S[u * ld + j];
      const double wi = (a_ui * sjj - a_uj * sij) / det, wj = (-a_ui * sij + a_uj * sii) / det;  // S21 S11^-1 (row u)
      x2[r] = x[u] - (mu[u] + wi * di + wj * dj);
      int cc = 0;
      for (int v = 0; v < m; v++) {
        if (v == i || v == j) continue;
        const double b_vi = S[v * ld + i], b_vj = S[v * ld + j];
        const double wvi = (b_vi * sjj - b_vj * sij) / det, wvj = (-b_vi * sij + b_vj * sii) / det;
        const double a = S[u * ld + v] - (wi * b_vi + wj * b_vj);
        const double b = S[v * ld + u] - (wvi * a_ui + wvj * a_uj);
        V[r * ldv + cc] = 0.5 * (a + b);
        cc++;
      }
      r++;
    }
  }
  if (r > 0) {
    double mn = V[0];
    for (int u = 1; u < r; u++) mn = fmin(mn, V[u * ldv + u]);
    if (mn <= 0.0)
      for (int u = 0; u < r; u++) V[u * ldv + u] = (V[u * ldv + u] >= 0.0 ? V[u * ldv + u] : 0.0) + 1e-9;
  }
  return dens;
}

struct GenGradShared {
  double kmean[QG], kcov[QG * QG];
  double mk[QG], bk[QG], xk[QG], zero[QG], Sigk[QG * QG];
  double gradpk[QG], hesspk[QG * QG];
  double phi_i[QG], dphi_dS[QG], dphi_dm[QG], Phi_i[QG], inv_sii[QG];
  double ck[QG], Sp[QG * QG];                       // conditional law of term B for the current i
  double Gik[QG * QG], Hik[QG * QG * QG];           // its GPhi (row i) and HPhi (block i)
  double px[QG], pV[QG * QG], pdens;                // the CDF problem being evaluated cooperatively
  double work[2 * QG + 2 * QG * QG + 8];
  double pk;
  int skip;
};

// entry e of the i < j pairs of an m-set, in the reference's loop order (i outer)
__device__ __forceinline__ void pair_of(int m, int e, int& i, int& j) {
  i = 0;
  while (e >= m - 1 - i) { e -= m - 1 - i; i++; }
  j = i + 1 + e;
}

// GPhi (j0 < 0 ... one index) / the off-diagonal part of HPhi (pairs), collectively by the CTA: out[e] = density x CDF.
// x, mu, S (leading dimension QG) in shared memory.  pairs = false: e = i, out[i]; pairs = true: out[i * QG + j] and [j * QG + i].
__device__ inline void decompo_all_cta(GenGradShared& sh, int m, const double* x, const double* mu, const double* S, bool pairs,
                                       double* out) {
  const int tid = threadIdx.x;
  const int cnt = pairs ? m * (m - 1) / 2 : m, ms = m - (pairs ? 2 : 1);
  __syncthreads();
  if (ms >= 5) {
#pragma unroll 1
    for (int e = 0; e < cnt; e++) {
      int i = e, j = -1;
      if (pairs) pair_of(m, e, i, j);
      if (tid == 0) sh.pdens = decompo_build(m, x, mu, S, QG, i, j, sh.px, sh.pV, QG);
      const double c = pmnorm_lattice_cta(ms, sh.px, sh.pV, QG, sh.work);   // synchronises before it reads
      if (tid == 0) {
        const double v = sh.pdens * c;
        if (pairs) { out[i * QG + j] = v; out[j * QG + i] = v; }
        else out[i] = v;
      }
      __syncthreads();
    }
  } else {
    // the deterministic rules are single-thread functions: one entry per thread, side by side
    if (tid < cnt) {
      int i = tid, j = -1;
      if (pairs) pair_of(m, tid, i, j);
      double x2[4], V[16];
      const double dens = decompo_build(m, x, mu, S, QG, i, j, x2, V, 4);
      const double v = (ms > 0) ? dens * pmnorm_dev(ms, x2, V) : dens;
      if (pairs) { out[i * QG + j] = v; out[j * QG + i] = v; }
      else out[i] = v;
    }
    __syncthreads();
  }
}

// HPhi(x, mu, S, gradient) (R/qEI.grad.R:39-60) into H (leading dimension QG)
__device__ inline void hphi_cta(GenGradShared& sh, int m, const double* x, const double* mu, const double* S, const double* g,
                                double* H) {
  const int tid = threadIdx.x;
  __syncthreads();
  if (m == 1) {
    if (tid == 0) H[0] = -(x[0] - mu[0]) / S[0] * dmnorm1(x[0], mu[0], S[0]);
    __syncthreads();
    return;
  }
  if (tid < m) H[tid * QG + tid] = 0.0;
  decompo_all_cta(sh, m, x, mu, S, true, H);
  if (tid < m) {
    const int i = tid;
    double sr = 0.0;
    for (int j = 0; j < m; j++)
      if (j != i) sr += S[i * QG + j] * H[j * QG + i];
    H[i * QG + i] = -((x[i] - mu[i]) * g[i] + sr) / S[i * QG + i];
  }
  __syncthreads();
}

// One CTA per (batch, k): the k-th summand of the gradient (R/qEI.grad.R:337-421 is a sum over k of independent terms) into
// partial[(b q + k)][q d]; k_qei_grad_sum_gen adds the q summands in order.  q CTAs per batch keep the latency of a single
// polish trip (a handful of batches) at one k's worth of CDFs.
__global__ void __launch_bounds__(CDF_THREADS)
k_qei_grad_cdf_gen(const double* __restrict__ moments, const double* __restrict__ deriv, int64_t B, int q, int d,
                   double plugin, int minimization, double sd2, double* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char gsm[];
  GenGradShared& sh = *reinterpret_cast<GenGradShared*>(gsm);
  const int tid = threadIdx.x;
  const int mstride = gen_mom_stride(q);
  const double sgn = minimization ? 1.0 : -1.0;
#pragma unroll 1
  for (int64_t item = blockIdx.x; item < B * q; item += gridDim.x) {
    const int64_t b = item / q;
    const int k = (int)(item - b * q);
    const double* mom = moments + b * mstride;
    const double* dv = deriv + b * qei_deriv_stride_gen(q, d);
    double* G = partial + item * (size_t)(q * d);
    __syncthreads();
    if (tid == 0) {
      const int mask = (int)mom[q + q * q];
      bool alltiny = true;
      for (int i = 0; i < q; i++) alltiny = alltiny && (mom[q + i * q + i] / sd2 < 1e-8);
      // zero matrix when any point is a duplicate (R/qEI.grad.R:293-295) or all variances vanish (:107)
      sh.skip = (mask != (1 << q) - 1 || alltiny) ? 1 : 0;
    }
    for (int e = tid; e < q * q; e += CDF_THREADS) sh.kcov[(e / q) * QG + e % q] = mom[q + e];
    if (tid < q) { sh.kmean[tid] = sgn * mom[tid]; sh.zero[tid] = 0.0; }
    __syncthreads();
    if (sh.skip) {
      for (int e = tid; e < q * d; e += CDF_THREADS) G[e] = 0.0;
      continue;
    }
    {
      __syncthreads();
      // Lk = -I with column k set to one: (Lk y)_a = y_k - y_a (a != k), y_k (a = k).  mk = Lk kmean, Sigk = Lk kcov Lk'.
      for (int e = tid; e < q * q; e += CDF_THREADS) {
        const int a = e / q, c = e % q;
        // raw[a][c] = sum over the one or two non-zero entries of rows a and c of Lk
        auto lk_kcov_lk = [&](int r1, int r2) {
          double s = sh.kcov[k * QG + k];
          if (r2 != k) s -= sh.kcov[k * QG + r2];
          if (r1 != k) {
            s -= sh.kcov[r1 * QG + k];
            if (r2 != k) s += sh.kcov[r1 * QG + r2];
          }
          return s;
        };
        sh.Sigk[a * QG + c] = 0.5 * (lk_kcov_lk(a, c) + lk_kcov_lk(c, a));
      }
      if (tid < q) {
        const int a = tid;
        sh.mk[a] = (a == k) ? sh.kmean[k] : sh.kmean[k] - sh.kmean[a];
        sh.bk[a] = (a == k) ? plugin : 0.0;
        sh.xk[a] = sh.bk[a] - sh.mk[a];
      }
      __syncthreads();
      const double pk = pmnorm_lattice_cta(q, sh.xk, sh.Sigk, QG, sh.work);
      decompo_all_cta(sh, q, sh.xk, sh.zero, sh.Sigk, false, sh.gradpk);        // GPhi(bk - mk, b = 0, Sigk)
      hphi_cta(sh, q, sh.bk, sh.mk, sh.Sigk, sh.gradpk, sh.hesspk);             // HPhi(bk, mk, Sigk, gradpk)

      // per-i quantities of the exact term B (R/qEI.grad.R:379-396)
#pragma unroll 1
      for (int i = 0; i < q; i++) {
        __syncthreads();
        const double S_ii = sh.Sigk[i * QG + i], dm = sh.bk[i] - sh.mk[i];
        if (tid == 0) {
          const double ph = dmnorm1(sh.bk[i], sh.mk[i], S_ii);
          sh.phi_i[i] = ph;
          sh.dphi_dS[i] = (dm * dm / (2.0 * S_ii * S_ii) - 0.5 / S_ii) * ph;
          sh.dphi_dm[i] = dm / S_ii * ph;
          sh.inv_sii[i] = 1.0 / S_ii;
        }
        for (int e = tid; e < (q - 1) * (q - 1); e += CDF_THREADS) {
          const int r = e / (q - 1), cc = e % (q - 1), u = r < i ? r : r + 1, v = cc < i ? cc : cc + 1;
          const double a1 = sh.Sigk[u * QG + v] - sh.Sigk[u * QG + i] * sh.Sigk[v * QG + i] / S_ii;
          const double a2 = sh.Sigk[v * QG + u] - sh.Sigk[v * QG + i] * sh.Sigk[u * QG + i] / S_ii;
          sh.Sp[r * QG + cc] = 0.5 * (a1 + a2);
          if (cc == 0) sh.ck[r] = sh.bk[u] - sh.mk[u] - dm / S_ii * sh.Sigk[u * QG + i];
        }
        __syncthreads();
        if (q - 1 >= 5) {
          const double ph = pmnorm_lattice_cta(q - 1, sh.ck, sh.Sp, QG, sh.work);
          if (tid == 0) sh.Phi_i[i] = ph;
        } else if (tid == 0) {
          double x4[4], S4[16];
          for (int a = 0; a < q - 1; a++) {
            x4[a] = sh.ck[a];
            for (int c = 0; c < q - 1; c++) S4[a * 4 + c] = sh.Sp[a * QG + c];
          }
          sh.Phi_i[i] = pmnorm_dev(q - 1, x4, S4);
        }
        decompo_all_cta(sh, q - 1, sh.ck, sh.zero, sh.Sp, false, sh.Gik + i * QG);
        hphi_cta(sh, q - 1, sh.ck, sh.zero, sh.Sp, sh.Gik + i * QG, sh.Hik + i * QG * QG);
      }
      __syncthreads();

      // A1 + A2 + B1 + B2 + B3 per (point l, dimension j): entries dealt over the threads
      const double fac = plugin - sh.kmean[k];
#pragma unroll 1
      for (int e = tid; e < q * d; e += CDF_THREADS) {
        const int l = e / d, j = e % d;
        // J = symmetrised cov.jacob[l, j, , ]: row / column l carry the one-sided derivatives, (l, l) twice.
        // Sdx = Lk J Lk', with (Lk y)_a = y_k - y_a: built from the rows l of J only
        double Jl[QG], Sdx[QG * QG], mdx[QG];
        for (int i = 0; i < q; i++) Jl[i] = dv[(size_t)(q + l * q + i) * d + j] * (i == l ? 2.0 : 1.0);
        auto Jat = [&](int a, int c) { return (a == l) ? Jl[c] : ((c == l) ? Jl[a] : 0.0); };
        for (int a = 0; a < q; a++)
          for (int c = 0; c < q; c++) {
            double s = Jat(k, k);
            if (c != k) s -= Jat(k, c);
            if (a != k) {
              s -= Jat(a, k);
              if (c != k) s += Jat(a, c);
            }
            Sdx[a * QG + c] = s;
          }
        const double mjl = sgn * dv[l * d + j];      // mean.jacob[l, j, l]
        for (int a = 0; a < q; a++) {
          // Lk[a][l]: column k is one, the diagonal elsewhere is minus one
          const double lk = (l == k) ? 1.0 : ((a == l) ? -1.0 : 0.0);
          mdx[a] = lk * mjl;
        }
        double hs = 0.0, gm = 0.0;
        for (int a = 0; a < q; a++) {
          gm += sh.gradpk[a] * mdx[a];
          for (int c = 0; c < q; c++) hs += sh.hesspk[a * QG + c] * Sdx[a * QG + c];
        }
        double acc = fac * (0.5 * hs - gm);
        if (l == k) acc -= mjl * pk;                 // A1
        for (int i = 0; i < q; i++) {
          const double S_ik = sh.Sigk[i * QG + k], S_ii = sh.Sigk[i * QG + i], dm = sh.bk[i] - sh.mk[i];
          const double sdx_ii = Sdx[i * QG + i];
          const double B1 = Sdx[i * QG + k] * sh.phi_i[i] * sh.Phi_i[i];
          const double B2 = S_ik * (mdx[i] * sh.dphi_dm[i] + sh.dphi_dS[i] * sdx_ii) * sh.Phi_i[i];
          const double is = sh.inv_sii[i], is2 = is * is;
          const double dc_a = (mdx[i] * S_ii + dm * sdx_ii) * is2, dc_b = dm * is, ds_a = sdx_ii * is2;
          double gd = 0.0, hd = 0.0;
          int r = 0;
          for (int u = 0; u < q; u++) {
            if (u == i) continue;
            const double S_ui = sh.Sigk[u * QG + i];
            const double dck = -mdx[u] + dc_a * S_ui - dc_b * Sdx[u * QG + i];
            gd += sh.Gik[i * QG + r] * dck;
            int cc = 0;
            for (int v = 0; v < q; v++) {
              if (v == i) continue;
              const double S_vi = sh.Sigk[v * QG + i];
              const double dS = Sdx[u * QG + v] + ds_a * S_ui * S_vi - (Sdx[u * QG + i] * S_vi + Sdx[v * QG + i] * S_ui) * is;
              hd += sh.Hik[(i * QG + r) * QG + cc] * dS;
              cc++;
            }
            r++;
          }
          acc += B1 + B2 + S_ik * sh.phi_i[i] * (gd + 0.5 * hd);
        }
        G[e] = acc;
      }
    }
  }
}

// out = sum over k of the q summands of a batch, in order; NaN anywhere -> zero matrix (R/qEI.grad.R:422).  One CTA per batch.
__global__ void __launch_bounds__(CDF_THREADS)
k_qei_grad_sum_gen(const double* __restrict__ partial, int64_t B, int q, int d, double* __restrict__ out /* (B*q) x d column-major */) {
  __shared__ double G[QG * DK_MAX_D];
  __shared__ int bad;
  const int tid = threadIdx.x;
  const int64_t N = B * q;
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    for (int e = tid; e < q * d; e += CDF_THREADS) {
      double s = 0.0;
      for (int k = 0; k < q; k++) s += partial[(b * q + k) * (size_t)(q * d) + e];
      G[e] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int e = 0; e < q * d; e++) s += G[e];
      bad = isnan(s) ? 1 : 0;
    }
    __syncthreads();
    for (int e = tid; e < q * d; e += CDF_THREADS) out[(b * q + e / d) + (int64_t)(e % d) * N] = bad ? 0.0 : G[e];
  }
}

}  // namespace dk
