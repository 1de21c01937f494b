// dk_qei.cuh — batched multipoint expected improvement (qEI) and its gradient,
// /root/reference/R/qEI.R:119-261 and R/qEI.grad.R:1-142,267-424.
//
// Value path, two kernels:
//   k_qei_moments : the fused DMMA forward solve of dk_kernels.cuh over tiles of 16 batches x 4 point
//                   slots, accumulating the Gram matrix V_a'V_b of each batch next to the solve, then
//                   predict.km(cov.compute=TRUE): mu (q) and the conditional covariance Sigma (q x q)
//                   = covMatrix(x) - V'V (+ S'S for UK) per batch (R/qEI.R:136-139), plus the duplicate
//                   mask of unique(round(rbind(X, x), 8)) (R/qEI.R:124-128).
//   k_qei_cdf     : per batch, the Chevalier-Ginsbourger formula: for each k the q-variate CDF p_k of
//                   Z^(k) (covZk, R/qEI.R:240-261) and the exact second term sum_i Sigma_k[i,k]
//                   phi_ik Phi_{q-1}(c_ik; Sigma_ik) (R/qEI.R:168-189).  Four lanes per batch (one per k),
//                   terms exchanged by warp shuffles, fixed summation order.
// Gradient path:
//   k_qei_grad_moments : the same tile, then krigingDeriv (R/qEI.grad.R:62-142) in adjoint form: one
//                   backward DMMA solve wt_l = T^-1 (v_l + M r_l) per point, and
//                     d m_k / d x_k        = dc_k' alpha + beta f.deltax
//                     d Sigma_kl / d x_k   = ker.grad(k, l) - dc_k' wt_l + f.deltax' r_l
//                   (the reference runs d triangular solves per point, W = backsolve(t(T), dc)).
//   k_qei_grad_cdf : R/qEI.grad.R:337-421 per batch: terms A1, A2 (GPhi / HPhi by decompo) and the exact
//                   B1 + B2 + B3; four lanes per batch, one per k, partial gradients summed in fixed order.
// The reference's fastCompute = TRUE replaces the second term (and term B) by finite differences
// (eps = 1e-5 / 1e-6) of sadmvn calls (abseps 1e-6); the engine returns the exact terms for both flags.
#pragma once
#include "dk_grad.cuh"
#include "dk_mvn.cuh"

namespace dk {

constexpr int QB = BN / 4;             // batches per tile (4 point slots each)
constexpr int MOM_STRIDE = 24;         // per batch: mu[4] | Sigma[16] row-major | keep mask | pad
constexpr int GRAM_DOUBLES = 4 * QB * 10;  // per CTA: [warp row][batch][10 products]
constexpr int QEI_AUX_DOUBLES = 1024;

__device__ __forceinline__ double round8(double v) { return rint(v * 1e8) / 1e8; }

// (a, b) -> slot in the 10 products a thread pair accumulates (see gram loop below)
__device__ __forceinline__ int gram_slot(int a, int b) {
  const int lo = a < b ? a : b, hi = a < b ? b : a;
  // (0,0)->0 (0,1)->1 (1,1)->2 (0,2)->3 (0,3)->4 (2,2)->5 (2,3)->6 (3,3)->7 (1,2)->8 (1,3)->9
  const int t[4][4] = {{0, 1, 3, 4}, {1, 2, 8, 9}, {3, 8, 5, 6}, {4, 9, 6, 7}};
  return t[lo][hi];
}

// covariance between two candidates of the tile (covMatrix off-diagonal entry)
__device__ __forceinline__ double cov_pair(const DevModel& M, const double* xs, int ca, int cb) {
  double s = 0.0, pr = 1.0;
  for (int k = 0; k < M.d; k++) {
    const double a = fabs(xs[k * BN + ca] - xs[k * BN + cb]) * M.w[k];
    switch (M.kernel) {
      case DK_GAUSS: s = fma(0.5 * a, a, s); break;
      case DK_EXP: s += a; break;
      case DK_MATERN3_2: pr *= (1.0 + a); s += a; break;
      case DK_MATERN5_2: pr *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0); s += a; break;
      default: s += pow(a, M.shape[k]); break;
    }
  }
  return M.sd2 * pr * exp(-s);
}

// d log k / d x_k for diff = x_k - X_k (covVector.dx)
__device__ __forceinline__ double dlogk_dev(const DevModel& M, int k, double diff) {
  const double wk = M.w[k], a = fabs(diff) * wk;
  const double sg = (diff > 0.0) ? 1.0 : ((diff < 0.0) ? -1.0 : 0.0);
  switch (M.kernel) {
    case DK_GAUSS: return -diff * wk * wk;
    case DK_EXP: return -sg * wk;
    case DK_MATERN3_2: return -sg * wk * a / (1.0 + a);
    case DK_MATERN5_2: return -sg * wk * (a * (1.0 / 3.0)) * (1.0 + a) / fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0);
    default: return (a > 0.0) ? -sg * M.shape[k] * pow(a, M.shape[k] - 1.0) * wk : 0.0;
  }
}

// Does candidate column c equal a design row after rounding to 8 digits (R/qEI.R:124-128)?  Rows slice, slice + 4, ... of the
// rounded design (Xrm_r8, rounded once per model); a row is dismissed on its first coordinate.  Eight rows' first coordinates
// are loaded together: one dependent L2 load per row made this scan a fifth of k_qei_moments' warp samples at n = 500.
__device__ __forceinline__ bool design_duplicate(const DevModel& M, const double* __restrict__ xs, int c, int slice) {
  const int n = M.n, d = M.d;
  const double x0 = round8(xs[c]);
  bool dup = false;
#pragma unroll 1
  for (int i0 = slice; i0 < n && !dup; i0 += 32) {
    double f[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = i0 + 4 * u;
      f[u] = (i < n) ? __ldg(M.Xrm_r8 + (size_t)i * d) : nan("");
    }
#pragma unroll
    for (int u = 0; u < 8; u++)
      if (f[u] == x0) {
        const int i = i0 + 4 * u;
        bool same = true;
        for (int k = 1; k < d && same; k++) same = (__ldg(M.Xrm_r8 + (size_t)i * d + k) == round8(xs[k * BN + c]));
        dup = dup || same;
      }
  }
  return dup;
}

// One tile of 16 batches: forward solve with Gram accumulation, then mu / Sigma / duplicate mask to
// `moments`.  When `coef` is non-NULL it also receives r = tuuinv (F.newdata - v'M)' per column (rows
// 2.. of the coefficient block, as the backward solve of the gradient kernels expects).
__device__ __forceinline__ void qei_tile_moments(const DevModel& M, int type, int q, const double* __restrict__ x,
                                                 int64_t B, int64_t b0, double* __restrict__ panel,
                                                 double* __restrict__ gram, double* __restrict__ moments,
                                                 double* __restrict__ coef, bool round_x, unsigned char* smem,
                                                 int tid, unsigned* stage_base, double* __restrict__ cpanel = nullptr) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  double* Cbuf = reinterpret_cast<double*>(smem);
  double* Xs = reinterpret_cast<double*>(smem + SM_XS_OFF);
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  const int d = M.d, n = M.n, p = M.p;
  const int64_t N = B * q;  // rows of x
  // column c = 4 g + a is point a of batch b0 + g; slots a >= q repeat point q - 1 (ignored later).
  // qEI evaluates the ROUNDED rows: x <- unique(round(rbind(X, x), 8)) (R/qEI.R:126-128); qEI.grad only
  // uses the rounding for its duplicate test (R/qEI.grad.R:292-295) and differentiates at x itself.
  for (int i = tid; i < d * BN; i += NT) {
    const int k = i / BN, c = i % BN;
    int64_t b = b0 + (c >> 2);
    if (b >= B) b = B - 1;
    const int a = min(c & 3, q - 1);
    const double xv = x[(b * q + a) + (int64_t)k * N];
    xs[i] = round_x ? round8(xv) : xv;
  }
  for (int i = tid; i < GRAM_DOUBLES; i += NT) gram[i] = 0.0;

#pragma unroll 1
  for (int I = 0; I < M.nblk; I++) {
    {
      const double* src = M.Xrm_s + (size_t)I * BM * d;  // design rows pre-scaled by w_k
      for (int i = tid; i < BM * d; i += NT) Xs[i] = __ldg(src + i);
    }
    __syncthreads();
    covgen_block(M, Xs, xs, Cbuf, I * BM, warp, lane);
    __syncthreads();
    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int ni = 0; ni < 4; ni++) {
        const double2 v = *reinterpret_cast<const double2*>(
            Cbuf + (32 * wm + 8 * mi + (lane >> 2)) * CB_STRIDE + 32 * wn + 8 * ni + 2 * (lane & 3));
        acc[mi][ni][0] = v.x;
        acc[mi][ni][1] = v.y;
        if (cpanel)
          *reinterpret_cast<double2*>(cpanel + (size_t)(I * BM + 32 * wm + 8 * mi + (lane >> 2)) * BN + 32 * wn + 8 * ni +
                                      2 * (lane & 3)) = v;
      }
    __syncthreads();
    solve_block_row<true>(acc, M.tiles_fwd, I, panel, smem, tid, stage_base);
    // ---- Gram products of each 4-column batch: a lane pair (lane ^ 1) covers one batch of an 8-column
    //      fragment; the even lane holds points (0,1), the odd lane (2,3); 5 products each. ----
    {
      const bool odd = lane & 1;
#pragma unroll
      for (int ni = 0; ni < 4; ni++) {
        double part[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          const bool live = (I * BM + solved_row(wm, mi, lane)) < n;
          const double o0 = live ? acc[mi][ni][0] : 0.0, o1 = live ? acc[mi][ni][1] : 0.0;
          const double p0 = __shfl_xor_sync(0xffffffffu, o0, 1), p1 = __shfl_xor_sync(0xffffffffu, o1, 1);
          part[0] = fma(o0, o0, part[0]);
          part[1] = fma(o0, o1, part[1]);
          part[2] = fma(o1, o1, part[2]);
          // even: (0,2) (0,3) = own0 * partner{0,1};  odd: (1,2) (1,3) = partner1 * own{0,1}
          part[3] = fma(odd ? p1 : o0, odd ? o0 : p0, part[3]);
          part[4] = fma(odd ? p1 : o0, odd ? o1 : p1, part[4]);
        }
#pragma unroll
        for (int j = 0; j < 5; j++) {
          double v = part[j];
          v += __shfl_xor_sync(0xffffffffu, v, 4);
          v += __shfl_xor_sync(0xffffffffu, v, 8);
          v += __shfl_xor_sync(0xffffffffu, v, 16);
          if (lane < 4) {
            const int g = 8 * wn + 2 * ni + (lane >> 1);
            gram[(wm * QB + g) * 10 + (odd ? 5 : 0) + j] += v;
          }
        }
      }
    }
    __syncthreads();
  }

  // ---- epilogue: duplicate test, mu, S, then Sigma per batch ----
  int* dupflag = reinterpret_cast<int*>(smem);                       // [64]
  double* mu_s = reinterpret_cast<double*>(smem + 1024);             // [64]
  double* S_s = reinterpret_cast<double*>(smem + 2048);              // [64][p]
  if (tid < BN) dupflag[tid] = 0;
  __syncthreads();
  {
    // 4 threads per candidate scan the design for a row equal after rounding to 8 digits
    const int c = tid & (BN - 1), slice = tid >> 6;
    if (design_duplicate(M, xs, c, slice)) atomicOr(&dupflag[c], 1);
  }
  if (tid < BN) {
    const int c = tid;
    double mean = 0.0, u[DK_MAX_P];
    for (int j = 0; j < p; j++) {
      const double fj = (j == 0) ? 1.0 : xs[(j - 1) * BN + c];
      mean = fma(fj, __ldg(M.beta + j), mean);
      u[j] = fj + panel_at(panel, n + 1 + j, c);
    }
    mu_s[c] = mean - panel_at(panel, n, c);
    if (type == DK_UK) {
      for (int j = 0; j < p; j++) {
        double t = u[j];
        for (int i = 0; i < j; i++) t -= __ldg(M.RM + i + j * p) * S_s[c * p + i];
        S_s[c * p + j] = t / __ldg(M.RM + j + j * p);
      }
    }
    if (coef) {
      for (int j = 0; j < p; j++) {
        double r = 0.0;
        if (type == DK_UK)
          for (int i = 0; i < p; i++) r = fma(__ldg(M.tuuinv + j + i * p), u[i], r);
        coef[(2 + j) * BN + c] = r;
      }
    }
  }
  __syncthreads();
  if (tid < BN) {
    const int c = tid, g = c >> 2, a = c & 3;
    const int64_t b = b0 + g;
    if (b < B && a < q) {
      double* mom = moments + b * MOM_STRIDE;
      mom[a] = mu_s[c];
      for (int bb = 0; bb < q; bb++) {
        const int cb = 4 * g + bb;
        const double kab = (a == bb) ? M.total_sd2 : cov_pair(M, xs, c, cb);
        const int sl = gram_slot(a, bb);
        const double gab = ((__ldcg(gram + (0 * QB + g) * 10 + sl) + __ldcg(gram + (1 * QB + g) * 10 + sl)) +
                            __ldcg(gram + (2 * QB + g) * 10 + sl)) + __ldcg(gram + (3 * QB + g) * 10 + sl);
        double sab = 0.0;
        if (type == DK_UK)
          for (int j = 0; j < p; j++) sab = fma(S_s[c * p + j], S_s[cb * p + j], sab);
        mom[4 + a * 4 + bb] = kab - gab + sab;
      }
      if (a == 0) {
        // unique(): a point is dropped if it equals a design row or an earlier point of the batch
        int mask = 0;
        for (int aa = 0; aa < q; aa++) {
          bool drop = dupflag[4 * g + aa] != 0;
          for (int e = 0; e < aa && !drop; e++) {
            bool same = true;
            for (int k = 0; k < d && same; k++) same = (round8(xs[k * BN + 4 * g + aa]) == round8(xs[k * BN + 4 * g + e]));
            drop = same;
          }
          if (!drop) mask |= 1 << aa;
        }
        mom[20] = (double)mask;
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(NT, 2)
k_qei_moments(const DevModel M, int type, int q, const double* __restrict__ x, int64_t B, int64_t ntiles,
              double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ auxscratch,
              double* __restrict__ moments) {
  extern __shared__ __align__(128) unsigned char smem[];
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* gram = auxscratch + (size_t)blockIdx.x * QEI_AUX_DOUBLES;
  pipeline_init(smem, threadIdx.x);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    qei_tile_moments(M, type, q, x, B, tile * QB, panel, gram, moments, nullptr, true, smem, threadIdx.x, &stage_base);
}

// covZk (R/qEI.R:240-261) on a q x q matrix stored with leading dimension 4
__device__ __forceinline__ void covzk_dev(int q, const double* S, int k, double* out) {
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++) {
      double v;
      if (i == k && j == k) v = S[k * 4 + k];
      else if (i == k) v = S[k * 4 + k] - S[k * 4 + j];
      else if (j == k) v = S[k * 4 + k] - S[k * 4 + i];
      else v = S[i * 4 + j] + S[k * 4 + k] - S[k * 4 + i] - S[k * 4 + j];
      out[i * 4 + j] = v;
    }
}

// Four lanes per batch (one per k): the throughput version, used when there are enough batches to fill the GPU.
#ifndef DK_CDF_MINB
#define DK_CDF_MINB 5
#endif
__global__ void __launch_bounds__(128, DK_CDF_MINB)
k_qei_cdf(const double* __restrict__ moments, int64_t B, int q, double plugin, int minimization,
          double* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t b = t >> 2;
  const int k = (int)(t & 3);
  const bool valid = b < B;
  const double* mom = moments + (valid ? b : 0) * MOM_STRIDE;
  const int mask = (int)mom[20];
  int idx[4], qq = 0;
  for (int a = 0; a < q; a++)
    if (mask & (1 << a)) idx[qq++] = a;
  for (int a = qq; a < 4; a++) idx[a] = 0;
  double mu[4], Sg[16];
  for (int i = 0; i < 4; i++) {
    mu[i] = mom[idx[i]];
    for (int j = 0; j < 4; j++) Sg[i * 4 + j] = mom[4 + idx[i] * 4 + idx[j]];
  }
  double term = 0.0;           // this lane's first_term[k] + sum_i sym[k,i] Sigma_k[i,k]
  double symrow[4] = {0.0, 0.0, 0.0, 0.0}, nonsym[4] = {0.0, 0.0, 0.0, 0.0};
  double first = 0.0, handed = 0.0;
  const bool active = valid && qq >= 2 && k < qq;
  // The second term needs sym[k][i] for i >= k: 4, 3, 2, 1 (q - 1)-variate CDFs on lanes 0..3.  Lane 0 hands its last
  // entry, sym[0][qq - 1], to the lane with the least work (qq - 1; lane 2 when qq = 2), so a group's lanes evaluate
  // 3, 3, 2, 2 of them: the warp runs three rounds instead of four.  The entry is computed from the same inputs by the same
  // code (one loop body), so the result does not depend on the lane that evaluates it.
  const int helper = (qq == 2) ? 2 : qq - 1;
  const bool helps = valid && qq >= 2 && k == helper;
  const int nown = active ? ((k == 0) ? qq - 1 : qq - k) : 0;
  if (active || helps) {
    double Sk[16], mk[4], bk[4], xk[4];
    auto setup = [&](int kk) {   // Z^(kk): covariance, mean and bounds (R/qEI.R:145-156)
      covzk_dev(qq, Sg, kk, Sk);
      for (int i = 0; i < qq; i++) {
        mk[i] = (i == kk) ? -mu[kk] : mu[i] - mu[kk];
        bk[i] = (i == kk) ? -plugin : 0.0;
        if (minimization) { mk[i] = -mk[i]; bk[i] = -bk[i]; }
        xk[i] = bk[i] - mk[i];
      }
    };
    if (active) {
      setup(k);
      const double pk = pmnorm_dev(qq, xk, Sk);
      first = (mu[k] - plugin) * pk;
      if (minimization) first = -first;
      for (int i = 0; i < qq; i++) nonsym[i] = Sk[i * 4 + k];
    }
    const int ntask = nown + (helps ? 1 : 0);
    for (int s = 0; s < ntask; s++) {
      const bool foreign = s >= nown;
      if (foreign) setup(0);
      const int i = foreign ? qq - 1 : k + s;
      const double sii = Sk[i * 4 + i], sdi = sqrt(sii);
      const double phi_ik = dnorm_std((bk[i] - mk[i]) / sdi) / sdi;
      // get_cik / get_sigmaik (R/qEI.R:196-227): condition Z^(k) on component i
      double cik[4], sik[16];
      int r = 0;
      for (int u = 0; u < qq; u++) {
        if (u == i) continue;
        cik[r] = (bk[u] - mk[u]) - (bk[i] - mk[i]) * (Sk[i * 4 + u] / sii);
        int cc = 0;
        for (int v = 0; v < qq; v++) {
          if (v == i) continue;
          sik[r * 4 + cc] = Sk[u * 4 + v] - Sk[u * 4 + i] * Sk[v * 4 + i] / sii;
          cc++;
        }
        r++;
      }
      const double val = phi_ik * pmnorm_dev(qq - 1, cik, sik);
      if (foreign) handed = val;
      else symrow[i] = val;
    }
  }
  {  // sym[0][qq - 1] back to lane 0 of the group
    const double hv = __shfl_sync(0xffffffffu, handed, (threadIdx.x & 28) + (helper & 3));
    if (active && k == 0) symrow[qq - 1] = hv;
  }
  // exchange the upper-triangular sym entries inside the 4-lane group: sym[k][i] = sym[i][k] for i < k
  const int base = threadIdx.x & 28;
  double full[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    // value held by lane i of the group for column k (k is per-lane: fetch all four columns, then pick)
    double col[4];
#pragma unroll
    for (int j = 0; j < 4; j++) col[j] = __shfl_sync(0xffffffffu, symrow[j], base + i);
    full[i] = (i < k) ? col[k & 3] : symrow[i];
  }
  if (active) {
    double second = 0.0;
    for (int i = 0; i < qq; i++) second = fma(full[i], nonsym[i], second);
    term = first + second;
  }
  const double t0 = __shfl_sync(0xffffffffu, term, base + 0), t1 = __shfl_sync(0xffffffffu, term, base + 1);
  const double t2 = __shfl_sync(0xffffffffu, term, base + 2), t3 = __shfl_sync(0xffffffffu, term, base + 3);
  if (valid && k == 0) {
    double res;
    if (qq == 0) {
      res = 0.0;                                   // every point duplicates the design (R/qEI.R:127)
    } else if (qq == 1) {
      // EI(x, model, plugin, type, envir): minimization takes its default TRUE (R/qEI.R:132-134)
      const double sd = sqrt(fmax(Sg[0], 0.0)), diff = plugin - mu[0], xcr = diff / sd;
      res = diff * normcdf(xcr) + sd * dnorm_std(xcr);
    } else {
      res = ((t0 + t1) + t2) + t3;
    }
    out[b] = res;
  }
}


// One warp per batch: lane = 8 k + s (the low-latency version for small B).  The 8 lanes of group k share the q-variate CDF p_k (path-integral nodes dealt
// round-robin, dk_mvn.cuh) and lane (k, s) evaluates the (q-1)-variate CDF of the second-term entry i = k + s.
__global__ void __launch_bounds__(128)
k_qei_cdf_warp(const double* __restrict__ moments, int64_t B, int q, double plugin, int minimization,
               double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int k = lane >> 3, s = lane & 7;
  const bool valid = b < B;
  const double* mom = moments + (valid ? b : 0) * MOM_STRIDE;
  const int mask = (int)mom[20];
  int idx[4], qq = 0;
  for (int a = 0; a < q; a++)
    if (mask & (1 << a)) idx[qq++] = a;
  for (int a = qq; a < 4; a++) idx[a] = 0;
  double mu[4], Sg[16];
  for (int i = 0; i < 4; i++) {
    mu[i] = mom[idx[i]];
    for (int j = 0; j < 4; j++) Sg[i * 4 + j] = mom[4 + idx[i] * 4 + idx[j]];
  }
  const bool active = valid && qq >= 2 && k < qq;
  double Sk[16], mk[4], bk[4], xk[4];
  for (int i = 0; i < 16; i++) Sk[i] = 0.0;
  for (int i = 0; i < 4; i++) mk[i] = bk[i] = xk[i] = 0.0;
  if (active) {
    covzk_dev(qq, Sg, k, Sk);
    for (int i = 0; i < qq; i++) {
      mk[i] = (i == k) ? -mu[k] : mu[i] - mu[k];
      bk[i] = (i == k) ? -plugin : 0.0;
      if (minimization) { mk[i] = -mk[i]; bk[i] = -bk[i]; }
      xk[i] = bk[i] - mk[i];
    }
  }
  // qq is the same for the whole warp (one batch): every lane runs the cooperative CDF, inactive groups idle inside
  const double pk = (valid && qq >= 2) ? pmnorm_coop(active, qq, xk, Sk, s) : 0.0;
  double symval = 0.0;  // sym[k][k + s]
  const int i = k + s;
  if (active && i < qq) {
    const double sii = Sk[i * 4 + i], sdi = sqrt(sii);
    const double phi_ik = dnorm_std((bk[i] - mk[i]) / sdi) / sdi;
    double cik[4], sik[16];
    int r = 0;
    for (int u = 0; u < qq; u++) {
      if (u == i) continue;
      cik[r] = (bk[u] - mk[u]) - (bk[i] - mk[i]) * (Sk[i * 4 + u] / sii);
      int cc = 0;
      for (int v = 0; v < qq; v++) {
        if (v == i) continue;
        sik[r * 4 + cc] = Sk[u * 4 + v] - Sk[u * 4 + i] * Sk[v * 4 + i] / sii;
        cc++;
      }
      r++;
    }
    symval = phi_ik * pmnorm_dev(qq - 1, cik, sik);
  }
  // group leader k gathers sym[k][i] (i >= k: lane (k, i - k)) and sym[i][k] (i < k: lane (i, k - i))
  double second = 0.0;
#pragma unroll
  for (int ii = 0; ii < 4; ii++) {
    const int lo = ii < k ? ii : k, hi = ii < k ? k : ii;
    const double v = __shfl_sync(0xffffffffu, symval, lo * 8 + (hi - lo));
    if (active && ii < qq) second = fma(v, Sk[ii * 4 + k], second);
  }
  double term = 0.0;
  if (active) {
    double first = (mu[k] - plugin) * pk;
    if (minimization) first = -first;
    term = first + second;
  }
  const double t0 = __shfl_sync(0xffffffffu, term, 0), t1 = __shfl_sync(0xffffffffu, term, 8);
  const double t2 = __shfl_sync(0xffffffffu, term, 16), t3 = __shfl_sync(0xffffffffu, term, 24);
  if (valid && lane == 0) {
    double res;
    if (qq == 0) {
      res = 0.0;                                   // every point duplicates the design (R/qEI.R:127)
    } else if (qq == 1) {
      // EI(x, model, plugin, type, envir): minimization takes its default TRUE (R/qEI.R:132-134)
      const double sd = sqrt(fmax(Sg[0], 0.0)), diff = plugin - mu[0], xcr = diff / sd;
      res = diff * normcdf(xcr) + sd * dnorm_std(xcr);
    } else {
      res = ((t0 + t1) + t2) + t3;
    }
    out[b] = res;
  }
}

// test hook kernel behind dk_mvn_cdf
__global__ void k_mvn_cdf_test(int q, int64_t B, const double* __restrict__ h, const double* __restrict__ S,
                               double* __restrict__ out) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double x[4], Sg[16];
  for (int i = 0; i < q; i++) {
    x[i] = h[b * q + i];
    for (int j = 0; j < q; j++) Sg[i * 4 + j] = S[(b * q + i) * q + j];
  }
  out[b] = pmnorm_dev(q, x, Sg);
}

// =========================================================================================
// qEI.grad
// =========================================================================================
// per batch: mj[k][j] = d m_k / d x_kj (4 x d) followed by cj[k][l][j] = one-sided d Sigma_kl / d x_kj (16 x d)
__host__ __device__ inline size_t qei_deriv_stride(int d) { return (size_t)20 * d; }

// The q + 1 derivative contractions of one batch point c (columns 4g .. 4g+3 hold its batch), dims k0 .. k0+3,
// in ONE sweep over the design rows i = slice, slice + 4, ...:
//   acc[0][kk]     += c_i alpha_i        dlogk_k(i)     (d m / d x, R/qEI.grad.R:128)
//   acc[1 + l][kk] += c_i wt_i[4g + l]   dlogk_k(i)     (dc' wt_l of d Sigma_kl / d x, R/qEI.grad.R:130-137)
// Same per-accumulator order of additions as a sweep per weight vector, so the results do not depend on the fusion.
template <int KERNEL>
__device__ __forceinline__ void qei_deriv_accumulate(const DevModel& M, const double* __restrict__ xs,
                                                     const double* __restrict__ cpanel,
                                                     const double* __restrict__ wpanel, int c, int slice, int q, int k0,
                                                     double (&acc)[5][4]) {
  const int n = M.n, d = M.d, g4 = c & ~3;
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
    om[0] = ci * __ldg(M.alpha + i);
#pragma unroll
    for (int l = 0; l < 4; l++) om[1 + l] = (l < q) ? ci * panel_at(wpanel, n - 1 - i, g4 + l) : 0.0;
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
k_qei_grad_moments(const DevModel M, int type, int q, const double* __restrict__ x, int64_t B, int64_t ntiles,
                   double* __restrict__ vscratch, double* __restrict__ wscratch, double* __restrict__ cscratch,
                   size_t panel_doubles, double* __restrict__ auxscratch, double* __restrict__ moments,
                   double* __restrict__ deriv) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;
  double* gram = auxscratch + (size_t)blockIdx.x * QEI_AUX_DOUBLES;
  double* cpanel = cscratch + (size_t)blockIdx.x * panel_doubles;
  const int n = M.n, d = M.d, p = M.p;
  const size_t dstride = qei_deriv_stride(d);
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b0 = tile * QB;
    qei_tile_moments(M, type, q, x, B, b0, vpanel, gram, moments, coef, false, smem, tid, &stage_base, cpanel);

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
      solve_block_row<true>(acc, M.tiles_bwd, I, wpanel, smem, tid, &stage_base);
      __syncthreads();
    }

    // ---- derivative contractions: weights alpha (mean) and wt of each point l of the batch, four dims per sweep ----
    const int c = tid & (BN - 1), slice = tid >> 6, g = c >> 2, kpt = c & 3;
    double* gbuf = reinterpret_cast<double*>(smem);  // [4 slices][5 weights][4 dims][64]
    const int64_t b = b0 + g;
#pragma unroll 1
    for (int k0 = 0; k0 < d; k0 += 4) {
      double acc[5][4];
#pragma unroll
      for (int ps = 0; ps < 5; ps++)
#pragma unroll
        for (int kk = 0; kk < 4; kk++) acc[ps][kk] = 0.0;
      switch (M.kernel) {
        case DK_GAUSS: qei_deriv_accumulate<DK_GAUSS>(M, xs, cpanel, wpanel, c, slice, q, k0, acc); break;
        case DK_EXP: qei_deriv_accumulate<DK_EXP>(M, xs, cpanel, wpanel, c, slice, q, k0, acc); break;
        case DK_MATERN3_2: qei_deriv_accumulate<DK_MATERN3_2>(M, xs, cpanel, wpanel, c, slice, q, k0, acc); break;
        case DK_MATERN5_2: qei_deriv_accumulate<DK_MATERN5_2>(M, xs, cpanel, wpanel, c, slice, q, k0, acc); break;
        default: qei_deriv_accumulate<DK_POWEXP>(M, xs, cpanel, wpanel, c, slice, q, k0, acc); break;
      }
#pragma unroll
      for (int ps = 0; ps < 5; ps++)
#pragma unroll
        for (int kk = 0; kk < 4; kk++) gbuf[((slice * 5 + ps) * 4 + kk) * BN + c] = acc[ps][kk];
      __syncthreads();
      // thread (c, slice) finishes dim k0 + slice for all weights
      const int k = k0 + slice;
      if (b < B && kpt < q && k < d) {
        double* dv = deriv + b * dstride;
        for (int ps = 0; ps <= q; ps++) {
          const double s = ((gbuf[((0 * 5 + ps) * 4 + slice) * BN + c] + gbuf[((1 * 5 + ps) * 4 + slice) * BN + c]) +
                            gbuf[((2 * 5 + ps) * 4 + slice) * BN + c]) + gbuf[((3 * 5 + ps) * 4 + slice) * BN + c];
          if (ps == 0) {
            double v = s;  // dc_k' alpha
            if (M.trend == DK_TREND_LINEAR) v += __ldg(M.beta + 1 + k);
            dv[kpt * d + k] = v;
          } else {
            // ker.grad(k, l) - dc_k' wt_l + f.deltax' r_l
            const int l = ps - 1;
            double v = -s;
            if (l != kpt) v += cov_pair(M, xs, c, 4 * g + l) * dlogk_dev(M, k, xs[k * BN + c] - xs[k * BN + 4 * g + l]);
            if (M.trend == DK_TREND_LINEAR) v += __ldcg(coef + (2 + 1 + k) * BN + 4 * g + l);
            dv[(4 + kpt * 4 + l) * d + k] = v;
          }
        }
      }
      __syncthreads();
    }
  }
}

// ---- small dense helpers on matrices with leading dimension 4 ----
__device__ __forceinline__ double dmnorm1(double x, double mu, double s2) {
  const double sd = sqrt(s2);
  return dnorm_std((x - mu) / sd) / sd;
}
// density of the components (i, j) at x (bivariate), mean mu, covariance from S
__device__ __forceinline__ double dmnorm2(double xi, double xj, double sii, double sij, double sjj) {
  const double det = sii * sjj - sij * sij;
  const double qf = (sjj * xi * xi - 2.0 * sij * xi * xj + sii * xj * xj) / det;
  return exp(-0.5 * qf) / (DK_TWOPI * sqrt(det));
}

// decompo(x, mu, Sigma, k = {i}) (R/qEI.grad.R:1-26): density of component i times the CDF of the rest given it
static __device__ __noinline__ double decompo1(int q, const double* x, const double* mu, const double* S, int i) {
  const double sii = S[i * 4 + i], dens = dmnorm1(x[i], mu[i], sii);
  if (q == 1) return dens;
  double x2[4], V[16];
  int r = 0;
#pragma unroll 1
  for (int u = 0; u < q; u++) {
    if (u == i) continue;
    const double moy = mu[u] + S[u * 4 + i] / sii * (x[i] - mu[i]);
    x2[r] = x[u] - moy;
    int cc = 0;
#pragma unroll 1
    for (int v = 0; v < q; v++) {
      if (v == i) continue;
      // 0.5 (varcov + t(varcov)) with varcov = S22 - S21 S11^-1 S12
      const double a = S[u * 4 + v] - S[u * 4 + i] * S[v * 4 + i] / sii;
      const double b = S[v * 4 + u] - S[v * 4 + i] * S[u * 4 + i] / sii;
      V[r * 4 + cc] = 0.5 * (a + b);
      cc++;
    }
    r++;
  }
  double mn = V[0];
#pragma unroll 1
  for (int u = 1; u < q - 1; u++) mn = fmin(mn, V[u * 4 + u]);
  if (mn <= 0.0)
#pragma unroll 1
    for (int u = 0; u < q - 1; u++) V[u * 4 + u] = (V[u * 4 + u] >= 0.0 ? V[u * 4 + u] : 0.0) + 1e-9;
  return dens * pmnorm_dev(q - 1, x2, V);
}

// decompo(x, mu, Sigma, k = {i, j}), i < j
static __device__ __noinline__ double decompo2(int q, const double* x, const double* mu, const double* S, int i, int j) {
  const double sii = S[i * 4 + i], sij = S[i * 4 + j], sjj = S[j * 4 + j];
  const double di = x[i] - mu[i], dj = x[j] - mu[j];
  const double dens = dmnorm2(di, dj, sii, sij, sjj);
  if (q == 2) return dens;
  const double det = sii * sjj - sij * sij;
  // S11^-1 = [[sjj, -sij], [-sij, sii]] / det
  double x2[4], V[16];
  int r = 0;
#pragma unroll 1
  for (int u = 0; u < q; u++) {
    if (u == i || u == j) continue;
    const double a_ui = S[u * 4 + i], a_uj = S[u * 4 + j];
    const double wi = (a_ui * sjj - a_uj * sij) / det, wj = (-a_ui * sij + a_uj * sii) / det;  // S21 S11^-1 (row u)
    x2[r] = x[u] - (mu[u] + wi * di + wj * dj);
    int cc = 0;
#pragma unroll 1
    for (int v = 0; v < q; v++) {
      if (v == i || v == j) continue;
      const double b_vi = S[v * 4 + i], b_vj = S[v * 4 + j];
      const double wvi = (b_vi * sjj - b_vj * sij) / det, wvj = (-b_vi * sij + b_vj * sii) / det;
      const double a = S[u * 4 + v] - (wi * b_vi + wj * b_vj);
      const double b = S[v * 4 + u] - (wvi * a_ui + wvj * a_uj);
      V[r * 4 + cc] = 0.5 * (a + b);
      cc++;
    }
    r++;
  }
  double mn = V[0];
#pragma unroll 1
  for (int u = 1; u < q - 2; u++) mn = fmin(mn, V[u * 4 + u]);
  if (mn <= 0.0)
#pragma unroll 1
    for (int u = 0; u < q - 2; u++) V[u * 4 + u] = (V[u * 4 + u] >= 0.0 ? V[u * 4 + u] : 0.0) + 1e-9;
  return dens * pmnorm_dev(q - 2, x2, V);
}

// GPhi / HPhi (R/qEI.grad.R:29-60): gradient and Hessian of the q-variate normal CDF
__device__ __forceinline__ void gphi_dev(int q, const double* x, const double* mu, const double* S, double* g) {
#pragma unroll 1
  for (int i = 0; i < q; i++) g[i] = decompo1(q, x, mu, S, i);
}
__device__ __forceinline__ void hphi_dev(int q, const double* x, const double* mu, const double* S, const double* g,
                                         double* H) {
  if (q == 1) {
    H[0] = -(x[0] - mu[0]) / S[0] * dmnorm1(x[0], mu[0], S[0]);
    return;
  }
#pragma unroll 1
  for (int i = 0; i < q; i++)
#pragma unroll 1
    for (int j = 0; j < q; j++) H[i * 4 + j] = 0.0;
#pragma unroll 1
  for (int i = 0; i < q - 1; i++)
#pragma unroll 1
    for (int j = i + 1; j < q; j++) H[i * 4 + j] = H[j * 4 + i] = decompo2(q, x, mu, S, i, j);
#pragma unroll 1
  for (int i = 0; i < q; i++) {
    double sr = 0.0;
#pragma unroll 1
    for (int j = 0; j < q; j++)
      if (j != i) sr += S[i * 4 + j] * H[j * 4 + i];
    H[i * 4 + i] = -((x[i] - mu[i]) * g[i] + sr) / S[i * 4 + i];
  }
}

__global__ void __launch_bounds__(128)
k_qei_grad_cdf(const double* __restrict__ moments, const double* __restrict__ deriv, int64_t B, int q, int d,
               double plugin, int minimization, double sd2, double* __restrict__ out /* (B*q) x d column-major */) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t b = t >> 2;
  const int k = (int)(t & 3);
  const bool valid = b < B;
  const double* mom = moments + (valid ? b : 0) * MOM_STRIDE;
  const double* dv = deriv + (valid ? b : 0) * qei_deriv_stride(d);
  const int mask = (int)mom[20];
  const double sgn = minimization ? 1.0 : -1.0;
  double kmean[4], kcov[16];
#pragma unroll 1
  for (int i = 0; i < q; i++) {
    kmean[i] = sgn * mom[i];
#pragma unroll 1
    for (int j = 0; j < q; j++) kcov[i * 4 + j] = mom[4 + i * 4 + j];
  }
  // zero matrix when any point is a duplicate (R/qEI.grad.R:293-295) or all variances vanish (:107)
  bool alltiny = true;
#pragma unroll 1
  for (int i = 0; i < q; i++) alltiny = alltiny && (kcov[i * 4 + i] / sd2 < 1e-8);
  const bool active = valid && k < q && mask == (1 << q) - 1 && !alltiny;

  double G[4][DK_MAX_D];
#pragma unroll 1
  for (int l = 0; l < 4; l++)
#pragma unroll 1
    for (int j = 0; j < DK_MAX_D; j++) G[l][j] = 0.0;

  if (active) {
    // Lk = -I with column k set to one; mk = Lk kmean; Sigk = Lk kcov Lk'
    double Lk[16], mk[4], Sigk[16], bk[4], zero4[4] = {0.0, 0.0, 0.0, 0.0}, xs_[4];
#pragma unroll 1
    for (int a = 0; a < q; a++)
#pragma unroll 1
      for (int c = 0; c < q; c++) Lk[a * 4 + c] = (c == k) ? 1.0 : ((a == c) ? -1.0 : 0.0);
#pragma unroll 1
    for (int a = 0; a < q; a++) {
      double s = 0.0;
#pragma unroll 1
      for (int c = 0; c < q; c++) s += Lk[a * 4 + c] * kmean[c];
      mk[a] = s;
      bk[a] = (a == k) ? plugin : 0.0;
    }
    {
      double tmp[16];
#pragma unroll 1
      for (int a = 0; a < q; a++)
#pragma unroll 1
        for (int c = 0; c < q; c++) {
          double s = 0.0;
#pragma unroll 1
          for (int e = 0; e < q; e++) s += Lk[a * 4 + e] * kcov[e * 4 + c];
          tmp[a * 4 + c] = s;
        }
      double raw[16];
#pragma unroll 1
      for (int a = 0; a < q; a++)
#pragma unroll 1
        for (int c = 0; c < q; c++) {
          double s = 0.0;
#pragma unroll 1
          for (int e = 0; e < q; e++) s += tmp[a * 4 + e] * Lk[c * 4 + e];
          raw[a * 4 + c] = s;
        }
#pragma unroll 1
      for (int a = 0; a < q; a++)
#pragma unroll 1
        for (int c = 0; c < q; c++) Sigk[a * 4 + c] = 0.5 * (raw[a * 4 + c] + raw[c * 4 + a]);
    }
#pragma unroll 1
    for (int a = 0; a < q; a++) xs_[a] = bk[a] - mk[a];
    const double pk = pmnorm_dev(q, xs_, Sigk);
    double gradpk[4], hesspk[16];
    gphi_dev(q, xs_, zero4, Sigk, gradpk);          // GPhi(bk - mk, b = 0, Sigk)
    hphi_dev(q, bk, mk, Sigk, gradpk, hesspk);      // HPhi(bk, mk, Sigk, gradpk)

    // per-i quantities of the exact term B (R/qEI.grad.R:379-396)
    double phi_i[4], dphi_dS[4], dphi_dm[4], Phi_i[4], Gik[4][4], Hik[4][16];
#pragma unroll 1
    for (int i = 0; i < q; i++) {
      const double S_ii = Sigk[i * 4 + i], dm = bk[i] - mk[i];
      phi_i[i] = dmnorm1(bk[i], mk[i], S_ii);
      dphi_dS[i] = (dm * dm / (2.0 * S_ii * S_ii) - 0.5 / S_ii) * phi_i[i];
      dphi_dm[i] = dm / S_ii * phi_i[i];
      double ck[4], Sp[16];
      int r = 0;
#pragma unroll 1
      for (int u = 0; u < q; u++) {
        if (u == i) continue;
        ck[r] = bk[u] - mk[u] - dm / S_ii * Sigk[u * 4 + i];
        int cc = 0;
#pragma unroll 1
        for (int v = 0; v < q; v++) {
          if (v == i) continue;
          const double a1 = Sigk[u * 4 + v] - Sigk[u * 4 + i] * Sigk[v * 4 + i] / S_ii;
          const double a2 = Sigk[v * 4 + u] - Sigk[v * 4 + i] * Sigk[u * 4 + i] / S_ii;
          Sp[r * 4 + cc] = 0.5 * (a1 + a2);
          cc++;
        }
        r++;
      }
      Phi_i[i] = pmnorm_dev(q - 1, ck, Sp);
      gphi_dev(q - 1, ck, zero4, Sp, Gik[i]);
      hphi_dev(q - 1, ck, zero4, Sp, Gik[i], Hik[i]);
    }

    const double fac = plugin - kmean[k];
    double inv_sii[4];
#pragma unroll 1
    for (int i = 0; i < q; i++) inv_sii[i] = 1.0 / Sigk[i * 4 + i];
#pragma unroll 1
    for (int l = 0; l < q; l++) {
#pragma unroll 1
      for (int j = 0; j < d; j++) {
        // J = symmetrised cov.jacob[l, j, , ]: row/column l carry the one-sided derivatives, (l, l) twice
        double J[16];
#pragma unroll 1
        for (int a = 0; a < q; a++)
#pragma unroll 1
          for (int c = 0; c < q; c++) J[a * 4 + c] = 0.0;
#pragma unroll 1
        for (int i = 0; i < q; i++) {
          const double v = dv[(4 + l * 4 + i) * d + j];
          if (i == l) J[l * 4 + l] = 2.0 * v;
          else { J[l * 4 + i] = v; J[i * 4 + l] = v; }
        }
        const double mjl = sgn * dv[l * d + j];      // mean.jacob[l, j, l]
        double Sdx[16], mdx[4];
        {
          double tmp[16];
#pragma unroll 1
          for (int a = 0; a < q; a++)
#pragma unroll 1
            for (int c = 0; c < q; c++) {
              double s = 0.0;
#pragma unroll 1
              for (int e = 0; e < q; e++) s += Lk[a * 4 + e] * J[e * 4 + c];
              tmp[a * 4 + c] = s;
            }
#pragma unroll 1
          for (int a = 0; a < q; a++)
#pragma unroll 1
            for (int c = 0; c < q; c++) {
              double s = 0.0;
#pragma unroll 1
              for (int e = 0; e < q; e++) s += tmp[a * 4 + e] * Lk[c * 4 + e];
              Sdx[a * 4 + c] = s;
            }
        }
#pragma unroll 1
        for (int a = 0; a < q; a++) mdx[a] = Lk[a * 4 + l] * mjl;
        // A2: Dpk = 0.5 sum(hesspk * Sdx) - gradpk' mdx
        double hs = 0.0, gm = 0.0;
#pragma unroll 1
        for (int a = 0; a < q; a++) {
          gm += gradpk[a] * mdx[a];
#pragma unroll 1
          for (int c = 0; c < q; c++) hs += hesspk[a * 4 + c] * Sdx[a * 4 + c];
        }
        double acc = fac * (0.5 * hs - gm);
        if (l == k) acc -= mjl * pk;                 // A1
        // B1 + B2 + B3
#pragma unroll 1
        for (int i = 0; i < q; i++) {
          const double S_ik = Sigk[i * 4 + k], S_ii = Sigk[i * 4 + i], dm = bk[i] - mk[i];
          const double sdx_ii = Sdx[i * 4 + i];
          const double B1 = Sdx[i * 4 + k] * phi_i[i] * Phi_i[i];
          const double B2 = S_ik * (mdx[i] * dphi_dm[i] + dphi_dS[i] * sdx_ii) * Phi_i[i];
          // the divisions by S_ii of R/qEI.grad.R:398-412 as products with 1 / S_ii (formed once per i above)
          const double is = inv_sii[i], is2 = is * is;
          const double dc_a = (mdx[i] * S_ii + dm * sdx_ii) * is2, dc_b = dm * is, ds_a = sdx_ii * is2;
          double gd = 0.0, hd = 0.0;
          int r = 0;
#pragma unroll 1
          for (int u = 0; u < q; u++) {
            if (u == i) continue;
            const double S_ui = Sigk[u * 4 + i];
            const double dck = -mdx[u] + dc_a * S_ui - dc_b * Sdx[u * 4 + i];
            gd += Gik[i][r] * dck;
            int cc = 0;
#pragma unroll 1
            for (int v = 0; v < q; v++) {
              if (v == i) continue;
              const double S_vi = Sigk[v * 4 + i];
              const double dS = Sdx[u * 4 + v] + ds_a * S_ui * S_vi - (Sdx[u * 4 + i] * S_vi + Sdx[v * 4 + i] * S_ui) * is;
              hd += Hik[i][r * 4 + cc] * dS;
              cc++;
            }
            r++;
          }
          const double B3 = S_ik * phi_i[i] * (gd + 0.5 * hd);
          acc += B1 + B2 + B3;
        }
        G[l][j] = acc;
      }
    }
  }
  // sum the four lanes' partial gradients in fixed order; NaN anywhere -> zero matrix (R/qEI.grad.R:422)
  const int base = threadIdx.x & 28;
  double nanacc = 0.0;
  for (int l = 0; l < q; l++)
    for (int j = 0; j < d; j++) {
      const double v = G[l][j];
      const double s = ((__shfl_sync(0xffffffffu, v, base) + __shfl_sync(0xffffffffu, v, base + 1)) +
                        __shfl_sync(0xffffffffu, v, base + 2)) + __shfl_sync(0xffffffffu, v, base + 3);
      G[l][j] = s;
      nanacc += s;
    }
  if (valid && k == 0) {
    const bool bad = nanacc != nanacc;
    const int64_t N = B * q;
    for (int l = 0; l < q; l++)
      for (int j = 0; j < d; j++) out[(b * q + l) + (int64_t)j * N] = bad ? 0.0 : G[l][j];
  }
}

}  // namespace dk
