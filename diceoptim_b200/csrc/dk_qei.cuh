// dk_qei.cuh — batched multipoint expected improvement (qEI), /root/reference/R/qEI.R:119-261.
//
// Two kernels:
//   k_qei_moments : the fused DMMA forward solve of dk_kernels.cuh over tiles of 16 batches x 4 point
//                   slots, accumulating the Gram matrix V_a'V_b of each batch next to the solve, then
//                   predict.km(cov.compute=TRUE): mu (q) and the conditional covariance Sigma (q x q)
//                   = covMatrix(x) - V'V (+ S'S for UK) per batch (R/qEI.R:136-139), plus the duplicate
//                   mask of unique(round(rbind(X, x), 8)) (R/qEI.R:124-128).
//   k_qei_cdf     : per batch, the Chevalier-Ginsbourger formula: for each k the q-variate CDF p_k of
//                   Z^(k) (covZk, R/qEI.R:240-261) and the exact second term sum_i Sigma_k[i,k]
//                   phi_ik Phi_{q-1}(c_ik; Sigma_ik) (R/qEI.R:168-189).  Four lanes per batch (one per k),
//                   terms exchanged by warp shuffles, fixed summation order.
// The reference's fastCompute = TRUE replaces the second term by a finite difference (eps = 1e-5) of
// two sadmvn calls (abseps 1e-6); the engine returns the exact term for both flag values.
#pragma once
#include "dk_kernels.cuh"
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

__global__ void __launch_bounds__(NT, 2)
k_qei_moments(const DevModel M, int type, int q, const double* __restrict__ x, int64_t B, int64_t ntiles,
              double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ auxscratch,
              double* __restrict__ moments) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  double* Cbuf = reinterpret_cast<double*>(smem);
  double* Xs = reinterpret_cast<double*>(smem + SM_XS_OFF);
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* gram = auxscratch + (size_t)blockIdx.x * QEI_AUX_DOUBLES;
  const int d = M.d, n = M.n, p = M.p;
  const int64_t N = B * q;  // rows of x

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b0 = tile * QB;
    // column c = 4 g + a is point a of batch b0 + g; slots a >= q repeat point q - 1 (ignored later).
    // The reference evaluates the ROUNDED rows: x <- unique(round(rbind(X, x), 8)) (R/qEI.R:126-128).
    for (int i = tid; i < d * BN; i += NT) {
      const int k = i / BN, c = i % BN;
      int64_t b = b0 + (c >> 2);
      if (b >= B) b = B - 1;
      const int a = min(c & 3, q - 1);
      xs[i] = round8(x[(b * q + a) + (int64_t)k * N]);
    }
    for (int i = tid; i < GRAM_DOUBLES; i += NT) gram[i] = 0.0;

#pragma unroll 1
    for (int I = 0; I < M.nblk; I++) {
      {
        const double* src = M.Xrm + (size_t)I * BM * d;
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
        }
      __syncthreads();
      solve_block_row(acc, M.tiles_fwd, I, panel, smem, tid);
      // ---- Gram products of each 4-column batch: a lane pair (lane ^ 1) covers one batch of an 8-column
      //      fragment; the even lane holds points (0,1), the odd lane (2,3); 5 products each. ----
      {
        const bool odd = lane & 1;
#pragma unroll
        for (int ni = 0; ni < 4; ni++) {
          double part[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
          for (int mi = 0; mi < 4; mi++) {
            const bool live = (I * BM + 32 * wm + 8 * mi + (lane >> 2)) < n;
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
      bool dup = false;
      for (int i = slice; i < n && !dup; i += 4) {
        bool same = true;
        for (int k = 0; k < d && same; k++) same = (round8(__ldg(M.Xrm + (size_t)i * d + k)) == xs[k * BN + c]);
        dup = same;
      }
      if (dup) atomicOr(&dupflag[c], 1);
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
              for (int k = 0; k < d && same; k++) same = (xs[k * BN + 4 * g + aa] == xs[k * BN + 4 * g + e]);
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

__global__ void __launch_bounds__(128)
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
  double first = 0.0;
  const bool active = valid && qq >= 2 && k < qq;
  if (active) {
    double Sk[16], mk[4], bk[4], xk[4];
    covzk_dev(qq, Sg, k, Sk);
    for (int i = 0; i < qq; i++) {
      mk[i] = (i == k) ? -mu[k] : mu[i] - mu[k];
      bk[i] = (i == k) ? -plugin : 0.0;
      if (minimization) { mk[i] = -mk[i]; bk[i] = -bk[i]; }
      xk[i] = bk[i] - mk[i];
    }
    const double pk = pmnorm_dev(qq, xk, Sk);
    first = (mu[k] - plugin) * pk;
    if (minimization) first = -first;
    for (int i = k; i < qq; i++) {
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
      symrow[i] = phi_ik * pmnorm_dev(qq - 1, cik, sik);
    }
    for (int i = 0; i < qq; i++) nonsym[i] = Sk[i * 4 + k];
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

}  // namespace dk
