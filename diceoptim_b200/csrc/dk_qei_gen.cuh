// dk_qei_gen.cuh — qEI for batches of 5 .. DK_MAX_Q_LATTICE points (/root/reference/R/qEI.R:119-261 takes any q; its own
// example runs q = 10 against Monte-Carlo, R/qEI.R:82-104).
//
// The q <= 4 path (dk_qei.cuh) packs four point slots per batch and evaluates the normal CDFs by deterministic
// bivariate / Plackett rules.  Beyond that:
//   k_qei_moments_gen : the fused forward solve over tiles of floor(64 / q) batches (the batch's rows are ordinary
//                       candidate columns), then predict.km(cov.compute = TRUE) per batch: mu (q) and
//                       Sigma = covMatrix(x) - V'V (+ S'S) (q x q, the Gram products read back from the solved panel),
//                       and the duplicate mask of unique(round(rbind(X, x), 8)).
//   k_qei_cdf_gen     : one CTA per batch, the Chevalier-Ginsbourger sums with every q'-variate normal CDF
//                       (mnormt::pmnorm -> Genz's sadmvn in the reference) evaluated by Genz's separation-of-variables
//                       transform on a FIXED lattice, all threads of the CTA sharing the lattice points; q' <= 4 falls
//                       back to the deterministic rules, one CDF per thread.  oracle/mvnorm.py runs the same rule.
#pragma once
#include "dk_qei.cuh"

namespace dk {

constexpr int QG = DK_MAX_Q_LATTICE;   // 10
constexpr int MVN_LATTICE_N = 8192;    // = oracle.mvnorm.LATTICE_N
constexpr int CDF_THREADS = 128;
__host__ __device__ inline int gen_mom_stride(int q) { return q + q * q + 2; }   // mu | Sigma row-major | mask | pad

static __device__ __constant__ double SQRT_PRIMES[9] = {1.4142135623730951, 1.7320508075688772, 2.23606797749979,
                                                          2.6457513110645907, 3.3166247903554, 3.605551275463989,
                                                          4.123105625617661, 4.358898943540674, 4.795831523312719};

// Step 1 of the rule (oracle.mvnorm.sov_factor), by ONE thread: standardise x / S (leading dimension ld), reorder the
// variables greedily by smallest conditional probability and factor the reordered correlation matrix.
// Out: b[q], C[q][QG] (lower).  Work arrays live in the caller's shared memory.
__device__ inline void sov_factor_dev(int q, const double* x, const double* S, int ld, double* b, double* C, double* R) {
  double y[QG], sd[QG];
  for (int i = 0; i < q; i++) {
    sd[i] = sqrt(S[i * ld + i]);
    b[i] = x[i] / sd[i];
  }
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++) {
      const double r = S[i * ld + j] / (sd[i] * sd[j]);
      R[i * QG + j] = (i == j) ? 1.0 : fmax(-1.0, fmin(1.0, r));
      C[i * QG + j] = 0.0;
    }
  for (int i = 0; i < q; i++) {
    double best = INFINITY;
    int bj = i;
    for (int j = i; j < q; j++) {
      double v = R[j * QG + j], m = b[j];
      for (int l = 0; l < i; l++) { v -= C[j * QG + l] * C[j * QG + l]; m -= C[j * QG + l] * y[l]; }
      // (the oracle forms the two dot products first and subtracts once; same values up to rounding of the order)
      const double e = normcdf(m / sqrt(fmax(v, 1e-280)));
      if (e < best) { best = e; bj = j; }
    }
    if (bj != i) {
      for (int c = 0; c < q; c++) { const double t = R[i * QG + c]; R[i * QG + c] = R[bj * QG + c]; R[bj * QG + c] = t; }
      for (int r = 0; r < q; r++) { const double t = R[r * QG + i]; R[r * QG + i] = R[r * QG + bj]; R[r * QG + bj] = t; }
      { const double t = b[i]; b[i] = b[bj]; b[bj] = t; }
      for (int c = 0; c < q; c++) { const double t = C[i * QG + c]; C[i * QG + c] = C[bj * QG + c]; C[bj * QG + c] = t; }
    }
    double v = R[i * QG + i];
    for (int l = 0; l < i; l++) v -= C[i * QG + l] * C[i * QG + l];
    const double cii = sqrt(fmax(v, 1e-280));
    C[i * QG + i] = cii;
    for (int j = i + 1; j < q; j++) {
      double s = R[j * QG + i];
      for (int l = 0; l < i; l++) s -= C[j * QG + l] * C[i * QG + l];
      C[j * QG + i] = s / cii;
    }
    double m = b[i];
    for (int l = 0; l < i; l++) m -= C[i * QG + l] * y[l];
    const double t = m / cii, e = normcdf(t);
    y[i] = -0.3989422804014326779 * exp(-0.5 * t * t) / fmax(e, 1e-300);
  }
}

// pmnorm(x, varcov = S) for 5 <= q <= QG, cooperatively by the CDF_THREADS threads of the CTA (all call it with the same
// arguments; x, S in shared memory).  work: >= 2 QG + 2 QG^2 + CDF_THREADS / 32 doubles of shared memory.
__device__ inline double pmnorm_lattice_cta(int q, const double* x, const double* S, int ld, double* work) {
  double* b = work;
  double* C = work + QG;
  double* R = C + QG * QG;
  double* red = R + QG * QG;
  const int tid = threadIdx.x;
  __syncthreads();   // the previous use of `work` is over
  if (tid == 0) sov_factor_dev(q, x, S, ld, b, C, R);
  __syncthreads();
  const double e0 = normcdf(b[0] / C[0]);
  double sum = 0.0;
  for (int k = tid + 1; k <= MVN_LATTICE_N; k += CDF_THREADS) {
    double y[QG - 1];
    double e = e0, f = e0;
#pragma unroll
    for (int i = 1; i < QG; i++) {
      if (i < q) {
        const double t = (double)k * SQRT_PRIMES[i - 1];
        const double w = fabs(2.0 * (t - floor(t)) - 1.0);
        y[i - 1] = normcdfinv(fmin(fmax(w * e, 1e-300), 0.99999999999999988898));
        double m = b[i];
#pragma unroll
        for (int l = 0; l < QG - 1; l++)
          if (l < i) m -= C[i * QG + l] * y[l];
        e = normcdf(m / C[i * QG + i]);
        f *= e;
      }
    }
    sum += f;
  }
  // fixed-order reduction: lanes by shuffles, then the warps in order
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  if ((tid & 31) == 0) red[tid >> 5] = sum;
  __syncthreads();
  double tot = 0.0;
  for (int w = 0; w < CDF_THREADS / 32; w++) tot += red[w];
  return tot / (double)MVN_LATTICE_N;
}

// ---------------------------------------------------------------------------------------------------------
// moments of batches of q points (5 <= q <= QG)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 2)
k_qei_moments_gen(const DevModel M, int type, int q, const double* __restrict__ x, int64_t B, int64_t ntiles,
                  double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ moments) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
  const int n = M.n, d = M.d, p = M.p;
  const int nb = BN / q;            // batches per tile
  const int64_t N = B * q;
  const int mstride = gen_mom_stride(q);
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b0 = tile * nb;
    // column c = q g + a is point a of batch b0 + g (rounded to 8 digits: qEI continues with the rounded rows,
    // R/qEI.R:126-128); columns beyond the tile's batches repeat the last row
    for (int i = tid; i < d * BN; i += NT) {
      const int k = i / BN, c = i % BN;
      int64_t row = b0 * q + c;
      if (c >= nb * q || row >= N) row = N - 1;
      xs[i] = round8(x[row + (int64_t)k * N]);
    }
    forward_tile<true, -1, false, true>(M, x, N, 0, panel, smem, tid, &stage_base);

    // ---- per column: duplicate flag, mean, S (UK) ----
    int* dupflag = reinterpret_cast<int*>(smem);            // [64]
    double* mu_s = reinterpret_cast<double*>(smem + 1024);  // [64]
    double* S_s = reinterpret_cast<double*>(smem + 2048);   // [64][p]
    if (tid < BN) dupflag[tid] = 0;
    __syncthreads();
    {
      const int c = tid & (BN - 1), slice = tid >> 6;
      if (design_duplicate(M, xs, c, slice)) atomicOr(&dupflag[c], 1);   // xs is rounded already: round8 is idempotent
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
      if (type == DK_UK)
        for (int j = 0; j < p; j++) {
          double t = u[j];
          for (int i = 0; i < j; i++) t -= __ldg(M.RM + i + j * p) * S_s[c * p + i];
          S_s[c * p + j] = t / __ldg(M.RM + j + j * p);
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
        for (int i = lane; i < n; i += 32) part = fma(panel_at(panel, i, ca), panel_at(panel, i, cb), part);
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
            for (int k = 0; k < d && same; k++) same = (xs[k * BN + q * g + aa] == xs[k * BN + q * g + e]);
            drop = same;
          }
          if (!drop) mask |= 1 << aa;
        }
        mom[q + q * q] = (double)mask;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// the Chevalier-Ginsbourger sums, one CTA per batch
// ---------------------------------------------------------------------------------------------------------
struct GenCdfShared {
  double mu[QG], Sg[QG * QG];       // the kept points
  double Sk[QG * QG], mk[QG], bk[QG], xk[QG];
  double cik[QG * QG], sik[QG * QG * QG];   // per second-term entry i: c_ik (row i) and Sigma_ik (block i), dimension qq - 1
  double sym[QG * QG];
  double pk[QG];
  double work[2 * QG + 2 * QG * QG + 8];
  int idx[QG];
  int qq;
};

// pmnorm for any dimension inside the CTA: >= 5 cooperative lattice; <= 4 the deterministic rules on thread 0's behalf of all
__device__ inline double pmnorm_any_cta(int q, const double* x, const double* S, int ld, double* work) {
  if (q >= 5) return pmnorm_lattice_cta(q, x, S, ld, work);
  double xx[4], S4[16];
  for (int i = 0; i < q; i++) {
    xx[i] = x[i];
    for (int j = 0; j < q; j++) S4[i * 4 + j] = S[i * ld + j];
  }
  return pmnorm_dev(q, xx, S4);
}

__global__ void __launch_bounds__(CDF_THREADS)
k_qei_cdf_gen(const double* __restrict__ moments, int64_t B, int q, double plugin, int minimization,
              double* __restrict__ out) {
  __shared__ GenCdfShared sh;
  const int tid = threadIdx.x;
  const int mstride = gen_mom_stride(q);
#pragma unroll 1
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const double* mom = moments + b * mstride;
    __syncthreads();
    if (tid == 0) {
      const int mask = (int)mom[q + q * q];
      int qq = 0;
      for (int a = 0; a < q; a++)
        if (mask & (1 << a)) sh.idx[qq++] = a;
      sh.qq = qq;
      for (int i = 0; i < qq; i++) {
        sh.mu[i] = mom[sh.idx[i]];
        for (int j = 0; j < qq; j++) sh.Sg[i * QG + j] = mom[q + sh.idx[i] * q + sh.idx[j]];
      }
    }
    __syncthreads();
    const int qq = sh.qq;
    if (qq <= 1) {
      if (tid == 0) {
        double res = 0.0;                       // every point duplicates the design (R/qEI.R:127)
        if (qq == 1) {                          // EI(x, model, plugin, type, envir): minimization = TRUE (R/qEI.R:132-134)
          const double sd = sqrt(fmax(sh.Sg[0], 0.0)), diff = plugin - sh.mu[0], xcr = diff / sd;
          res = diff * normcdf(xcr) + sd * dnorm_std(xcr);
        }
        out[b] = res;
      }
      continue;
    }
    double total = 0.0;   // thread 0's running sum_k first[k]; the second term is summed from sh.sym at the end
    double second = 0.0;
#pragma unroll 1
    for (int k = 0; k < qq; k++) {
      __syncthreads();
      // covZk (R/qEI.R:240-261), mu_k, b_k: element-parallel
      for (int e = tid; e < qq * qq; e += CDF_THREADS) {
        const int i = e / qq, j = e % qq;
        const double* S = sh.Sg;
        double v;
        if (i == k && j == k) v = S[k * QG + k];
        else if (i == k) v = S[k * QG + k] - S[k * QG + j];
        else if (j == k) v = S[k * QG + k] - S[k * QG + i];
        else v = S[i * QG + j] + S[k * QG + k] - S[k * QG + i] - S[k * QG + j];
        sh.Sk[i * QG + j] = v;
      }
      if (tid < qq) {
        const int i = tid;
        double m = (i == k) ? -sh.mu[k] : sh.mu[i] - sh.mu[k];
        double bb = (i == k) ? -plugin : 0.0;
        if (minimization) { m = -m; bb = -bb; }
        sh.mk[i] = m; sh.bk[i] = bb; sh.xk[i] = bb - m;
      }
      __syncthreads();
      // get_cik / get_sigmaik (R/qEI.R:196-227) for every i >= k
      for (int e = tid; e < (qq - k) * (qq - 1) * (qq - 1); e += CDF_THREADS) {
        const int i = k + e / ((qq - 1) * (qq - 1)), rc = e % ((qq - 1) * (qq - 1)), r = rc / (qq - 1), cc = rc % (qq - 1);
        const int u = r < i ? r : r + 1, v = cc < i ? cc : cc + 1;
        const double sii = sh.Sk[i * QG + i];
        sh.sik[(i * QG + r) * QG + cc] = sh.Sk[u * QG + v] - sh.Sk[u * QG + i] * sh.Sk[v * QG + i] / sii;
        if (cc == 0) sh.cik[i * QG + r] = (sh.bk[u] - sh.mk[u]) - (sh.bk[i] - sh.mk[i]) * (sh.Sk[i * QG + u] / sii);
      }
      __syncthreads();
      // p_k: the qq-variate CDF
      double pk;
      if (qq >= 5) pk = pmnorm_lattice_cta(qq, sh.xk, sh.Sk, QG, sh.work);
      else { if (tid == 0) sh.pk[k] = pmnorm_any_cta(qq, sh.xk, sh.Sk, QG, sh.work); __syncthreads(); pk = sh.pk[k]; }
      // sym[k][i], i >= k: (qq - 1)-variate CDFs
      if (qq - 1 >= 5) {
        for (int i = k; i < qq; i++) {
          const double c = pmnorm_lattice_cta(qq - 1, sh.cik + i * QG, sh.sik + i * QG * QG, QG, sh.work);
          if (tid == 0) {
            const double sdi = sqrt(sh.Sk[i * QG + i]);
            sh.sym[k * QG + i] = dnorm_std((sh.bk[i] - sh.mk[i]) / sdi) / sdi * c;
          }
        }
      } else {
        // one thread per entry (the deterministic rules are single-thread functions; the lanes run them side by side)
        if (tid < qq - k) {
          const int i = k + tid;
          const double sdi = sqrt(sh.Sk[i * QG + i]);
          sh.sym[k * QG + i] = dnorm_std((sh.bk[i] - sh.mk[i]) / sdi) / sdi *
                               pmnorm_any_cta(qq - 1, sh.cik + i * QG, sh.sik + i * QG * QG, QG, nullptr);
        }
      }
      __syncthreads();
      if (tid == 0) {
        double first = (sh.mu[k] - plugin) * pk;
        if (minimization) first = -first;
        total += first;
        // second term contribution of row k needs sym[i][k] for i < k (already there) and sym[k][i] for i >= k
        double s2 = 0.0;
        for (int i = 0; i < qq; i++) {
          const double sv = i < k ? sh.sym[i * QG + k] : sh.sym[k * QG + i];
          s2 = fma(sv, sh.Sk[i * QG + k], s2);
        }
        second += s2;
      }
    }
    if (tid == 0) out[b] = total + second;
  }
}

// test hook: pmnorm for dimension 5 .. QG, one CTA per problem
__global__ void __launch_bounds__(CDF_THREADS)
k_mvn_lattice_test(int q, int64_t B, const double* __restrict__ h, const double* __restrict__ S, double* __restrict__ out) {
  __shared__ double xs[QG], Ss[QG * QG], work[2 * QG + 2 * QG * QG + 8];
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    for (int e = threadIdx.x; e < q * q; e += CDF_THREADS) Ss[(e / q) * QG + e % q] = S[b * q * q + e];
    if (threadIdx.x < q) xs[threadIdx.x] = h[b * q + threadIdx.x];
    __syncthreads();
    const double v = pmnorm_lattice_cta(q, xs, Ss, QG, work);
    if (threadIdx.x == 0) out[b] = v;
  }
}

}  // namespace dk
