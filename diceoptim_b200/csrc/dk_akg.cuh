// dk_akg.cuh — batched approximate knowledge gradient, /root/reference/R/AKG.R:93-229.
//
// The reference, for ONE candidate x: predicts at x, re-predicts at the whole design (R/AKG.R:121, an
// n^3 job repeated on every call), forms  cn = c.x - V.X'V.x + mu.x  (R/AKG.R:136), sorts the n+1 lines
// (a, b) = (-(mk.X, mk.x), (cn, sk.x^2)/sqrt(tau2 + sk.x^2)) by (b, a), walks the upper envelope and sums
// a_k (Phi(c_{k+1}) - Phi(c_k)) + b_k (phi(c_k) - phi(c_{k+1}))  (R/AKG.R:147-205).
//
// Candidate-independent algebra, hoisted into dk_model_create (delta = noise.var, or 0 for nugget /
// noise-free models, where covMat1Mat2(X, X) already equals K):
//     V.X'V.x = C_X T^-1 v = c.x - delta .* (T^-1 v)               since C_X = K - diag(delta), K T^-1 v = c.x
//     mu.x    = (F - V.X'M) tuuinv (F.x - v'M)' = delta .* (T^-1 M r)       r = tuuinv (F.x - v'M)'
//  => cn      = delta .* T^-1 (v + M r)         the SAME back substitution the gradient kernel runs
//     mk.X    = y - delta .* alpha              alpha = T^-1 z
// so one candidate costs two blocked DMMA solves (2 n^2 flops instead of the reference's 3 n^2 + n^3) plus
// a sort of n+1 keys.  Where the reference's cn is pure rounding noise (delta = 0) this gives exact zeros;
// AKG values agree, the order of the noise-level slopes is not reproduced.
//
// Per candidate tile: forward solve, epilogue (mk.x, sk.x, r), backward solve, then for each of the 64
// candidates a convex-hull pre-filter (16 extreme directions) and a shared-memory bitonic sort of the surviving
// (b, rank-of-a) keys by the whole CTA, and finally 64 threads walk their candidates' envelopes in parallel
// with stacks in per-CTA global scratch.
#pragma once
#include "dk_grad.cuh"

namespace dk {

constexpr int AKG_MAX_SORT = 8192;  // n + 1 <= 8192: keys (8 B) + ranks (4 B) fill region 0 exactly
constexpr int AKG_DIRS = 16;        // directions of the hull pre-filter (two per warp)

constexpr int AKG_GRAD_MAX_HULL = 256;  // envelope members whose K^-1 rows the gradient kernel contracts

// per-CTA scratch: sorted a, sorted b, stack a, stack b, stack cend, sorted rank, stack rank : 7 arrays [n+1][64]
__host__ __device__ inline size_t akg_scratch_doubles(int n) { return (size_t)7 * (n + 1) * BN; }
// per-CTA scratch of the gradient passes: one d x 64 block per weight vector (alpha, wt, p columns of T^-1 M, hull rows)
__host__ __device__ inline size_t akg_grad_scratch_doubles(int d) {
  return (size_t)(2 + DK_MAX_P + AKG_GRAD_MAX_HULL) * d * BN;
}

__device__ __forceinline__ bool akg_less(double b1, int r1, double b2, int r2) {
  return b1 < b2 || (b1 == b2 && r1 < r2);
}

// GRAD = true adds AKG.grad (R/AKG.grad.R:65-281, envir = NULL path).  In the engine's terms, with
//   G1 = dc'alpha + beta f.deltax (gradient of mk.x), G2 = dc'wt, sk2.grad = -2 G2 + 2 r'f.deltax,
//   G3 = (T^-1 M)'dc (UK), q_k = tuuinv (f.deltax[,k] - G3[,k]), H_t = (K^-1 dc)_{i_t} for envelope member i_t,
// the per-line gradients are  A.grad = -G1 (own line) or 0,  B.grad = (delta_i H_t + 2 delta_i (T^-1 M)_i q)/den
// - sk2.grad cn_i / (2 den^3) (design line i; the factor 2 is the reference's, :148) or
// sk^2 (2 tau2 + sk^2)/den^4 sk2.grad / (2 sQ) (own line), and the sum of :258-264 follows, ct.grad as coded at :256.
// m_min.grad is identically zero in the reference (:268-276).  dc = c .* dlogk is contracted once per weight vector.
template <bool GRAD>
__global__ void __launch_bounds__(NT, 2)
k_akg_fused(const DevModel M, int type, double tau2, const double* __restrict__ x, int64_t N, int64_t ntiles,
            double* __restrict__ vscratch, double* __restrict__ wscratch, size_t panel_doubles,
            double* __restrict__ akgscratch, double* __restrict__ out, double* __restrict__ cscratch,
            const double* __restrict__ Kinv, double* __restrict__ gscratch, double* __restrict__ out_grad) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;  // rows: 0 mk.x, 1 sk.x, 2.. r
  const int n = M.n, p = M.p, m0 = n + 1;
  double* sa = akgscratch + (size_t)blockIdx.x * akg_scratch_doubles(n);  // sorted a   [pos][c]
  double* sb = sa + (size_t)m0 * BN;                                      // sorted b
  double* ka = sb + (size_t)m0 * BN;                                      // stack a
  double* kb = ka + (size_t)m0 * BN;                                      // stack b
  double* kc = kb + (size_t)m0 * BN;                                      // stack cend
  double* sr = kc + (size_t)m0 * BN;                                      // sorted rank (as double)
  double* kr = sr + (size_t)m0 * BN;                                      // stack rank
  double* cpanel = GRAD ? cscratch + (size_t)blockIdx.x * panel_doubles : nullptr;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;

#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BN;
    forward_tile<true>(M, x, N, c0, vpanel, smem, tid, &stage_base, cpanel);

    if (tid < BN) {
      const int c = tid;
      double mean, sd, u[DK_MAX_P];
      cand_predict(M, type, vpanel, xs, s2acc, c, mean, sd, u);
      coef[c] = mean;
      coef[BN + c] = sd;
      for (int j = 0; j < p; j++) {
        double r = 0.0;
        if (type == DK_UK)
          for (int i = 0; i < p; i++) r = fma(__ldg(M.tuuinv + j + i * p), u[i], r);
        coef[(2 + j) * BN + c] = r;
      }
    }
    __syncthreads();

    // ---- backward blocked solve: wt = T^-1 (v + M r), flipped row order in wpanel ----
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

    // ---- per candidate: slopes b -> hull pre-filter -> bitonic sort of the survivors -> sorted (a, b) ----
    // Only lines on the upper envelope (the upper convex hull of the points (b, a)) contribute.  The extreme
    // points in AKG_DIRS directions are hull vertices; a point clearly below a chord between two of them
    // cannot be on the hull, so it is dropped before the sort (the reference's stack scan would pop or skip
    // it without effect on the breakpoints of the surviving lines).  Survivors are sorted by (b, rank of a)
    // exactly as order(B, A) does and go through the reference's scan unchanged.
    double* keyb = reinterpret_cast<double*>(smem);                        // [8192] all slopes, then survivor keys
    int* keyr = reinterpret_cast<int*>(smem + (size_t)AKG_MAX_SORT * 8);   // [8192] survivor ranks
    double* red = reinterpret_cast<double*>(smem + (size_t)AKG_MAX_SORT * 8);       // reduction scratch (aliases keyr)
    double* chb = red + 64;                                                // chain vertices b [AKG_DIRS]
    double* cha = chb + AKG_DIRS;                                          // chain vertices a
    int* s_cnt = reinterpret_cast<int*>(smem + SM_S2_OFF);                 // survivor counter, then cnt[64] at +8
    int* cnt = s_cnt + 2;
#pragma unroll 1
    for (int c = 0; c < BN; c++) {
      const double mkx = __ldcg(coef + c), skx = __ldcg(coef + BN + c);
      const double den = sqrt(tau2 + skx * skx);
      const double ax = -mkx;
      // rank of the candidate's own line among the design lines sorted by a (ties: it comes last)
      int lo = 0, hi = n;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(M.akg_a_sorted + mid) <= ax) lo = mid + 1; else hi = mid;
      }
      const int rank_x = 2 * lo - 1;
      // 1. slopes of all n + 1 lines; running min / max of b
      double bmin = INFINITY, bmax = -INFINITY;
      for (int j = tid; j < m0; j += NT) {
        double b;
        if (j < n) {
          const int i = __ldg(M.akg_perm + j);  // design row with the j-th smallest a
          b = __ldg(M.akg_delta + i) * panel_at(wpanel, n - 1 - i, c) / den;
        } else {
          b = skx * skx / den;
        }
        keyb[j] = b;
        bmin = fmin(bmin, b); bmax = fmax(bmax, b);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        bmin = fmin(bmin, __shfl_xor_sync(0xffffffffu, bmin, o));
        bmax = fmax(bmax, __shfl_xor_sync(0xffffffffu, bmax, o));
      }
      if (lane == 0) { red[warp] = bmin; red[8 + warp] = bmax; }
      if (tid == 0) *s_cnt = 0;
      __syncthreads();
      bmin = red[0]; bmax = red[8];
#pragma unroll
      for (int w = 1; w < 8; w++) { bmin = fmin(bmin, red[w]); bmax = fmax(bmax, red[8 + w]); }
      const double amin = fmin(__ldg(M.akg_a_sorted), ax), amax = fmax(__ldg(M.akg_a_sorted + n - 1), ax);
      const double isb = (bmax > bmin) ? 1.0 / (bmax - bmin) : 0.0, isa = (amax > amin) ? 1.0 / (amax - amin) : 0.0;
      // 2. extreme point per direction: warp w takes directions 2w, 2w + 1 (angles from 0 to pi)
      {
        double cs[2], sn[2], best[2] = {-INFINITY, -INFINITY};
        int bi[2] = {0, 0};
#pragma unroll
        for (int u = 0; u < 2; u++) sincospi((double)(2 * warp + u) / (AKG_DIRS - 1), &sn[u], &cs[u]);
        for (int j = lane; j < m0; j += 32) {
          const double bs = (keyb[j] - bmin) * isb;
          const double as = (((j < n) ? __ldg(M.akg_a_sorted + j) : ax) - amin) * isa;
#pragma unroll
          for (int u = 0; u < 2; u++) {
            const double pv = fma(cs[u], bs, sn[u] * as);
            if (pv > best[u]) { best[u] = pv; bi[u] = j; }
          }
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best[u], o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi[u], o);
            if (ob > best[u] || (ob == best[u] && oi < bi[u])) { best[u] = ob; bi[u] = oi; }
          }
          if (lane == 0) {
            chb[2 * warp + u] = keyb[bi[u]];
            cha[2 * warp + u] = (bi[u] < n) ? __ldg(M.akg_a_sorted + bi[u]) : ax;
          }
        }
      }
      __syncthreads();
      // 3. drop points clearly below a chord of the chain; survivors (b, rank) -> global staging (kc, ka).
      //    A chord between ANY two data points lies inside the hull, so the test is safe whatever the vertices.
      double* clo = cha + AKG_DIRS;   // per chord: b-span [clo, chi], value = cic + csl * b
      double* chi = clo + AKG_DIRS;
      double* csl = chi + AKG_DIRS;
      double* cic = csl + AKG_DIRS;
      if (tid < AKG_DIRS - 1) {
        const double b0 = chb[tid], b1 = chb[tid + 1];
        const double blo = fmin(b0, b1), bhi = fmax(b0, b1);
        const bool ok = bhi > blo;
        const double sl = ok ? (cha[tid + 1] - cha[tid]) / (b1 - b0) : 0.0;
        clo[tid] = ok ? blo : INFINITY;   // empty span: never tested
        chi[tid] = ok ? bhi : -INFINITY;
        csl[tid] = sl;
        cic[tid] = cha[tid] - sl * b0;
      }
      __syncthreads();
      {
        const double tol = 1e-9 * (amax - amin) + 1e-300;
        for (int j = tid; j < m0; j += NT) {
          const double bj = keyb[j];
          const double aj = (j < n) ? __ldg(M.akg_a_sorted + j) : ax;
          bool keep = true;
#pragma unroll
          for (int s = 0; s < AKG_DIRS - 1; s++)
            if (bj >= clo[s] && bj <= chi[s] && aj < fma(csl[s], bj, cic[s]) - tol) keep = false;
          if (keep) {
            const int pos = atomicAdd(s_cnt, 1);
            kc[(size_t)pos * BN + c] = bj;
            ka[(size_t)pos * BN + c] = (double)((j < n) ? 2 * j : rank_x);
          }
        }
      }
      __syncthreads();
      const int ms = *s_cnt;  // survivors
      int P2 = 2;
      while (P2 < ms) P2 <<= 1;
      for (int j = tid; j < P2; j += NT) {
        if (j < ms) {
          keyb[j] = __ldcg(kc + (size_t)j * BN + c);
          keyr[j] = (int)__ldcg(ka + (size_t)j * BN + c);
        } else {
          keyb[j] = INFINITY; keyr[j] = 0x7fffffff;
        }
      }
      if (tid == 0) cnt[c] = ms;
      __syncthreads();
      for (int k2 = 2; k2 <= P2; k2 <<= 1)
        for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
          // pair pi <-> (t, t | j2) with bit j2 of t clear: every lane has work, and the four pairs a thread
          // takes per step are loaded before any is stored (they are disjoint), so the loads overlap
          const int npairs = P2 >> 1;
          for (int p0 = tid; p0 < npairs; p0 += 4 * NT) {
            double b1[4], b2[4];
            int r1[4], r2[4], tt[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const int pi = p0 + u * NT;
              tt[u] = -1;
              if (pi < npairs) {
                const int t = ((pi & ~(j2 - 1)) << 1) | (pi & (j2 - 1));
                tt[u] = t;
                b1[u] = keyb[t]; b2[u] = keyb[t | j2];
                r1[u] = keyr[t]; r2[u] = keyr[t | j2];
              }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
              if (tt[u] >= 0) {
                const int t = tt[u];
                const bool up = (t & k2) == 0;
                const bool swap = up ? akg_less(b2[u], r2[u], b1[u], r1[u]) : akg_less(b1[u], r1[u], b2[u], r2[u]);
                if (swap) { keyb[t] = b2[u]; keyb[t | j2] = b1[u]; keyr[t] = r2[u]; keyr[t | j2] = r1[u]; }
              }
            }
          }
          __syncthreads();
        }
      for (int j = tid; j < ms; j += NT) {
        const int r = keyr[j];
        sb[(size_t)j * BN + c] = keyb[j];
        sa[(size_t)j * BN + c] = (r & 1) ? ax : __ldg(M.akg_a_sorted + (r >> 1));
        if (GRAD) sr[(size_t)j * BN + c] = (double)r;
      }
      __syncthreads();
    }

    // ---- 64 threads: drop equal slopes (keep the last), upper-envelope stack, Phi/phi sum ----
    int* hlen = cnt + BN;  // envelope length per candidate (gradient kernel)
    if (GRAD && tid < BN) hlen[tid] = 0;
    if (tid < BN && c0 + tid < N) {
      const int c = tid;
      const double mkx = __ldcg(coef + c), skx = __ldcg(coef + BN + c);
      double res = 0.0;
      if (!(skx < sqrt(M.sd2) / 1e6 || M.sd2 < 1e-20)) {  // R/AKG.R:115
        const int m0 = cnt[c];    // lines that survived the hull pre-filter, sorted by (b, a)
        int len = 0;              // stack length
        double ta = 0.0, tb = 0.0;  // top of stack (a, b)
        double cbelow = -1e36;      // cend of the element below the top
        for (int j = 0; j < m0; j++) {
          const double bj = __ldcg(sb + (size_t)j * BN + c);
          if (j + 1 < m0 && __ldcg(sb + (size_t)(j + 1) * BN + c) == bj) continue;  // R/AKG.R:153-160
          const double aj = __ldcg(sa + (size_t)j * BN + c);
          if (len == 0) {
            ka[c] = aj; kb[c] = bj; ta = aj; tb = bj; len = 1;
            if (GRAD) kr[c] = __ldcg(sr + (size_t)j * BN + c);
            continue;
          }
          int count = 0;
          while (count < 1000) {  // R/AKG.R:179-192
            count++;
            const double cu = (ta - aj) / (bj - tb);
            if (len > 1 && cu <= cbelow) {
              len--;               // pop; reload the new top and the cend below it
              ta = __ldcg(ka + (size_t)(len - 1) * BN + c);
              tb = __ldcg(kb + (size_t)(len - 1) * BN + c);
              cbelow = (len > 1) ? __ldcg(kc + (size_t)(len - 2) * BN + c) : -1e36;
            } else {
              kc[(size_t)(len - 1) * BN + c] = cu;
              cbelow = cu;
              ka[(size_t)len * BN + c] = aj; kb[(size_t)len * BN + c] = bj;
              if (GRAD) kr[(size_t)len * BN + c] = __ldcg(sr + (size_t)j * BN + c);
              ta = aj; tb = bj; len++;
              break;
            }
          }
        }
        kc[(size_t)(len - 1) * BN + c] = 1e36;
        __threadfence_block();
        double lo = -1e36, maxnew = 0.0;
        for (int t = 0; t < len; t++) {
          const double hi = __ldcg(kc + (size_t)t * BN + c);
          const double at = __ldcg(ka + (size_t)t * BN + c), bt = __ldcg(kb + (size_t)t * BN + c);
          maxnew += at * (normcdf(hi) - normcdf(lo)) + bt * (dnorm_std(lo) - dnorm_std(hi));
          lo = hi;
        }
        const double m_min = fmin(M.akg_min_mkX, mkx);
        res = maxnew + m_min;  // AKG <- maxNew - (-m_min)
        if (GRAD) hlen[c] = len;
      }
      if (out) out[c0 + c] = res;
    }
    __syncthreads();

    if constexpr (GRAD) {
      const int d = M.d;
      const int pu = (type == DK_UK) ? p : 0;
      int maxlen = 0;
      for (int c2 = 0; c2 < BN; c2++) maxlen = max(maxlen, hlen[c2]);
      maxlen = min(maxlen, AKG_GRAD_MAX_HULL);
      const int npass = 2 + pu + maxlen;
      double* gbuf = reinterpret_cast<double*>(smem);  // [4][DK_MAX_D][64]
      double* G = gscratch + (size_t)blockIdx.x * akg_grad_scratch_doubles(d);  // [pass][k][64]
      {
        const int c = tid & (BN - 1), slice = tid >> 6;
#pragma unroll 1
        for (int pass = 0; pass < npass; pass++) {
          double g[DK_MAX_D];
#pragma unroll
          for (int k = 0; k < DK_MAX_D; k++) g[k] = 0.0;
          const double* wvec = nullptr;
          bool use_panel = false;
          if (pass == 0) wvec = M.alpha;
          else if (pass == 1) use_panel = true;
          else if (pass < 2 + pu) wvec = M.KinvF + (size_t)(pass - 2) * n;
          else {
            const int t = pass - 2 - pu;
            if (t < hlen[c]) {
              const int r = (int)kr[(size_t)t * BN + c];
              if (!(r & 1)) wvec = Kinv + (size_t)__ldg(M.akg_perm + (r >> 1)) * n;  // row of K^-1 (symmetric)
            }
          }
          if (use_panel) {
            switch (M.kernel) {
              case DK_GAUSS: grad_accumulate<DK_GAUSS>(M, xs, cpanel, wpanel, c, slice, 0.0, 1.0, g); break;
              case DK_EXP: grad_accumulate<DK_EXP>(M, xs, cpanel, wpanel, c, slice, 0.0, 1.0, g); break;
              case DK_MATERN3_2: grad_accumulate<DK_MATERN3_2>(M, xs, cpanel, wpanel, c, slice, 0.0, 1.0, g); break;
              case DK_MATERN5_2: grad_accumulate<DK_MATERN5_2>(M, xs, cpanel, wpanel, c, slice, 0.0, 1.0, g); break;
              default: grad_accumulate<DK_POWEXP>(M, xs, cpanel, wpanel, c, slice, 0.0, 1.0, g); break;
            }
          } else if (wvec) {
            switch (M.kernel) {
              case DK_GAUSS: grad_accumulate_vec<DK_GAUSS>(M, xs, cpanel, wvec, c, slice, g); break;
              case DK_EXP: grad_accumulate_vec<DK_EXP>(M, xs, cpanel, wvec, c, slice, g); break;
              case DK_MATERN3_2: grad_accumulate_vec<DK_MATERN3_2>(M, xs, cpanel, wvec, c, slice, g); break;
              case DK_MATERN5_2: grad_accumulate_vec<DK_MATERN5_2>(M, xs, cpanel, wvec, c, slice, g); break;
              default: grad_accumulate_vec<DK_POWEXP>(M, xs, cpanel, wvec, c, slice, g); break;
            }
          }
#pragma unroll
          for (int k = 0; k < DK_MAX_D; k++)
            if (k < d) gbuf[(slice * DK_MAX_D + k) * BN + c] = g[k];
          __syncthreads();
          for (int k = slice; k < d; k += 4)
            G[((size_t)pass * d + k) * BN + c] = ((gbuf[(0 * DK_MAX_D + k) * BN + c] + gbuf[(1 * DK_MAX_D + k) * BN + c]) +
                                                  gbuf[(2 * DK_MAX_D + k) * BN + c]) + gbuf[(3 * DK_MAX_D + k) * BN + c];
          __syncthreads();
        }
      }
      // ---- assembly: one thread per candidate walks its envelope ----
      if (tid < BN && c0 + tid < N) {
        const int c = tid, len = hlen[c];
        const int64_t gi = c0 + c;
        if (len == 0) {
          for (int k = 0; k < d; k++) out_grad[gi + (int64_t)k * N] = 0.0;          // R/AKG.grad.R:126-127
        } else if (len > AKG_GRAD_MAX_HULL) {
          for (int k = 0; k < d; k++) out_grad[gi + (int64_t)k * N] = nan("");      // envelope longer than the scratch
        } else {
          const double skx = __ldcg(coef + BN + c);
          const double den2 = tau2 + skx * skx, den = sqrt(den2), sQ = skx * skx / den;
          const bool lin = M.trend == DK_TREND_LINEAR;
          double grad[DK_MAX_D], sk2g[DK_MAX_D], meang[DK_MAX_D];
          double atg_p[DK_MAX_D], btg_p[DK_MAX_D];
          for (int k = 0; k < d; k++) {
            grad[k] = 0.0;
            meang[k] = __ldcg(G + ((size_t)0 * d + k) * BN + c) + (lin ? __ldg(M.beta + 1 + k) : 0.0);
            sk2g[k] = -2.0 * __ldcg(G + ((size_t)1 * d + k) * BN + c) + (lin ? 2.0 * __ldcg(coef + (2 + 1 + k) * BN + c) : 0.0);
          }
          double at_p = 0.0, bt_p = 0.0, lo = -1e36;
          for (int t = 0; t < len; t++) {
            const double at = ka[(size_t)t * BN + c], bt = kb[(size_t)t * BN + c], hi = kc[(size_t)t * BN + c];
            const int r = (int)kr[(size_t)t * BN + c];
            const bool own = r & 1;
            const int row = own ? 0 : __ldg(M.akg_perm + (r >> 1));
            const double delta = own ? 0.0 : __ldg(M.akg_delta + row);
            const double cn_i = bt * den;
            const double pl = normcdf(hi) - normcdf(lo), dl = dnorm_std(lo) - dnorm_std(hi);
            const double philo = dnorm_std(lo);
            for (int k = 0; k < d; k++) {
              double atg, btg;
              if (own) {
                atg = -meang[k];
                btg = (skx * skx * (2.0 * tau2 + skx * skx) / (den2 * den2)) * sk2g[k] / (2.0 * sQ);
              } else {
                double cng = delta * __ldcg(G + ((size_t)(2 + pu + t) * d + k) * BN + c);
                if (pu) {
                  // mu.x.grad[k, i] = 2 delta_i (T^-1 M)[i, ] tuuinv (f.deltax[, k] - G3[, k])
                  double mu = 0.0;
                  for (int j = 0; j < p; j++) {
                    double qj = 0.0;
                    for (int j2 = 0; j2 < p; j2++) {
                      const double fd = (lin && j2 == 1 + k) ? 1.0 : 0.0;
                      qj = fma(__ldg(M.tuuinv + j + j2 * p), fd - __ldcg(G + ((size_t)(2 + j2) * d + k) * BN + c), qj);
                    }
                    mu = fma(__ldg(M.KinvF + row + (size_t)j * n), qj, mu);
                  }
                  cng = fma(2.0 * delta, mu, cng);
                }
                atg = 0.0;
                btg = cng / den - sk2g[k] * cn_i * (0.5 / (den2 * den));
              }
              double gk = grad[k] + atg * pl + btg * dl;
              if (t > 0) {
                const double ctg = (atg_p[k] - atg) - (at_p - at) * (btg - btg_p[k]) / (bt - bt_p);
                gk += ctg * philo * ((at_p + bt_p * lo) - (at + bt * lo));
              }
              grad[k] = gk;
              atg_p[k] = atg; btg_p[k] = btg;
            }
            at_p = at; bt_p = bt; lo = hi;
          }
          for (int k = 0; k < d; k++) out_grad[gi + (int64_t)k * N] = grad[k];
        }
      }
      __syncthreads();
    }
  }
}

// K^-1 = T^-1 T^-T, 64 columns per tile: forward and backward blocked solves of identity columns with the same
// DMMA code (model set-up for AKG.grad; candidate independent).
__global__ void __launch_bounds__(NT, 2)
k_kinv(const DevModel M, int64_t ntiles, double* __restrict__ vscratch, double* __restrict__ wscratch,
       size_t panel_doubles, double* __restrict__ Kinv) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  const int n = M.n;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int j0 = (int)tile * BN;
    for (int pass = 0; pass < 2; pass++) {
      const int nb = pass == 0 ? M.nblk : M.nblk_b;
#pragma unroll 1
      for (int I = 0; I < nb; I++) {
        double acc[4][4][2];
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          const int ip = I * BM + 32 * wm + 8 * mi + (lane >> 2);
#pragma unroll
          for (int ni = 0; ni < 4; ni++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
              const int c = 32 * wn + 8 * ni + 2 * (lane & 3) + e;
              double v = 0.0;
              if (pass == 0) v = (ip == j0 + c && ip < n) ? 1.0 : 0.0;            // identity column
              else if (ip < n) v = panel_at(vpanel, n - 1 - ip, c);               // flipped forward solution
              acc[mi][ni][e] = v;
            }
        }
        solve_block_row<true>(acc, pass == 0 ? M.tiles_fwd : M.tiles_bwd, I, pass == 0 ? vpanel : wpanel, smem, tid,
                              &stage_base);
        __syncthreads();
      }
    }
    for (int idx = tid; idx < n * BN; idx += NT) {
      const int i = idx % n, c = idx / n;
      if (j0 + c < n) Kinv[(size_t)i + (size_t)(j0 + c) * n] = panel_at(wpanel, n - 1 - i, c);
    }
    __syncthreads();
  }
}

}  // namespace dk
