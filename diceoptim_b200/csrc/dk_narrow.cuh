// dk_narrow.cuh — narrow-tile (16 candidates per CTA) variants of the fused forward and gradient kernels, for
// optimiser-sized batches: the polish iterations of max_crit / max_EI (optim(fn = EI, gr = EI.grad),
// /root/reference/R/max_crit.R:118-121) evaluate a few dozen points per call.  The wide kernels give every CTA 64
// candidates; a call with fewer than gridDim tiles leaves SMs idle and still pays a full 64-wide tile's latency
// (2.7 ms forward at n = 2000).  Here a CTA owns 16 candidates: the 8 warps stand as 8 x 1 over the 128 x 16 block, the
// DMMA work per tile is a quarter, four times as many CTAs share a small batch.
//
// Same data as the wide path (factor tiles, panel format, shared-memory map); per candidate the covariance values, the
// GEMM accumulation order (k-chunks ascending) and the diagonal phase (chunks 0..q for row pair q) are those of the wide
// kernel, so the solved panel V is bit-identical; only the reductions over rows (sum of v^2: 8 partials instead of 4;
// gradient: 16 row slices instead of 4) group their terms differently, i.e. sd and the gradients differ from the wide
// path in the last bits (tests: 1e-13), not beyond.
#pragma once
#include "dk_grad.cuh"

namespace dk {

constexpr int BNN = 16;   // candidates per narrow tile; the panel / B-tile layout stays that of BN = 64 (columns 0..15 used)

// 4 DMMAs per k4 step on the warp's 16 x 16 tile: warp w owns row pair w (rows 16w .. 16w + 15)
__device__ __forceinline__ void mma_stage_narrow(double (&acc)[2][2][2], const double* __restrict__ As,
                                                 const double* __restrict__ Bs, int warp, int lane) {
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const double2 a = *reinterpret_cast<const double2*>(As + ((warp * 4 + k4) * 32 + lane) * 2);
    const double2 b = *reinterpret_cast<const double2*>(Bs + ((k4 * 4) * 32 + lane) * 2);
    dmma884(acc[0][0], a.x, b.x); dmma884(acc[0][1], a.x, b.y);
    dmma884(acc[1][0], a.y, b.x); dmma884(acc[1][1], a.y, b.y);
  }
}

// covariance block: 128 design rows x 16 candidates.  Warp w owns rows 16w .. 16w + 15; lane = 16 h + col handles column
// col of rows 8h .. 8h + 7: eight independent chains per thread, the arithmetic per pair is covgen_rows' own.
template <int KERNEL, bool NUG>
__device__ __forceinline__ void covgen_rows_narrow(const DevModel& M, const double* __restrict__ Xs,
                                                   const double* __restrict__ xs, double* __restrict__ Cbuf,
                                                   int row0_global, int warp, int lane) {
  constexpr int RB = 8;
  const int d = M.d, col = lane & 15, rr = 8 * (lane >> 4);
  if (row0_global + 16 * warp + rr >= M.n) {   // eight rows beyond the design: zeros
#pragma unroll
    for (int j = 0; j < RB; j++) Cbuf[(16 * warp + rr + j) * CB_STRIDE + col] = 0.0;
    return;
  }
  double s[RB], pr[RB], hs[NUG ? RB : 1];
#pragma unroll
  for (int j = 0; j < RB; j++) {
    s[j] = 0.0;
    pr[j] = 1.0;
    if (NUG) hs[j] = 0.0;
  }
  const double* Xr = Xs + (16 * warp + rr) * d;
#pragma unroll 1
  for (int k = 0; k < d; k++) {
    const double wk = M.w[k];
    const double c0 = __dmul_rn(xs[k * BN + col], wk);
    const double iwk = NUG ? M.iw[k] : 0.0;
#pragma unroll
    for (int j = 0; j < RB; j++) {
      const double a = fabs(__dsub_rn(Xr[j * d + k], c0));
      if (NUG) hs[j] = fma(a, iwk, hs[j]);
      if (KERNEL == DK_GAUSS) {
        s[j] = fma(0.5 * a, a, s[j]);
      } else if (KERNEL == DK_EXP) {
        s[j] += a;
      } else if (KERNEL == DK_MATERN3_2) {
        pr[j] *= (1.0 + a);
        s[j] += a;
      } else if (KERNEL == DK_MATERN5_2) {
        pr[j] *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0);
        s[j] += a;
      } else {
        s[j] += pow(a, M.shape[k]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < RB; j++) {
    const int r = 16 * warp + rr + j;
    const bool live = (row0_global + r) < M.n;
    double c = M.sd2 * pr[j] * exp(-s[j]);
    if (NUG && hs[j] < 1e-15) c += M.nugget;
    Cbuf[r * CB_STRIDE + col] = live ? c : 0.0;
  }
}

__device__ __forceinline__ void covgen_block_narrow(const DevModel& M, const double* Xs, const double* xs, double* Cbuf,
                                                    int row0_global, int warp, int lane) {
#define DK_NCOV(K)                                                                                   \
  case K:                                                                                            \
    if (M.nugget_flag) covgen_rows_narrow<K, true>(M, Xs, xs, Cbuf, row0_global, warp, lane);        \
    else covgen_rows_narrow<K, false>(M, Xs, xs, Cbuf, row0_global, warp, lane);                     \
    break;
  switch (M.kernel) {
    DK_NCOV(DK_GAUSS) DK_NCOV(DK_EXP) DK_NCOV(DK_MATERN3_2) DK_NCOV(DK_MATERN5_2)
    default: DK_NCOV(DK_POWEXP)
  }
#undef DK_NCOV
}

// One block row of the blocked substitution, narrow tile (bulk-copy + mbarrier pipeline as solve_block_row<true>).
// acc[mi][ni][e]: row 16 warp + 8 mi + (lane >> 2), column 8 ni + 2 (lane & 3) + e.
__device__ __forceinline__ void solve_block_row_narrow(double (&acc)[2][2][2], const double* __restrict__ tiles, int I,
                                                       double* __restrict__ panel, unsigned char* smem, int tid,
                                                       unsigned* stage_base, int live_rows = BM) {
  const int warp = tid >> 5, lane = tid & 31;
  const bool pair_live = 16 * warp < live_rows;   // a row pair made of padding alone has nothing to accumulate
  double* ring = reinterpret_cast<double*>(smem);
  const double* Atiles = tiles + tile_base(I) * A_TILE;
  const int nk = CHUNKS_PER_BLOCK * I;
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smem + SM_TOTAL);
  unsigned long long* empty = full + STAGES;
  const unsigned base = *stage_base;
  // of a B tile (16 x 64, fragment order) the narrow tile reads the column pair np = 0 of each k4 group: 4 x 512 B
  constexpr unsigned STAGE_BYTES = A_TILE * 8 + 4 * 512;
  asm volatile("fence.proxy.async;" ::: "memory");
  __syncthreads();
  auto issue = [&](int idx) {  // thread 0 only
    const unsigned gi = base + idx, slot = gi % STAGES, par = (gi / STAGES) & 1;
    mbar_wait(&empty[slot], par ^ 1);
    double* dst = ring + slot * (A_TILE + B_TILE);
    mbar_arrive_expect_tx(&full[slot], STAGE_BYTES);
    bulk_load(dst, Atiles + (size_t)idx * A_TILE, A_TILE * 8, &full[slot]);
    const double* bsrc = panel + (size_t)idx * B_TILE;
#pragma unroll
    for (int k4 = 0; k4 < 4; k4++) bulk_load(dst + A_TILE + k4 * 256, bsrc + k4 * 256, 512, &full[slot]);
  };
  if (tid == 0)
    for (int s = 0; s < STAGES - 1 && s < nk; s++) issue(s);
#pragma unroll 1
  for (int kc = 0; kc < nk; kc++) {
    if (tid == 0 && kc + STAGES - 1 < nk) issue(kc + STAGES - 1);
    const unsigned gi = base + kc, slot = gi % STAGES, par = (gi / STAGES) & 1;
    mbar_wait(&full[slot], par);
    const double* s = ring + slot * (A_TILE + B_TILE);
    if (pair_live) mma_stage_narrow(acc, s, s + A_TILE, warp, lane);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[slot]);
  }
  *stage_base = base + nk;
  __syncthreads();

  // ---- V_I = inv(L_II) * acc ----
  double* Rbuf = reinterpret_cast<double*>(smem);
  double* Abuf = reinterpret_cast<double*>(smem + SM_ABUF_OFF);
  const double* Dtiles = Atiles + (size_t)nk * A_TILE;
  load_tile_async(Abuf, Dtiles, A_TILE, tid);
  cp_async_commit();
#pragma unroll
  for (int mi = 0; mi < 2; mi++) {
    const int k = 8 * mi + (lane >> 2);
#pragma unroll
    for (int ni = 0; ni < 2; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int nn = 8 * ni + 2 * (lane & 3) + e;
        Rbuf[warp * B_TILE + b_tile_index(k, nn)] = acc[mi][ni][e];
        acc[mi][ni][e] = 0.0;
      }
  }
#pragma unroll 1
  for (int j = 0; j < CHUNKS_PER_BLOCK; j++) {
    cp_async_wait<0>();
    __syncthreads();
    if (j + 1 < CHUNKS_PER_BLOCK) {
      load_tile_async(Abuf + ((j + 1) & 1) * A_TILE, Dtiles + (size_t)(j + 1) * A_TILE, A_TILE, tid);
      cp_async_commit();
    }
    if (j <= warp && pair_live) mma_stage_narrow(acc, Abuf + (j & 1) * A_TILE, Rbuf + j * B_TILE, warp, lane);  // pair q: chunks 0..q
  }
  // ---- publish V_I ----
#pragma unroll
  for (int mi = 0; mi < 2; mi++) {
    const int k = 8 * mi + (lane >> 2);
#pragma unroll
    for (int ni = 0; ni < 2; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int nn = 8 * ni + 2 * (lane & 3) + e;
        __stcg(panel + (size_t)(CHUNKS_PER_BLOCK * I + warp) * B_TILE + b_tile_index(k, nn), acc[mi][ni][e]);
      }
  }
  __threadfence();
  asm volatile("fence.proxy.async.global;" ::: "memory");
}

// s2acc layout of the narrow tile: [8 warps][16 columns]
__device__ __forceinline__ void cand_predict_narrow(const DevModel& M, int type, const double* panel, const double* xs,
                                                    const double* s2acc, int c, double& mean_out, double& sd_out,
                                                    double* u) {
  const int p = M.p, n = M.n;
  double S[DK_MAX_P];
  double mean = 0.0;
  for (int j = 0; j < p; j++) {
    const double fj = (j == 0) ? 1.0 : xs[(j - 1) * BN + c];
    mean = fma(fj, __ldg(M.beta + j), mean);
    u[j] = fj + panel_at(panel, n + 1 + j, c);
  }
  mean += -panel_at(panel, n, c);
  double sum = 0.0;
#pragma unroll
  for (int w = 0; w < 8; w++) sum += s2acc[w * BNN + c];
  double s2 = M.total_sd2 - sum;
  if (type == DK_UK) {
    double s22 = 0.0;
    for (int j = 0; j < p; j++) {
      double t = u[j];
      for (int i = 0; i < j; i++) t -= __ldg(M.RM + i + j * p) * S[i];
      S[j] = t / __ldg(M.RM + j + j * p);
      s22 = fma(S[j], S[j], s22);
    }
    s2 += s22;
  }
  s2 = fmax(s2, 0.0);
  mean_out = mean;
  sd_out = sqrt(s2);
}

// Forward pass of one narrow tile (candidates c0 .. c0 + 15); cpanel as in forward_tile (row-major [i][64], columns 0..15).
__device__ __forceinline__ void forward_tile_narrow(const DevModel& M, const double* __restrict__ x, int64_t N, int64_t c0,
                                                    double* __restrict__ panel, unsigned char* smem, int tid,
                                                    unsigned* stage_base, double* __restrict__ cpanel = nullptr) {
  const int warp = tid >> 5, lane = tid & 31;
  double* Cbuf = reinterpret_cast<double*>(smem);
  double* Xs = reinterpret_cast<double*>(smem + SM_XS_OFF);
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  double* s2acc = reinterpret_cast<double*>(smem + SM_S2_OFF);
  const int d = M.d;
  for (int i = tid; i < d * BNN; i += NT) {
    const int k = i / BNN, j = i % BNN;
    const int64_t g = c0 + j;
    xs[k * BN + j] = (g < N) ? x[g + (int64_t)k * N] : x[(N - 1) + (int64_t)k * N];
  }
  if (tid < 8 * BNN) s2acc[tid] = 0.0;
#pragma unroll 1
  for (int I = 0; I < M.nblk; I++) {
    {
      const double* src = M.Xrm_s + (size_t)I * BM * d;
      for (int i = tid; i < BM * d; i += NT) Xs[i] = __ldg(src + i);
    }
    __syncthreads();
    covgen_block_narrow(M, Xs, xs, Cbuf, I * BM, warp, lane);
    __syncthreads();
    double acc[2][2][2];
#pragma unroll
    for (int mi = 0; mi < 2; mi++)
#pragma unroll
      for (int ni = 0; ni < 2; ni++) {
        const int row = 16 * warp + 8 * mi + (lane >> 2), col = 8 * ni + 2 * (lane & 3);
        const double2 v = *reinterpret_cast<const double2*>(Cbuf + row * CB_STRIDE + col);
        acc[mi][ni][0] = v.x;
        acc[mi][ni][1] = v.y;
        if (cpanel) *reinterpret_cast<double2*>(cpanel + (size_t)(I * BM + row) * BN + col) = v;
      }
    __syncthreads();
    solve_block_row_narrow(acc, M.tiles_fwd, I, panel, smem, tid, stage_base, min(BM, M.n_aug - I * BM));
    {
      double part[2][2];
#pragma unroll
      for (int ni = 0; ni < 2; ni++) part[ni][0] = part[ni][1] = 0.0;
#pragma unroll
      for (int mi = 0; mi < 2; mi++) {
        const bool live = (I * BM + 16 * warp + 8 * mi + (lane >> 2)) < M.n;
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const double v = live ? acc[mi][ni][e] : 0.0;
            part[ni][e] = fma(v, v, part[ni][e]);
          }
      }
#pragma unroll
      for (int ni = 0; ni < 2; ni++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          double v = part[ni][e];
          v += __shfl_xor_sync(0xffffffffu, v, 4);
          v += __shfl_xor_sync(0xffffffffu, v, 8);
          v += __shfl_xor_sync(0xffffffffu, v, 16);
          if (lane < 4) s2acc[warp * BNN + 8 * ni + 2 * lane + e] += v;
        }
    }
    __syncthreads();
  }
}

// ---- forward kernel: one CTA per narrow tile (grid-strided when there are more tiles than CTAs) ----
__global__ void __launch_bounds__(NT, 2)
k_predict_narrow(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
                 double* __restrict__ vscratch, size_t panel_doubles, double* __restrict__ out_main,
                 double* __restrict__ out_mean, double* __restrict__ out_sd) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* panel = vscratch + (size_t)blockIdx.x * panel_doubles;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BNN;
    forward_tile_narrow(M, x, N, c0, panel, smem, tid, &stage_base);
    if (tid < BNN && c0 + tid < N) {
      double mean, sd, u[DK_MAX_P];
      cand_predict_narrow(M, P.type, panel, xs, s2acc, tid, mean, sd, u);
      const int64_t g = c0 + tid;
      if (out_mean) out_mean[g] = mean;
      if (out_sd) out_sd[g] = sd;
      if (out_main) {
        double v = crit_value(P, mean, sd);
        if (P.combine == COMBINE_MUL) v *= out_main[g];
        else if (P.combine == COMBINE_MIN) v = fmin(v, out_main[g]);
        out_main[g] = v;
      }
    }
    __syncthreads();
  }
}

// outer-derivative coefficients of the gradient kernels (k_grad_fused's epilogue 1), shared with the narrow kernel
__device__ __forceinline__ void grad_coefficients(const DevModel& M, const CritParams& P, double mean, double sd, double& cA,
                                                  double& cW, bool& drop_r) {
  drop_r = false;
  if (P.mode == MODE_QUANTILE) {
    cA = 1.0;
    if (sd < sqrt(M.sd2) / 1e6) { cW = 0.0; drop_r = true; }
    else cW = -P.qnorm_beta / sd;
  } else if (P.mode == MODE_AEI) {
    if (sd < sqrt(P.sd2) / 1e6) { cA = 0.0; cW = 0.0; drop_r = true; }
    else {
      const double diff = P.plugin - mean, xcr = diff / sd, Phi = normcdf(xcr), phi = dnorm_std(xcr);
      const double v2 = P.tau2 + sd * sd, pen = 1.0 - sqrt(P.tau2) / sqrt(v2);
      const double eival = diff * Phi + sd * phi;
      const double kappa = phi * pen / (2.0 * sd) + eival * sqrt(P.tau2) / (2.0 * v2 * sqrt(v2));
      cA = -Phi * pen;
      cW = -2.0 * kappa;
    }
  } else if (P.mode == MODE_EQI) {
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
  } else {
    const double sgn = P.minimization ? 1.0 : -1.0;
    if (sd / sqrt(M.sd2) < 1e-6) { cA = 0.0; cW = 0.0; drop_r = true; }
    else if (P.proxy) { cA = -sgn; cW = 0.0; drop_r = true; }
    else {
      const double xcr = (P.plugin - sgn * mean) / sd;
      cA = -sgn * normcdf(xcr);
      cW = -dnorm_std(xcr) / sd;
    }
  }
}

// g_k += sum over rows i = slice, slice + 16, ... of c_i omega_i dlogk_k(x - X_i)   (16 row slices x 16 candidates)
template <int KERNEL>
__device__ __forceinline__ void grad_accumulate_narrow(const DevModel& M, const double* __restrict__ xs,
                                                       const double* __restrict__ cpanel, const double* __restrict__ wpanel,
                                                       int c, int slice, double cA, double cW, double (&g)[DK_MAX_D]) {
  const int n = M.n, d = M.d;
  int i = slice;
#pragma unroll 1
  for (; i + 16 < n; i += 32) {
    const double om0 = cA * __ldg(M.alpha + i) + cW * panel_at(wpanel, n - 1 - i, c);
    const double om1 = cA * __ldg(M.alpha + i + 16) + cW * panel_at(wpanel, n - 1 - (i + 16), c);
    const double c0 = __ldcg(cpanel + (size_t)i * BN + c), c1 = __ldcg(cpanel + (size_t)(i + 16) * BN + c);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)i * d, c, c0 * om0, g);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)(i + 16) * d, c, c1 * om1, g);
  }
  if (i < n) {
    const double omega = cA * __ldg(M.alpha + i) + cW * panel_at(wpanel, n - 1 - i, c);
    accum_dlogk<KERNEL>(M, xs, M.Xrm + (size_t)i * d, c, __ldcg(cpanel + (size_t)i * BN + c) * omega, g);
  }
}

__global__ void __launch_bounds__(NT, 2)
k_grad_narrow(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
              double* __restrict__ vscratch, double* __restrict__ wscratch, double* __restrict__ cscratch,
              size_t panel_doubles, double* __restrict__ out_grad, double* __restrict__ out_main) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const double* xs = reinterpret_cast<const double*>(smem + SM_CAND_OFF);
  const double* s2acc = reinterpret_cast<const double*>(smem + SM_S2_OFF);
  double* vpanel = vscratch + (size_t)blockIdx.x * panel_doubles;
  double* wpanel = wscratch + (size_t)blockIdx.x * (panel_doubles + COEF_DOUBLES);
  double* coef = wpanel + panel_doubles;
  double* cpanel = cscratch + (size_t)blockIdx.x * panel_doubles;
  const int n = M.n, d = M.d, p = M.p;
  pipeline_init(smem, tid);
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * BNN;
    forward_tile_narrow(M, x, N, c0, vpanel, smem, tid, &stage_base, cpanel);
    if (tid < BNN) {
      const int c = tid;
      double mean, sd, u[DK_MAX_P], cA, cW;
      bool drop_r;
      cand_predict_narrow(M, P.type, vpanel, xs, s2acc, c, mean, sd, u);
      if (out_main && c0 + c < N) out_main[c0 + c] = crit_value(P, mean, sd);
      grad_coefficients(M, P, mean, sd, cA, cW, drop_r);
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
    // backward blocked solve with the flipped factor
#pragma unroll 1
    for (int I = 0; I < M.nblk_b; I++) {
      double acc[2][2][2];
#pragma unroll
      for (int mi = 0; mi < 2; mi++) {
        const int ip = I * BM + 16 * warp + 8 * mi + (lane >> 2);
        const int i = n - 1 - ip;
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int c = 8 * ni + 2 * (lane & 3) + e;
            double v = 0.0;
            if (ip < n) {
              v = panel_at(vpanel, i, c);
              if (P.type == DK_UK)
                for (int j = 0; j < p; j++) v = fma(__ldg(M.Mcm + i + (size_t)j * n), __ldcg(coef + (2 + j) * BN + c), v);
            }
            acc[mi][ni][e] = v;
          }
      }
      solve_block_row_narrow(acc, M.tiles_bwd, I, wpanel, smem, tid, &stage_base, min(BM, n - I * BM));
      __syncthreads();
    }
    // gradient reduction: 16 row slices x 16 candidates
    {
      const int c = tid & (BNN - 1), slice = tid >> 4;
      const double cA = __ldcg(coef + c), cW = __ldcg(coef + BN + c);
      double g[DK_MAX_D];
#pragma unroll
      for (int k = 0; k < DK_MAX_D; k++) g[k] = 0.0;
      switch (M.kernel) {
        case DK_GAUSS: grad_accumulate_narrow<DK_GAUSS>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_EXP: grad_accumulate_narrow<DK_EXP>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_MATERN3_2: grad_accumulate_narrow<DK_MATERN3_2>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        case DK_MATERN5_2: grad_accumulate_narrow<DK_MATERN5_2>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
        default: grad_accumulate_narrow<DK_POWEXP>(M, xs, cpanel, wpanel, c, slice, cA, cW, g); break;
      }
      double* gbuf = reinterpret_cast<double*>(smem);  // [16 slices][DK_MAX_D][16] in region 0 (48 KB)
#pragma unroll
      for (int k = 0; k < DK_MAX_D; k++)
        if (k < d) gbuf[(slice * DK_MAX_D + k) * BNN + c] = g[k];
      __syncthreads();
      if (c0 + c < N) {
        for (int k = slice; k < d; k += 16) {
          double s = 0.0;
#pragma unroll
          for (int sl = 0; sl < 16; sl++) s += gbuf[(sl * DK_MAX_D + k) * BNN + c];
          if (M.trend == DK_TREND_LINEAR) s += cA * __ldg(M.beta + 1 + k) - cW * __ldcg(coef + (2 + 1 + k) * BN + c);
          out_grad[(c0 + c) + (int64_t)k * N] = s;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace dk
