// dk_small.cuh — forward kernel for SMALL designs (n + p + 1 <= 256), e.g. BASELINE cfg1 (n = 20) and cfg2 (n = 200).
//
// With one or two 128-row block rows per tile the blocked kernel (dk_kernels.cuh) is barrier- and latency-bound: ~14 CTA
// barriers and 8 serialised inverse-tile fetches per block row, FP64 pipe 50 % busy at cfg2.  Here the WHOLE padded factor
// is one diagonal block: its explicit inverse (N x N, N = n_pad <= 256, formed in long double on the host like the 128 x 128
// block inverses, and rounded once) is applied as a lower-triangular GEMM  V = inv(L_aug) c  straight from the covariance
// tile, so a tile is: covariance generation (written directly as B fragments), then up to 24 GEMM stages on one
// bulk-copy / mbarrier pipeline with no CTA barrier between them, then the epilogue.  No V panel: the sum of v^2 and the
// augmented rows (-z'v, -M'v) are taken from the accumulators, so the kernel uses NO global scratch and DRAM traffic is
// the algorithmic 8 d + 8 bytes per candidate.  One CTA per SM: 64 candidates, 16 warps as 4 x 4 (warp tile 32 x 16), the
// covariance tile (up to 128 KB) resident in shared memory next to a 4-stage ring of 16 KB factor tiles, so that an L2
// round trip per stage is hidden (a first version with two 32-candidate CTAs per SM and 2 stages was latency-bound).
// The explicit inverse of the whole factor is less accurate than the blocked substitution on ill-conditioned factors
// (the off-diagonal block of the inverse cancels against the diagonal one), so dk_model_create enables this kernel only
// when a probe against long-double substitution shows it loses nothing (small_inverse_is_accurate, dk_api.cu).
#pragma once
#include "dk_kernels.cuh"

namespace dk {

constexpr int SBN = BN;                 // 64 candidates per tile; B tiles in the wide kernel's 16 x 64 fragment order
constexpr int NTS = 512;                // 16 warps: 4 (rows) x 4 (columns), warp tile 32 x 16
constexpr int S_STAGES = 4;
constexpr int SMALL_MAX_PAD = 256;
// shared-memory map (bytes), one CTA per SM
constexpr int SS_RBUF = 0;                                   // covariance tile as B fragments: N x 64 doubles (<= 128 KB)
constexpr int SS_RING = SMALL_MAX_PAD * SBN * 8;             // 131072: 4 A-tile stages; the design rows of cov-gen alias it
constexpr int SS_CAND = SS_RING + S_STAGES * A_TILE * 8;     // 196608: xs[d][64]
constexpr int SS_S2 = SS_CAND + DK_MAX_D * SBN * 8;          // 208896: s2acc[4][64]
constexpr int SS_AUG = SS_S2 + 4 * SBN * 8;                  // 210944: aug[p + 1][64]
constexpr int SS_BAR = SS_AUG + (DK_MAX_P + 1) * SBN * 8;    // 224256: 4 full + 4 empty mbarriers
constexpr int SS_TOTAL = SS_BAR + 64;                        // 224320 (<= 227 KB)
// first tile of row block rb in the small tile array: row block rb' < rb holds 8 (rb' + 1) tiles
__host__ __device__ __forceinline__ size_t small_tile_base(int rb) { return (size_t)4 * rb * (rb + 1); }

// covariance of ALL design rows with the 64 candidates, written as B fragments into Rbuf.  Warp w (of 16) owns rows
// 16 w .. 16 w + 15 (= k-chunk w), lane owns candidates lane and lane + 32; two passes of 8 rows x 2 columns = 16 chains
// (the arithmetic per pair is covgen_rows' own).
template <int KERNEL, bool NUG>
__device__ __forceinline__ void covgen_small(const DevModel& M, const double* __restrict__ Xs, const double* __restrict__ xs,
                                             double* __restrict__ Rbuf, int warp, int lane) {
  constexpr int RB = 8;
  const int d = M.d;
  double* chunk = Rbuf + (size_t)warp * B_TILE;
#pragma unroll 1
  for (int rr = 0; rr < 16; rr += RB) {
    const int r0 = 16 * warp + rr;
    if (r0 >= M.n) {
#pragma unroll
      for (int j = 0; j < RB; j++) {
        chunk[b_tile_index(rr + j, lane)] = 0.0;
        chunk[b_tile_index(rr + j, lane + 32)] = 0.0;
      }
      continue;
    }
    double s[RB][2], pr[RB][2], hs[NUG ? RB : 1][2];
#pragma unroll
    for (int j = 0; j < RB; j++) {
      s[j][0] = s[j][1] = 0.0;
      pr[j][0] = pr[j][1] = 1.0;
      if (NUG) hs[j][0] = hs[j][1] = 0.0;
    }
    const double* Xr = Xs + (size_t)r0 * d;
#pragma unroll 1
    for (int k = 0; k < d; k++) {
      const double wk = M.w[k];
      const double c0 = __dmul_rn(xs[k * SBN + lane], wk), c1 = __dmul_rn(xs[k * SBN + lane + 32], wk);
      const double iwk = NUG ? M.iw[k] : 0.0;
#pragma unroll
      for (int j = 0; j < RB; j++) {
        const double xr = Xr[j * d + k];
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const double a = fabs(__dsub_rn(xr, e ? c1 : c0));
          if (NUG) hs[j][e] = fma(a, iwk, hs[j][e]);
          if (KERNEL == DK_GAUSS) s[j][e] = fma(0.5 * a, a, s[j][e]);
          else if (KERNEL == DK_EXP) s[j][e] += a;
          else if (KERNEL == DK_MATERN3_2) { pr[j][e] *= (1.0 + a); s[j][e] += a; }
          else if (KERNEL == DK_MATERN5_2) { pr[j][e] *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0); s[j][e] += a; }
          else s[j][e] += pow(a, M.shape[k]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < RB; j++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        double c = M.sd2 * pr[j][e] * exp(-s[j][e]);
        if (NUG && hs[j][e] < 1e-15) c += M.nugget;
        chunk[b_tile_index(rr + j, lane + 32 * e)] = (r0 + j < M.n) ? c : 0.0;
      }
  }
}

__device__ __forceinline__ void covgen_small_any(const DevModel& M, const double* Xs, const double* xs, double* Rbuf, int warp,
                                                 int lane) {
#define DK_SCOV(K)                                                                   \
  case K:                                                                            \
    if (M.nugget_flag) covgen_small<K, true>(M, Xs, xs, Rbuf, warp, lane);           \
    else covgen_small<K, false>(M, Xs, xs, Rbuf, warp, lane);                        \
    break;
  switch (M.kernel) {
    DK_SCOV(DK_GAUSS) DK_SCOV(DK_EXP) DK_SCOV(DK_MATERN3_2) DK_SCOV(DK_MATERN5_2)
    default: DK_SCOV(DK_POWEXP)
  }
#undef DK_SCOV
}

// one k-chunk on the warp's 32 x 16 tile; lo / hi: which of the warp's two 16-row pairs takes part
__device__ __forceinline__ void mma_stage_small(double (&acc)[4][2][2], const double* __restrict__ As,
                                                const double* __restrict__ Bs, int wm, int wn, int lane, bool lo, bool hi) {
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const double2 b = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + wn) * 32 + lane) * 2);
    if (lo) {
      const double2 a = *reinterpret_cast<const double2*>(As + (((2 * wm) * 4 + k4) * 32 + lane) * 2);
      dmma884(acc[0][0], a.x, b.x); dmma884(acc[0][1], a.x, b.y);
      dmma884(acc[1][0], a.y, b.x); dmma884(acc[1][1], a.y, b.y);
    }
    if (hi) {
      const double2 a = *reinterpret_cast<const double2*>(As + (((2 * wm + 1) * 4 + k4) * 32 + lane) * 2);
      dmma884(acc[2][0], a.x, b.x); dmma884(acc[2][1], a.x, b.y);
      dmma884(acc[3][0], a.y, b.x); dmma884(acc[3][1], a.y, b.y);
    }
  }
}

__global__ void __launch_bounds__(NTS, 1)
k_predict_small(const DevModel M, const CritParams P, const double* __restrict__ x, int64_t N, int64_t ntiles,
                double* __restrict__ out_main, double* __restrict__ out_mean, double* __restrict__ out_sd) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  double* Rbuf = reinterpret_cast<double*>(smem + SS_RBUF);
  double* ring = reinterpret_cast<double*>(smem + SS_RING);
  double* Xs = ring;   // ALL design rows (<= 256 x d doubles = 48 KB), staged here while the ring is idle
  double* xs = reinterpret_cast<double*>(smem + SS_CAND);
  double* s2acc = reinterpret_cast<double*>(smem + SS_S2);
  double* aug = reinterpret_cast<double*>(smem + SS_AUG);
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smem + SS_BAR);
  unsigned long long* empty = full + S_STAGES;
  const int n = M.n, d = M.d, p = M.p, nrb = M.nblk;
  const int live16 = (M.n_aug + 15) >> 4;        // 16-row pairs that hold real rows (design + augmented)
  const int nchunks = (n + 15) >> 4;             // k-chunks with a non-zero right-hand side (covariance rows)
  if (tid == 0) {
    for (int s = 0; s < S_STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], NTS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned stage_base = 0;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * SBN;
    __syncthreads();   // the previous tile's epilogue has read xs / aug / s2acc
    for (int i = tid; i < d * SBN; i += NTS) {
      const int k = i / SBN, j = i % SBN;
      const int64_t g = c0 + j;
      xs[i] = (g < N) ? x[g + (int64_t)k * N] : x[(N - 1) + (int64_t)k * N];
    }
    // ---- covariance tile of all row blocks at once, as B fragments ----
    __syncthreads();   // Xs free (the previous tile's ring reads are over), xs visible
    for (int i = tid; i < nrb * BM * d; i += NTS) Xs[i] = __ldg(M.Xrm_s + i);
    __syncthreads();
    if (warp < nrb * CHUNKS_PER_BLOCK) covgen_small_any(M, Xs, xs, Rbuf, warp, lane);
    // Xs (generic-proxy writes) gives the ring back to the bulk copies; Rbuf must be complete before any warp reads it
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncthreads();
    double s2p[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    // ---- V = inv(L_aug) c, row block by row block, all stages on one pipeline ----
#pragma unroll 1
    for (int rb = 0; rb < nrb; rb++) {
      double acc[4][2][2];
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
      const int nk = min(CHUNKS_PER_BLOCK * (rb + 1), nchunks);
      const double* Atiles = M.tiles_small + small_tile_base(rb) * A_TILE;
      const int plo = CHUNKS_PER_BLOCK * rb + 2 * wm, phi = plo + 1;   // this warp's global pair indices
      auto issue = [&](int idx) {   // thread 0 only
        const unsigned gi = stage_base + idx, slot = gi % S_STAGES, par = (gi / S_STAGES) & 1;
        mbar_wait(&empty[slot], par ^ 1);
        mbar_arrive_expect_tx(&full[slot], A_TILE * 8);
        bulk_load(ring + slot * A_TILE, Atiles + (size_t)idx * A_TILE, A_TILE * 8, &full[slot]);
      };
      if (tid == 0)
        for (int s = 0; s < S_STAGES - 1 && s < nk; s++) issue(s);
#pragma unroll 1
      for (int kc = 0; kc < nk; kc++) {
        if (tid == 0 && kc + S_STAGES - 1 < nk) issue(kc + S_STAGES - 1);
        const unsigned gi = stage_base + kc, slot = gi % S_STAGES, par = (gi / S_STAGES) & 1;
        mbar_wait(&full[slot], par);
        // the inverse is lower triangular: pair q needs chunks 0..q; pairs made of padding take no part
        const bool lo = kc <= plo && plo < live16, hi = kc <= phi && phi < live16;
        if (lo || hi) mma_stage_small(acc, ring + slot * A_TILE, Rbuf + (size_t)kc * B_TILE, wm, wn, lane, lo, hi);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
      }
      stage_base += nk;
      // ---- sum of v^2 over the design rows; the augmented rows go to `aug` ----
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const int r = rb * BM + 32 * wm + 8 * mi + (lane >> 2);
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const double v = acc[mi][ni][e];
            if (r < n) s2p[ni][e] = fma(v, v, s2p[ni][e]);
            else if (r < n + p + 1) aug[(r - n) * SBN + 16 * wn + 8 * ni + 2 * (lane & 3) + e] = v;
          }
      }
    }
#pragma unroll
    for (int ni = 0; ni < 2; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        double v = s2p[ni][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (lane < 4) s2acc[wm * SBN + 16 * wn + 8 * ni + 2 * lane + e] = v;
      }
    __syncthreads();
    // ---- per-candidate predict.km epilogue (cand_predict with the augmented rows read from shared memory) ----
    if (tid < SBN && c0 + tid < N) {
      const int c = tid;
      double S[DK_MAX_P], u[DK_MAX_P];
      double mean = 0.0;
      for (int j = 0; j < p; j++) {
        const double fj = (j == 0) ? 1.0 : xs[(j - 1) * SBN + c];
        mean = fma(fj, __ldg(M.beta + j), mean);
        u[j] = fj + aug[(1 + j) * SBN + c];
      }
      mean += -aug[c];
      double s2 = M.total_sd2 - (((s2acc[c] + s2acc[SBN + c]) + s2acc[2 * SBN + c]) + s2acc[3 * SBN + c]);
      if (P.type == DK_UK) {
        double s22 = 0.0;
        for (int j = 0; j < p; j++) {
          double t = u[j];
          for (int i = 0; i < j; i++) t -= __ldg(M.RM + i + j * p) * S[i];
          S[j] = t / __ldg(M.RM + j + j * p);
          s22 = fma(S[j], S[j], s22);
        }
        s2 += s22;
      }
      const double sd = sqrt(fmax(s2, 0.0));
      const int64_t g = c0 + c;
      if (out_mean) out_mean[g] = mean;
      if (out_sd) out_sd[g] = sd;
      if (out_main) {
        double v = crit_value(P, mean, sd);
        if (P.combine == COMBINE_MUL) v *= out_main[g];
        else if (P.combine == COMBINE_MIN) v = fmin(v, out_main[g]);
        out_main[g] = v;
      }
    }
  }
}

}  // namespace dk
