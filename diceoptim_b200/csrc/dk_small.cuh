// dk_small.cuh — forward kernel for SMALL designs (n + p + 1 <= 256), e.g. BASELINE cfg1 (n = 20) and cfg2 (n = 200).
//
// With one or two 128-row block rows per tile the blocked kernel (dk_kernels.cuh) is barrier- and latency-bound: ~14 CTA
// barriers and 8 serialised inverse-tile fetches per block row, FP64 pipe 50 % busy at cfg2.  Here the WHOLE padded factor
// is one diagonal block: its explicit inverse (N x N, N = n_pad <= 256, formed in long double on the host like the 128 x 128
// block inverses, and rounded once) is applied as a lower-triangular GEMM  V = inv(L_aug) c  straight from the covariance
// tile.  No V panel: the sum of v^2 and the augmented rows (-z'v, -M'v) are taken from the accumulators, so the kernel uses
// NO global scratch and DRAM traffic is the algorithmic 8 d + 8 bytes per candidate.
//
// One 512-thread CTA per SM, 64 candidates per tile, 16 warps as 4 x 4 (warp tile 32 x 16).  Per tile: covariance
// generation straight into B fragments (one 16-row chunk per warp), then up to 24 GEMM stages with no CTA barrier between
// them, then the epilogue — which warps 0 and 1 run for tile t - 1 while the other warps already generate tile t.
// The factor tiles a candidate tile needs are the same for every tile, so thread 0 streams that list round and round
// through a ring of 16 KB stages (bulk copies + mbarriers) that never drains: not between row blocks, not between
// tiles, not during covariance generation.  The ring takes whatever shared memory the model leaves (small_layout).
// Measured on the way (cfg2 / cfg1, 4e6 candidates, ms): blocked kernel 13.1 / 3.9; two 256-thread CTAs per SM with 32
// candidates and a 2-stage ring 12.4 / 2.9; this shape with a ring that drains per row block and design rows re-staged
// per tile 11.2 / 3.7; warp-specialised (8 producer warps generating the covariance of the next 32 candidates while 8
// consumer warps multiply) 13.9 / 3.0 — eight warps with eight accumulators each do not keep the DMMA pipe busy; both
// row blocks advancing together (16 accumulators per warp, the B fragments read once) 16.2 / 3.2: over the 128-register
// limit of a 512-thread CTA; this kernel as two 256-thread CTAs of 32 candidates per SM (3-stage ring each) 13.2 / 2.5:
// twice the factor-tile traffic from L2 (3.2 TB/s) costs more than the overlap of the two CTAs' phases gains; a dataflow
// form without CTA barriers (per-chunk mbarriers, covariance units taken by idle warps) 14.2 / 2.3.  What did pay: several
// k-chunks per ring stage (12.1 / 11.2 / 10.9 ms for one / two / three at cfg2): the hand-shakes, not the DMMA pipe or the
// copies, were the cost (what-if builds -DDK_X_SMALL_NOMMA / NOCOV / NOLOAD: skeleton 2.2, covariance 3.3, DMMA 6.5 ms).
// The explicit inverse of the whole factor is less accurate than the blocked substitution on ill-conditioned factors
// (the off-diagonal block of the inverse cancels against the diagonal one), so dk_model_create enables this kernel only
// when a probe against long-double substitution shows it loses nothing (small_inverse_is_accurate, dk_api.cu).
#pragma once
#include "dk_kernels.cuh"

namespace dk {

constexpr int SBN = BN;                 // 64 candidates per tile; B tiles in the wide kernel's 16 x 64 fragment order
constexpr int NTS = 512;                // 16 warps: 4 (rows) x 4 (columns), warp tile 32 x 16
constexpr int SMALL_MAX_PAD = 256;
constexpr int SMALL_MAX_STAGES = 8;
constexpr int SMALL_MAX_TPS = 4;        // most k-chunks (16 KB factor tiles) per ring stage
constexpr int SMALL_SMEM_MAX = 232448;  // 227 KB of dynamic shared memory per CTA

// Shared-memory map, sized for the model: what the fixed parts leave goes to the ring (cfg2: 6 stages).  Designs whose
// rows x d leave fewer than 3 stages stay on the blocked kernel.
__host__ inline SmallLayout small_layout(int n, int n_pad, int d, int p) {
  SmallLayout L{};
  if (n_pad > SMALL_MAX_PAD) return L;
  const int nchunks = (n + 15) / 16;
  L.off_aug = 256 + 4 * SBN * 8;
  L.off_cand = L.off_aug + (p + 1) * SBN * 8;
  L.off_xs = L.off_cand + 4 * d * SBN * 8;   // [2 buffers][plain | scaled by w][d][64]
  L.off_rbuf = L.off_xs + n_pad * d * 8;
  L.off_ring = L.off_rbuf + nchunks * B_TILE * 8;
  // several k-chunks per ring stage: every stage costs each warp two mbarrier hand-shakes, and the factor tiles of a row
  // block are consecutive in memory, so a stage is still ONE bulk copy.  The most chunks per stage that leave two stages
  // (the copies themselves are cheap: a build that never fetches them is 1.4 % faster); DK_SMALL_TPS overrides (A/B runs).
  if (L.off_ring > SMALL_SMEM_MAX) return L;
  static const int forced = [] { const char* e = std::getenv("DK_SMALL_TPS"); return e ? std::atoi(e) : 0; }();
  for (int tps = SMALL_MAX_TPS; tps >= 1; tps--) {
    if (forced >= 1 && forced <= SMALL_MAX_TPS && tps != forced) continue;
    const int room = (SMALL_SMEM_MAX - L.off_ring) / (tps * A_TILE * 8);
    if (room < 2) continue;
    L.tps = tps;
    L.stages = room < SMALL_MAX_STAGES ? room : SMALL_MAX_STAGES;
    L.total = L.off_ring + L.stages * tps * A_TILE * 8;
    break;
  }
  return L;
}
// first tile of row block rb in the small tile array: row block rb' < rb holds 8 (rb' + 1) tiles
__host__ __device__ __forceinline__ size_t small_tile_base(int rb) { return (size_t)4 * rb * (rb + 1); }

// mbarrier wait that cannot hang the GPU: a protocol error traps (the launch fails) instead of spinning for ever
__device__ __forceinline__ void mbar_wait_guarded(unsigned long long* bar, unsigned parity) {
  unsigned done = 0;
  long long t0 = 0;
  for (unsigned long long spin = 0; !done; spin++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (!done && (spin & 1023) == 1023) {        // ~2 s of waiting is a protocol error, not a slow copy
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000ll) __trap();
    }
  }
}

// covariance of the 16 design rows of k-chunk `chunk` with candidates 32 half .. 32 half + 31, written as B fragments.
// The lane is the fragment's own (lane = 4 nl + kl): it owns row 4 k4 + kl of each of the chunk's four row groups and the
// candidates nl + 8 g, so a warp's stores are whole conflict-free 16-byte fragment entries; 4 rows x 4 candidates = 16
// chains per thread.  xsw holds the candidates already scaled by w_k (the arithmetic per pair is covgen_rows' own).
template <int KERNEL, bool NUG>
__device__ __forceinline__ void covgen_small(const DevModel& M, const double* __restrict__ Xs, const double* __restrict__ xsw,
                                             double* __restrict__ Rbuf, int chunk, int half, int lane) {
  const int d = M.d, kl = lane & 3, nl = lane >> 2;
  const int nk4 = min(4, (M.n - 16 * chunk + 3) >> 2);      // row groups of this chunk that hold design rows
  double* out = Rbuf + (size_t)chunk * B_TILE;
  double s[4][4], pr[4][4], hs[NUG ? 4 : 1][4];
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++)
#pragma unroll
    for (int g = 0; g < 4; g++) { s[k4][g] = 0.0; pr[k4][g] = 1.0; if (NUG) hs[k4][g] = 0.0; }
  const double* Xr = Xs + (size_t)(16 * chunk + kl) * d;
  const double* xc = xsw + 32 * half + nl;
#pragma unroll 1
  for (int k = 0; k < d; k++) {
    const double iwk = NUG ? M.iw[k] : 0.0;
    double c[4];
#pragma unroll
    for (int g = 0; g < 4; g++) c[g] = xc[k * SBN + 8 * g];
#pragma unroll
    for (int k4 = 0; k4 < 4; k4++) {
      if (k4 < nk4) {
        const double xr = Xr[(4 * k4) * d + k];
#pragma unroll
        for (int g = 0; g < 4; g++) {
          const double a = fabs(__dsub_rn(xr, c[g]));
          if (NUG) hs[k4][g] = fma(a, iwk, hs[k4][g]);
          if (KERNEL == DK_GAUSS) s[k4][g] = fma(0.5 * a, a, s[k4][g]);
          else if (KERNEL == DK_EXP) s[k4][g] += a;
          else if (KERNEL == DK_MATERN3_2) { pr[k4][g] *= (1.0 + a); s[k4][g] += a; }
          else if (KERNEL == DK_MATERN5_2) { pr[k4][g] *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0); s[k4][g] += a; }
          else s[k4][g] += pow(a, M.shape[k]);
        }
      }
    }
  }
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const bool live = k4 < nk4 && 16 * chunk + 4 * k4 + kl < M.n;
    double v[4];
#pragma unroll
    for (int g = 0; g < 4; g++) {
      double c = 0.0;
      if (live) {
        c = M.sd2 * pr[k4][g] * exp(-s[k4][g]);
        if (NUG && hs[k4][g] < 1e-15) c += M.nugget;
      }
      v[g] = c;
    }
#pragma unroll
    for (int q = 0; q < 2; q++)   // fragment (k4, np = 2 half + q): candidates 8 (2 np) + nl and 8 (2 np + 1) + nl
      *reinterpret_cast<double2*>(out + ((k4 * 4 + 2 * half + q) * 32 + lane) * 2) = make_double2(v[2 * q], v[2 * q + 1]);
  }
}

__device__ __forceinline__ void covgen_small_any(const DevModel& M, const double* Xs, const double* xsw, double* Rbuf, int chunk,
                                                 int half, int lane) {
#define DK_SCOV(K)                                                                          \
  case K:                                                                                   \
    if (M.nugget_flag) covgen_small<K, true>(M, Xs, xsw, Rbuf, chunk, half, lane);          \
    else covgen_small<K, false>(M, Xs, xsw, Rbuf, chunk, half, lane);                       \
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

// per-candidate predict.km epilogue of one tile (cand_predict with the augmented rows read from shared memory)
__device__ __forceinline__ void small_epilogue(const DevModel& M, const CritParams& P, const double* __restrict__ xs,
                                               const double* __restrict__ aug, const double* __restrict__ s2acc, int c,
                                               int64_t g, double* __restrict__ out_main, double* __restrict__ out_mean,
                                               double* __restrict__ out_sd) {
  const int p = M.p;
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
  if (out_mean) out_mean[g] = mean;
  if (out_sd) out_sd[g] = sd;
  if (out_main) {
    double v = crit_value(P, mean, sd);
    if (P.combine == COMBINE_MUL) v *= out_main[g];
    else if (P.combine == COMBINE_MIN) v = fmin(v, out_main[g]);
    out_main[g] = v;
  }
}

__global__ void __launch_bounds__(NTS, 1)
k_predict_small(const DevModel M, const CritParams P, const SmallLayout L, const double* __restrict__ x, int64_t N,
                int64_t ntiles, double* __restrict__ out_main, double* __restrict__ out_mean, double* __restrict__ out_sd) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smem);            // ring: tile landed
  unsigned long long* empty = full + SMALL_MAX_STAGES;                                // ring: slot consumed (16 warps)
  double* s2acc = reinterpret_cast<double*>(smem + 256);
  double* aug = reinterpret_cast<double*>(smem + L.off_aug);
  double* cand = reinterpret_cast<double*>(smem + L.off_cand);                        // [2][plain | scaled][d][64]
  double* Xs = reinterpret_cast<double*>(smem + L.off_xs);
  double* Rbuf = reinterpret_cast<double*>(smem + L.off_rbuf);
  double* ring = reinterpret_cast<double*>(smem + L.off_ring);
  const int n = M.n, d = M.d, p = M.p, nrb = M.nblk;
  const int live16 = (M.n_aug + 15) >> 4;        // 16-row pairs that hold real rows (design + augmented)
  const int nchunks = (n + 15) >> 4;             // k-chunks with a non-zero right-hand side (covariance rows)
  const int nstages = L.stages;
  const int nunits = 2 * nchunks;                // covariance work: (chunk, candidate half), dealt from warp 15 downwards
  if (tid == 0) {
    for (int s = 0; s < nstages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], NTS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < M.n_pad * d; i += NTS) Xs[i] = __ldg(M.Xrm_s + i);   // the design does not depend on the tile
  auto load_candidates = [&](int64_t tile, double* xs) {
    const int64_t c0 = tile * SBN;
    for (int i = tid; i < d * SBN; i += NTS) {
      const int k = i / SBN, j = i % SBN;
      const int64_t g = c0 + j;
      const double v = (g < N) ? x[g + (int64_t)k * N] : x[(N - 1) + (int64_t)k * N];
      xs[i] = v;
      xs[d * SBN + i] = __dmul_rn(v, M.w[k]);   // never contracted into an FMA: a pair's value cannot depend on its tile position
    }
  };
  load_candidates(blockIdx.x, cand);
  __syncthreads();

  // factor-tile stream: row block 0's nk0 chunks, then row block 1's, L.tps chunks (consecutive tiles, one bulk copy) per
  // ring stage; the same list of stages for every candidate tile
  const int nk0 = min(CHUNKS_PER_BLOCK, nchunks);
  const int tps = L.tps;
  const int ns0 = (nk0 + tps - 1) / tps, ns1 = nrb > 1 ? (nchunks + tps - 1) / tps : 0;
  const int per_tile = ns0 + ns1;
  const int64_t my_tiles = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const int64_t total = my_tiles * per_tile;
  int64_t issued = 0;
  int i_pos = 0, i_slot = 0; unsigned i_par = 0;     // thread 0: next list position / slot / parity to issue
  int c_slot = 0; unsigned c_par = 0;                // every warp: next slot / parity to consume
  auto issue = [&]() {   // thread 0 only
    mbar_wait_guarded(&empty[i_slot], i_par ^ 1);
    const int rb = i_pos < ns0 ? 0 : 1, st = i_pos - (rb ? ns0 : 0), kc = st * tps;
    const int ntl = min(tps, (rb ? nchunks : nk0) - kc);
#ifdef DK_X_SMALL_NOLOAD   // what-if build (timing only): the factor tiles are never fetched
    mbar_arrive(&full[i_slot]);
#else
    mbar_arrive_expect_tx(&full[i_slot], ntl * A_TILE * 8);
    const double* src = M.tiles_small + (small_tile_base(rb) + (size_t)kc) * A_TILE;
    bulk_load(ring + (size_t)i_slot * tps * A_TILE, src, ntl * A_TILE * 8, &full[i_slot]);
#endif
    if (++i_pos == per_tile) i_pos = 0;
    if (++i_slot == nstages) { i_slot = 0; i_par ^= 1; }
    issued++;
  };
  if (tid == 0)
    while (issued < total && issued < nstages - 1) issue();

  int it = 0;
  int64_t prev = -1;
#pragma unroll 1
  for (int64_t tile = blockIdx.x; tile < ntiles; prev = tile, tile += gridDim.x, it++) {
    const double* xs = cand + (size_t)(it & 1) * 2 * d * SBN;
    // ---- epilogue of the previous tile (warps 0, 1) beside the covariance tile of this one ----
    if (prev >= 0 && tid < SBN && prev * SBN + tid < N)
      small_epilogue(M, P, cand + (size_t)((it - 1) & 1) * 2 * d * SBN, aug, s2acc, tid, prev * SBN + tid, out_main, out_mean, out_sd);
#ifndef DK_X_SMALL_NOCOV   // what-if build (timing only): the covariance tile is never written
    for (int u = NTS / 32 - 1 - warp; u < nunits; u += NTS / 32)
#else
    for (int u = nunits; u < nunits; u++)
#endif
      covgen_small_any(M, Xs, xs + d * SBN, Rbuf, u >> 1, u & 1, lane);
    __syncthreads();   // Rbuf complete; aug / s2acc free
    // the next tile's candidates: asynchronous copies issued now land in shared memory while the GEMM runs (no registers held)
    const bool have_next = tile + gridDim.x < ntiles;
    double* xn = cand + (size_t)((it + 1) & 1) * 2 * d * SBN;
    if (have_next) {
      const int64_t c1 = (tile + gridDim.x) * SBN;
      for (int i = tid; i < d * SBN; i += NTS) {
        const int64_t g = c1 + (i % SBN);
        const double* src = x + min(g, N - 1) + (int64_t)(i / SBN) * N;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(xn + i)), "l"(src) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    double s2p[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    // ---- V = inv(L_aug) c, row block by row block, all stages on one pipeline ----
#pragma unroll 1
    for (int rb = 0; rb < nrb; rb++) {
      double acc[4][2][2];
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
      const int nk = rb == 0 ? nk0 : nchunks;
      const int plo = CHUNKS_PER_BLOCK * rb + 2 * wm, phi = plo + 1;   // this warp's global pair indices
#pragma unroll 1
      for (int kc0 = 0; kc0 < nk; kc0 += tps) {
        if (tid == 0 && issued < total) issue();   // nstages - 1 stages ahead: the slot this warp consumed one step ago
        mbar_wait_guarded(&full[c_slot], c_par);
#pragma unroll 1
        for (int t = 0; t < tps; t++) {
          const int kc = kc0 + t;
          // the inverse is lower triangular: pair q needs chunks 0..q; pairs made of padding take no part
          const bool lo = kc < nk && kc <= plo && plo < live16, hi = kc < nk && kc <= phi && phi < live16;
#ifndef DK_X_SMALL_NOMMA   // what-if build (timing only): no DMMA work
          if (lo || hi)
#else
          if (false)
#endif
            mma_stage_small(acc, ring + ((size_t)c_slot * tps + t) * A_TILE, Rbuf + (size_t)kc * B_TILE, wm, wn, lane, lo, hi);
        }
        // the slot is refilled through the async proxy: order this lane's generic-proxy reads of it before that (the blocked
        // kernel's ring raced without this fence: about one corrupted tile in a thousand, dk_kernels.cuh)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[c_slot]);
        if (++c_slot == nstages) { c_slot = 0; c_par ^= 1; }
      }
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
    // the next tile's candidates, into the buffer the epilogue above has finished with
    // ... and their copy scaled by w_k, by the thread that fetched them
    if (have_next) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      for (int i = tid; i < d * SBN; i += NTS) xn[d * SBN + i] = __dmul_rn(xn[i], M.w[i / SBN]);   // never contracted into an FMA
    }
    __syncthreads();   // aug / s2acc / candidates visible; every warp is done with Rbuf
  }
  if (prev >= 0 && tid < SBN && prev * SBN + tid < N)
    small_epilogue(M, P, cand + (size_t)((it - 1) & 1) * 2 * d * SBN, aug, s2acc, tid, prev * SBN + tid, out_main, out_mean, out_sd);
}

}  // namespace dk
