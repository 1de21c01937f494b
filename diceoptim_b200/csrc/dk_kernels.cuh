// dk_kernels.cuh — device code of the fused kriging-prediction kernel for sm_100a.
//
// One CTA owns a tile of BN = 64 candidates and walks the block rows of the augmented
// Cholesky factor top to bottom (left-looking blocked forward substitution):
//
//   for block row I:
//     P1  c_I   = sd2 * prod_k k1d(|X_ik - x_jk|)                 fused covariance generation
//                                                                 (replaces covMat1Mat2)
//     P3  acc   = c_I - sum_{J<I} L_IJ V_J                        FP64 DMMA m8n8k4, A = pre-tiled -L from L2,
//                                                                 B = this CTA's V panel (3-stage cp.async ring)
//     P4  V_I   = inv(L_II) acc                                   block inverse applied as a lower-triangular GEMM
//                                                                 (replaces backsolve(t(T), c))
//     P5  s2   += colsum(V_I^2); V_I -> per-CTA panel in HBM/L2   (apply(Tinv.c, 2, crossprod))
//   epilogue: rows n..n+p of the augmented system hold -z'v and -M'v  -> mean, sd, EI / quantile.
//
// The V panel is the kernel's scratch: n_pad x 64 doubles per CTA, written once and re-read
// (n_pad / 256) times on average; at n = 2000 it is 1 MB per CTA and mostly L2 traffic.
#pragma once
#include "dk_common.cuh"

namespace dk {

// Optional per-phase cycle accounting (build with -DDK_PHASE_TIMING; tools/phase_timing.py): thread 0 of every CTA
// adds the cycles between marks to g_phase_cycles[phase].  Not compiled into the shipped library.
#ifdef DK_PHASE_TIMING
static __device__ unsigned long long g_phase_cycles[8];
#define DK_MARK(ph)                                                                  \
  do {                                                                               \
    if (threadIdx.x == 0) {                                                          \
      const long long now__ = clock64();                                             \
      atomicAdd(&g_phase_cycles[ph], (unsigned long long)(now__ - dk_t0__));         \
      dk_t0__ = now__;                                                               \
    }                                                                                \
  } while (0)
#define DK_MARK_INIT long long dk_t0__ = clock64()
#else
#define DK_MARK(ph) do { } while (0)
#define DK_MARK_INIT do { } while (0)
#endif

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c[0]), "+d"(c[1])
      : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// 16 DMMAs per k4 step on the warp's 32 x 32 tile; As/Bs are fragment-ordered tiles in smem.
__device__ __forceinline__ void mma_stage(double (&acc)[4][4][2], const double* __restrict__ As,
                                          const double* __restrict__ Bs, int wm, int wn, int lane) {
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const double2 a01 = *reinterpret_cast<const double2*>(As + (((2 * wm) * 4 + k4) * 32 + lane) * 2);
    const double2 a23 = *reinterpret_cast<const double2*>(As + (((2 * wm + 1) * 4 + k4) * 32 + lane) * 2);
    const double2 b01 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn) * 32 + lane) * 2);
    const double2 b23 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn + 1) * 32 + lane) * 2);
    const double a[4] = {a01.x, a01.y, a23.x, a23.y};
    const double b[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int ni = 0; ni < 4; ni++) dmma884(acc[mi][ni], a[mi], b[ni]);
  }
}
// The same for ONE of the warp's two 16-row pairs (half = 0: rows 32 wm .. +15, acc[0..1]; half = 1: acc[2..3]): the
// last block row of a small design is mostly padding, and a pair that is all padding has nothing to accumulate.
__device__ __forceinline__ void mma_stage_half(double (&acc)[4][4][2], const double* __restrict__ As,
                                               const double* __restrict__ Bs, int wm, int wn, int lane, int half) {
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const double2 a = *reinterpret_cast<const double2*>(As + (((2 * wm + half) * 4 + k4) * 32 + lane) * 2);
    const double2 b01 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn) * 32 + lane) * 2);
    const double2 b23 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn + 1) * 32 + lane) * 2);
    const double b[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
    for (int ni = 0; ni < 4; ni++) { dmma884(acc[2 * half][ni], a.x, b[ni]); dmma884(acc[2 * half + 1][ni], a.y, b[ni]); }
  }
}
// live16 = number of 16-row pairs of this block row that hold real rows (8 everywhere but in the last block row)
__device__ __forceinline__ void mma_stage_live(double (&acc)[4][4][2], const double* __restrict__ As,
                                               const double* __restrict__ Bs, int wm, int wn, int lane, int live16) {
  if (2 * wm + 1 < live16) mma_stage(acc, As, Bs, wm, wn, lane);
  else if (2 * wm < live16) mma_stage_half(acc, As, Bs, wm, wn, lane, 0);
}

// Row ownership AFTER the diagonal phase of solve_block_row.  The inverse of a diagonal block is lower triangular: the
// 16-row pair q of the block needs the k-chunks 0..q only.  With the GEMM-phase ownership (warp row wm owns pairs 2wm,
// 2wm + 1) warp row 3 would do 15 pair-chunks and warp row 0 three; instead the diagonal phase gives warp row wm the
// pairs wm and 7 - wm (9 pair-chunks each), the right-hand side passing through shared memory anyway.  acc[0..1] then
// hold pair wm, acc[2..3] pair 7 - wm; everything downstream of the solve (panel store, column sums) uses solved_row.
__device__ __forceinline__ int solved_pair(int wm, int mi) { return mi < 2 ? wm : 7 - wm; }
__device__ __forceinline__ int solved_row(int wm, int mi, int lane) { return 16 * solved_pair(wm, mi) + 8 * (mi & 1) + (lane >> 2); }

// One k-chunk of the diagonal phase: acc[0..1] += inv[pair wm] * R, acc[2..3] += inv[pair 7 - wm] * R (when needed)
__device__ __forceinline__ void mma_stage_diag(double (&acc)[4][4][2], const double* __restrict__ As,
                                               const double* __restrict__ Bs, int wm, int wn, int lane, bool do_lo,
                                               bool do_hi) {
#pragma unroll
  for (int k4 = 0; k4 < 4; k4++) {
    const double2 b01 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn) * 32 + lane) * 2);
    const double2 b23 = *reinterpret_cast<const double2*>(Bs + ((k4 * 4 + 2 * wn + 1) * 32 + lane) * 2);
    const double b[4] = {b01.x, b01.y, b23.x, b23.y};
    if (do_lo) {
      const double2 a = *reinterpret_cast<const double2*>(As + ((wm * 4 + k4) * 32 + lane) * 2);
#pragma unroll
      for (int ni = 0; ni < 4; ni++) { dmma884(acc[0][ni], a.x, b[ni]); dmma884(acc[1][ni], a.y, b[ni]); }
    }
    if (do_hi) {
      const double2 a = *reinterpret_cast<const double2*>(As + (((7 - wm) * 4 + k4) * 32 + lane) * 2);
#pragma unroll
      for (int ni = 0; ni < 4; ni++) { dmma884(acc[2][ni], a.x, b[ni]); dmma884(acc[3][ni], a.y, b[ni]); }
    }
  }
}

__device__ __forceinline__ void load_tile_async(double* dst, const double* __restrict__ src, int ndoubles, int tid) {
  // straight copy, 16 B per cp.async
  for (int i = tid * 2; i < ndoubles; i += NT * 2) cp_async16(dst + i, src + i);
}

// ---------------------------------------------------------------------------------------
// P1: covariance block c[r][j] for the 128 design rows of block I and the tile's 64 candidates
// into Cbuf (row stride CB_STRIDE).  Warp w owns rows 16w..16w+15; lane owns columns lane, lane+32.
// 1-D kernels as DiceKriging codes them; matern kernels in product form
// prod(1 + a + a^2/3) * exp(-sum a), algebraically the reference's exp(-sum(a - log(1+a+a^2/3))).
// ---------------------------------------------------------------------------------------
// Xs holds the design rows PRE-SCALED by w_k (DevModel::Xrm_s); the candidate coordinate is scaled once per k, so a
// pair costs one DADD per dimension for a = |X_ik w_k - x_jk w_k|.  Eight rows x two columns = 16 independent
// chains per thread keep the FP64 pipe fed (the phase is latency-bound with fewer).
template <int KERNEL, bool NUG>
__device__ __forceinline__ void covgen_rows(const DevModel& M, const double* __restrict__ Xs,
                                            const double* __restrict__ xs, double* __restrict__ Cbuf,
                                            int row0_global, int warp, int lane) {
  constexpr int RB = 8;
  const int d = M.d;
#pragma unroll 1
  for (int rr = 0; rr < 16; rr += RB) {
#ifndef DK_X_NOSKIP
    if (row0_global + 16 * warp + rr >= M.n) {   // eight rows beyond the design (padding, augmented rows): zeros
#pragma unroll
      for (int j = 0; j < RB; j++) {
        Cbuf[(16 * warp + rr + j) * CB_STRIDE + lane] = 0.0;
        Cbuf[(16 * warp + rr + j) * CB_STRIDE + lane + 32] = 0.0;
      }
      continue;
    }
#endif
    double s[RB][2], pr[RB][2], hs[NUG ? RB : 1][2];
#pragma unroll
    for (int j = 0; j < RB; j++) {
      s[j][0] = s[j][1] = 0.0;
      pr[j][0] = pr[j][1] = 1.0;
      if (NUG) hs[j][0] = hs[j][1] = 0.0;
    }
    const double* Xr = Xs + (16 * warp + rr) * d;
#pragma unroll 1
    for (int k = 0; k < d; k++) {
      const double wk = M.w[k];
      // __dmul_rn / __dsub_rn: never contracted into an FMA, so a pair's value cannot depend on its tile position
      const double c0 = __dmul_rn(xs[k * BN + lane], wk), c1 = __dmul_rn(xs[k * BN + lane + 32], wk);
      const double iwk = NUG ? M.iw[k] : 0.0;
#pragma unroll
      for (int j = 0; j < RB; j++) {
        const double xr = Xr[j * d + k];
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const double a = fabs(__dsub_rn(xr, e ? c1 : c0));
          if (NUG) hs[j][e] = fma(a, iwk, hs[j][e]);
          if (KERNEL == DK_GAUSS) {
            s[j][e] = fma(0.5 * a, a, s[j][e]);
          } else if (KERNEL == DK_EXP) {
            s[j][e] += a;
          } else if (KERNEL == DK_MATERN3_2) {
            pr[j][e] *= (1.0 + a);
            s[j][e] += a;
          } else if (KERNEL == DK_MATERN5_2) {
            pr[j][e] *= fma(a, fma(a, 1.0 / 3.0, 1.0), 1.0);
            s[j][e] += a;
          } else {  // powexp
            s[j][e] += pow(a, M.shape[k]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < RB; j++) {
      const int r = 16 * warp + rr + j;
      const bool live = (row0_global + r) < M.n;
#pragma unroll
      for (int e = 0; e < 2; e++) {
        double c = M.sd2 * pr[j][e] * exp(-s[j][e]);
        if (NUG && hs[j][e] < 1e-15) c += M.nugget;  // covMat1Mat2(nugget.flag=TRUE) white-noise term
        Cbuf[r * CB_STRIDE + lane + 32 * e] = live ? c : 0.0;
      }
    }
  }
}

template <bool NUG>
__device__ __forceinline__ void covgen_block_n(const DevModel& M, const double* Xs, const double* xs, double* Cbuf,
                                               int row0_global, int warp, int lane) {
  switch (M.kernel) {
    case DK_GAUSS: covgen_rows<DK_GAUSS, NUG>(M, Xs, xs, Cbuf, row0_global, warp, lane); break;
    case DK_EXP: covgen_rows<DK_EXP, NUG>(M, Xs, xs, Cbuf, row0_global, warp, lane); break;
    case DK_MATERN3_2: covgen_rows<DK_MATERN3_2, NUG>(M, Xs, xs, Cbuf, row0_global, warp, lane); break;
    case DK_MATERN5_2: covgen_rows<DK_MATERN5_2, NUG>(M, Xs, xs, Cbuf, row0_global, warp, lane); break;
    default: covgen_rows<DK_POWEXP, NUG>(M, Xs, xs, Cbuf, row0_global, warp, lane); break;
  }
}
__device__ __forceinline__ void covgen_block(const DevModel& M, const double* Xs, const double* xs, double* Cbuf,
                                             int row0_global, int warp, int lane) {
  if (M.nugget_flag) covgen_block_n<true>(M, Xs, xs, Cbuf, row0_global, warp, lane);
  else covgen_block_n<false>(M, Xs, xs, Cbuf, row0_global, warp, lane);
}

// ---- mbarrier + bulk-copy (TMA 1-D) helpers for the warp-synchronous variant of the GEMM phase ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// one bulk asynchronous copy global -> shared, completion counted in bytes on `bar` (SASS: UBLKCP)
__device__ __forceinline__ void bulk_load(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// Barriers of the bulk-copy pipeline live right after the fixed shared-memory map (SM_TOTAL + PIPE_SMEM bytes).
constexpr int PIPE_SMEM = 64;
__device__ __forceinline__ void pipeline_init(unsigned char* smem, int tid) {
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(smem + SM_TOTAL);
  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&bars[s], 1);                 // full: one arrive.expect_tx by the issuing thread
      mbar_init(&bars[STAGES + s], NT / 32);  // empty: one arrive per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------
// One block row of the blocked substitution given its right-hand side already in `acc`
// (C-fragment layout).  On return acc holds V_I; the panel chunk for rows of block I has been
// written to `panel` (fragment-ordered B tiles).  `tiles` is the tile array to use (forward or
// backward factor).  Contains the __syncthreads the phase changes need; callers must sync once
// more before touching region 0 again.
// ---------------------------------------------------------------------------------------
// TMA = true: the GEMM phase is fed by bulk asynchronous copies (one 16 KB A tile + one 8 KB B tile per stage,
// issued by thread 0) and synchronised per stage by full/empty mbarriers instead of __syncthreads; `stage_base`
// counts the stages this CTA has consumed so far (slot = index % STAGES, parity = (index / STAGES) & 1).
// live_rows: rows of this block row that are real (< 128 only in the last block row of the system); 16-row pairs made of
// padding alone are skipped in both phases (their V is zero: zero right-hand side, identity diagonal).
template <bool TMA = false>
__device__ __forceinline__ void solve_block_row(double (&acc)[4][4][2], const double* __restrict__ tiles, int I,
                                                double* __restrict__ panel, unsigned char* smem, int tid,
                                                unsigned* stage_base = nullptr, int live_rows = BM) {
  DK_MARK_INIT;
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
#ifdef DK_X_NOSKIP   // what-if build: execute the padded pairs as round 1 did
  const int live16 = 8;
  (void)live_rows;
#else
  const int live16 = (live_rows + 15) >> 4;
#endif
  double* ring = reinterpret_cast<double*>(smem);
  const double* Atiles = tiles + tile_base(I) * A_TILE;
  const int nk = CHUNKS_PER_BLOCK * I;

  if constexpr (TMA) {
    unsigned long long* full = reinterpret_cast<unsigned long long*>(smem + SM_TOTAL);
    unsigned long long* empty = full + STAGES;
    const unsigned base = *stage_base;
    constexpr unsigned STAGE_BYTES = (A_TILE + B_TILE) * 8;
    // Region 0 was last written through the generic proxy (covariance tile, design rows, diagonal-phase buffers) and
    // the panel by generic global stores; every thread orders its own accesses against the async proxy, then the
    // CTA meets, and only then does thread 0 start issuing bulk copies into / from those locations.
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncthreads();
    // DK_TPS k-chunks per ring stage (consecutive tiles of the factor and of the panel: still two bulk copies per stage)
    // and STAGES / DK_TPS stages: fewer mbarrier hand-shakes per block row for the same shared memory
#ifndef DK_TPS
#define DK_TPS 2
#endif
    constexpr int TPS = DK_TPS, NS = STAGES / TPS;
    static_assert(NS >= 2 && NS * TPS == STAGES && CHUNKS_PER_BLOCK % TPS == 0, "DK_TPS must divide the ring into >= 2 stages");
    const int nst = nk / TPS;
    auto issue = [&](int idx) {  // thread 0 only
      const unsigned gi = base + idx, slot = gi % NS, par = (gi / NS) & 1;
      mbar_wait(&empty[slot], par ^ 1);  // every warp has released the slot's previous contents
      double* dst = ring + slot * TPS * (A_TILE + B_TILE);
      mbar_arrive_expect_tx(&full[slot], TPS * STAGE_BYTES);
      bulk_load(dst, Atiles + (size_t)idx * TPS * A_TILE, TPS * A_TILE * 8, &full[slot]);
      bulk_load(dst + TPS * A_TILE, panel + (size_t)idx * TPS * B_TILE, TPS * B_TILE * 8, &full[slot]);
    };
    if (tid == 0)
      for (int s = 0; s < NS - 1 && s < nst; s++) issue(s);
#pragma unroll 1
    for (int kc = 0; kc < nst; kc++) {
      if (tid == 0 && kc + NS - 1 < nst) issue(kc + NS - 1);
      const unsigned gi = base + kc, slot = gi % NS, par = (gi / NS) & 1;
      mbar_wait(&full[slot], par);
      const double* s = ring + slot * TPS * (A_TILE + B_TILE);
#pragma unroll
      for (int t = 0; t < TPS; t++) mma_stage_live(acc, s + t * A_TILE, s + TPS * A_TILE + t * B_TILE, wm, wn, lane, live16);
      // this slot is refilled through the async proxy: order this lane's generic-proxy reads of it before that
      // (without this fence the refill raced the LDS reads: about one corrupted tile in a thousand)
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
    }
    *stage_base = base + nst;
    __syncthreads();  // every warp is done with the ring: region 0 can change role
    DK_MARK(2);
  } else {
  // ---- P3: acc += (-L_IJ) V_J over all k-chunks left of the diagonal block -----------------
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nk) {
      double* slot = ring + s * (A_TILE + B_TILE);
      load_tile_async(slot, Atiles + (size_t)s * A_TILE, A_TILE, tid);
      load_tile_async(slot + A_TILE, panel + (size_t)s * B_TILE, B_TILE, tid);
    }
    cp_async_commit();
  }
#pragma unroll 1
  for (int kc = 0; kc < nk; kc++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    {
      const int nx = kc + STAGES - 1;
      if (nx < nk) {
        double* slot = ring + (nx % STAGES) * (A_TILE + B_TILE);
        load_tile_async(slot, Atiles + (size_t)nx * A_TILE, A_TILE, tid);
        load_tile_async(slot + A_TILE, panel + (size_t)nx * B_TILE, B_TILE, tid);
      }
      cp_async_commit();
    }
    const double* slot = ring + (kc % STAGES) * (A_TILE + B_TILE);
    mma_stage_live(acc, slot, slot + A_TILE, wm, wn, lane, live16);
  }
  cp_async_wait<0>();
  __syncthreads();  // every warp is done with the ring: region 0 can change role
  }

  // ---- P4: V_I = inv(L_II) * acc, the right-hand side goes through smem as B fragments ------
  double* Rbuf = reinterpret_cast<double*>(smem);
  double* Abuf = reinterpret_cast<double*>(smem + SM_ABUF_OFF);
  const double* Dtiles = Atiles + (size_t)nk * A_TILE;
  load_tile_async(Abuf, Dtiles, A_TILE, tid);
  cp_async_commit();
#pragma unroll
  for (int mi = 0; mi < 4; mi++) {
    const int chunk = 2 * wm + (mi >> 1);
    const int k = 8 * (mi & 1) + (lane >> 2);
#pragma unroll
    for (int ni = 0; ni < 4; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int nn = 32 * wn + 8 * ni + 2 * (lane & 3) + e;
        Rbuf[chunk * B_TILE + b_tile_index(k, nn)] = acc[mi][ni][e];
        acc[mi][ni][e] = 0.0;
      }
  }
#ifdef DK_X_NODIAG  // what-if build: one diagonal stage instead of eight (wrong results, timing only)
  const int ndiag = 1;
#else
  const int ndiag = CHUNKS_PER_BLOCK;
#endif
#pragma unroll 1
  for (int j = 0; j < ndiag; j++) {
    cp_async_wait<0>();
    __syncthreads();
    if (j + 1 < ndiag) {
      load_tile_async(Abuf + ((j + 1) & 1) * A_TILE, Dtiles + (size_t)(j + 1) * A_TILE, A_TILE, tid);
      cp_async_commit();
    }
    // inverse is lower triangular: pair q needs chunks 0..q (see solved_pair)
    if (j <= 7 - wm && (wm < live16 || 7 - wm < live16))
      mma_stage_diag(acc, Abuf + (j & 1) * A_TILE, Rbuf + j * B_TILE, wm, wn, lane, j <= wm && wm < live16, 7 - wm < live16);
  }

  DK_MARK(3);
  // ---- P5: publish V_I to the panel (fragment-ordered B tiles) -----------------------------
#pragma unroll
  for (int mi = 0; mi < 4; mi++) {
    const int chunk = CHUNKS_PER_BLOCK * I + solved_pair(wm, mi);
    const int k = 8 * (mi & 1) + (lane >> 2);
#pragma unroll
    for (int ni = 0; ni < 4; ni++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int nn = 32 * wn + 8 * ni + 2 * (lane & 3) + e;
        __stcg(panel + (size_t)chunk * B_TILE + b_tile_index(k, nn), acc[mi][ni][e]);  // straight to L2
      }
  }
  // The panel is next read by bulk copies, i.e. through the async proxy and straight from L2.  The block-level
  // barrier that follows only makes these generic-proxy stores visible at CTA scope (L1), so they are first
  // pushed to GPU scope (__threadfence) and then ordered against the async proxy.
  if constexpr (TMA) {
    __threadfence();
    asm volatile("fence.proxy.async.global;" ::: "memory");
  }
  DK_MARK(4);
}

__device__ __forceinline__ double panel_at(const double* panel, int row, int col) {
  return __ldcg(panel + (size_t)(row >> 4) * B_TILE + b_tile_index(row & 15, col));
}

__device__ __forceinline__ double dnorm_std(double x) { return 0.3989422804014326779 * exp(-0.5 * x * x); }

// Per-candidate predict.km epilogue: mean and sd from the solved panel.  The panel rows
// n .. n+p hold -z'v and -M'v (augmented system).  u[j] = (F.newdata - t(Tinv.c) %*% M)_j is returned
// for the gradient kernels.
__device__ __forceinline__ void cand_predict(const DevModel& M, int type, const double* panel, const double* xs,
                                             const double* s2acc, int c, double& mean_out, double& sd_out,
                                             double* u) {
  const int p = M.p, n = M.n;
  double S[DK_MAX_P];
  double mean = 0.0;
  for (int j = 0; j < p; j++) {
    const double fj = (j == 0) ? 1.0 : xs[(j - 1) * BN + c];  // model.matrix: ~1 -> 1 ; ~. -> (1, x)
    mean = fma(fj, __ldg(M.beta + j), mean);
    u[j] = fj + panel_at(panel, n + 1 + j, c);
  }
  mean += -panel_at(panel, n, c);  // + Tinv.c' z
  double s2 = M.total_sd2 - (((s2acc[c] + s2acc[BN + c]) + s2acc[2 * BN + c]) + s2acc[3 * BN + c]);
  if (type == DK_UK) {
    // s2.predict.2 = colSums(backsolve(t(T.M), t(F.newdata - t(Tinv.c) %*% M))^2)
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

// EI (R/EI.R:118-132), kriging.quantile (R/kriging.quantile.R:102-105) or the bare mean.
__device__ __forceinline__ double crit_value(const CritParams& P, double mean, double sd) {
  if (P.mode == MODE_PREDICT) return mean;
  if (P.mode == MODE_QUANTILE) return mean + P.qnorm_beta * sd;
  if (P.mode == MODE_EI_WEIGHT) {
    // sampleFromEI's sampling weight (R/sampleFromEI.R:119-125): sn (u Phi(u) + phi(u)), NaN -> +-(T - mn), clipped at 0
    double uu = (mean - P.plugin) / sd;
    if (P.minimization) uu = -uu;
    double ei = sd * (uu * normcdf(uu) + dnorm_std(uu));
    if (ei != ei) ei = P.minimization ? (P.plugin - mean) : (mean - P.plugin);
    return ei > 0.0 ? ei : 0.0;
  }
  if (P.mode == MODE_AEI) {
    // R/AEI.R:128-143: EI against y.min times the noise penalty 1 - tau / sqrt(tau^2 + sk^2); 0 when sk is negligible
    if (sd < sqrt(P.sd2) / 1e6) return 0.0;
    const double diff = P.plugin - mean, xcr = diff / sd;
    return (diff * normcdf(xcr) + sd * dnorm_std(xcr)) * (1.0 - sqrt(P.tau2) / sqrt(P.tau2 + sd * sd));
  }
  if (P.mode == MODE_EQI) {
    // R/EQI.R:121-148: improvement of the future beta-quantile; mq, sq from the old mean / sd and new.noise.var
    if (sd < sqrt(P.sd2) / 1e6) return 0.0;
    const double mq = mean + P.qnorm_beta * sqrt((P.tau2 * sd * sd) / (P.tau2 + sd * sd));
    const double sq = sd * sd / sqrt(P.tau2 + sd * sd);
    const double xcr = (P.plugin - mq) / sq;
    return sq * (xcr * normcdf(xcr) + dnorm_std(xcr));
  }
  if (P.mode == MODE_FEAS) {
    // Feas (R/crit_EFI.R:231-238): pnorm(q = threshold, mean, sd), minus pnorm(-threshold, ...) for an equality
    // constraint (P.proxy carries the flag).  sd = 0 follows R's pnorm: 1 when q >= mean, else 0.
    const double thr = P.plugin;
    const double hi = (sd > 0.0) ? normcdf((thr - mean) / sd) : ((thr >= mean) ? 1.0 : 0.0);
    if (!P.proxy) return hi;
    const double lo = (sd > 0.0) ? normcdf((-thr - mean) / sd) : ((-thr >= mean) ? 1.0 : 0.0);
    return hi - lo;
  }
  if (P.mode == MODE_COVRATIO) return sd / sqrt(P.sd2);  // checkPredict "covratio" (R/checkPredict.R:43-48)
  const double km = P.minimization ? mean : -mean;
  const double diff = P.plugin - km;
  if (P.proxy) return diff;
  const double xcr = diff / sd;
  return diff * normcdf(xcr) + sd * dnorm_std(xcr);
}

static __device__ __noinline__ void crit_epilogue(const DevModel& M, const CritParams& P, const double* panel,
                                           const double* xs, const double* s2acc, int c, int64_t gidx,
                                           double* out_main, double* out_mean, double* out_sd) {
  double mean, sd, u[DK_MAX_P];
  cand_predict(M, P.type, panel, xs, s2acc, c, mean, sd, u);
  if (out_mean) out_mean[gidx] = mean;
  if (out_sd) out_sd[gidx] = sd;
  if (out_main) {
    double v = crit_value(P, mean, sd);
    if (P.combine == COMBINE_MUL) v *= out_main[gidx];
    else if (P.combine == COMBINE_MIN) v = fmin(v, out_main[gidx]);  // pmin(mindist, .) of R/checkPredict.R:47
    out_main[gidx] = v;
  }
}

// ---------------------------------------------------------------------------------------
// Forward pass of one candidate tile: stage the candidates, then for every block row
// cov-gen -> solve -> s2.  On return (after a __syncthreads) the panel holds V (and the
// augmented rows), s2acc the per-warp-row partial sums of v^2, xs the candidate tile.
// ---------------------------------------------------------------------------------------
// cpanel (optional): the covariance columns c(x_j) themselves are kept, row-major [i][64], for the gradient
// kernels (covVector.dx needs c again: dc = c .* dlogk).
// KSEL >= 0 fixes the covariance kernel (and NUGSEL the nugget path) at compile time, so the instantiation carries
// only the code it runs (the generic version inlines all five kernels: 240 KB of SASS).
// STAGED: the caller has already written the candidate tile xs[k][64] (qEI's general-q kernel lays its batches out itself).
template <bool TMA = false, int KSEL = -1, bool NUGSEL = false, bool STAGED = false>
__device__ __forceinline__ void forward_tile(const DevModel& M, const double* __restrict__ x, int64_t N, int64_t c0,
                                             double* __restrict__ panel, unsigned char* smem, int tid,
                                             unsigned* stage_base = nullptr, double* __restrict__ cpanel = nullptr) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 1, wn = warp & 1;
  double* Cbuf = reinterpret_cast<double*>(smem);
  double* Xs = reinterpret_cast<double*>(smem + SM_XS_OFF);
  double* xs = reinterpret_cast<double*>(smem + SM_CAND_OFF);
  double* s2acc = reinterpret_cast<double*>(smem + SM_S2_OFF);
  const int d = M.d;
  // candidate tile: xs[k][j], coalesced along j (R's column-major N x d is already SoA)
  if constexpr (!STAGED)
    for (int i = tid; i < d * BN; i += NT) {
      const int k = i / BN, j = i % BN;
      const int64_t g = c0 + j;
      xs[i] = (g < N) ? x[g + (int64_t)k * N] : x[(N - 1) + (int64_t)k * N];
    }
  if (tid < 4 * BN) s2acc[tid] = 0.0;

  DK_MARK_INIT;
#pragma unroll 1
  for (int I = 0; I < M.nblk; I++) {
    DK_MARK(6);
    // ---- P1: design rows of this block -> smem, then covariance block -> Cbuf ----
    {
      const double* src = M.Xrm_s + (size_t)I * BM * d;  // design rows pre-scaled by w_k
      for (int i = tid; i < BM * d; i += NT) Xs[i] = __ldg(src + i);
    }
    __syncthreads();
#ifndef DK_X_NOCOV  // what-if build (tools/ab_lib.py): cost of cov-gen = time(normal) - time(-DDK_X_NOCOV)
    if constexpr (KSEL >= 0) covgen_rows<KSEL, NUGSEL>(M, Xs, xs, Cbuf, I * BM, warp, lane);
    else covgen_block(M, Xs, xs, Cbuf, I * BM, warp, lane);
#endif
    __syncthreads();
    // ---- P2: Cbuf -> accumulator fragments ----
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
    DK_MARK(0);  // P1 + P2: cov-gen and fragment load
    // ---- P3..P5 ----
    solve_block_row<TMA>(acc, M.tiles_fwd, I, panel, smem, tid, stage_base, min(BM, M.n_aug - I * BM));
    DK_MARK(7);  // (solve_block_row accounts 2 = GEMM, 3 = diagonal, 4 = panel store itself)
    // ---- s2 += sum over this block's real rows of v^2 (fixed order: rows, then lanes, then warps) ----
    {
      double part[4][2];
#pragma unroll
      for (int ni = 0; ni < 4; ni++) part[ni][0] = part[ni][1] = 0.0;
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const bool live = (I * BM + solved_row(wm, mi, lane)) < M.n;
#pragma unroll
        for (int ni = 0; ni < 4; ni++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const double v = live ? acc[mi][ni][e] : 0.0;
            part[ni][e] = fma(v, v, part[ni][e]);
          }
      }
#pragma unroll
      for (int ni = 0; ni < 4; ni++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          double v = part[ni][e];
          v += __shfl_xor_sync(0xffffffffu, v, 4);
          v += __shfl_xor_sync(0xffffffffu, v, 8);
          v += __shfl_xor_sync(0xffffffffu, v, 16);
          if (lane < 4) s2acc[wm * BN + 32 * wn + 8 * ni + 2 * lane + e] += v;
        }
    }
    __syncthreads();  // panel writes + s2acc visible; region 0 free for the next block row
    DK_MARK(5);  // s2 reduction + barrier
  }
}

}  // namespace dk
