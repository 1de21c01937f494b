// dk_api.cu — host side of the C ABI (include/dk_b200.h): model upload + pre-tiling, the forward / gradient launches
// and the entry points built on them.  The qEI, AKG and constrained families live in dk_api_qei.cu, dk_api_akg.cu
// and dk_api_cst.cu (shared host helpers: dk_host.h).
#include "dk_host.h"
#include "dk_kernels.cuh"
#include "dk_predict_kernels.cuh"
#include "dk_grad.cuh"
#include "dk_narrow.cuh"
#include "dk_small.cuh"
#include "dk_reduce.cuh"

namespace dk {

std::string& last_error() {
  static thread_local std::string e;
  return e;
}
int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  last_error() = buf;
  return code;
}

namespace {

// Lower-triangular augmented factor  L_aug = [[T', 0, 0], [[z M]', I, 0], [0, 0, I]]  (n_pad square).
struct AugFactor {
  const double *T, *z, *M;
  int n, p, n_aug;
  size_t ldT, ldM;   // leading dimensions of T and of M (z is a plain vector)
  double operator()(int i, int j) const {  // j <= i
    if (i < n) return T[(size_t)j + (size_t)i * ldT];
    if (i < n_aug) {
      const int r = i - n;
      if (j < n) return r == 0 ? z[j] : M[(size_t)j + (size_t)(r - 1) * ldM];
    }
    return i == j ? 1.0 : 0.0;
  }
};
// Flipped upper factor for the backward solve  T w = v :  Lb[i'][j'] = T[n-1-i'][n-1-j'] (lower), identity pad.
struct FlipFactor {
  const double* T;
  int n;
  size_t ldT;
  double operator()(int i, int j) const {  // j <= i
    if (i < n) return T[(size_t)(n - 1 - i) + (size_t)(n - 1 - j) * ldT];
    return i == j ? 1.0 : 0.0;
  }
};

// Build the fragment-ordered tile array of a lower-triangular factor: for each block row I the
// 8 I tiles of -L_IJ (J < I, k-chunks of 16 columns) followed by the 8 tiles of inv(L_II).
// The 128 x 128 diagonal inverses are formed in long double and rounded once.
// Block rows [I0, nblk) only when I0 > 0 (dk_model_append: a new design row changes the last block rows only); `tiles`
// then starts at block row I0.
template <class Factor>
void build_tiles(const Factor& L, int nblk, std::vector<double>& tiles, int I0 = 0) {
  const size_t first = tile_base(I0) * A_TILE;
  tiles.assign(tile_base(nblk) * A_TILE - first, 0.0);
  std::vector<long double> inv((size_t)BM * BM);
#pragma omp parallel for schedule(dynamic) firstprivate(inv)
  for (int I = I0; I < nblk; I++) {
    double* base = tiles.data() + (tile_base(I) * A_TILE - first);
    const int r0 = I * BM;
    for (int kc = 0; kc < CHUNKS_PER_BLOCK * I; kc++) {
      double* t = base + (size_t)kc * A_TILE;
      for (int r = 0; r < BM; r++)
        for (int k = 0; k < KC; k++) t[a_tile_index(r, k)] = -L(r0 + r, kc * KC + k);
    }
    // inverse of the diagonal block by forward substitution, column by column
    std::fill(inv.begin(), inv.end(), 0.0L);
    for (int j = 0; j < BM; j++) {
      inv[(size_t)j * BM + j] = 1.0L / (long double)L(r0 + j, r0 + j);
      for (int i = j + 1; i < BM; i++) {
        long double s = 0.0L;
        for (int k = j; k < i; k++) s += (long double)L(r0 + i, r0 + k) * inv[(size_t)k * BM + j];
        inv[(size_t)i * BM + j] = -s / (long double)L(r0 + i, r0 + i);
      }
    }
    double* dt = base + (size_t)CHUNKS_PER_BLOCK * I * A_TILE;
    for (int kc = 0; kc < CHUNKS_PER_BLOCK; kc++)
      for (int r = 0; r < BM; r++)
        for (int k = 0; k < KC; k++)
          dt[(size_t)kc * A_TILE + a_tile_index(r, k)] = (double)inv[(size_t)r * BM + kc * KC + k];
  }
}

// Small designs (n_pad <= 256): the explicit inverse of the WHOLE padded factor, formed in long double and rounded once,
// as A tiles [row block rb][k-chunk kc <= 8 rb + 7] for k_predict_small (dk_small.cuh).
template <class Factor>
void invert_factor(const Factor& L, int N, std::vector<long double>& inv) {
  inv.assign((size_t)N * N, 0.0L);
  for (int j = 0; j < N; j++) {
    inv[(size_t)j * N + j] = 1.0L / (long double)L(j, j);
    for (int i = j + 1; i < N; i++) {
      long double s = 0.0L;
      for (int k = j; k < i; k++) s += (long double)L(i, k) * inv[(size_t)k * N + j];
      inv[(size_t)i * N + j] = -s / (long double)L(i, i);
    }
  }
}

void host_cov_vector(const dk_model* m, const double* x, double* c);

// Applying the rounded whole-factor inverse is less accurate than substitution when the factor is ill-conditioned (the
// off-diagonal block of the inverse cancels against the diagonal one).  Probe: 64 points of the design's bounding box,
// the mean term z'v and sum v^2 from the rounded inverse (double, ascending k, as the kernel accumulates) against
// long-double substitution; the kernel is enabled only if both stay within 1e-11 of their scales (a tenth of the
// tolerance: the largest error over a screen of thousands of candidates is a few times the largest over 64 probes).
template <class Factor>
bool small_inverse_is_accurate(const dk_model* m, const Factor& L, const std::vector<long double>& inv, int N) {
  const int n = m->n, d = m->d;
  std::vector<double> lo(d, 1e300), hi(d, -1e300), x(d), c(n), vs(N);
  std::vector<long double> vr(N);
  for (int i = 0; i < n; i++)
    for (int k = 0; k < d; k++) {
      lo[k] = std::min(lo[k], m->h_X[(size_t)i * d + k]);
      hi[k] = std::max(hi[k], m->h_X[(size_t)i * d + k]);
    }
  double ymax = 0.0;
  if (m->h_y)
    for (int i = 0; i < n; i++) ymax = std::max(ymax, std::fabs(m->h_y[i]));
  const double scale = ymax + std::sqrt(m->dm.sd2);
  unsigned long long rng = 0x9E3779B97F4A7C15ULL;
  for (int t = 0; t < 64; t++) {
    for (int k = 0; k < d; k++) {
      rng = rng * 6364136223846793005ULL + 1442695040888963407ULL;
      x[k] = lo[k] + (hi[k] - lo[k]) * ((double)(rng >> 11) * 1.1102230246251565e-16);
    }
    host_cov_vector(m, x.data(), c.data());
    long double s2r = 0.0L;
    for (int i = 0; i < N; i++) {          // long-double substitution, right-hand side [c; 0]
      long double s = i < n ? (long double)c[i] : 0.0L;
      for (int k = 0; k < i; k++) s -= (long double)L(i, k) * vr[k];
      vr[i] = s / (long double)L(i, i);
      if (i < n) s2r += vr[i] * vr[i];
    }
    double s2s = 0.0;
    for (int i = 0; i <= n; i++) {         // rows 0..n: the design rows and the z'v row
      double a = 0.0;
      for (int k = 0; k <= std::min(i, n - 1); k++) a = std::fma((double)inv[(size_t)i * N + k], c[k], a);
      vs[i] = a;
      if (i < n) s2s = std::fma(a, a, s2s);
    }
    if (std::fabs((double)((long double)vs[n] - vr[n])) > 1e-11 * scale) return false;
    if (std::fabs((double)((long double)s2s - s2r)) > 1e-11 * m->dm.total_sd2) return false;
  }
  return true;
}

}  // namespace

// (re)build the small-design tiles of a handle from its host copies; no-op for n_pad > 256 and for factors on which the
// whole-factor inverse would lose accuracy (the blocked kernels then serve the model as before)
int refresh_small_tiles(dk_model* m) {
  auto drop = [&] {
    dev_release(m, &m->d_tiles_small, m->tiles_small_doubles);
    m->tiles_small_doubles = 0;
    m->dm.tiles_small = nullptr;
    return DK_OK;
  };
  const SmallLayout lay = small_layout(m->n, m->dm.n_pad, m->d, m->p);
  if (lay.stages == 0) return drop();   // more than 256 padded rows, or rows x d too large to keep the design in shared memory
  const int N = m->dm.n_pad, nrb = N / BM;
  const size_t ld = (size_t)m->ld;
  AugFactor L{m->h_T, m->h_zM, m->h_zM + ld, m->n, m->p, m->dm.n_aug, ld, ld};
  std::vector<long double> inv;
  std::vector<double> tiles;
  try {
    invert_factor(L, N, inv);
    if (!small_inverse_is_accurate(m, L, inv, N)) return drop();
    tiles.assign(small_tile_base(nrb) * A_TILE, 0.0);
    for (int rb = 0; rb < nrb; rb++)
      for (int kc = 0; kc < CHUNKS_PER_BLOCK * (rb + 1); kc++) {
        double* t = tiles.data() + (small_tile_base(rb) + kc) * A_TILE;
        for (int r = 0; r < BM; r++)
          for (int k = 0; k < KC; k++) t[a_tile_index(r, k)] = (double)inv[(size_t)(rb * BM + r) * N + kc * KC + k];
      }
  } catch (const std::bad_alloc&) {
    return fail(DK_ERR_NOMEM, "host allocation failed while inverting the factor");
  }
  if (m->tiles_small_doubles < tiles.size()) {
    drop();
    int rc = dev_alloc(m, &m->d_tiles_small, tiles.size());
    if (rc) return rc;
    m->tiles_small_doubles = tiles.size();
  }
  DK_CUDA_TRY(cudaMemcpy(m->d_tiles_small, tiles.data(), tiles.size() * 8, cudaMemcpyHostToDevice));
  m->dm.tiles_small = m->d_tiles_small;
  m->small_lay = lay;
  return DK_OK;
}

int ensure_io(dk_model* m, size_t doubles) {
  if (m->io_doubles >= doubles) return DK_OK;
  if (m->d_io) {
    cudaFree(m->d_io);
    m->device_bytes -= m->io_doubles * 8;
    m->d_io = nullptr;
    m->io_doubles = 0;
  }
  int rc = dev_alloc(m, &m->d_io, doubles);
  if (rc) return rc;
  m->io_doubles = doubles;
  return DK_OK;
}

int ensure_red(dk_model* m, size_t bytes) {
  if (m->red_bytes >= bytes) return DK_OK;
  if (m->d_red) {
    cudaFree(m->d_red);
    m->device_bytes -= m->red_bytes;
    m->d_red = nullptr;
  }
  unsigned char* p = nullptr;
  int rc = dev_alloc(m, &p, bytes);
  if (rc) return rc;
  m->d_red = p;
  m->red_bytes = bytes;
  return DK_OK;
}

int ensure_bwd(dk_model* m) {
  if (m->d_tiles_bwd && m->bwd_valid) return DK_OK;
  std::vector<double> tiles;
  try {
    FlipFactor L{m->h_T, m->n, (size_t)m->ld};
    build_tiles(L, m->dm.nblk_b, tiles);
  } catch (const std::bad_alloc&) {
    return fail(DK_ERR_NOMEM, "host allocation for the backward tiles failed");
  }
  if (!m->d_tiles_bwd) {
    // allocate into locals and publish every pointer only once all steps succeeded: a failed call leaves the handle
    // exactly as it was (a later gradient call retries instead of launching with half of the buffers).  The tile buffer
    // is sized for the handle's row capacity, so dk_model_append only has to refill it.
    const int cap_blk = (int)((m->scratch_doubles_per_cta / BN) / BM);
    double *t = nullptr, *w = nullptr, *c = nullptr;
    const size_t tcount = tile_base(cap_blk) * A_TILE;
    const size_t wcount = (size_t)m->grid * (m->scratch_doubles_per_cta + COEF_DOUBLES);
    const size_t ccount = (size_t)m->grid * m->scratch_doubles_per_cta;
    int rc = dev_alloc(m, &t, tcount);
    if (!rc) rc = dev_alloc(m, &w, wcount);
    if (!rc) rc = dev_alloc(m, &c, ccount);  // covariance panels (gradients)
    if (rc) {
      dev_release(m, &t, tcount);
      dev_release(m, &w, wcount);
      dev_release(m, &c, ccount);
      return rc;
    }
    m->d_wscratch = w;
    m->d_cscratch = c;
    m->d_tiles_bwd = t;
    m->bwd_tile_doubles = tcount;
  }
  cudaError_t e = cudaMemcpy(m->d_tiles_bwd, tiles.data(), tiles.size() * 8, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return fail(DK_ERR_CUDA, "upload of the backward tiles failed: %s", cudaGetErrorString(e));
  m->dm.tiles_bwd = m->d_tiles_bwd;
  m->bwd_valid = true;
  return DK_OK;
}

namespace {
bool g_attr_done[64] = {};
int set_kernel_attrs(int device) {
  if (device < 64 && g_attr_done[device]) return DK_OK;
  DK_CUDA_TRY(cudaFuncSetAttribute(k_predict_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL));
#define DK_ATTR_TMA(K)                                                                                                  \
  DK_CUDA_TRY(cudaFuncSetAttribute(k_predict_fused_tma<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM)); \
  DK_CUDA_TRY(cudaFuncSetAttribute(k_predict_fused_tma<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
  DK_ATTR_TMA(DK_GAUSS) DK_ATTR_TMA(DK_EXP) DK_ATTR_TMA(DK_MATERN3_2) DK_ATTR_TMA(DK_MATERN5_2) DK_ATTR_TMA(DK_POWEXP)
#undef DK_ATTR_TMA
  DK_CUDA_TRY(cudaFuncSetAttribute(k_predict_small, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_MAX));
  DK_CUDA_TRY(cudaFuncSetAttribute(k_predict_narrow, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
  DK_CUDA_TRY(cudaFuncSetAttribute(k_grad_narrow, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
  DK_CUDA_TRY(cudaFuncSetAttribute(k_grad_fused<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
  DK_CUDA_TRY(cudaFuncSetAttribute(k_grad_fused<DK_MATERN5_2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL + PIPE_SMEM));
  if (device < 64) g_attr_done[device] = true;
  return DK_OK;
}

}  // namespace

int check_call(const dk_model* m, const void* x, int64_t N) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  if (!x) return fail(DK_ERR_ARG, "x is NULL");
  if (N < 0) return fail(DK_ERR_ARG, "N = %lld is negative", (long long)N);
  return DK_OK;
}

// Narrow tiles (16 candidates per CTA, dk_narrow.cuh) when the batch would leave at least three quarters of the
// persistent grid without a 64-wide tile: optimiser-sized calls.  DK_NARROW=0 / 1 forces the choice (A/B runs).
static bool use_narrow(const dk_model* m, int64_t N) {
  static const int forced = [] { const char* e = std::getenv("DK_NARROW"); return e ? std::atoi(e) : -1; }();
  if (m->tile_mode == DK_TILE_WIDE) return false;
  if (m->tile_mode == DK_TILE_NARROW) return true;
  if (forced >= 0) return forced != 0;
  return ((N + BN - 1) / BN) * 4 <= m->grid;
}

// Small designs (n_pad <= 256) take the single-block kernel of dk_small.cuh for every batch the narrow tiles do not serve.
// DK_SMALL=0 keeps them on the blocked kernel (A/B runs); DK_TILE_WIDE / DK_TILE_NARROW also bypass it.
static bool use_small(const dk_model* m) {
  static const bool off = [] { const char* e = std::getenv("DK_SMALL"); return e && std::atoi(e) == 0; }();
  return !off && m->dm.tiles_small != nullptr && m->tile_mode == DK_TILE_AUTO;
}

// forward pass over device-resident candidates
int launch_predict(dk_model* m, const CritParams& P, const double* d_x, int64_t N, double* d_main, double* d_mean,
                   double* d_sd, cudaStream_t st) {
  if (N == 0) return DK_OK;
  if (use_narrow(m, N)) {
    const int64_t nt = (N + BNN - 1) / BNN;
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
    k_predict_narrow<<<(int)std::min<int64_t>(nt, m->grid), NT, SM_TOTAL + PIPE_SMEM, st>>>(
        m->dm, P, d_x, N, nt, m->d_vscratch, m->scratch_doubles_per_cta, d_main, d_mean, d_sd);
    if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
    m->launches++;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  if (use_small(m)) {
    const int64_t nt = (N + SBN - 1) / SBN;
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
    k_predict_small<<<(int)std::min<int64_t>(nt, m->num_sms), NTS, m->small_lay.total, st>>>(m->dm, P, m->small_lay, d_x, N, nt, d_main, d_mean, d_sd);
    if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
    m->launches++;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  const int64_t ntiles = (N + BN - 1) / BN;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  // default: bulk-copy + mbarrier pipeline (+8.6 % at cfg4); DK_PIPELINE=cpasync selects the cp.async / __syncthreads ring
  static const bool use_tma = [] { const char* e = std::getenv("DK_PIPELINE"); return !(e && std::strcmp(e, "cpasync") == 0); }();
  if (use_tma) {
    // one instantiation per covariance kernel and nugget path: each carries only the code it runs
#define DK_LAUNCH_TMA(K)                                                                                               \
  case K:                                                                                                              \
    if (m->dm.nugget_flag)                                                                                             \
      k_predict_fused_tma<K, true><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, P, d_x, N, ntiles, m->d_vscratch,     \
                                                                         m->scratch_doubles_per_cta, d_main, d_mean, d_sd); \
    else                                                                                                               \
      k_predict_fused_tma<K, false><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, P, d_x, N, ntiles, m->d_vscratch,    \
                                                                          m->scratch_doubles_per_cta, d_main, d_mean, d_sd); \
    break;
    switch (m->dm.kernel) {
      DK_LAUNCH_TMA(DK_GAUSS) DK_LAUNCH_TMA(DK_EXP) DK_LAUNCH_TMA(DK_MATERN3_2) DK_LAUNCH_TMA(DK_MATERN5_2)
      default: DK_LAUNCH_TMA(DK_POWEXP)
    }
#undef DK_LAUNCH_TMA
  } else
    k_predict_fused<<<grid, NT, SM_TOTAL, st>>>(m->dm, P, d_x, N, ntiles, m->d_vscratch, m->scratch_doubles_per_cta,
                                                 d_main, d_mean, d_sd);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

int launch_grad(dk_model* m, const CritParams& P, const double* d_x, int64_t N, double* d_grad, double* d_main,
                cudaStream_t st) {
  if (N == 0) return DK_OK;
  int rc = ensure_bwd(m);
  if (rc) return rc;
  if (use_narrow(m, N)) {
    const int64_t nt = (N + BNN - 1) / BNN;
    if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
    k_grad_narrow<<<(int)std::min<int64_t>(nt, m->grid), NT, SM_TOTAL + PIPE_SMEM, st>>>(
        m->dm, P, d_x, N, nt, m->d_vscratch, m->d_wscratch, m->d_cscratch, m->scratch_doubles_per_cta, d_grad, d_main);
    if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
    m->launches++;
    DK_CUDA_TRY(cudaGetLastError());
    return DK_OK;
  }
  const int64_t ntiles = (N + BN - 1) / BN;
  const int grid = (int)std::min<int64_t>(ntiles, m->grid);
  if (m->timing) DK_CUDA_TRY(cudaEventRecord(m->ev0, st));
  if (m->dm.kernel == DK_MATERN5_2 && !m->dm.nugget_flag)
    k_grad_fused<DK_MATERN5_2><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, P, d_x, N, ntiles, m->d_vscratch, m->d_wscratch,
                                                                     m->d_cscratch, m->scratch_doubles_per_cta, d_grad, d_main);
  else
    k_grad_fused<-1><<<grid, NT, SM_TOTAL + PIPE_SMEM, st>>>(m->dm, P, d_x, N, ntiles, m->d_vscratch, m->d_wscratch,
                                                           m->d_cscratch, m->scratch_doubles_per_cta, d_grad, d_main);
  if (m->timing) { DK_CUDA_TRY(cudaEventRecord(m->ev1, st)); m->ev_valid = true; m->ev_mid_valid = false; }
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

// ------------------------------------------------------------------------------------------
// chunked host <-> device pipeline (dk_host.h)
// ------------------------------------------------------------------------------------------
int64_t chunk_units_for(const dk_model* m, int unit, int tile_units) {
  const int64_t wave = (int64_t)tile_units * m->grid;
  const double bytes_per_wave = (double)wave * unit * m->d * 8.0;
  int64_t k = (int64_t)std::llround(24.0e6 / bytes_per_wave);
  k = std::max<int64_t>(1, std::min<int64_t>(k, 256));
  return wave * k;
}

namespace {
int ensure_pipeline(dk_model* m, size_t in_doubles, size_t out_doubles) {
  if (!m->copy_in) {
    DK_CUDA_TRY(cudaStreamCreateWithFlags(&m->copy_in, cudaStreamNonBlocking));
    DK_CUDA_TRY(cudaStreamCreateWithFlags(&m->copy_out, cudaStreamNonBlocking));
    for (int s = 0; s < 2; s++) {
      DK_CUDA_TRY(cudaEventCreateWithFlags(&m->ev_h2d[s], cudaEventDisableTiming));
      DK_CUDA_TRY(cudaEventCreateWithFlags(&m->ev_kern[s], cudaEventDisableTiming));
      DK_CUDA_TRY(cudaEventCreateWithFlags(&m->ev_d2h[s], cudaEventDisableTiming));
    }
  }
  in_doubles = (in_doubles + 31) & ~(size_t)31;
  out_doubles = (out_doubles + 31) & ~(size_t)31;
  if (m->stage_in_doubles >= in_doubles && m->stage_out_doubles >= out_doubles) return DK_OK;
  in_doubles = std::max(in_doubles, m->stage_in_doubles);
  out_doubles = std::max(out_doubles, m->stage_out_doubles);
  if (m->h_stage) { cudaFreeHost(m->h_stage); m->h_stage = nullptr; }
  dev_release(m, &m->d_chunk, 2 * (m->stage_in_doubles + m->stage_out_doubles));
  m->stage_in_doubles = m->stage_out_doubles = 0;
  const size_t total = 2 * (in_doubles + out_doubles);
  cudaError_t e = cudaHostAlloc((void**)&m->h_stage, total * 8, cudaHostAllocDefault);
  if (e != cudaSuccess) {
    cudaGetLastError();
    m->h_stage = nullptr;
    return fail(DK_ERR_NOMEM, "cudaHostAlloc of %zu bytes of pinned staging failed: %s", total * 8, cudaGetErrorString(e));
  }
  int rc = dev_alloc(m, &m->d_chunk, total);
  if (rc) { cudaFreeHost(m->h_stage); m->h_stage = nullptr; return rc; }
  m->stage_in_doubles = in_doubles;
  m->stage_out_doubles = out_doubles;
  return DK_OK;
}
}  // namespace

int run_chunked(dk_model* m, const double* x, int64_t ldx, int64_t row0, int64_t nunits, int unit, int64_t chunk_units,
                const HostOut* outs, int nouts, const ChunkLaunch& launch) {
  if (nunits <= 0) return DK_OK;
  if (nouts > 8) return fail(DK_ERR_ARG, "run_chunked: too many outputs");
  const int d = m->d;
  chunk_units = std::max<int64_t>(1, std::min(chunk_units, nunits));
  const int64_t crows = chunk_units * unit;
  size_t out_per_chunk = 0;
  for (int j = 0; j < nouts; j++)
    if (outs[j].host) out_per_chunk += (size_t)outs[j].cols * (size_t)(outs[j].per_unit ? chunk_units : crows);
  int rc = ensure_pipeline(m, (size_t)crows * d, out_per_chunk);
  if (rc) return rc;
  const size_t in_sz = m->stage_in_doubles, out_sz = m->stage_out_doubles;
  double* h_in[2] = {m->h_stage, m->h_stage + in_sz};
  double* h_out[2] = {m->h_stage + 2 * in_sz, m->h_stage + 2 * in_sz + out_sz};
  double* d_in[2] = {m->d_chunk, m->d_chunk + in_sz};
  double* d_out[2] = {m->d_chunk + 2 * in_sz, m->d_chunk + 2 * in_sz + out_sz};
  const int64_t unit_base = row0 / unit;
  const int64_t nchunks = (nunits + chunk_units - 1) / chunk_units;

  // copy a finished chunk's outputs from pinned staging into the caller's arrays
  auto scatter = [&](int64_t ci) -> int {
    const int s = (int)(ci & 1);
    DK_CUDA_TRY(cudaEventSynchronize(m->ev_d2h[s]));
    const int64_t u0 = ci * chunk_units, nu = std::min(chunk_units, nunits - u0);
    size_t off = 0;
    for (int j = 0; j < nouts; j++) {
      if (!outs[j].host) continue;
      const int64_t cnt = outs[j].per_unit ? nu : nu * unit;
      const int64_t first = outs[j].per_unit ? unit_base + u0 : row0 + u0 * unit;
      for (int c = 0; c < outs[j].cols; c++)
        std::memcpy(outs[j].host + (size_t)c * outs[j].ld + first, h_out[s] + off + (size_t)c * cnt, (size_t)cnt * 8);
      off += (size_t)outs[j].cols * cnt;
    }
    return DK_OK;
  };

  for (int64_t ci = 0; ci < nchunks; ci++) {
    const int s = (int)(ci & 1);
    if (ci >= 2 && (rc = scatter(ci - 2))) return rc;   // also: slot s of every buffer is free again
    const int64_t u0 = ci * chunk_units, nu = std::min(chunk_units, nunits - u0), rows = nu * unit;
    const double* src = x + row0 + u0 * unit;
    for (int k = 0; k < d; k++) std::memcpy(h_in[s] + (size_t)k * rows, src + (size_t)k * ldx, (size_t)rows * 8);
    DK_CUDA_TRY(cudaMemcpyAsync(d_in[s], h_in[s], (size_t)rows * d * 8, cudaMemcpyHostToDevice, m->copy_in));
    DK_CUDA_TRY(cudaEventRecord(m->ev_h2d[s], m->copy_in));
    DK_CUDA_TRY(cudaStreamWaitEvent(m->stream, m->ev_h2d[s], 0));
    double* dptr[8];
    size_t off = 0;
    for (int j = 0; j < nouts; j++) {
      const int64_t cnt = outs[j].per_unit ? nu : rows;
      dptr[j] = outs[j].host ? d_out[s] + off : nullptr;
      if (outs[j].host) off += (size_t)outs[j].cols * cnt;
    }
    if ((rc = launch(d_in[s], rows, dptr, u0, m->stream))) return rc;
    DK_CUDA_TRY(cudaEventRecord(m->ev_kern[s], m->stream));
    DK_CUDA_TRY(cudaStreamWaitEvent(m->copy_out, m->ev_kern[s], 0));
    if (off) DK_CUDA_TRY(cudaMemcpyAsync(h_out[s], d_out[s], off * 8, cudaMemcpyDeviceToHost, m->copy_out));
    DK_CUDA_TRY(cudaEventRecord(m->ev_d2h[s], m->copy_out));
  }
  for (int64_t ci = std::max<int64_t>(0, nchunks - 2); ci < nchunks; ci++)
    if ((rc = scatter(ci))) return rc;
  return DK_OK;
}

}  // namespace dk

using namespace dk;

extern "C" {

const char* dk_last_error(void) { return dk::last_error().c_str(); }
const char* dk_version(void) { return "dk_b200 0.1 sm_100a (FP64 DMMA m8n8k4 blocked TRSM)"; }

int dk_model_create(const dk_model_desc* D, dk_model** out) {
  if (!D || !out) return fail(DK_ERR_ARG, "dk_model_create: NULL argument");
  *out = nullptr;
  if (D->n < 1 || D->d < 1 || D->p < 1) return fail(DK_ERR_ARG, "bad dimensions n=%d d=%d p=%d", D->n, D->d, D->p);
  if (D->d > DK_MAX_D) return fail(DK_ERR_ARG, "d=%d exceeds DK_MAX_D=%d", D->d, DK_MAX_D);
  if (D->p > DK_MAX_P) return fail(DK_ERR_ARG, "p=%d exceeds DK_MAX_P=%d", D->p, DK_MAX_P);
  if (D->kernel < DK_GAUSS || D->kernel > DK_POWEXP)
    return fail(DK_ERR_ARG, "unsupported covariance kernel id %d (covTensorProduct gauss/exp/matern3_2/matern5_2/powexp only)", D->kernel);
  if (D->trend != DK_TREND_CONST && D->trend != DK_TREND_LINEAR)
    return fail(DK_ERR_ARG, "unsupported trend id %d (~1 and ~. only)", D->trend);
  if (D->trend == DK_TREND_CONST && D->p != 1) return fail(DK_ERR_ARG, "trend ~1 needs p=1, got %d", D->p);
  if (D->trend == DK_TREND_LINEAR && D->p != D->d + 1) return fail(DK_ERR_ARG, "trend ~. needs p=d+1, got p=%d d=%d", D->p, D->d);
  if (!D->theta || !D->T || !D->z || !D->M || !D->X || !D->beta) return fail(DK_ERR_ARG, "a required model array is NULL");
  if (D->kernel == DK_POWEXP && !D->shape) return fail(DK_ERR_ARG, "powexp needs shape.val");
  if (!(D->sd2 > 0)) return fail(DK_ERR_ARG, "sd2 must be positive");
  for (int k = 0; k < D->d; k++)
    if (!(D->theta[k] > 0)) return fail(DK_ERR_ARG, "range.val[%d] must be positive", k);
  for (int i = 0; i < D->n; i++)
    if (!(D->T[(size_t)i + (size_t)i * D->n] > 0)) return fail(DK_ERR_ARG, "T[%d,%d] is not positive: not a Cholesky factor", i, i);

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(DK_ERR_UNSUPPORTED, "no CUDA device visible");
  if (D->device < 0 || D->device >= ndev) return fail(DK_ERR_ARG, "device %d out of range (%d visible)", D->device, ndev);
  cudaDeviceProp prop;
  DK_CUDA_TRY(cudaGetDeviceProperties(&prop, D->device));
  if (prop.major != 10) return fail(DK_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", D->device, prop.major, prop.minor);
  DeviceGuard guard(D->device);
  int rc = set_kernel_attrs(D->device);
  if (rc) return rc;

  dk_model* m = new (std::nothrow) dk_model();
  if (!m) return fail(DK_ERR_NOMEM, "host allocation failed");
  m->device = D->device;
  m->num_sms = prop.multiProcessorCount;
  const int n = D->n, d = D->d, p = D->p;
  m->n = n; m->d = d; m->p = p;
  DevModel& dm = m->dm;
  dm.n = n; dm.d = d; dm.p = p;
  dm.n_aug = n + p + 1;
  dm.nblk = (dm.n_aug + BM - 1) / BM;
  dm.n_pad = dm.nblk * BM;
  dm.nblk_b = (n + BM - 1) / BM;
  dm.kernel = D->kernel; dm.trend = D->trend; dm.nugget_flag = D->nugget_flag ? 1 : 0;
  dm.sd2 = D->sd2; dm.nugget = D->nugget;
  dm.total_sd2 = D->sd2 + (D->nugget_flag ? D->nugget : 0.0);
  for (int k = 0; k < DK_MAX_D; k++) { dm.w[k] = 0; dm.iw[k] = 0; dm.shape[k] = 1; dm.theta[k] = 1; }
  for (int k = 0; k < d; k++) {
    const double th = D->theta[k];
    dm.theta[k] = th;
    dm.w[k] = D->kernel == DK_MATERN3_2 ? std::sqrt(3.0) / th : D->kernel == DK_MATERN5_2 ? std::sqrt(5.0) / th : 1.0 / th;
    dm.iw[k] = 1.0 / dm.w[k];
    if (D->shape) dm.shape[k] = D->shape[k];
  }

  auto bail = [&](int code) { dk_model_destroy(m); return code; };
  try {
    // ---- tiles of the augmented factor ----
    std::vector<double> tiles;
    AugFactor L{D->T, D->z, D->M, n, p, dm.n_aug, (size_t)n, (size_t)n};
    build_tiles(L, dm.nblk, tiles);
    if ((rc = dev_alloc(m, &m->d_tiles_fwd, tiles.size()))) return bail(rc);
    m->tiles_fwd_doubles = tiles.size();
    auto upload = [&](double* dst, const double* src, size_t count, const char* what) {
      const cudaError_t e = cudaMemcpy(dst, src, count * 8, cudaMemcpyHostToDevice);
      return e == cudaSuccess ? DK_OK : fail(DK_ERR_CUDA, "upload of %s failed: %s", what, cudaGetErrorString(e));
    };
    if ((rc = upload(m->d_tiles_fwd, tiles.data(), tiles.size(), "the factor tiles"))) return bail(rc);
    dm.tiles_fwd = m->d_tiles_fwd;
    tiles.clear(); tiles.shrink_to_fit();

    // ---- design, row-major, zero padded ----
    std::vector<double> Xrm((size_t)dm.n_pad * d, 0.0);
    for (int i = 0; i < n; i++)
      for (int k = 0; k < d; k++) Xrm[(size_t)i * d + k] = D->X[(size_t)i + (size_t)k * n];
    if ((rc = dev_alloc(m, &m->d_Xrm, 3 * Xrm.size()))) return bail(rc);
    m->xrm_doubles = 3 * Xrm.size();
    if ((rc = upload(m->d_Xrm, Xrm.data(), Xrm.size(), "the design"))) return bail(rc);
    dm.Xrm = m->d_Xrm;
    {
      std::vector<double> Xs(Xrm.size());
      for (size_t i = 0; i < Xrm.size(); i++) Xs[i] = Xrm[i] * dm.w[i % d];
      if ((rc = upload(m->d_Xrm + Xrm.size(), Xs.data(), Xs.size(), "the scaled design"))) return bail(rc);
      dm.Xrm_s = m->d_Xrm + Xrm.size();
    }
    {
      // round(X, 8) once per model: the qEI kernels compare it with the rounded batch points (R/qEI.R:124-134)
      std::vector<double> Xr(Xrm.size());
      for (size_t i = 0; i < Xrm.size(); i++) Xr[i] = std::nearbyint(Xrm[i] * 1e8) / 1e8;
      if ((rc = upload(m->d_Xrm + 2 * Xrm.size(), Xr.data(), Xr.size(), "the rounded design"))) return bail(rc);
      dm.Xrm_r8 = m->d_Xrm + 2 * Xrm.size();
    }

    // ---- small candidate-independent arrays (SURVEY 8(a1) "derived once per model") ----
    // G = M'M ; RM = chol(G) upper (predict.km: T.M <- chol(t(M)%*%M)) ; tuuinv = solve(G) (R/EI.grad.R:169)
    std::vector<double> G((size_t)p * p, 0.0), RM((size_t)p * p, 0.0), Ginv((size_t)p * p, 0.0);
    for (int a = 0; a < p; a++)
      for (int b = 0; b < p; b++) {
        long double s = 0;
        for (int i = 0; i < n; i++) s += (long double)D->M[(size_t)i + (size_t)a * n] * D->M[(size_t)i + (size_t)b * n];
        G[a + (size_t)b * p] = (double)s;
      }
    for (int j = 0; j < p; j++) {  // upper Cholesky, column by column: G = RM' RM
      for (int i = 0; i <= j; i++) {
        double s = G[i + (size_t)j * p];
        for (int k = 0; k < i; k++) s -= RM[k + (size_t)i * p] * RM[k + (size_t)j * p];
        if (i < j) RM[i + (size_t)j * p] = s / RM[i + (size_t)i * p];
        else {
          if (!(s > 0)) return bail(fail(DK_ERR_ARG, "M'M is not positive definite (trend matrix rank deficient)"));
          RM[j + (size_t)j * p] = std::sqrt(s);
        }
      }
    }
    {  // Ginv = RM^-1 RM^-T  via two triangular solves per unit vector
      std::vector<double> col(p);
      for (int e = 0; e < p; e++) {
        for (int i = 0; i < p; i++) {  // RM' y = e_e (forward)
          double s = (i == e) ? 1.0 : 0.0;
          for (int k = 0; k < i; k++) s -= RM[k + (size_t)i * p] * col[k];
          col[i] = s / RM[i + (size_t)i * p];
        }
        for (int i = p - 1; i >= 0; i--) {  // RM x = y (backward)
          double s = col[i];
          for (int k = i + 1; k < p; k++) s -= RM[i + (size_t)k * p] * col[k];
          col[i] = s / RM[i + (size_t)i * p];
        }
        for (int i = 0; i < p; i++) Ginv[i + (size_t)e * p] = col[i];
      }
    }
    // alpha = T^-1 z, KinvF = T^-1 M (back substitution with the upper factor)
    std::vector<double> aK((size_t)n * (p + 1));
    for (int c = 0; c <= p; c++) {
      const double* rhs = c == 0 ? D->z : D->M + (size_t)(c - 1) * n;
      double* o = aK.data() + (size_t)c * n;
      for (int i = n - 1; i >= 0; i--) {
        long double s = rhs[i];
        for (int k = i + 1; k < n; k++) s -= (long double)D->T[(size_t)i + (size_t)k * n] * o[k];
        o[i] = (double)(s / D->T[(size_t)i + (size_t)i * n]);
      }
    }
    const size_t small_count = (size_t)p + 2 * (size_t)p * p + 2 * (size_t)n * (p + 1) + (D->noise_var ? n : 0);
    std::vector<double> small(small_count);
    size_t off = 0;
    const size_t o_beta = off; std::memcpy(&small[off], D->beta, p * 8); off += p;
    const size_t o_RM = off; std::memcpy(&small[off], RM.data(), (size_t)p * p * 8); off += (size_t)p * p;
    const size_t o_tu = off; std::memcpy(&small[off], Ginv.data(), (size_t)p * p * 8); off += (size_t)p * p;
    const size_t o_al = off; std::memcpy(&small[off], aK.data(), (size_t)n * (p + 1) * 8); off += (size_t)n * (p + 1);
    const size_t o_zM = off;
    std::memcpy(&small[off], D->z, (size_t)n * 8);
    std::memcpy(&small[off + n], D->M, (size_t)n * p * 8);
    off += (size_t)n * (p + 1);
    const size_t o_nv = off;
    if (D->noise_var) std::memcpy(&small[off], D->noise_var, (size_t)n * 8);
    // capacity: the rows the padded panels can hold (dk_model_append fills them without reallocating)
    const size_t cap_n = (size_t)dm.n_pad - (size_t)(p + 1);
    const size_t small_cap = (size_t)p + 2 * (size_t)p * p + 2 * cap_n * (p + 1) + cap_n;
    if ((rc = dev_alloc(m, &m->d_small, small_cap))) return bail(rc);
    m->small_doubles = small_cap;
    if ((rc = upload(m->d_small, small.data(), small.size(), "the trend / weight arrays"))) return bail(rc);
    dm.beta = m->d_small + o_beta; dm.RM = m->d_small + o_RM; dm.tuuinv = m->d_small + o_tu;
    dm.alpha = m->d_small + o_al; dm.KinvF = m->d_small + o_al + n;
    dm.noise_var = D->noise_var ? m->d_small + o_nv : nullptr;
    dm.Mcm = m->d_small + o_zM + n;
    m->h_zM = (double*)std::malloc((size_t)n * (p + 1) * 8);
    if (!m->h_zM) return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
    std::memcpy(m->h_zM, &small[o_zM], (size_t)n * (p + 1) * 8);
    if (D->y) {
      m->h_y = (double*)std::malloc((size_t)n * 8);
      if (!m->h_y) return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
      std::memcpy(m->h_y, D->y, (size_t)n * 8);
    }
    if (D->noise_var) {
      m->h_nv = (double*)std::malloc((size_t)n * 8);
      if (!m->h_nv) return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
      std::memcpy(m->h_nv, D->noise_var, (size_t)n * 8);
    }

    // host copy of T for the lazily built backward tiles
    m->h_T = (double*)std::malloc((size_t)n * n * 8);
    if (!m->h_T) return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
    std::memcpy(m->h_T, D->T, (size_t)n * n * 8);
    // what dk_model_append needs to grow the model without re-factorising it
    m->ld = n;
    m->h_X = (double*)std::malloc((size_t)n * d * 8);
    m->h_aK = (double*)std::malloc((size_t)n * (p + 1) * 8);
    if (!m->h_X || !m->h_aK) return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
    std::memcpy(m->h_X, Xrm.data(), (size_t)n * d * 8);
    std::memcpy(m->h_aK, aK.data(), (size_t)n * (p + 1) * 8);
    std::memcpy(m->h_G, G.data(), (size_t)p * p * 8);
    std::memcpy(m->h_beta, D->beta, (size_t)p * 8);
  } catch (const std::bad_alloc&) {
    return bail(fail(DK_ERR_NOMEM, "host allocation failed while tiling the factor"));
  }

  if ((rc = refresh_small_tiles(m))) return bail(rc);

  // ---- per-CTA V panels: two resident CTAs per SM ----
  m->grid = 2 * m->num_sms;
  if (const char* e = std::getenv("DK_CTAS_PER_SM")) {  // experiment knob: 1 = one resident CTA per SM
    if (std::atoi(e) == 1) m->grid = m->num_sms;
  }
  m->scratch_doubles_per_cta = (size_t)dm.n_pad * BN;
  if ((rc = dev_alloc(m, &m->d_vscratch, (size_t)m->grid * m->scratch_doubles_per_cta))) return bail(rc);
  if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail(DK_ERR_CUDA, "cudaStreamCreate failed"));
  if (cudaDeviceSynchronize() != cudaSuccess)
    return bail(fail(DK_ERR_CUDA, "model upload failed: %s", cudaGetErrorString(cudaGetLastError())));
  *out = m;
  return DK_OK;
}

// A replica of `src` on another device: the three uploads (factor tiles, design, small arrays) are copied device to
// device, nothing is re-tiled on the host; the lazily built pieces (backward tiles, AKG constants, K^-1) are rebuilt
// per replica on first use from the host copies.
int dk_model_clone(const dk_model* src, int device, dk_model** out) {
  if (!src || !out) return fail(DK_ERR_ARG, "dk_model_clone: NULL argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(DK_ERR_UNSUPPORTED, "no CUDA device visible");
  if (device < 0 || device >= ndev) return fail(DK_ERR_ARG, "device %d out of range (%d visible)", device, ndev);
  cudaDeviceProp prop;
  DK_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(DK_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  DeviceGuard guard(device);
  int rc = set_kernel_attrs(device);
  if (rc) return rc;
  dk_model* m = new (std::nothrow) dk_model();
  if (!m) return fail(DK_ERR_NOMEM, "host allocation failed");
  auto bail = [&](int code) { dk_model_destroy(m); return code; };
  m->device = device;
  m->num_sms = prop.multiProcessorCount;
  m->n = src->n; m->d = src->d; m->p = src->p;
  m->dm = src->dm;
  m->dm.tiles_bwd = nullptr; m->dm.akg_a_sorted = nullptr; m->dm.akg_delta = nullptr; m->dm.akg_perm = nullptr;
  m->tiles_fwd_doubles = src->tiles_fwd_doubles; m->xrm_doubles = src->xrm_doubles; m->small_doubles = src->small_doubles;
  if ((rc = dev_alloc(m, &m->d_tiles_fwd, m->tiles_fwd_doubles))) return bail(rc);
  if ((rc = dev_alloc(m, &m->d_Xrm, m->xrm_doubles))) return bail(rc);
  if ((rc = dev_alloc(m, &m->d_small, m->small_doubles))) return bail(rc);
  auto peer = [&](double* dst, const double* s, size_t count) {
    const cudaError_t e = cudaMemcpyPeer(dst, device, s, src->device, count * 8);
    return e == cudaSuccess ? DK_OK : fail(DK_ERR_CUDA, "device-to-device copy of the model failed: %s", cudaGetErrorString(e));
  };
  if ((rc = peer(m->d_tiles_fwd, src->d_tiles_fwd, m->tiles_fwd_doubles))) return bail(rc);
  if ((rc = peer(m->d_Xrm, src->d_Xrm, m->xrm_doubles))) return bail(rc);
  if ((rc = peer(m->d_small, src->d_small, m->small_doubles))) return bail(rc);
  auto rebase = [](const double* p, const double* old_base, double* new_base) -> const double* {
    return p ? new_base + (p - old_base) : nullptr;
  };
  DevModel& dm = m->dm;
  dm.tiles_fwd = m->d_tiles_fwd;
  dm.Xrm = rebase(src->dm.Xrm, src->d_Xrm, m->d_Xrm);
  dm.Xrm_s = rebase(src->dm.Xrm_s, src->d_Xrm, m->d_Xrm);
  dm.Xrm_r8 = rebase(src->dm.Xrm_r8, src->d_Xrm, m->d_Xrm);
  dm.beta = rebase(src->dm.beta, src->d_small, m->d_small);
  dm.RM = rebase(src->dm.RM, src->d_small, m->d_small);
  dm.tuuinv = rebase(src->dm.tuuinv, src->d_small, m->d_small);
  dm.alpha = rebase(src->dm.alpha, src->d_small, m->d_small);
  dm.KinvF = rebase(src->dm.KinvF, src->d_small, m->d_small);
  dm.Mcm = rebase(src->dm.Mcm, src->d_small, m->d_small);
  dm.noise_var = rebase(src->dm.noise_var, src->d_small, m->d_small);
  const size_t n = (size_t)src->n, p = (size_t)src->p;
  auto dup = [&](const double* h, size_t count, double** dst) {
    if (!h) return true;
    *dst = (double*)std::malloc(count * 8);
    if (!*dst) return false;
    std::memcpy(*dst, h, count * 8);
    return true;
  };
  const size_t ld = (size_t)src->ld;
  m->ld = src->ld;
  if (!dup(src->h_T, ld * ld, &m->h_T) || !dup(src->h_zM, ld * (p + 1), &m->h_zM) || !dup(src->h_y, ld, &m->h_y) ||
      !dup(src->h_nv, ld, &m->h_nv) || !dup(src->h_X, ld * src->d, &m->h_X) || !dup(src->h_aK, ld * (p + 1), &m->h_aK))
    return bail(fail(DK_ERR_NOMEM, "host allocation failed"));
  std::memcpy(m->h_G, src->h_G, sizeof m->h_G);
  std::memcpy(m->h_beta, src->h_beta, sizeof m->h_beta);
  (void)n;
  if (src->d_tiles_small) {
    if ((rc = dev_alloc(m, &m->d_tiles_small, src->tiles_small_doubles))) return bail(rc);
    m->tiles_small_doubles = src->tiles_small_doubles;
    if ((rc = peer(m->d_tiles_small, src->d_tiles_small, src->tiles_small_doubles))) return bail(rc);
    dm.tiles_small = m->d_tiles_small;
    m->small_lay = src->small_lay;
  } else dm.tiles_small = nullptr;
  m->grid = src->grid * m->num_sms / std::max(1, src->num_sms);
  m->scratch_doubles_per_cta = src->scratch_doubles_per_cta;
  if ((rc = dev_alloc(m, &m->d_vscratch, (size_t)m->grid * m->scratch_doubles_per_cta))) return bail(rc);
  if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess)
    return bail(fail(DK_ERR_CUDA, "cudaStreamCreate failed"));
  if (cudaDeviceSynchronize() != cudaSuccess)
    return bail(fail(DK_ERR_CUDA, "model replication failed: %s", cudaGetErrorString(cudaGetLastError())));
  *out = m;
  return DK_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// dk_model_append: one more observation, hyper-parameters and trend coefficients unchanged
// ------------------------------------------------------------------------------------------
namespace dk { namespace {
// 1-D kernels as DiceKriging codes them: k(x, X_i) for the n design rows held on the host
void host_cov_vector(const dk_model* m, const double* x, double* c) {
  const DevModel& dm = m->dm;
  const int n = m->n, d = m->d;
  for (int i = 0; i < n; i++) {
    double s = 0.0, hs = 0.0;
    for (int k = 0; k < d; k++) {
      const double h = std::fabs(m->h_X[(size_t)i * d + k] - x[k]);
      hs += h;
      switch (dm.kernel) {
        case DK_GAUSS: { const double a = h / dm.theta[k]; s += 0.5 * a * a; break; }
        case DK_EXP: s += h / dm.theta[k]; break;
        case DK_MATERN3_2: { const double a = std::sqrt(3.0) * h / dm.theta[k]; s += a - std::log1p(a); break; }
        case DK_MATERN5_2: { const double a = std::sqrt(5.0) * h / dm.theta[k]; s += a - std::log(1.0 + a + a * a / 3.0); break; }
        default: s += std::pow(h / dm.theta[k], dm.shape[k]); break;
      }
    }
    c[i] = dm.sd2 * std::exp(-s);
    if (dm.nugget_flag && hs < 1e-15) c[i] += dm.nugget;
  }
}

void free_lazy(dk_model* m) {
  // everything built lazily from the factor: rebuilt on next use (ensure_bwd refills its buffers in place when they fit)
  m->bwd_valid = false;
  const size_t n = (size_t)m->n;
  auto drop = [&](double** q) { if (*q) { cudaFree(*q); *q = nullptr; } };
  if (m->d_akg_const) { m->device_bytes -= 2 * n * 8; drop(&m->d_akg_const); }
  if (m->d_akg_perm) { m->device_bytes -= n * 4; cudaFree(m->d_akg_perm); m->d_akg_perm = nullptr; }
  if (m->d_akg_scratch) { cudaFree(m->d_akg_scratch); m->d_akg_scratch = nullptr; }
  if (m->d_Kinv) { m->device_bytes -= n * n * 8; drop(&m->d_Kinv); }
  if (m->d_delta) { m->device_bytes -= n * 8; drop(&m->d_delta); }
  for (int t = 0; t < 2; t++)
    if (m->d_knyy[t]) { m->device_bytes -= n * 8; drop(&m->d_knyy[t]); }
  m->dm.akg_a_sorted = nullptr; m->dm.akg_delta = nullptr; m->dm.akg_perm = nullptr;
}

// host leading dimension and device buffers for at least want_n design points (+ 256 rows of head-room)
int grow_capacity(dk_model* m, int want_n) {
  const int p = m->p, d = m->d, n = m->n;
  const int old_pad = (int)(m->scratch_doubles_per_cta / BN);
  const int new_pad = BM * ((want_n + 256 + p + 1 + BM - 1) / BM);
  // ---- host: re-stride to the new leading dimension ----
  const size_t ld0 = (size_t)m->ld, ld1 = (size_t)new_pad;
  auto restride = [&](double** a, int cols, size_t rows_used) -> bool {
    if (!*a) return true;
    double* b = (double*)std::calloc(ld1 * (size_t)cols, 8);
    if (!b) return false;
    for (int c = 0; c < cols; c++) std::memcpy(b + (size_t)c * ld1, *a + (size_t)c * ld0, rows_used * 8);
    std::free(*a);
    *a = b;
    return true;
  };
  bool ok = restride(&m->h_T, n, (size_t)n);
  if (ok) {  // h_T must be ld1 x ld1, not ld1 x n
    double* b = (double*)std::realloc(m->h_T, ld1 * ld1 * 8);
    ok = b != nullptr;
    if (ok) { m->h_T = b; std::memset(m->h_T + ld1 * (size_t)n, 0, ld1 * (ld1 - (size_t)n) * 8); }
  }
  ok = ok && restride(&m->h_zM, p + 1, (size_t)n) && restride(&m->h_aK, p + 1, (size_t)n);
  auto regrow = [&](double** a, size_t per_row) -> bool {
    if (!*a) return true;
    double* b = (double*)std::realloc(*a, ld1 * per_row * 8);
    if (!b) return false;
    std::memset(b + (size_t)n * per_row, 0, (ld1 - (size_t)n) * per_row * 8);
    *a = b;
    return true;
  };
  ok = ok && regrow(&m->h_y, 1) && regrow(&m->h_nv, 1) && regrow(&m->h_X, (size_t)d);
  if (!ok) return fail(DK_ERR_NOMEM, "host allocation failed while growing the model");
  m->ld = new_pad;
  // ---- device ----
  if (new_pad > old_pad) {
    DK_CUDA_TRY(cudaDeviceSynchronize());
    double *t = nullptr, *X = nullptr, *sm = nullptr, *vs = nullptr;
    const size_t tcount = tile_base(new_pad / BM) * A_TILE, xcount = 3 * (size_t)new_pad * d;
    const size_t cap_n = (size_t)new_pad - (size_t)(p + 1);
    const size_t scount = (size_t)p + 2 * (size_t)p * p + 2 * cap_n * (p + 1) + cap_n;
    const size_t vcount = (size_t)m->grid * (size_t)new_pad * BN;
    // the V panels go first: freeing the old ones makes room for the new ones
    dev_release(m, &m->d_vscratch, (size_t)m->grid * m->scratch_doubles_per_cta);
    if (m->d_wscratch) dev_release(m, &m->d_wscratch, (size_t)m->grid * (m->scratch_doubles_per_cta + COEF_DOUBLES));
    if (m->d_cscratch) dev_release(m, &m->d_cscratch, (size_t)m->grid * m->scratch_doubles_per_cta);
    if (m->d_tiles_bwd) dev_release(m, &m->d_tiles_bwd, m->bwd_tile_doubles);
    m->dm.tiles_bwd = nullptr;
    int rc = dev_alloc(m, &vs, vcount);
    if (!rc) rc = dev_alloc(m, &t, tcount);
    if (!rc) rc = dev_alloc(m, &X, xcount);
    if (!rc) rc = dev_alloc(m, &sm, scount);
    if (rc) return rc;   // the handle is unusable after this (out of device memory): the caller destroys it
    DK_CUDA_TRY(cudaMemcpy(t, m->d_tiles_fwd, m->tiles_fwd_doubles * 8, cudaMemcpyDeviceToDevice));
    DK_CUDA_TRY(cudaMemset(X, 0, xcount * 8));
    for (int a = 0; a < 3; a++)
      DK_CUDA_TRY(cudaMemcpy(X + (size_t)a * new_pad * d, m->d_Xrm + (size_t)a * old_pad * d, (size_t)old_pad * d * 8,
                             cudaMemcpyDeviceToDevice));
    dev_release(m, &m->d_tiles_fwd, m->tiles_fwd_doubles);
    dev_release(m, &m->d_Xrm, m->xrm_doubles);
    dev_release(m, &m->d_small, m->small_doubles);
    m->d_vscratch = vs; m->d_tiles_fwd = t; m->d_Xrm = X; m->d_small = sm;
    m->tiles_fwd_doubles = tcount; m->xrm_doubles = xcount; m->small_doubles = scount;
    m->scratch_doubles_per_cta = (size_t)new_pad * BN;
    m->dm.tiles_fwd = t;
    m->dm.Xrm = X; m->dm.Xrm_s = X + (size_t)new_pad * d; m->dm.Xrm_r8 = X + 2 * (size_t)new_pad * d;
  }
  return DK_OK;
}

// pack beta | RM | tuuinv | [alpha | T^-1 M] | [z | M] | noise.var from the host state, upload, set the kernel pointers
int upload_small(dk_model* m) {
  const int n = m->n, p = m->p;
  const size_t ld = (size_t)m->ld;
  std::vector<double> RM((size_t)p * p, 0.0), Ginv((size_t)p * p, 0.0);
  const double* G = m->h_G;
  for (int j = 0; j < p; j++)
    for (int i = 0; i <= j; i++) {
      double s = G[i + (size_t)j * p];
      for (int k = 0; k < i; k++) s -= RM[k + (size_t)i * p] * RM[k + (size_t)j * p];
      if (i < j) RM[i + (size_t)j * p] = s / RM[i + (size_t)i * p];
      else {
        if (!(s > 0)) return fail(DK_ERR_ARG, "M'M is not positive definite (trend matrix rank deficient)");
        RM[j + (size_t)j * p] = std::sqrt(s);
      }
    }
  std::vector<double> col(p);
  for (int e = 0; e < p; e++) {
    for (int i = 0; i < p; i++) {
      double s = (i == e) ? 1.0 : 0.0;
      for (int k = 0; k < i; k++) s -= RM[k + (size_t)i * p] * col[k];
      col[i] = s / RM[i + (size_t)i * p];
    }
    for (int i = p - 1; i >= 0; i--) {
      double s = col[i];
      for (int k = i + 1; k < p; k++) s -= RM[i + (size_t)k * p] * col[k];
      col[i] = s / RM[i + (size_t)i * p];
    }
    for (int i = 0; i < p; i++) Ginv[i + (size_t)e * p] = col[i];
  }
  const bool nv = m->h_nv != nullptr;
  std::vector<double> small((size_t)p + 2 * (size_t)p * p + 2 * (size_t)n * (p + 1) + (nv ? n : 0));
  if (small.size() > m->small_doubles) return fail(DK_ERR_NOMEM, "internal: small-array capacity exceeded");
  size_t off = 0;
  const size_t o_beta = off; std::memcpy(&small[off], m->h_beta, p * 8); off += p;
  const size_t o_RM = off; std::memcpy(&small[off], RM.data(), (size_t)p * p * 8); off += (size_t)p * p;
  const size_t o_tu = off; std::memcpy(&small[off], Ginv.data(), (size_t)p * p * 8); off += (size_t)p * p;
  const size_t o_al = off;
  for (int c = 0; c <= p; c++) std::memcpy(&small[off + (size_t)c * n], m->h_aK + (size_t)c * ld, (size_t)n * 8);
  off += (size_t)n * (p + 1);
  const size_t o_zM = off;
  for (int c = 0; c <= p; c++) std::memcpy(&small[off + (size_t)c * n], m->h_zM + (size_t)c * ld, (size_t)n * 8);
  off += (size_t)n * (p + 1);
  const size_t o_nv = off;
  if (nv) std::memcpy(&small[off], m->h_nv, (size_t)n * 8);
  DK_CUDA_TRY(cudaMemcpy(m->d_small, small.data(), small.size() * 8, cudaMemcpyHostToDevice));
  DevModel& dm = m->dm;
  dm.beta = m->d_small + o_beta; dm.RM = m->d_small + o_RM; dm.tuuinv = m->d_small + o_tu;
  dm.alpha = m->d_small + o_al; dm.KinvF = m->d_small + o_al + n;
  dm.noise_var = nv ? m->d_small + o_nv : nullptr;
  dm.Mcm = m->d_small + o_zM + n;
  return DK_OK;
}

}  // namespace
}  // namespace dk

// The last column of the new Cholesky factor needs one forward substitution against the old one:
//   v = T'^-1 c(x),  t = sqrt(k(x, x) - v'v),  T+ = [[T, v], [0, t]],  z+ = [z, (y - f'beta - v'z) / t],
//   M+ = [M; (f - v'M) / t],  alpha+ = [alpha - a u; a] with a = z+_n / t and u = T^-1 v (likewise T^-1 M),
// and on the device only the block rows of the tiled factor that hold row n or later change (the new row of T', and the
// augmented rows [z M]' one row further down).
extern "C" int dk_model_append(dk_model* m, const double* x_new, double y_new, double new_noise_var) {
  if (!m || !x_new) return fail(DK_ERR_ARG, "dk_model_append: NULL argument");
  if (!m->h_y) return fail(DK_ERR_ARG, "dk_model_append needs model@y (dk_model_desc.y was NULL)");
  DeviceGuard guard(m->device);
  DK_CUDA_TRY(cudaDeviceSynchronize());   // nothing of this handle may be in flight while its factor changes
  const int n = m->n, p = m->p, d = m->d;
  int rc;
  if (n + 1 + p + 1 > (int)(m->scratch_doubles_per_cta / BN) || n + 1 > m->ld)
    if ((rc = grow_capacity(m, n + 1))) return rc;
  const size_t ld = (size_t)m->ld;
  DevModel& dm = m->dm;
  std::vector<double> v((size_t)n), u((size_t)n);
  host_cov_vector(m, x_new, v.data());
  // v <- T'^-1 c: row i of T' is column i of T (contiguous)
  for (int i = 0; i < n; i++) {
    const double* col = m->h_T + (size_t)i * ld;
    long double s = v[i];
    for (int j = 0; j < i; j++) s -= (long double)col[j] * v[j];
    v[i] = (double)(s / col[i]);
  }
  long double vv = 0.0L;
  for (int i = 0; i < n; i++) vv += (long double)v[i] * v[i];
  const double knn = dm.sd2 + (dm.nugget_flag ? dm.nugget : (m->h_nv ? new_noise_var : 0.0));
  const double t2 = (double)((long double)knn - vv);
  if (!(t2 > 0.0))
    return fail(DK_ERR_ARG, "the new point makes the covariance matrix numerically singular (k(x,x) - v'v = %.3e); "
                            "checkPredict guards against this (R/checkPredict.R)", t2);
  const double t = std::sqrt(t2);
  // trend row f(x), z+ and M+
  double f[DK_MAX_P], mnew[DK_MAX_P];
  f[0] = 1.0;
  if (dm.trend == DK_TREND_LINEAR)
    for (int k = 0; k < d; k++) f[1 + k] = x_new[k];
  long double fb = 0.0L, vz = 0.0L;
  for (int j = 0; j < p; j++) fb += (long double)f[j] * m->h_beta[j];
  for (int i = 0; i < n; i++) vz += (long double)v[i] * m->h_zM[i];
  const double znew = (double)(((long double)y_new - fb - vz) / t);
  for (int j = 0; j < p; j++) {
    long double vm = 0.0L;
    const double* Mj = m->h_zM + (size_t)(1 + j) * ld;
    for (int i = 0; i < n; i++) vm += (long double)v[i] * Mj[i];
    mnew[j] = (double)(((long double)f[j] - vm) / t);
  }
  // u = T^-1 v (column-oriented back substitution: column k of T is contiguous)
  u = v;
  for (int k = n - 1; k >= 0; k--) {
    const double* col = m->h_T + (size_t)k * ld;
    u[k] /= col[k];
    const double uk = u[k];
    for (int i = 0; i < k; i++) u[i] -= uk * col[i];
  }
  // ---- host state ----
  std::memcpy(m->h_T + (size_t)n * ld, v.data(), (size_t)n * 8);
  m->h_T[(size_t)n * ld + n] = t;
  m->h_zM[n] = znew;
  for (int j = 0; j < p; j++) m->h_zM[(size_t)(1 + j) * ld + n] = mnew[j];
  for (int c = 0; c <= p; c++) {
    double* col = m->h_aK + (size_t)c * ld;
    const double a = (c == 0 ? znew : mnew[c - 1]) / t;
    for (int i = 0; i < n; i++) col[i] -= a * u[i];
    col[n] = a;
  }
  for (int a = 0; a < p; a++)
    for (int b = 0; b < p; b++) m->h_G[a + (size_t)b * p] += mnew[a] * mnew[b];
  m->h_y[n] = y_new;
  if (m->h_nv) m->h_nv[n] = new_noise_var;
  std::memcpy(m->h_X + (size_t)n * d, x_new, (size_t)d * 8);
  // ---- device: lazily built pieces go stale, sizes move on ----
  free_lazy(m);
  const int I0 = n / BM;                        // first block row that holds row n of L_aug
  m->n = n + 1;
  dm.n = n + 1;
  dm.n_aug = n + 1 + p + 1;
  dm.nblk = (dm.n_aug + BM - 1) / BM;
  dm.n_pad = dm.nblk * BM;
  dm.nblk_b = (n + 1 + BM - 1) / BM;
  std::vector<double> tiles;
  try {
    AugFactor L{m->h_T, m->h_zM, m->h_zM + ld, n + 1, p, dm.n_aug, ld, ld};
    build_tiles(L, dm.nblk, tiles, I0);
  } catch (const std::bad_alloc&) {
    return fail(DK_ERR_NOMEM, "host allocation failed while tiling the new rows");
  }
  if (tile_base(dm.nblk) * A_TILE > m->tiles_fwd_doubles) return fail(DK_ERR_NOMEM, "internal: tile capacity exceeded");
  DK_CUDA_TRY(cudaMemcpy(m->d_tiles_fwd + tile_base(I0) * A_TILE, tiles.data(), tiles.size() * 8, cudaMemcpyHostToDevice));
  // design row n in its three forms (plain, scaled by w_k, rounded to 8 digits)
  const size_t cap_pad = m->scratch_doubles_per_cta / BN;
  std::vector<double> row(3 * (size_t)d);
  for (int k = 0; k < d; k++) {
    row[k] = x_new[k];
    row[d + k] = x_new[k] * dm.w[k];
    row[2 * d + k] = std::nearbyint(x_new[k] * 1e8) / 1e8;
  }
  for (int a = 0; a < 3; a++)
    DK_CUDA_TRY(cudaMemcpy(m->d_Xrm + (size_t)a * cap_pad * d + (size_t)n * d, row.data() + (size_t)a * d, (size_t)d * 8,
                           cudaMemcpyHostToDevice));
  if ((rc = upload_small(m))) return rc;
  return refresh_small_tiles(m);   // small designs: the whole-factor inverse (dropped once n_pad exceeds 256)
}

extern "C" {

int dk_model_n(const dk_model* m) { return m ? m->n : 0; }
int dk_model_small_kernel(const dk_model* m) { return (m && m->dm.tiles_small) ? 1 : 0; }

void dk_model_destroy(dk_model* m) {
  if (!m) return;
  DeviceGuard guard(m->device);
  cudaDeviceSynchronize();
  cudaFree(m->d_tiles_fwd); cudaFree(m->d_tiles_bwd); cudaFree(m->d_tiles_small); cudaFree(m->d_Xrm); cudaFree(m->d_small);
  cudaFree(m->d_vscratch); cudaFree(m->d_wscratch); cudaFree(m->d_cscratch); cudaFree(m->d_io); cudaFree(m->d_red);
  cudaFree(m->d_aux); cudaFree(m->d_mom); cudaFree(m->d_deriv);
  if (m->stream) cudaStreamDestroy(m->stream);
  if (m->copy_in) cudaStreamDestroy(m->copy_in);
  if (m->copy_out) cudaStreamDestroy(m->copy_out);
  for (int s = 0; s < 2; s++) {
    if (m->ev_h2d[s]) cudaEventDestroy(m->ev_h2d[s]);
    if (m->ev_kern[s]) cudaEventDestroy(m->ev_kern[s]);
    if (m->ev_d2h[s]) cudaEventDestroy(m->ev_d2h[s]);
  }
  if (m->h_stage) cudaFreeHost(m->h_stage);
  cudaFree(m->d_chunk); cudaFree(m->d_vec); cudaFree(m->d_gen);
  if (m->ev0) cudaEventDestroy(m->ev0);
  if (m->ev1) cudaEventDestroy(m->ev1);
  if (m->ev_mid) cudaEventDestroy(m->ev_mid);
  std::free(m->h_T);
  std::free(m->h_X);
  std::free(m->h_aK);
  std::free(m->h_zM);
  std::free(m->h_y);
  std::free(m->h_nv);
  cudaFree(m->d_akg_const); cudaFree(m->d_akg_perm); cudaFree(m->d_akg_scratch);
  cudaFree(m->d_Kinv); cudaFree(m->d_akg_gscratch);
  cudaFree(m->d_delta); cudaFree(m->d_knyy[0]); cudaFree(m->d_knyy[1]);
  delete m;
}

size_t dk_model_device_bytes(const dk_model* m) { return m ? m->device_bytes : 0; }
long long dk_model_launches(const dk_model* m) { return m ? m->launches : 0; }

int dk_model_set_timing(dk_model* m, int on) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  DeviceGuard guard(m->device);
  if (on && !m->ev0) {
    DK_CUDA_TRY(cudaEventCreate(&m->ev0));
    DK_CUDA_TRY(cudaEventCreate(&m->ev1));
    DK_CUDA_TRY(cudaEventCreate(&m->ev_mid));
  }
  m->timing = on ? 1 : 0;
  m->ev_valid = false;
  m->ev_mid_valid = false;
  return DK_OK;
}

double dk_model_last_kernel_ms(dk_model* m) {
  if (!m || !m->ev_valid) return -1.0;
  DeviceGuard guard(m->device);
  if (cudaEventSynchronize(m->ev1) != cudaSuccess) return -1.0;
  float ms = -1.0f;
  if (cudaEventElapsedTime(&ms, m->ev0, m->ev1) != cudaSuccess) return -1.0;
  return (double)ms;
}

int dk_model_prepare(dk_model* m, int what) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  DeviceGuard guard(m->device);
  int rc = DK_OK;
  if (what & (DK_PREP_GRADIENTS | DK_PREP_AKG | DK_PREP_AKG_GRAD)) rc = ensure_bwd(m);
  if (!rc && (what & (DK_PREP_AKG | DK_PREP_AKG_GRAD))) rc = prepare_akg(m, (what & DK_PREP_AKG_GRAD) != 0, m->stream);
  if (!rc) DK_CUDA_TRY(cudaDeviceSynchronize());
  return rc;
}

int dk_model_set_tile_mode(dk_model* m, int mode) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  if (mode != DK_TILE_AUTO && mode != DK_TILE_WIDE && mode != DK_TILE_NARROW) return fail(DK_ERR_ARG, "tile mode %d", mode);
  m->tile_mode = mode;
  return DK_OK;
}

double dk_model_last_phase_ms(dk_model* m) {
  if (!m || !m->ev_valid || !m->ev_mid_valid) return -1.0;
  DeviceGuard guard(m->device);
  if (cudaEventSynchronize(m->ev_mid) != cudaSuccess) return -1.0;
  float ms = -1.0f;
  if (cudaEventElapsedTime(&ms, m->ev0, m->ev_mid) != cudaSuccess) return -1.0;
  return (double)ms;
}

// ------------------------------------------------------------------------------------------
// device-pointer entry points
// ------------------------------------------------------------------------------------------
int dk_predict_dev(dk_model* m, const double* d_x, int64_t N, int type, double* d_mean, double* d_sd, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_mean) return fail(DK_ERR_ARG, "d_mean is NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_PREDICT, type, 1, 0, 0.0, 0.0};
  return launch_predict(m, P, d_x, N, nullptr, d_mean, d_sd, (cudaStream_t)stream);
}

int dk_ei_dev(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization, int proxy,
              double* d_ei, double* d_mean, double* d_sd, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_ei) return fail(DK_ERR_ARG, "d_ei is NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return launch_predict(m, P, d_x, N, d_ei, d_mean, d_sd, (cudaStream_t)stream);
}

int dk_quantile_dev(dk_model* m, const double* d_x, int64_t N, double qnorm_beta, int type, double* d_q,
                    double* d_grad, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_q && !d_grad) return fail(DK_ERR_ARG, "both outputs are NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_QUANTILE, type, 1, 0, 0.0, qnorm_beta};
  if (d_grad) return launch_grad(m, P, d_x, N, d_grad, d_q, (cudaStream_t)stream);
  return launch_predict(m, P, d_x, N, d_q, nullptr, nullptr, (cudaStream_t)stream);
}

int dk_ei_grad_dev(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization, int proxy,
                   double* d_grad, double* d_ei, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_grad) return fail(DK_ERR_ARG, "d_grad is NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return launch_grad(m, P, d_x, N, d_grad, d_ei, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------
// host-pointer entry points: chunked H2D -> kernel -> D2H pipeline (run_chunked), then synchronise.
// The host_* forms work on rows [row0, row0 + n) of the caller's matrix and write the same rows of the caller's
// outputs: a whole call is (0, N); the multi-GPU entry points (dk_api_multi.cu) hand every device its own range.
// ------------------------------------------------------------------------------------------
}  // extern "C"

namespace dk {

int host_predict(dk_model* m, const CritParams& P, const double* x, int64_t ldx, int64_t row0, int64_t n,
                 double* out_main, double* out_mean, double* out_sd) {
  DeviceGuard guard(m->device);
  const HostOut outs[3] = {{out_main, 1, ldx, false}, {out_mean, 1, ldx, false}, {out_sd, 1, ldx, false}};
  int rc = run_chunked(m, x, ldx, row0, n, 1, chunk_units_for(m, 1, BN), outs, 3,
                       [&](const double* d_x, int64_t rows, double* const* o, int64_t, cudaStream_t st) {
                         return launch_predict(m, P, d_x, rows, o[0], o[1], o[2], st);
                       });
  if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); cudaStreamSynchronize(m->copy_out); }
  return rc;
}

int host_grad(dk_model* m, const CritParams& P, const double* x, int64_t ldx, int64_t row0, int64_t n, double* out_grad,
              double* out_main) {
  DeviceGuard guard(m->device);
  int rc = ensure_bwd(m);
  if (rc) return rc;
  const HostOut outs[2] = {{out_grad, m->d, ldx, false}, {out_main, 1, ldx, false}};
  rc = run_chunked(m, x, ldx, row0, n, 1, chunk_units_for(m, 1, BN), outs, 2,
                   [&](const double* d_x, int64_t rows, double* const* o, int64_t, cudaStream_t st) {
                     return launch_grad(m, P, d_x, rows, o[0], o[1], st);
                   });
  if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); cudaStreamSynchronize(m->copy_out); }
  return rc;
}

namespace {
int ensure_vec(dk_model* m, size_t doubles) {
  if (m->vec_doubles >= doubles) return DK_OK;
  dev_release(m, &m->d_vec, m->vec_doubles);
  m->vec_doubles = 0;
  int rc = dev_alloc(m, &m->d_vec, doubles);
  if (rc) return rc;
  m->vec_doubles = doubles;
  return DK_OK;
}
}  // namespace

namespace {
int ensure_gen(dk_model* m, size_t doubles) {
  if (m->gen_doubles >= doubles) return DK_OK;
  dev_release(m, &m->d_gen, m->gen_doubles);
  m->gen_doubles = 0;
  int rc = dev_alloc(m, &m->d_gen, doubles);
  if (rc) return rc;
  m->gen_doubles = doubles;
  return DK_OK;
}

// one criterion over device-resident rows
int crit_launch(dk_model* m, const ScreenSpec& S, const double* d_x, int64_t rows, double* d_out, cudaStream_t st) {
  if (S.criterion == DK_CRIT_AKG) return dk_akg_dev(m, d_x, rows, S.tau2, S.P.type, d_out, st);
  return launch_predict(m, S.P, d_x, rows, d_out, nullptr, nullptr, st);
}
}  // namespace

// The screen of R/max_crit.R:104-110 over rows [row0, row0 + n): the criterion into a device vector (never copied
// back), then the k best (EI, AKG: largest; kriging.quantile: smallest; ties to the lower index; indices relative to
// row0, -1 when fewer than k are comparable).  Candidates: the caller's host matrix (chunked pipeline), a device matrix,
// or rows [row0, row0 + n) of the counter-based generator `gen` (generated chunk by chunk on the device, never uploaded).
int host_screen(dk_model* m, const ScreenSpec& S, const double* x, int x_on_device, const dk_uniform_spec* gen, int64_t ldx,
                int64_t row0, int64_t n, int k, int64_t* out_idx, double* out_val, cudaStream_t user_stream) {
  DeviceGuard guard(m->device);
  int rc = ensure_vec(m, (size_t)std::max<int64_t>(n, 1));
  if (rc) return rc;
  if ((rc = ensure_red(m, topk_workspace_bytes(k)))) return rc;
  cudaStream_t st = user_stream ? user_stream : ((x && x_on_device) ? (cudaStream_t) nullptr : m->stream);
  if (!x) {
    if (!gen) return fail(DK_ERR_ARG, "neither candidates nor a generator were given");
    GenSpec G;
    G.seed = gen->seed;
    for (int c = 0; c < m->d; c++) {
      const double lo = gen->lower ? gen->lower[c] : 0.0, up = gen->upper ? gen->upper[c] : 1.0;
      if (!(up >= lo)) return fail(DK_ERR_ARG, "generator bounds: upper[%d] < lower[%d]", c, c);
      G.lower[c] = lo;
      G.width[c] = up - lo;
    }
    const int64_t chunk = std::min<int64_t>(std::max<int64_t>(n, 1), chunk_units_for(m, 1, BN));
    if ((rc = ensure_gen(m, (size_t)chunk * m->d))) return rc;
    for (int64_t c0 = 0; c0 < n; c0 += chunk) {
      const int64_t rows = std::min(chunk, n - c0);
      k_gen_uniform<<<(unsigned)std::min<int64_t>((rows * m->d + 255) / 256, 4 * 148), 256, 0, st>>>(G, m->d, row0 + c0, rows, m->d_gen);
      m->launches++;
      DK_CUDA_TRY(cudaGetLastError());
      if ((rc = crit_launch(m, S, m->d_gen, rows, m->d_vec + c0, st))) return rc;
    }
  } else if (x_on_device) {
    if (ldx != n) return fail(DK_ERR_ARG, "device-resident candidates must be contiguous (ld = N)");
    if ((rc = crit_launch(m, S, x, n, m->d_vec, st))) return rc;
  } else {
    if (user_stream) {  // the pipeline runs on the handle's streams: order it after the caller's stream
      DK_CUDA_TRY(cudaStreamSynchronize(user_stream));
      st = m->stream;
    }
    rc = run_chunked(m, x, ldx, row0, n, 1, chunk_units_for(m, 1, BN), nullptr, 0,
                     [&](const double* d_x, int64_t rows, double* const*, int64_t u0, cudaStream_t s) {
                       return crit_launch(m, S, d_x, rows, m->d_vec + u0, s);
                     });
    if (rc) { cudaStreamSynchronize(m->copy_in); cudaStreamSynchronize(m->stream); return rc; }
  }
  const double sign = S.criterion == DK_CRIT_QUANTILE ? -1.0 : 1.0;
  if ((rc = launch_topk(m->d_vec, n, k, m->d_red, st, sign))) return rc;
  m->launches += k;
  std::vector<TopkEntry> host(k);
  DK_CUDA_TRY(cudaMemcpyAsync(host.data(), topk_result_ptr(m->d_red), sizeof(TopkEntry) * k, cudaMemcpyDeviceToHost, st));
  DK_CUDA_TRY(cudaStreamSynchronize(st));
  for (int i = 0; i < k; i++) {
    out_idx[i] = host[i].idx;
    out_val[i] = host[i].idx >= 0 ? sign * host[i].val : 0.0;
  }
  return DK_OK;
}

// criterion S over the caller's host matrix into the handle's device vector d_vec (dk_sample_from_ei ranks keys made
// from it); synchronises nothing: the values are complete in stream order on m->stream
int host_weights_to_vec(dk_model* m, const ScreenSpec& S, const double* x, int64_t N) {
  DeviceGuard guard(m->device);
  int rc = ensure_vec(m, (size_t)std::max<int64_t>(N, 1));
  if (rc) return rc;
  rc = run_chunked(m, x, N, 0, N, 1, chunk_units_for(m, 1, BN), nullptr, 0,
                   [&](const double* d_x, int64_t rows, double* const*, int64_t u0, cudaStream_t s) {
                     return crit_launch(m, S, d_x, rows, m->d_vec + u0, s);
                   });
  if (rc) cudaStreamSynchronize(m->stream);
  return rc;
}

// ScreenSpec of the public dk_crit_spec (argument checks included)
int make_screen_spec(const dk_model* m, const dk_crit_spec* c, ScreenSpec* S) {
  if (!c) return fail(DK_ERR_ARG, "criterion spec is NULL");
  S->criterion = c->criterion;
  S->tau2 = c->new_noise_var;
  switch (c->criterion) {
    case DK_CRIT_EI: S->P = CritParams{MODE_EI, c->type, c->minimization, c->proxy, c->plugin, 0.0}; break;
    case DK_CRIT_QUANTILE: S->P = CritParams{MODE_QUANTILE, c->type, 1, 0, 0.0, c->qnorm_beta}; break;
    case DK_CRIT_AKG:
      S->P = CritParams{MODE_PREDICT, c->type, 1, 0, 0.0, 0.0};
      if (!(c->new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
      break;
    default: return fail(DK_ERR_ARG, "criterion %d cannot be screened (DK_CRIT_EI / DK_CRIT_QUANTILE / DK_CRIT_AKG can)", c->criterion);
  }
  (void)m;
  return DK_OK;
}

// coordinates of generated candidates idx[0..k) (host arithmetic: the generator is a pure function of (seed, j, k))
void generated_rows(const dk_uniform_spec* gen, int d, const int64_t* idx, int k, double* out_x) {
  GenSpec G;
  G.seed = gen->seed;
  for (int c = 0; c < d; c++) {
    G.lower[c] = gen->lower ? gen->lower[c] : 0.0;
    G.width[c] = (gen->upper ? gen->upper[c] : 1.0) - G.lower[c];
  }
  for (int i = 0; i < k; i++)
    for (int c = 0; c < d; c++)
      out_x[i + (size_t)c * k] = idx[i] >= 0 ? gen_coordinate(G, idx[i], c) : std::nan("");
}

}  // namespace dk

extern "C" {

int dk_predict(dk_model* m, const double* x, int64_t N, int type, double* out_mean, double* out_sd) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_mean) return fail(DK_ERR_ARG, "out_mean is NULL");
  CritParams P{MODE_PREDICT, type, 1, 0, 0.0, 0.0};
  return host_predict(m, P, x, N, 0, N, nullptr, out_mean, out_sd);
}

int dk_ei(dk_model* m, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
          double* out_ei, double* out_mean, double* out_sd) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_ei) return fail(DK_ERR_ARG, "out_ei is NULL");
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return host_predict(m, P, x, N, 0, N, out_ei, out_mean, out_sd);
}

int dk_ei_grad(dk_model* m, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
               double* out_grad, double* out_ei) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  return host_grad(m, P, x, N, 0, N, out_grad, out_ei);
}

int dk_quantile(dk_model* m, const double* x, int64_t N, double qnorm_beta, int type, double* out_q, double* out_grad) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_q) return fail(DK_ERR_ARG, "out_q is NULL");
  CritParams P{MODE_QUANTILE, type, 1, 0, 0.0, qnorm_beta};
  if (out_grad) return host_grad(m, P, x, N, 0, N, out_grad, out_q);
  return host_predict(m, P, x, N, 0, N, out_q, nullptr, nullptr);
}

int dk_screen_ei(dk_model* m, const double* x, int x_on_device, int64_t N, double plugin, int type, int minimization,
                 int proxy, int k, int64_t* out_idx, double* out_val, void* stream) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (k < 1 || k > DK_TOPK_MAX) return fail(DK_ERR_ARG, "k=%d outside 1..%d", k, DK_TOPK_MAX);
  if (!out_idx || !out_val) return fail(DK_ERR_ARG, "output pointer is NULL");
  ScreenSpec S{DK_CRIT_EI, CritParams{MODE_EI, type, minimization, proxy, plugin, 0.0}, 0.0};
  return host_screen(m, S, x, x_on_device, nullptr, N, 0, N, k, out_idx, out_val, (cudaStream_t)stream);
}

int dk_screen(dk_model* m, const double* x, int x_on_device, const dk_uniform_spec* gen, int64_t N, const dk_crit_spec* crit,
              int k, int64_t* out_idx, double* out_val, double* out_x, void* stream) {
  if (!m) return fail(DK_ERR_ARG, "model handle is NULL");
  if (!x && !gen) return fail(DK_ERR_ARG, "neither candidates nor a generator were given");
  if (N < 0) return fail(DK_ERR_ARG, "N = %lld is negative", (long long)N);
  if (k < 1 || k > DK_TOPK_MAX) return fail(DK_ERR_ARG, "k=%d outside 1..%d", k, DK_TOPK_MAX);
  if (!out_idx || !out_val) return fail(DK_ERR_ARG, "output pointer is NULL");
  if (out_x && x) return fail(DK_ERR_ARG, "out_x is for generated candidates (the caller already holds x)");
  ScreenSpec S;
  int rc = make_screen_spec(m, crit, &S);
  if (rc) return rc;
  if ((rc = host_screen(m, S, x, x_on_device, x ? nullptr : gen, N, 0, N, k, out_idx, out_val, (cudaStream_t)stream))) return rc;
  if (out_x) generated_rows(gen, m->d, out_idx, k, out_x);
  return DK_OK;
}

/* Debug aid (not part of the drop-in surface): the same criterion through the plain
 * one-thread-per-candidate substitution kernel; tests use it to localise tiling bugs. */
int dk_debug_plain(dk_model* m, const double* x, int64_t N, int mode, double plugin, double qnorm_beta, int type,
                   int minimization, int proxy, double* out_main, double* out_mean, double* out_sd) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  DeviceGuard guard(m->device);
  const int n = m->n, p = m->p;
  DevBuf bT, bzM, bv, bxo;
  DK_CUDA_TRY(bT.alloc((size_t)n * n));
  DK_CUDA_TRY(bzM.alloc((size_t)n * (p + 1)));
  DK_CUDA_TRY(bv.alloc((size_t)n * N));
  DK_CUDA_TRY(bxo.alloc((size_t)N * (m->d + 3)));
  double *d_T = bT.p, *d_zM = bzM.p, *d_v = bv.p, *d_xo = bxo.p;
  DK_CUDA_TRY(cudaMemcpy2D(d_T, (size_t)n * 8, m->h_T, (size_t)m->ld * 8, (size_t)n * 8, n, cudaMemcpyHostToDevice));
  DK_CUDA_TRY(cudaMemcpy2D(d_zM, (size_t)n * 8, m->h_zM, (size_t)m->ld * 8, (size_t)n * 8, p + 1, cudaMemcpyHostToDevice));
  DK_CUDA_TRY(cudaMemcpy(d_xo, x, (size_t)N * m->d * 8, cudaMemcpyHostToDevice));
  double* d_main = d_xo + (size_t)N * m->d;
  CritParams P{mode, type, minimization, proxy, plugin, qnorm_beta};
  k_predict_plain<<<(unsigned)((N + 127) / 128), 128>>>(m->dm, P, d_xo, N, d_T, d_zM, d_v, d_main, d_main + N, d_main + 2 * N);
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  DK_CUDA_TRY(cudaDeviceSynchronize());
  if (out_main) DK_CUDA_TRY(cudaMemcpy(out_main, d_main, N * 8, cudaMemcpyDeviceToHost));
  if (out_mean) DK_CUDA_TRY(cudaMemcpy(out_mean, d_main + N, N * 8, cudaMemcpyDeviceToHost));
  if (out_sd) DK_CUDA_TRY(cudaMemcpy(out_sd, d_main + 2 * N, N * 8, cudaMemcpyDeviceToHost));
  return DK_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// Allocation-free EI entry point.
//
// Some hosts of this library may not allocate device memory once they are running: a serving
// loop whose memory budget was fixed at start-up, code that runs under another framework's
// memory pool (a raw cudaMalloc there synchronises the whole device and can fail although the
// pool has room), or a step replayed at a steady cadence where an allocation stall is not
// acceptable.  dk_ei_dev() is fine for them except for one thing: its V panels (the solve's
// scratch) were cudaMalloc'ed by dk_model_create().  dk_ei_dev_noalloc() runs the same fused
// kernel with the panels placed in the fixed array below instead, so a handle is not even
// needed to own scratch and nothing is allocated or freed during the call.
//
// The array lives in every CUDA context that loads this library, from load until exit.  All
// calls (all handles, all streams) on one device share it and it keeps whatever the previous
// call left: the kernel writes every panel element of a tile before reading it, so nothing has
// to be cleared, but two calls that overlap on the GPU would overwrite each other's panels.
// ------------------------------------------------------------------------------------------
#define DK_NOALLOC_SCRATCH_DOUBLES ((size_t)DK_NOALLOC_SCRATCH_BYTES / 8)
static __device__ __align__(256) double g_noalloc_scratch[DK_NOALLOC_SCRATCH_DOUBLES];

namespace {
// g_noalloc_scratch is only a host-side stand-in for the symbol here; the kernel needs the device
// address, looked up once per device (with that device current) and kept.
double* g_noalloc_ptr[64] = {};

int noalloc_scratch_address(int device, double** out) {
  if (device < 0 || device >= 64) return dk::fail(DK_ERR_ARG, "device %d out of range for the fixed scratch table", device);
  if (!g_noalloc_ptr[device]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return dk::fail(DK_ERR_CUDA, "cudaGetSymbolAddress(fixed scratch) on device %d: %s", device, cudaGetErrorString(e));
    g_noalloc_ptr[device] = p;
  }
  *out = g_noalloc_ptr[device];
  return DK_OK;
}
}  // namespace

extern "C" int dk_ei_dev_noalloc(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization,
                                 int proxy, double* d_ei, double* d_mean, double* d_sd, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_ei) return fail(DK_ERR_ARG, "d_ei is NULL");
  const size_t panel = m->scratch_doubles_per_cta;  // n_pad x 64 doubles: one CTA's V panel
  if (panel > DK_NOALLOC_SCRATCH_DOUBLES)
    return fail(DK_ERR_ARG,
                "dk_ei_dev_noalloc: this model needs %zu bytes of scratch per CTA (n + p + 1 = %d rows padded to %d), "
                "more than the fixed %zu-byte array; use dk_ei_dev",
                panel * 8, m->dm.n_aug, m->dm.n_pad, (size_t)DK_NOALLOC_SCRATCH_BYTES);
  DeviceGuard guard(m->device);
  double* scratch = nullptr;
  if ((rc = noalloc_scratch_address(m->device, &scratch))) return rc;
  if (N == 0) return DK_OK;
  const int64_t fit = (int64_t)(DK_NOALLOC_SCRATCH_DOUBLES / panel);  // CTAs whose panels fit the array
  CritParams P{MODE_EI, type, minimization, proxy, plugin, 0.0};
  if (!use_narrow(m, N) && use_small(m)) {
    // small designs: the single-block kernel needs no scratch at all
    const int64_t nt = (N + SBN - 1) / SBN;
    k_predict_small<<<(int)std::min<int64_t>(nt, m->num_sms), NTS, m->small_lay.total, (cudaStream_t)stream>>>(m->dm, P, m->small_lay, d_x, N, nt, d_ei, d_mean, d_sd);
  } else if (use_narrow(m, N)) {
    // the same tile shape dk_ei_dev picks for this batch size, so the two entry points return the same numbers
    const int64_t nt = (N + BNN - 1) / BNN;
    const int grid = (int)std::min<int64_t>(std::min<int64_t>(nt, m->grid), fit);
    k_predict_narrow<<<grid, NT, SM_TOTAL + PIPE_SMEM, (cudaStream_t)stream>>>(m->dm, P, d_x, N, nt, scratch, panel, d_ei,
                                                                               d_mean, d_sd);
  } else {
    const int64_t ntiles = (N + BN - 1) / BN;
    const int grid = (int)std::min<int64_t>(std::min<int64_t>(ntiles, m->grid), fit);
    k_predict_fused<<<grid, NT, SM_TOTAL, (cudaStream_t)stream>>>(m->dm, P, d_x, N, ntiles, scratch, panel, d_ei, d_mean,
                                                                   d_sd);
  }
  m->launches++;
  DK_CUDA_TRY(cudaGetLastError());
  return DK_OK;
}

// ------------------------------------------------------------------------------------------
// sampleFromEI's candidate screen (R/sampleFromEI.R:109-125): EI-proportional sampling weights
// ------------------------------------------------------------------------------------------
extern "C" int dk_sample_weights_dev(dk_model* m, const double* d_x, int64_t N, double T, int minimization, double* d_w,
                                     void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_w) return fail(DK_ERR_ARG, "d_w is NULL");
  DeviceGuard guard(m->device);
  CritParams P{MODE_EI_WEIGHT, DK_UK, minimization, 0, T, 0.0};  // the reference hard-codes type = "UK" here
  return launch_predict(m, P, d_x, N, d_w, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int dk_sample_weights(dk_model* m, const double* x, int64_t N, double T, int minimization, double* out_w) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_w) return fail(DK_ERR_ARG, "out_w is NULL");
  CritParams P{MODE_EI_WEIGHT, DK_UK, minimization, 0, T, 0.0};  // the reference hard-codes type = "UK" here
  return host_predict(m, P, x, N, 0, N, out_w, nullptr, nullptr);
}

#ifdef DK_PHASE_TIMING
/* Debug build only (tools/phase_timing.py): copy out the per-phase cycle counters. */
extern "C" int dk_debug_phase_cycles(unsigned long long* out8) {
  DK_CUDA_TRY(cudaMemcpyFromSymbol(out8, dk::g_phase_cycles, 8 * sizeof(unsigned long long)));
  return DK_OK;
}
#endif

// ------------------------------------------------------------------------------------------
// AEI (R/AEI.R:105-146) and EQI (R/EQI.R:100-151): noisy-criteria epilogues on the same predict kernel
// ------------------------------------------------------------------------------------------
extern "C" int dk_aei_dev(dk_model* m, const double* d_x, int64_t N, double y_min, double new_noise_var, int type,
                          double* d_out, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  DeviceGuard guard(m->device);
  CritParams P{MODE_AEI, type, 1, 0, y_min, 0.0, new_noise_var, m->dm.sd2};
  return launch_predict(m, P, d_x, N, d_out, nullptr, nullptr, (cudaStream_t)stream);
}
extern "C" int dk_eqi_dev(dk_model* m, const double* d_x, int64_t N, double q_min, double new_noise_var, double qnorm_beta,
                          int type, double* d_out, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_out) return fail(DK_ERR_ARG, "d_out is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  DeviceGuard guard(m->device);
  CritParams P{MODE_EQI, type, 1, 0, q_min, qnorm_beta, new_noise_var, m->dm.sd2};
  return launch_predict(m, P, d_x, N, d_out, nullptr, nullptr, (cudaStream_t)stream);
}
extern "C" int dk_aei(dk_model* m, const double* x, int64_t N, double y_min, double new_noise_var, int type, double* out) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  CritParams P{MODE_AEI, type, 1, 0, y_min, 0.0, new_noise_var, m->dm.sd2};
  return host_predict(m, P, x, N, 0, N, out, nullptr, nullptr);
}
extern "C" int dk_eqi(dk_model* m, const double* x, int64_t N, double q_min, double new_noise_var, double qnorm_beta, int type,
                      double* out) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out) return fail(DK_ERR_ARG, "out is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  CritParams P{MODE_EQI, type, 1, 0, q_min, qnorm_beta, new_noise_var, m->dm.sd2};
  return host_predict(m, P, x, N, 0, N, out, nullptr, nullptr);
}

// AEI.grad (R/AEI.grad.R:65-154) and EQI.grad (R/EQI.grad.R:66-146): the gradient kernel with other outer derivatives
extern "C" int dk_aei_grad_dev(dk_model* m, const double* d_x, int64_t N, double y_min, double new_noise_var, int type,
                               double* d_grad, double* d_aei, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_grad) return fail(DK_ERR_ARG, "d_grad is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  DeviceGuard guard(m->device);
  CritParams P{MODE_AEI, type, 1, 0, y_min, 0.0, new_noise_var, m->dm.sd2};
  return launch_grad(m, P, d_x, N, d_grad, d_aei, (cudaStream_t)stream);
}
extern "C" int dk_eqi_grad_dev(dk_model* m, const double* d_x, int64_t N, double q_min, double new_noise_var,
                               double qnorm_beta, int type, double* d_grad, double* d_eqi, void* stream) {
  int rc = check_call(m, d_x, N);
  if (rc) return rc;
  if (!d_grad) return fail(DK_ERR_ARG, "d_grad is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  DeviceGuard guard(m->device);
  CritParams P{MODE_EQI, type, 1, 0, q_min, qnorm_beta, new_noise_var, m->dm.sd2};
  return launch_grad(m, P, d_x, N, d_grad, d_eqi, (cudaStream_t)stream);
}
extern "C" int dk_aei_grad(dk_model* m, const double* x, int64_t N, double y_min, double new_noise_var, int type,
                           double* out_grad) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  CritParams P{MODE_AEI, type, 1, 0, y_min, 0.0, new_noise_var, m->dm.sd2};
  return host_grad(m, P, x, N, 0, N, out_grad, nullptr);
}
extern "C" int dk_eqi_grad(dk_model* m, const double* x, int64_t N, double q_min, double new_noise_var, double qnorm_beta,
                           int type, double* out_grad) {
  int rc = check_call(m, x, N);
  if (rc) return rc;
  if (!out_grad) return fail(DK_ERR_ARG, "out_grad is NULL");
  if (!(new_noise_var >= 0)) return fail(DK_ERR_ARG, "new.noise.var must be >= 0");
  CritParams P{MODE_EQI, type, 1, 0, q_min, qnorm_beta, new_noise_var, m->dm.sd2};
  return host_grad(m, P, x, N, 0, N, out_grad, nullptr);
}
