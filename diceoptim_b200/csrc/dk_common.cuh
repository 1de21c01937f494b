// dk_common.cuh — shared definitions of the dk_b200 engine (host + device).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>

#include "dk_b200.h"

namespace dk {

// ---------------------------------------------------------------------------------------
// Tile geometry of the blocked triangular solve (DESIGN.md "kernels").
//   block row  : BM = 128 rows of the augmented factor  L_aug = [[T', 0], [ [z M]', I ]]
//   candidates : BN = 64 columns per CTA tile
//   k-chunk    : KC = 16 (four DMMA m8n8k4 steps per pipeline stage)
// A and B tiles are stored in DMMA *fragment order* (a_tile_index / b_tile_index) both in
// HBM and in shared memory, so a stage is a straight 16-byte copy and every fragment read
// is one conflict-free LDS.128.
// ---------------------------------------------------------------------------------------
constexpr int BM = 128;
constexpr int BN = 64;
constexpr int KC = 16;
constexpr int NT = 256;          // threads per CTA: 8 warps as 4 (rows) x 2 (cols), warp tile 32 x 32
constexpr int STAGES = 4;   // 4 x 24 KB = region 0 exactly (the design-row staging is dead during the GEMM phase)
constexpr int A_TILE = BM * KC;  // doubles
constexpr int B_TILE = KC * BN;  // doubles
constexpr int CB_STRIDE = 72;    // row stride (doubles) of the covariance staging tile
constexpr int CHUNKS_PER_BLOCK = BM / KC;  // 8

// Shared-memory map (bytes).  Region 0 is time-multiplexed inside one block row:
//   cov-gen : [Cbuf 128 x 72 doubles = 73728][Xs 128 x d doubles <= 24576]
//   GEMM    : [3 stages x (A 16384 + B 8192) = 73728][unused]
//   diag    : [R 8 chunks x 8192 = 65536][Abuf 2 x 16384 = 32768]
constexpr int SM_REGION0 = 98304;
constexpr int SM_XS_OFF = CB_STRIDE * BM * 8;             // 73728
constexpr int SM_ABUF_OFF = CHUNKS_PER_BLOCK * B_TILE * 8;  // 65536
constexpr int SM_CAND_OFF = SM_REGION0;                   // xs[d][64]
constexpr int SM_CAND_BYTES = DK_MAX_D * BN * 8;          // 12288
constexpr int SM_S2_OFF = SM_CAND_OFF + SM_CAND_BYTES;    // s2acc[4][64]
constexpr int SM_S2_BYTES = 4 * BN * 8;                   // 2048
constexpr int SM_TOTAL = SM_S2_OFF + SM_S2_BYTES;         // 112640 (two CTAs per SM fit in 227 KB)

__host__ __device__ __forceinline__ int a_tile_index(int r, int k) {  // r < 128, k < 16
  const int g = r >> 3, mp = g >> 1, e = g & 1, rl = r & 7, k4 = k >> 2, kl = k & 3;
  return (((mp * 4 + k4) * 32) + (rl * 4 + kl)) * 2 + e;
}
__host__ __device__ __forceinline__ int b_tile_index(int k, int n) {  // k < 16, n < 64
  const int k4 = k >> 2, kl = k & 3, g = n >> 3, np = g >> 1, e = g & 1, nl = n & 7;
  return (((k4 * 4 + np) * 32) + (nl * 4 + kl)) * 2 + e;
}
// first tile of block row I in a tile array: rows I' < I hold 8 I' off-diagonal tiles + 8 inverse tiles
__host__ __device__ __forceinline__ size_t tile_base(int I) { return (size_t)4 * I * (I + 1); }

// Shared-memory map of k_predict_small (dk_small.cuh), sized for the model by small_layout()
struct SmallLayout {
  int off_aug, off_cand, off_xs, off_rbuf, off_ring;   // bytes; barriers at 0, s2acc at 256
  int stages;                                          // ring depth (2 .. 8); 0: the design does not fit
  int tps;                                             // factor tiles (k-chunks) per ring stage, 1 .. 4
  int total;                                           // bytes of dynamic shared memory
};

// Model constants handed to kernels by value (constant bank).
struct DevModel {
  int n, d, p, n_aug, nblk, n_pad;
  int nblk_b;               // block rows of the backward (flipped T) system: ceil(n / 128)
  int kernel, trend, nugget_flag;
  double sd2, nugget, total_sd2;
  const double* tiles_fwd;  // forward tiles: -L_aug off-diagonal, block inverses on the diagonal
  const double* tiles_bwd;  // backward (flipped T) tiles, same format; NULL until a gradient call
  const double* tiles_small; // n_pad <= 256 only: the explicit inverse of the whole padded factor as A tiles (dk_small.cuh)
  const double* Xrm;        // [n_pad][d] row-major design (pad rows zero)
  const double* Xrm_s;      // the same rows scaled by w_k (cov-gen operand)
  const double* Xrm_r8;     // the same rows rounded to 8 digits, rint(v 1e8) / 1e8 (qEI's duplicate rule)
  const double* beta;       // p
  const double* RM;         // chol(M'M) upper, p x p column-major
  const double* tuuinv;     // (M'M)^-1, p x p
  const double* alpha;      // T^-1 z, n
  const double* KinvF;      // T^-1 M, n x p column-major
  const double* Mcm;        // model@M, n x p column-major
  const double* noise_var;  // n or NULL
  // AKG constants (built lazily): design lines sorted by a_i = -mk.X_i
  const double* akg_a_sorted;  // n, ascending
  const double* akg_delta;     // n: noise.var (0 for nugget / noise-free models)
  const int* akg_perm;         // n: design row with the j-th smallest a
  double akg_min_mkX;          // min_i mk.X_i
  double w[DK_MAX_D];       // per-dimension scale of |h|: 1/theta (gauss, exp, powexp), sqrt3/theta, sqrt5/theta
  double iw[DK_MAX_D];      // 1 / w_k
  double shape[DK_MAX_D];   // powexp exponents
  double theta[DK_MAX_D];
};

enum Mode { MODE_PREDICT = 0, MODE_EI = 1, MODE_QUANTILE = 2, MODE_EI_WEIGHT = 3, MODE_AEI = 4, MODE_EQI = 5,
            MODE_FEAS = 6, MODE_COVRATIO = 7 };
enum Combine { COMBINE_STORE = 0, COMBINE_MUL = 1, COMBINE_MIN = 2 };

struct CritParams {
  int mode, type, minimization, proxy;
  double plugin, qnorm_beta;
  double tau2;  // new.noise.var (AEI, EQI)
  double sd2;   // covariance@sd2 for the negligible-sd guards of AEI / EQI
  int combine = COMBINE_STORE;  // how the criterion meets what out[] already holds (products / minima over several models)
};

// ---------------------------------------------------------------------------------------
// error channel
// ---------------------------------------------------------------------------------------
std::string& last_error();
int fail(int code, const char* fmt, ...);

#define DK_CUDA_TRY(expr)                                                                         \
  do {                                                                                            \
    cudaError_t e__ = (expr);                                                                     \
    if (e__ != cudaSuccess)                                                                       \
      return dk::fail(DK_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

}  // namespace dk

// The opaque handle.
struct dk_model {
  int device = 0;
  int num_sms = 0;
  dk::DevModel dm{};
  // device allocations owned by the handle
  double* d_tiles_fwd = nullptr;
  double* d_tiles_bwd = nullptr;
  double* d_tiles_small = nullptr;   // small designs: inverse of the whole factor (dk_small.cuh)
  size_t tiles_small_doubles = 0;
  dk::SmallLayout small_lay{};           // its shared-memory map (valid while dm.tiles_small != nullptr)
  double* d_Xrm = nullptr;
  double* d_small = nullptr;     // beta | RM | tuuinv | alpha | KinvF | noise_var
  size_t tiles_fwd_doubles = 0, xrm_doubles = 0, small_doubles = 0, bwd_tile_doubles = 0;  // sizes of the three uploads (dk_model_clone)
  double* d_vscratch = nullptr;  // per-CTA V panels (forward solve), grid x n_pad x 64 doubles
  double* d_wscratch = nullptr;  // per-CTA W panels (backward solve)
  double* d_cscratch = nullptr;  // per-CTA covariance panels kept for the gradient kernels
  double* d_aux = nullptr;       // per-CTA Gram accumulators of the qEI kernel
  double* d_mom = nullptr;       // per-batch (mu, Sigma, mask) of the qEI kernels, grow-only
  size_t mom_doubles = 0;
  double* d_deriv = nullptr;     // per-batch kriging derivatives of the qEI.grad kernels, grow-only
  size_t deriv_doubles = 0;
  size_t scratch_doubles_per_cta = 0;
  int grid = 0;
  // grow-only staging for the host entry points
  double* d_io = nullptr;
  size_t io_doubles = 0;
  // chunked host<->device pipeline of the host entry points (run_chunked): two slots of pinned staging and of device
  // chunk buffers, one stream per direction
  double* h_stage = nullptr;     // pinned, 2 x (in + out)
  size_t stage_in_doubles = 0, stage_out_doubles = 0;
  double* d_chunk = nullptr;     // device, 2 x (in + out)
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_kern[2] = {nullptr, nullptr}, ev_d2h[2] = {nullptr, nullptr};
  double* d_vec = nullptr;       // device vector of one criterion value per candidate (screens), grow-only
  size_t vec_doubles = 0;
  double* d_gen = nullptr;       // one chunk of generated candidates (dk_screen with a generator), grow-only
  size_t gen_doubles = 0;
  void* d_red = nullptr;         // argmax partials
  size_t red_bytes = 0;
  cudaStream_t stream = nullptr;
  size_t device_bytes = 0;
  // host copies kept for lazily built pieces
  int n = 0, d = 0, p = 0;
  // Host copies use the leading dimension `ld` (= n after dk_model_create, the capacity after dk_model_append grew it).
  int ld = 0;                    // leading dimension of h_T, h_zM, h_aK; rows held by h_y, h_nv, h_X
  bool bwd_valid = false;        // d_tiles_bwd matches the current factor (dk_model_append invalidates it)
  double* h_X = nullptr;         // design, ld x d row-major
  double* h_aK = nullptr;        // [alpha | T^-1 M], ld x (p+1) column-major
  double h_G[DK_MAX_P * DK_MAX_P] = {};   // M'M, p x p
  double h_beta[DK_MAX_P] = {};
  double* h_T = nullptr;         // ld x ld column-major copy of the upper factor (for the backward tiles)
  double* h_zM = nullptr;        // [z | M], ld x (p+1) column-major
  double* h_y = nullptr;         // n
  double* h_nv = nullptr;        // n noise.var or NULL
  double* d_akg_const = nullptr; // a_sorted | delta
  int* d_akg_perm = nullptr;
  double* d_akg_scratch = nullptr;
  double* d_Kinv = nullptr;        // K^-1, n x n (AKG.grad only, built lazily on the device)
  double* d_akg_gscratch = nullptr; // per-CTA blocks of the AKG.grad contraction passes
  double* d_delta = nullptr;        // n: noise.var or zeros (checkPredict covdist when AKG never ran)
  double* d_knyy[2] = {nullptr, nullptr};  // sd(X_j)^2 per type (SK, UK), checkPredict covdist
  int tile_mode = 0;             // DK_TILE_AUTO / DK_TILE_WIDE / DK_TILE_NARROW (dk_model_set_tile_mode)
  long long launches = 0;        // kernels launched through this handle
  int timing = 0;                // bracket the solve kernel with events
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_mid = nullptr;  // ev_mid: between the two kernels of a two-kernel call
  bool ev_valid = false, ev_mid_valid = false;
};
