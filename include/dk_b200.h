/*
 * dk_b200.h — C ABI of the B200-native batched acquisition engine for DiceOptim.
 *
 * This is the R-free layer (1) of SURVEY.md 8(b): plain pointers and sizes, int status
 * return (0 = ok), message via dk_last_error().  The R `.Call` shim (r-package/src/dk_rcall.c,
 * registered next to the reference's own table, /root/reference/src/packagename_init.c:12-21)
 * is a mechanical wrapper over these entry points; INTEGRATION.md shows the binding.
 *
 * Conventions (R's): all matrices are COLUMN-MAJOR doubles.  Candidates `x` are N x d
 * (coordinate k of candidate j at x[j + k*N]); gradients are N x d likewise.
 * Host entry points take HOST pointers (pageable memory is fine) and stream the candidates through the device in
 * chunks: the host<->device copies overlap the kernels and the device holds two chunks whatever N;
 * the *_dev entry points take DEVICE pointers on the model's device and enqueue on `stream`
 * (a cudaStream_t passed as void*; NULL = the default stream) without synchronising.
 * One model handle is bound to one CUDA device; calls on one handle must not overlap
 * (R is single-threaded; the handle owns per-call scratch).
 */
#ifndef DK_B200_H
#define DK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* covariance@name of a DiceKriging covTensorProduct */
enum { DK_GAUSS = 0, DK_EXP = 1, DK_MATERN3_2 = 2, DK_MATERN5_2 = 3, DK_POWEXP = 4 };
/* trend.formula: ~1 or ~. */
enum { DK_TREND_CONST = 0, DK_TREND_LINEAR = 1 };
/* predict type */
enum { DK_SK = 0, DK_UK = 1 };
/* criteria of dk_screen (the first three) and dk_polish (all): kriging.quantile is MINIMISED, the others MAXIMISED */
enum { DK_CRIT_EI = 0, DK_CRIT_QUANTILE = 1, DK_CRIT_AKG = 2, DK_CRIT_AEI = 3, DK_CRIT_EQI = 4, DK_CRIT_QEI = 5 };

enum {
  DK_OK = 0,
  DK_ERR_ARG = 1,        /* bad dimension / NULL / unsupported kernel, trend, q */
  DK_ERR_CUDA = 2,       /* CUDA runtime error (message has the string) */
  DK_ERR_NOMEM = 3,      /* host or device allocation failed */
  DK_ERR_UNSUPPORTED = 4 /* no usable sm_100 device */
};

#define DK_MAX_D 24   /* input dimension limit of the fused kernels */
#define DK_MAX_P 25   /* trend columns (1 for ~1, d+1 for ~.) */
#define DK_MAX_Q 4    /* qEI / qEI.grad batch size served by the deterministic CDF kernels */
#define DK_MAX_Q_LATTICE 10 /* qEI / qEI.grad batch size served by the fixed-lattice Genz CDF (5 .. 10 points) */

typedef struct dk_model dk_model;

/* The km S4 slots the hot path reads — replaces the reads at
 * /root/reference/R/EI.grad.R:118-122, R/AKG.R:110-113,124, R/qEI.grad.R:78-83,
 * R/kriging.quantile.grad.R:92-97, R/EI.R:84,86,109. */
typedef struct {
  int n, d, p;
  int kernel;              /* DK_GAUSS ... DK_POWEXP  (covariance@name) */
  int trend;               /* DK_TREND_CONST | DK_TREND_LINEAR (trend.formula) */
  double sd2;              /* covariance@sd2 */
  const double* theta;     /* covariance@range.val, d */
  const double* shape;     /* covariance@shape.val, d (powexp) or NULL */
  double nugget;           /* covariance@nugget */
  int nugget_flag;         /* covariance@nugget.flag */
  const double* noise_var; /* model@noise.var, n, or NULL */
  const double* T;         /* model@T, n x n, upper Cholesky factor (only j<=i of column i read) */
  const double* z;         /* model@z, n */
  const double* M;         /* model@M, n x p */
  const double* X;         /* model@X, n x d */
  const double* y;         /* model@y, n */
  const double* F;         /* model@F, n x p */
  const double* beta;      /* model@trend.coef, p */
  int device;              /* CUDA device ordinal the handle lives on */
} dk_model_desc;

/* Thread-local message of the last failing call. */
const char* dk_last_error(void);
/* Library / build identification string ("dk_b200 <ver> sm_100a"). */
const char* dk_version(void);

/* Upload the model, pre-tile the Cholesky factor for the DMMA solve, precompute the
 * candidate-independent arrays (alpha = T^-1 z, T^-1 M, (M'M)^-1, chol(M'M), block inverses). */
int dk_model_create(const dk_model_desc* desc, dk_model** out);
void dk_model_destroy(dk_model* m);
/* One more observation (x_new[d], y_new) with the covariance parameters and trend.coef unchanged — replaces the
 * rbind(X, x) / computeAuxVariables refresh of the constant-liar loop, /root/reference/R/max_qEI.R:305-316 (and update()
 * with cov.reestim = FALSE, R/fastEGO.nsteps.R:231-235), which re-factorises the whole model (n^3 / 3 flops) and, with a
 * device handle, re-tiles and re-uploads it.  Here the new Cholesky column is one forward substitution (n^2 flops on
 * the host), and on the device only the block rows of the tiled factor that hold the new row change (one or two of
 * n / 128), plus one design row and the O(n p) weight vectors.  new_noise_var is the new point's noise.var for models
 * that have one (the reference's loop uses 0).  Buffers grow by 256 rows at a time when the padded capacity runs out.
 * The pieces built lazily from the factor (backward tiles for gradients, AKG constants, K^-1) are rebuilt on next use.
 * Not to be called while another call on the handle is in flight.  dk_model_n: the handle's current design size. */
int dk_model_append(dk_model* m, const double* x_new, double y_new, double new_noise_var);
int dk_model_n(const dk_model* m);
/* 1 when the handle is a small design (n + p + 1 <= 256) whose large forward batches run the single-block kernel (the
 * explicit inverse of the whole factor applied as one triangular GEMM, no scratch).  dk_model_create enables it only when
 * a probe against long-double substitution shows that inverse loses no accuracy (ill-conditioned gauss designs do not
 * qualify and stay on the blocked kernels). */
int dk_model_small_kernel(const dk_model* m);
/* Device-memory bytes held by the handle (tiles + scratch). */
size_t dk_model_device_bytes(const dk_model* m);
/* Number of kernels launched through the handle so far (bench.py's gpu_launches). */
long long dk_model_launches(const dk_model* m);
/* Bracket the dominant (solve) kernel of every call with CUDA events on its launch stream (on = 0/1);
 * dk_model_last_kernel_ms returns the device time of the most recent completed one, < 0 if none. */
int dk_model_set_timing(dk_model* m, int on);
double dk_model_last_kernel_ms(dk_model* m);
/* For calls made of two kernels (dk_qei_dev: moments, then CDFs): device time of the FIRST kernel of the most recent
 * timed call, < 0 if that call had a single kernel. */
double dk_model_last_phase_ms(dk_model* m);

/* Pieces of a handle that are built on first use — the flipped factor's tiles and the second panels (every gradient,
 * AKG, covdist, crit_SUR_cst), the AKG constants, K^-1 for AKG.grad — allocate device memory, copy and (K^-1) launch on
 * the FIRST call of their family, so that first *_dev call is not "enqueue only".  dk_model_prepare builds them up front
 * (and synchronises), e.g. before a stream capture or a latency-sensitive loop.  what: OR of DK_PREP_*. */
enum { DK_PREP_GRADIENTS = 1, DK_PREP_AKG = 2, DK_PREP_AKG_GRAD = 4 };
int dk_model_prepare(dk_model* m, int what);

/* Tile shape of the forward / gradient kernels (EI, quantile, predict, AEI, EQI and their gradients).  AUTO (default):
 * 16-candidate tiles when the call would leave three quarters of the persistent grid without a 64-candidate tile
 * (N <= 4736 on a B200: the optimiser-sized batches of optim(fn = EI, gr = EI.grad), /root/reference/R/max_crit.R:118-121),
 * 64-candidate tiles otherwise.  Both solve the same systems in the same order; the row reductions group their terms
 * differently, so sd and gradients agree to the last bits, not bit for bit.  WIDE / NARROW force one shape. */
enum { DK_TILE_AUTO = 0, DK_TILE_WIDE = 1, DK_TILE_NARROW = 2 };
int dk_model_set_tile_mode(dk_model* m, int mode);

/* predict.km(light.return=TRUE): mean and sd — replaces the DiceKriging call at
 * /root/reference/R/EI.R:112-116 and R/sampleFromEI.R:109.  out_sd may be NULL. */
int dk_predict(dk_model* m, const double* x, int64_t N, int type, double* out_mean, double* out_sd);
int dk_predict_dev(dk_model* m, const double* d_x, int64_t N, int type, double* d_mean, double* d_sd,
                   void* stream);

/* EI(x, model, plugin, type, minimization, proxy) over an N x d matrix — replaces
 * /root/reference/R/EI.R:80-145.  out_mean / out_sd (the envir values kriging.mean before
 * the sign flip, kriging.sd) may be NULL. */
int dk_ei(dk_model* m, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
          double* out_ei, double* out_mean, double* out_sd);
int dk_ei_dev(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization,
              int proxy, double* d_ei, double* d_mean, double* d_sd, void* stream);

/* EI.grad for every row of x (the reference evaluates one row per call) — replaces
 * /root/reference/R/EI.grad.R:97-186.  out_grad is N x d; out_ei may be NULL. */
int dk_ei_grad(dk_model* m, const double* x, int64_t N, double plugin, int type, int minimization,
               int proxy, double* out_grad, double* out_ei);
int dk_ei_grad_dev(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization,
                   int proxy, double* d_grad, double* d_ei, void* stream);

/* kriging.quantile / kriging.quantile.grad — replaces /root/reference/R/kriging.quantile.R:92-115
 * and R/kriging.quantile.grad.R:83-138.  qnorm_beta = qnorm(beta) is computed by the caller
 * (R's qnorm).  out_grad (N x d) may be NULL. */
int dk_quantile(dk_model* m, const double* x, int64_t N, double qnorm_beta, int type, double* out_q,
                double* out_grad);
int dk_quantile_dev(dk_model* m, const double* d_x, int64_t N, double qnorm_beta, int type, double* d_q,
                    double* d_grad, void* stream);

/* The batched screen of /root/reference/R/max_crit.R:104-110: EI over N candidates and
 * which.max (first maximal index, NaN skipped).  Returns the k best (value desc, index asc);
 * out_idx are 0-based, -1 when fewer than k comparable values exist.  d_x is a DEVICE or HOST
 * pointer according to x_on_device. */
int dk_screen_ei(dk_model* m, const double* x, int x_on_device, int64_t N, double plugin, int type,
                 int minimization, int proxy, int k, int64_t* out_idx, double* out_val, void* stream);

/* The same screen for any criterion, over given candidates OR over candidates generated on the device (SURVEY 8(b)
 * "dk_screen(model, x_or_generator_spec, N, criterion, k, ...)"): the random search of R/max_crit.R:104 is
 * `lower + runif(d N) (upper - lower)`, which the device can draw itself, so the N x d matrix (800 MB at cfg4) is never
 * built on the host nor uploaded.  The generator is counter-based (Philox4x32-10): coordinate c of candidate j is a pure
 * function of (seed, j, c) (counter = (j lo, j hi, c / 2, 0), key = seed; 53-bit uniform u in [0, 1), value =
 * fma(u, upper - lower, lower)), so any GPU generates any range and the result does not depend on the device count.
 * It is NOT R's Mersenne-Twister stream: a caller that needs runif()'s exact draws passes x instead.
 * x == NULL selects the generator.  out_x (k x d column-major, generated candidates only, may be NULL) receives the
 * coordinates of the k winners. */
typedef struct {
  unsigned long long seed;
  const double* lower; /* d, NULL = 0 */
  const double* upper; /* d, NULL = 1 */
} dk_uniform_spec;
typedef struct {
  int criterion;        /* DK_CRIT_* */
  int type;             /* DK_SK | DK_UK */
  int minimization;     /* EI, qEI */
  int proxy;            /* EI */
  double plugin;        /* EI, qEI: plugin;  AEI: y.min;  EQI: q.min */
  double qnorm_beta;    /* kriging.quantile, EQI: qnorm(beta) */
  double new_noise_var; /* AKG, AEI, EQI */
} dk_crit_spec;
int dk_screen(dk_model* m, const double* x, int x_on_device, const dk_uniform_spec* gen, int64_t N, const dk_crit_spec* crit,
              int k, int64_t* out_idx, double* out_val, double* out_x, void* stream);

/* The polish step of the multistart maximisers ON THE DEVICE — replaces the restart-by-restart
 * optim(par, fn = EI, gr = EI.grad, method = "L-BFGS-B", lower, upper) of /root/reference/R/max_crit.R:114-122, the
 * optim(c(x), qEI, qEI.grad, ...) loop of R/max_qEI.R:204-222 and the BFGS stage of genoud in R/max_EI.R:207-219,
 * R/max_AKG.R, R/max_AEI.R, R/max_EQI.R, R/min_quantile.R.  All S starts advance together: one criterion launch per trip
 * evaluates every start's current trial point (value + gradient), the optimiser state (iterates, gradients, (s, y)
 * history, step lengths) stays in device memory, the host reads one int per trip (starts still running).
 * Projected L-BFGS with Armijo backtracking along the projection arc; same stopping rules as optim's L-BFGS-B (factr,
 * pgtol, maxit).  x0 / out_x: (S q) x d column-major, rows s q .. s q + q - 1 = start s (q = 1 except for DK_CRIT_QEI,
 * 2 <= q <= DK_MAX_Q_LATTICE); lower / upper: d (NULL = 0 / 1); out_val[S] = criterion at out_x; out_iters[S] may be NULL. */
typedef struct {
  int maxit;          /* iterations per start (optim's maxit; default 100) */
  int history;        /* (s, y) pairs kept, 1..16 (default 8; optim's lmm is 5) */
  double pgtol;       /* stop when the projected gradient's max-norm is <= pgtol (default 0, as optim) */
  double factr;       /* stop when the decrease is <= factr * eps * max(|f|, |f'|, 1) (default 1e7, as optim) */
  int max_backtracks; /* halvings per line search (default 25) */
} dk_lbfgs_options;
int dk_polish(dk_model* m, const dk_crit_spec* crit, int q, const double* x0, int S, const double* lower,
              const double* upper, const dk_lbfgs_options* opt, double* out_x, double* out_val, int* out_iters);

/* sampleFromEI's draw on the device — replaces /root/reference/R/sampleFromEI.R:109-135: the EI weights of the N rows of x
 * (dk_sample_weights), the reference's fix-up when fewer than n weights are positive (:126-133), then n_draw <= 64 rows
 * drawn WITHOUT replacement with probability proportional to the weights: out_idx[n_draw], 0-based, in draw order.
 * Successive sampling is realised as "the n largest keys log(u_i) / w_i" (Efraimidis & Spirakis), u_i from the
 * counter-based generator with `seed`: the same law as R's sample(prob =, replace = FALSE), not R's random stream. */
int dk_sample_from_ei(dk_model* m, const double* x, int64_t N, double T, int minimization, int n_draw,
                      unsigned long long seed, int64_t* out_idx);

/* The candidate screen of sampleFromEI — replaces /root/reference/R/sampleFromEI.R:109-125: the
 * EI-proportional sampling weight sn (u Phi(u) + phi(u)) of every row (type "UK" as in the reference,
 * NaN -> +-(T - mn), clipped at 0).  The weighted draw itself (R's sample()) stays on the host. */
int dk_sample_weights(dk_model* m, const double* x, int64_t N, double T, int minimization, double* out_w);
int dk_sample_weights_dev(dk_model* m, const double* d_x, int64_t N, double T, int minimization, double* d_w, void* stream);

/* AEI(x, model, new.noise.var, y.min, type) and EQI(x, model, new.noise.var, beta, q.min, type) for every row of x
 * — replace /root/reference/R/AEI.R:105-146 and R/EQI.R:100-151.  y_min / q_min must be given (the R defaults come
 * from one dk_predict over model@X: mk[which.min(mk + qnorm(0.75) sk)] and min(mk + qnorm(beta) sk)). */
int dk_aei(dk_model* m, const double* x, int64_t N, double y_min, double new_noise_var, int type, double* out);
int dk_aei_dev(dk_model* m, const double* d_x, int64_t N, double y_min, double new_noise_var, int type, double* d_out, void* stream);
int dk_eqi(dk_model* m, const double* x, int64_t N, double q_min, double new_noise_var, double qnorm_beta, int type, double* out);
int dk_eqi_dev(dk_model* m, const double* d_x, int64_t N, double q_min, double new_noise_var, double qnorm_beta, int type,
               double* d_out, void* stream);

/* AEI.grad / EQI.grad for every row of x — replace /root/reference/R/AEI.grad.R:65-154 and R/EQI.grad.R:66-146
 * (N x d outputs; same y_min / q_min conventions as dk_aei / dk_eqi). */
int dk_aei_grad(dk_model* m, const double* x, int64_t N, double y_min, double new_noise_var, int type, double* out_grad);
int dk_aei_grad_dev(dk_model* m, const double* d_x, int64_t N, double y_min, double new_noise_var, int type, double* d_grad,
                    double* d_aei, void* stream);
int dk_eqi_grad(dk_model* m, const double* x, int64_t N, double q_min, double new_noise_var, double qnorm_beta, int type,
                double* out_grad);
int dk_eqi_grad_dev(dk_model* m, const double* d_x, int64_t N, double q_min, double new_noise_var, double qnorm_beta, int type,
                    double* d_grad, double* d_eqi, void* stream);

/* AKG(x, model, new.noise.var, type) for every row of x (the reference evaluates one row per call and
 * re-predicts at the whole design each time) — replaces /root/reference/R/AKG.R:93-229.  Needs n + 1 <= 8192. */
int dk_akg(dk_model* m, const double* x, int64_t N, double new_noise_var, int type, double* out);
int dk_akg_dev(dk_model* m, const double* d_x, int64_t N, double new_noise_var, int type, double* d_out, void* stream);

/* AKG.grad(x, model, new.noise.var, type) for every row of x — replaces /root/reference/R/AKG.grad.R:65-281
 * (envir = NULL path; the reference's quirks are kept: factor 2 in mu.x.grad, m_min.grad never taken).
 * out_grad is N x d; out_akg (N, the AKG values from the same pass) may be NULL.  Builds K^-1 on the device at
 * the first call (n^2 doubles).  A candidate whose envelope has more than 256 lines gets NaN. */
int dk_akg_grad(dk_model* m, const double* x, int64_t N, double new_noise_var, int type, double* out_grad, double* out_akg);
int dk_akg_grad_dev(dk_model* m, const double* d_x, int64_t N, double new_noise_var, int type, double* d_grad,
                    double* d_akg, void* stream);

/* qEI(x, model, plugin, type, minimization) for B batches of q points — replaces
 * /root/reference/R/qEI.R:119-261 (incl. covZk / get_cik / get_sigmaik and the duplicate test of
 * R/qEI.R:124-128).  x is (B*q) x d column-major; rows b*q .. b*q+q-1 form batch b.  1 <= q <= DK_MAX_Q_LATTICE.
 * plugin must be given (R default: min(y), or max(y) when !minimization, R/qEI.R:129-130).
 * The q-variate normal CDFs (mnormt::pmnorm in the reference) are deterministic Genz bivariate /
 * Plackett trivariate and quadrivariate rules; the exact second term is returned for fastCompute
 * TRUE and FALSE alike. */
int dk_qei(dk_model* m, const double* x, int64_t B, int q, double plugin, int type, int minimization, double* out);
int dk_qei_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization,
               double* d_out, void* stream);
/* qEI.grad for B batches of q points (2 <= q <= DK_MAX_Q_LATTICE; q = 1 is dk_ei_grad) — replaces
 * /root/reference/R/qEI.grad.R:267-424 with krigingDeriv / decompo / GPhi / HPhi (:1-142).  out_grad is
 * (B*q) x d column-major, row b*q + l = gradient with respect to point l of batch b.  plugin must be given
 * (R default: min(y), or -max(y) when !minimization, R/qEI.grad.R:272-276,297).  Exact term B for both
 * values of fastCompute.  q > DK_MAX_Q: the CDFs of dimension >= 5 inside GPhi / HPhi / term B come from the fixed
 * lattice (one CTA per batch; 5160 of them at q = 10, about 0.15 s per batch). */
int dk_qei_grad(dk_model* m, const double* x, int64_t B, int q, double plugin, int type, int minimization, double* out_grad);
int dk_qei_grad_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization,
                    double* d_grad, void* stream);
/* qEI and qEI.grad of the same batches from ONE pass over the model — what qEI(x, envir) followed by qEI.grad(x, envir) saves in
 * the reference (R/qEI.R:190-198 stores the kriging mean / covariance and the CDF terms, R/qEI.grad.R:299-316 reads them back;
 * the optim loop of R/max_qEI.R:205-222 relies on it).  d_val[B], d_grad as dk_qei_grad_dev.  The value is taken at x itself,
 * as the gradient is (dk_qei rounds x to 8 digits first, R/qEI.R:126-128). */
int dk_qei_value_grad_dev(dk_model* m, const double* d_x, int64_t B, int q, double plugin, int type, int minimization,
                          double* d_val, double* d_grad, void* stream);
/* pmnorm(h_b, varcov = S_b) for B independent q-variate problems (q <= DK_MAX_Q_LATTICE): the device CDFs behind
 * dk_qei, exposed for parity tests against mnormt-class references.  h: B x q, S: B x q x q, row-major, host. */
int dk_mvn_cdf(int q, int64_t B, const double* h, const double* S, double* out);

/* EI over device-resident candidates WITHOUT any device allocation (no counterpart in the reference:
 * it exists for hosts that must not allocate once running, e.g. a serving loop with a memory budget fixed
 * at start-up, or code living under another framework's memory pool where a raw cudaMalloc stalls the
 * device and may fail).  Same arguments and results as dk_ei_dev; the solve's scratch (the per-CTA V
 * panels) sits in a fixed DK_NOALLOC_SCRATCH_BYTES array inside the library instead of the handle's
 * allocation.  Limits and rules:
 *  - the model must satisfy  128*ceil((n+p+1)/128) * 64 * 8 <= DK_NOALLOC_SCRATCH_BYTES, else DK_ERR_ARG;
 *    only as many CTAs run as have a panel in the array (n = 2000: 64 of 296), so large models run slower
 *    here than through dk_ei_dev;
 *  - every context that loads the library holds the array until it exits;
 *  - the first call on a device looks up the array's device address; make that first call before any
 *    stream capture begins;
 *  - all calls on a device share the array: calls that overlap on the GPU (two streams, two handles) race.
 *    Nothing needs clearing between calls (every panel element is written before it is read). */
#define DK_NOALLOC_SCRATCH_BYTES (64u << 20)
int dk_ei_dev_noalloc(dk_model* m, const double* d_x, int64_t N, double plugin, int type, int minimization,
                      int proxy, double* d_ei, double* d_mean, double* d_sd, void* stream);

/* ---- Constrained criteria over several km models on one device (objective + constraints) ----
 * dk_feas_dev replaces Feas, /root/reference/R/crit_EFI.R:231-238: pnorm(threshold, mean, sd) of every row, minus
 * pnorm(-threshold, mean, sd) when `equality`.  multiply != 0 multiplies the result into d_out instead of storing it,
 * so a product over constraint models needs no pass of its own. */
int dk_feas_dev(dk_model* m, const double* d_x, int64_t N, int type, int equality, double threshold, int multiply,
                double* d_out, void* stream);

/* Distances of checkPredict, /root/reference/R/checkPredict.R:35-66. */
enum { DK_DIST_EUCLIDEAN = 0, DK_DIST_COVDIST = 1, DK_DIST_COVRATIO = 2 };
/* One model's distance for every row; running_min != 0 keeps pmin(d_out, distance) (lists of models, :47, :64). */
int dk_check_distance_dev(dk_model* m, const double* d_x, int64_t N, int type, int distance, int running_min,
                          double* d_out, void* stream);
/* checkPredict(x, model = list(...), distance, type) for every row of x (host buffers): out_mindist[i] is the `mindist`
 * the reference compares with `threshold` (:74); "euclidean" looks at models[0] only, as the reference does. */
int dk_check_predict(dk_model* const* models, int n_models, const double* x, int64_t N, int type, int distance,
                     double* out_mindist);

/* crit_EFI's product (kind = DK_FI_EXPECTED; /root/reference/R/crit_EFI.R:158-191, without the checkPredict gate, which
 * is dk_check_predict) and prob.of.feasible.improvement (kind = DK_FI_PROBABILITY;
 * /root/reference/R/prob.of.feasible.improvement.R:39-79) for every row of x: objective model `fun`, n_cst constraint
 * models, plugin = current feasible best (+Inf: none known), equality[n_cst] / tol[n_cst] may be NULL (all FALSE / 0). */
enum { DK_FI_EXPECTED = 0, DK_FI_PROBABILITY = 1 };
int dk_feasible_improvement(dk_model* fun, dk_model* const* cst, int n_cst, const double* x, int64_t N, double plugin,
                            int type, int kind, const int* equality, const double* tol, double* out);

/* crit_SUR_cst for every row of x — replaces /root/reference/R/crit_SUR_cst.R:120-378 together with its helpers
 * precomputeUpdateData (R/precomputeUpdateData.R:1-22) and computeQuickKrigcov2 (R/computeQuickKrigcov2.R:22-33); the
 * checkPredict gate of :133-136 is dk_check_predict.  integration_points: Mi x d column-major (critcontrol$integration.points),
 * integration_weights: Mi or NULL (1 / Mi each), obj_min: best feasible observation (+Inf: none), n_cst in 1..3,
 * equality[n_cst] / tol[n_cst] may be NULL.  out[i] = mean((oldcrit - newcrit) * weights, na.rm = TRUE) of row i. */
int dk_crit_sur_cst(dk_model* fun, dk_model* const* cst, int n_cst, const double* x, int64_t N,
                    const double* integration_points, int64_t Mi, const double* integration_weights, double obj_min,
                    int type, const int* equality, const double* tol, double* out);

/* ---- One model on several GPUs of this process (SURVEY 8(b) "copies ... to every selected GPU", 8(e)) ----
 * R cannot fork CUDA contexts, so the reference's only parallelism (mclapply over restarts,
 * /root/reference/R/max_crit.R:125) is replaced by a split INSIDE the library: candidates (qEI: batches) are cut into
 * contiguous ranges, one per device (sizes differ by at most one), one host thread per device streams its range of the
 * caller's arrays through that device and writes the same range of the caller's outputs.  Results are bit-identical
 * to the single-device entry points (every candidate is evaluated by the same kernel; the top-k merge keeps the lower
 * index on ties, i.e. which.max's first maximum).  The model is tiled once on the host and copied device to device. */
typedef struct dk_multi dk_multi;
/* devices: ndev CUDA ordinals, or NULL for the first ndev visible devices (all of them when ndev <= 0);
 * desc->device is ignored.  An ordinal may appear more than once (the replicas then share that device: no speed-up,
 * but the split and merge run as they would on separate GPUs, which is how the single-GPU tests cover them). */
int dk_multi_create(const dk_model_desc* desc, const int* devices, int ndev, dk_multi** out);
void dk_multi_destroy(dk_multi* mm);
int dk_multi_count(const dk_multi* mm);
/* The single-device handle of replica i (owned by mm), e.g. for the *_dev entry points. */
dk_model* dk_multi_model(dk_multi* mm, int i);
/* dk_model_append on every replica. */
int dk_multi_append(dk_multi* mm, const double* x_new, double y_new, double new_noise_var);
/* A replica of an existing handle on another device (device-to-device copy of the tiled model). */
int dk_model_clone(const dk_model* src, int device, dk_model** out);
/* Host-pointer criteria, same arguments and results as dk_predict / dk_ei / dk_ei_grad / dk_quantile / dk_akg / dk_qei
 * (+ dk_qei_grad: out or out_grad may be NULL) / dk_screen_ei. */
int dk_multi_predict(dk_multi* mm, const double* x, int64_t N, int type, double* out_mean, double* out_sd);
int dk_multi_ei(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                double* out_ei, double* out_mean, double* out_sd);
int dk_multi_ei_grad(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                     double* out_grad, double* out_ei);
int dk_multi_quantile(dk_multi* mm, const double* x, int64_t N, double qnorm_beta, int type, double* out_q,
                      double* out_grad);
int dk_multi_akg(dk_multi* mm, const double* x, int64_t N, double new_noise_var, int type, double* out);
int dk_multi_qei(dk_multi* mm, const double* x, int64_t B, int q, double plugin, int type, int minimization,
                 double* out, double* out_grad);
int dk_multi_screen_ei(dk_multi* mm, const double* x, int64_t N, double plugin, int type, int minimization, int proxy,
                       int k, int64_t* out_idx, double* out_val);
int dk_multi_screen(dk_multi* mm, const double* x, const dk_uniform_spec* gen, int64_t N, const dk_crit_spec* crit, int k,
                    int64_t* out_idx, double* out_val, double* out_x);

#ifdef __cplusplus
}
#endif
#endif /* DK_B200_H */
