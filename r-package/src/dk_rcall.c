/* dk_rcall.c — the R `.Call` shim over include/dk_b200.h (layer (2) of SURVEY.md 8(b)).
 *
 * Mechanical wrappers: unwrap REAL()/INTEGER() pointers (R matrices are column-major already, no
 * copies), allocate results with allocVector, hold the model as an external pointer whose finalizer
 * calls dk_model_destroy.  Errors are raised with Rf_error AFTER unprotecting, so the callers' try()
 * chains (e.g. /root/reference/R/fastEGO.nsteps.R:137) keep working.
 *
 * R and its headers are not installed in the build image: this file is compiled only where
 * DK_HAVE_R is defined (R CMD SHLIB adds it through Makevars) and is not exercised by tests/.
 */
#ifdef DK_HAVE_R
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <stdio.h>
#include <string.h>

#include "dk_b200.h"

/* The R-side handle is ALWAYS a multi-device handle, on one or several devices: R is a single process that cannot fork CUDA
 * contexts (the reference's mclapply, /root/reference/R/max_crit.R:125), so the split over GPUs happens inside the
 * library.  multi_of() gives the multi-GPU handle, handle_of() replica 0 for the single-device-only entry points. */
static void dkR_finalize(SEXP ptr) {
  dk_multi* mm = (dk_multi*)R_ExternalPtrAddr(ptr);
  if (mm) dk_multi_destroy(mm);
  R_ClearExternalPtr(ptr);
}
static dk_multi* multi_of(SEXP h) {
  dk_multi* mm = (dk_multi*)R_ExternalPtrAddr(h);
  if (!mm) Rf_error("dk_b200: model handle is NULL (released or restored from a saved workspace)");
  return mm;
}
static dk_model* handle_of(SEXP h) { return dk_multi_model(multi_of(h), 0); }
static SEXP el(SEXP list, const char* name) {
  SEXP names = Rf_getAttrib(list, R_NamesSymbol);
  for (R_xlen_t i = 0; i < Rf_xlength(list); i++)
    if (strcmp(CHAR(STRING_ELT(names, i)), name) == 0) return VECTOR_ELT(list, i);
  return R_NilValue;
}
static const double* opt_real(SEXP x) { return (x == R_NilValue || Rf_xlength(x) == 0) ? NULL : REAL(x); }

/* slots: list(T, z, M, X, y, F, trend.coef, range.val, shape.val, noise.var, n, d, p, kernel, trend, sd2, nugget, nugget.flag) */
SEXP dkR_model_create(SEXP slots, SEXP devices) {
  dk_model_desc D;
  memset(&D, 0, sizeof D);
  D.n = Rf_asInteger(el(slots, "n")); D.d = Rf_asInteger(el(slots, "d")); D.p = Rf_asInteger(el(slots, "p"));
  D.kernel = Rf_asInteger(el(slots, "kernel")); D.trend = Rf_asInteger(el(slots, "trend"));
  D.sd2 = Rf_asReal(el(slots, "sd2")); D.nugget = Rf_asReal(el(slots, "nugget"));
  D.nugget_flag = Rf_asLogical(el(slots, "nugget.flag"));
  D.theta = REAL(el(slots, "range.val")); D.shape = opt_real(el(slots, "shape.val"));
  D.noise_var = opt_real(el(slots, "noise.var"));
  D.T = REAL(el(slots, "T")); D.z = REAL(el(slots, "z")); D.M = REAL(el(slots, "M")); D.X = REAL(el(slots, "X"));
  D.y = REAL(el(slots, "y")); D.F = REAL(el(slots, "F")); D.beta = REAL(el(slots, "trend.coef"));
  /* devices: integer vector of CUDA ordinals; a single NA / negative value = every visible device */
  int ndev = (int)Rf_xlength(devices);
  const int* dev = INTEGER(devices);
  int all = (ndev == 1 && (dev[0] == NA_INTEGER || dev[0] < 0));
  dk_multi* m = NULL;
  if (dk_multi_create(&D, all ? NULL : dev, all ? 0 : ndev, &m)) Rf_error("dk_multi_create: %s", dk_last_error());
  SEXP ptr = PROTECT(R_MakeExternalPtr(m, R_NilValue, R_NilValue));
  R_RegisterCFinalizerEx(ptr, dkR_finalize, TRUE);
  UNPROTECT(1);
  return ptr;
}

SEXP dkR_ei(SEXP h, SEXP x, SEXP plugin, SEXP type, SEXP minimization, SEXP proxy, SEXP want_envir) {
  dk_multi* m = multi_of(h);
  R_xlen_t N = Rf_nrows(x);
  int env = Rf_asLogical(want_envir);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 3));
  SEXP ei = PROTECT(Rf_allocVector(REALSXP, N));
  SEXP mean = PROTECT(Rf_allocVector(REALSXP, env ? N : 0));
  SEXP sd = PROTECT(Rf_allocVector(REALSXP, env ? N : 0));
  int rc = dk_multi_ei(m, REAL(x), (int64_t)N, Rf_asReal(plugin), Rf_asInteger(type), Rf_asLogical(minimization),
                 Rf_asLogical(proxy), REAL(ei), env ? REAL(mean) : NULL, env ? REAL(sd) : NULL);
  SET_VECTOR_ELT(out, 0, ei); SET_VECTOR_ELT(out, 1, mean); SET_VECTOR_ELT(out, 2, sd);
  UNPROTECT(4);
  if (rc) Rf_error("dk_ei: %s", dk_last_error());
  return out;
}

SEXP dkR_ei_grad(SEXP h, SEXP x, SEXP plugin, SEXP type, SEXP minimization, SEXP proxy) {
  dk_multi* m = multi_of(h);
  R_xlen_t N = Rf_nrows(x), d = Rf_ncols(x);
  SEXP g = PROTECT(Rf_allocMatrix(REALSXP, N, d));
  int rc = dk_multi_ei_grad(m, REAL(x), (int64_t)N, Rf_asReal(plugin), Rf_asInteger(type), Rf_asLogical(minimization),
                      Rf_asLogical(proxy), REAL(g), NULL);
  UNPROTECT(1);
  if (rc) Rf_error("dk_ei_grad: %s", dk_last_error());
  return g;
}

SEXP dkR_quantile(SEXP h, SEXP x, SEXP qnorm_beta, SEXP type, SEXP want_grad) {
  dk_multi* m = multi_of(h);
  R_xlen_t N = Rf_nrows(x), d = Rf_ncols(x);
  int wg = Rf_asLogical(want_grad);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));
  SEXP q = PROTECT(Rf_allocVector(REALSXP, N));
  SEXP g = PROTECT(wg ? Rf_allocMatrix(REALSXP, N, d) : Rf_allocVector(REALSXP, 0));
  int rc = dk_multi_quantile(m, REAL(x), (int64_t)N, Rf_asReal(qnorm_beta), Rf_asInteger(type), REAL(q), wg ? REAL(g) : NULL);
  SET_VECTOR_ELT(out, 0, q); SET_VECTOR_ELT(out, 1, g);
  UNPROTECT(3);
  if (rc) Rf_error("dk_quantile: %s", dk_last_error());
  return out;
}

SEXP dkR_akg(SEXP h, SEXP x, SEXP new_noise_var, SEXP type) {
  dk_multi* m = multi_of(h);
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_multi_akg(m, REAL(x), (int64_t)N, Rf_asReal(new_noise_var), Rf_asInteger(type), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_akg: %s", dk_last_error());
  return out;
}

/* x: (B*q) x d matrix, rows b*q .. b*q+q-1 = batch b */
SEXP dkR_qei(SEXP h, SEXP x, SEXP q, SEXP plugin, SEXP type, SEXP minimization, SEXP want_grad) {
  dk_multi* m = multi_of(h);
  int qq = Rf_asInteger(q), wg = Rf_asLogical(want_grad);
  R_xlen_t rows = Rf_nrows(x), d = Rf_ncols(x), B = rows / qq;
  SEXP out = PROTECT(wg ? Rf_allocMatrix(REALSXP, rows, d) : Rf_allocVector(REALSXP, B));
  int rc = dk_multi_qei(m, REAL(x), (int64_t)B, qq, Rf_asReal(plugin), Rf_asInteger(type), Rf_asLogical(minimization),
                        wg ? NULL : REAL(out), wg ? REAL(out) : NULL);
  UNPROTECT(1);
  if (rc) Rf_error("dk_qei: %s", dk_last_error());
  return out;
}

SEXP dkR_screen_ei(SEXP h, SEXP x, SEXP plugin, SEXP type, SEXP minimization, SEXP proxy, SEXP k) {
  dk_multi* m = multi_of(h);
  int kk = Rf_asInteger(k);
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));
  SEXP idx = PROTECT(Rf_allocVector(REALSXP, kk));
  SEXP val = PROTECT(Rf_allocVector(REALSXP, kk));
  int64_t* tmp = (int64_t*)R_alloc(kk, sizeof(int64_t));
  int rc = dk_multi_screen_ei(m, REAL(x), (int64_t)N, Rf_asReal(plugin), Rf_asInteger(type), Rf_asLogical(minimization),
                              Rf_asLogical(proxy), kk, tmp, REAL(val));
  for (int i = 0; i < kk; i++) REAL(idx)[i] = tmp[i] < 0 ? NA_REAL : (double)(tmp[i] + 1); /* 1-based like which.max */
  SET_VECTOR_ELT(out, 0, idx); SET_VECTOR_ELT(out, 1, val);
  UNPROTECT(3);
  if (rc) Rf_error("dk_screen_ei: %s", dk_last_error());
  return out;
}

SEXP dkR_sample_weights(SEXP h, SEXP x, SEXP T, SEXP minimization) {
  dk_model* m = handle_of(h);
  R_xlen_t N = Rf_nrows(x);
  SEXP w = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_sample_weights(m, REAL(x), (int64_t)N, Rf_asReal(T), Rf_asLogical(minimization), REAL(w));
  UNPROTECT(1);
  if (rc) Rf_error("dk_sample_weights: %s", dk_last_error());
  return w;
}

/* handles: list of external pointers (one per km model, all on one device) */
static dk_model** handles_of(SEXP lst, int* n) {
  *n = (int)Rf_xlength(lst);
  dk_model** hs = (dk_model**)R_alloc(*n > 0 ? *n : 1, sizeof(dk_model*));
  for (int i = 0; i < *n; i++) hs[i] = handle_of(VECTOR_ELT(lst, i));
  return hs;
}

/* checkPredict's mindist for every row of x (R/checkPredict.R:35-66); distance: 0 euclidean, 1 covdist, 2 covratio */
SEXP dkR_check_predict(SEXP handles, SEXP x, SEXP type, SEXP distance) {
  int n = 0;
  dk_model** hs = handles_of(handles, &n);
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_check_predict(hs, n, REAL(x), (int64_t)N, Rf_asInteger(type), Rf_asInteger(distance), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_check_predict: %s", dk_last_error());
  return out;
}

/* kind 0: EI x prod Feas (R/crit_EFI.R:158-191); kind 1: prob.of.feasible.improvement */
SEXP dkR_feasible_improvement(SEXP hfun, SEXP hcst, SEXP x, SEXP plugin, SEXP type, SEXP kind, SEXP equality, SEXP tol) {
  int n = 0;
  dk_model** hs = handles_of(hcst, &n);
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_feasible_improvement(handle_of(hfun), hs, n, REAL(x), (int64_t)N, Rf_asReal(plugin), Rf_asInteger(type),
                                   Rf_asInteger(kind), LOGICAL(equality), opt_real(tol), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_feasible_improvement: %s", dk_last_error());
  return out;
}

/* crit_SUR_cst for every row of x (R/crit_SUR_cst.R:120-378); pts: M x d matrix, weights: numeric(M) or NULL */
SEXP dkR_crit_sur_cst(SEXP hfun, SEXP hcst, SEXP x, SEXP pts, SEXP weights, SEXP obj_min, SEXP type, SEXP equality, SEXP tol) {
  int n = 0;
  dk_model** hs = handles_of(hcst, &n);
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_crit_sur_cst(handle_of(hfun), hs, n, REAL(x), (int64_t)N, REAL(pts), (int64_t)Rf_nrows(pts), opt_real(weights),
                           Rf_asReal(obj_min), Rf_asInteger(type), LOGICAL(equality), opt_real(tol), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_crit_sur_cst: %s", dk_last_error());
  return out;
}
/* Explicit release (the cache in gpu_engine.R evicts with it); the finalizer then finds a cleared pointer. */
SEXP dkR_release(SEXP h) {
  dkR_finalize(h);
  return R_NilValue;
}

/* Cache key of a km object: FNV-1a over the bytes of every array that, with the scalars, determines the model
 * (X, y, z, trend.coef, range.val, shape.val, noise.var, diag(T)) and over the scalars themselves.  O(n d), not
 * O(n^2): T itself is a function of X and the covariance parameters. */
static unsigned long long fnv(unsigned long long hsh, const void* p, size_t bytes) {
  const unsigned char* c = (const unsigned char*)p;
  for (size_t i = 0; i < bytes; i++) { hsh ^= c[i]; hsh *= 1099511628211ULL; }
  return hsh;
}
SEXP dkR_model_key(SEXP slots) {
  unsigned long long hsh = 1469598103934665603ULL;
  const char* arrays[] = {"X", "y", "z", "M", "trend.coef", "range.val", "shape.val", "noise.var"};
  for (int i = 0; i < 8; i++) {
    SEXP a = el(slots, arrays[i]);
    R_xlen_t len = a == R_NilValue ? 0 : Rf_xlength(a);
    hsh = fnv(hsh, &len, sizeof len);
    if (len) hsh = fnv(hsh, REAL(a), (size_t)len * sizeof(double));
  }
  SEXP T = el(slots, "T");
  int n = Rf_asInteger(el(slots, "n"));
  for (int i = 0; i < n; i++) hsh = fnv(hsh, REAL(T) + (size_t)i * n + i, sizeof(double));
  const char* scalars[] = {"n", "d", "p", "kernel", "trend", "sd2", "nugget", "nugget.flag"};
  for (int i = 0; i < 8; i++) { double v = Rf_asReal(el(slots, scalars[i])); hsh = fnv(hsh, &v, sizeof v); }
  char buf[32];
  snprintf(buf, sizeof buf, "%016llx", hsh);
  return Rf_mkString(buf);
}

/* predict.km(light.return = TRUE): list(mean, sd) */
SEXP dkR_predict(SEXP h, SEXP x, SEXP type) {
  R_xlen_t N = Rf_nrows(x);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));
  SEXP mean = PROTECT(Rf_allocVector(REALSXP, N));
  SEXP sd = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_multi_predict(multi_of(h), REAL(x), (int64_t)N, Rf_asInteger(type), REAL(mean), REAL(sd));
  SET_VECTOR_ELT(out, 0, mean); SET_VECTOR_ELT(out, 1, sd);
  UNPROTECT(3);
  if (rc) Rf_error("dk_predict: %s", dk_last_error());
  return out;
}

/* AKG.grad for every row of x: list(grad N x d, akg N) */
SEXP dkR_akg_grad(SEXP h, SEXP x, SEXP new_noise_var, SEXP type) {
  R_xlen_t N = Rf_nrows(x), d = Rf_ncols(x);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 2));
  SEXP g = PROTECT(Rf_allocMatrix(REALSXP, N, d));
  SEXP v = PROTECT(Rf_allocVector(REALSXP, N));
  int rc = dk_akg_grad(handle_of(h), REAL(x), (int64_t)N, Rf_asReal(new_noise_var), Rf_asInteger(type), REAL(g), REAL(v));
  SET_VECTOR_ELT(out, 0, g); SET_VECTOR_ELT(out, 1, v);
  UNPROTECT(3);
  if (rc) Rf_error("dk_akg_grad: %s", dk_last_error());
  return out;
}

/* AEI / AEI.grad (want_grad) for every row of x; ref = y.min */
SEXP dkR_aei(SEXP h, SEXP x, SEXP y_min, SEXP new_noise_var, SEXP type, SEXP want_grad) {
  R_xlen_t N = Rf_nrows(x), d = Rf_ncols(x);
  int wg = Rf_asLogical(want_grad);
  SEXP out = PROTECT(wg ? Rf_allocMatrix(REALSXP, N, d) : Rf_allocVector(REALSXP, N));
  int rc = wg ? dk_aei_grad(handle_of(h), REAL(x), (int64_t)N, Rf_asReal(y_min), Rf_asReal(new_noise_var), Rf_asInteger(type), REAL(out))
              : dk_aei(handle_of(h), REAL(x), (int64_t)N, Rf_asReal(y_min), Rf_asReal(new_noise_var), Rf_asInteger(type), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_aei: %s", dk_last_error());
  return out;
}

/* EQI / EQI.grad (want_grad) for every row of x; ref = q.min, qnorm_beta = qnorm(beta) */
SEXP dkR_eqi(SEXP h, SEXP x, SEXP q_min, SEXP new_noise_var, SEXP qnorm_beta, SEXP type, SEXP want_grad) {
  R_xlen_t N = Rf_nrows(x), d = Rf_ncols(x);
  int wg = Rf_asLogical(want_grad);
  SEXP out = PROTECT(wg ? Rf_allocMatrix(REALSXP, N, d) : Rf_allocVector(REALSXP, N));
  int rc = wg ? dk_eqi_grad(handle_of(h), REAL(x), (int64_t)N, Rf_asReal(q_min), Rf_asReal(new_noise_var), Rf_asReal(qnorm_beta),
                            Rf_asInteger(type), REAL(out))
              : dk_eqi(handle_of(h), REAL(x), (int64_t)N, Rf_asReal(q_min), Rf_asReal(new_noise_var), Rf_asReal(qnorm_beta),
                       Rf_asInteger(type), REAL(out));
  UNPROTECT(1);
  if (rc) Rf_error("dk_eqi: %s", dk_last_error());
  return out;
}
/* ---- round-2 entry points: on-device candidates, on-device polish, on-device draw, in-place model growth ---- */

static void fill_crit(dk_crit_spec* c, SEXP crit) {
  /* crit: list(criterion, type, minimization, proxy, plugin, qnorm.beta, new.noise.var) */
  c->criterion = Rf_asInteger(el(crit, "criterion"));
  c->type = Rf_asInteger(el(crit, "type"));
  c->minimization = Rf_asLogical(el(crit, "minimization"));
  c->proxy = Rf_asLogical(el(crit, "proxy"));
  c->plugin = Rf_asReal(el(crit, "plugin"));
  c->qnorm_beta = Rf_asReal(el(crit, "qnorm.beta"));
  c->new_noise_var = Rf_asReal(el(crit, "new.noise.var"));
}

/* The random search of max_crit (R/max_crit.R:104-107) with the candidates drawn ON the device: N uniform points in
 * [lower, upper] from the counter-based generator seeded by `seed` (two integers: low and high 32 bits), criterion `crit`,
 * the k best: list(index (1-based), value, x (k x d)). */
SEXP dkR_screen_generated(SEXP h, SEXP N, SEXP lower, SEXP upper, SEXP seed, SEXP crit, SEXP k) {
  dk_multi* m = multi_of(h);
  int kk = Rf_asInteger(k), d = (int)Rf_xlength(lower);
  dk_crit_spec c;
  fill_crit(&c, crit);
  dk_uniform_spec g;
  g.seed = ((unsigned long long)(unsigned)INTEGER(seed)[1] << 32) | (unsigned)INTEGER(seed)[0];
  g.lower = REAL(lower); g.upper = REAL(upper);
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 3));
  SEXP idx = PROTECT(Rf_allocVector(REALSXP, kk));
  SEXP val = PROTECT(Rf_allocVector(REALSXP, kk));
  SEXP xs = PROTECT(Rf_allocMatrix(REALSXP, kk, d));
  int64_t* tmp = (int64_t*)R_alloc(kk, sizeof(int64_t));
  int rc = dk_multi_screen(m, NULL, &g, (int64_t)Rf_asReal(N), &c, kk, tmp, REAL(val), REAL(xs));
  for (int i = 0; i < kk; i++) REAL(idx)[i] = tmp[i] < 0 ? NA_REAL : (double)(tmp[i] + 1);
  SET_VECTOR_ELT(out, 0, idx); SET_VECTOR_ELT(out, 1, val); SET_VECTOR_ELT(out, 2, xs);
  UNPROTECT(4);
  if (rc) Rf_error("dk_screen: %s", dk_last_error());
  return out;
}

/* The L-BFGS-B polish of all starts together on the device (replaces the optim() calls of R/max_crit.R:114-122 and
 * R/max_qEI.R:204-222): x0 is (S q) x d, rows s q .. s q + q - 1 = start s; list(par, value, iterations). */
SEXP dkR_polish(SEXP h, SEXP crit, SEXP q, SEXP x0, SEXP lower, SEXP upper, SEXP maxit) {
  dk_model* m = handle_of(h);
  dk_crit_spec c;
  fill_crit(&c, crit);
  int qq = Rf_asInteger(q), rows = Rf_nrows(x0), d = Rf_ncols(x0), S = rows / qq;
  dk_lbfgs_options o = {Rf_asInteger(maxit), 8, 0.0, 1e7, 25};
  SEXP out = PROTECT(Rf_allocVector(VECSXP, 3));
  SEXP par = PROTECT(Rf_allocMatrix(REALSXP, rows, d));
  SEXP val = PROTECT(Rf_allocVector(REALSXP, S));
  SEXP its = PROTECT(Rf_allocVector(INTSXP, S));
  int rc = dk_polish(m, &c, qq, REAL(x0), S, REAL(lower), REAL(upper), &o, REAL(par), REAL(val), INTEGER(its));
  SET_VECTOR_ELT(out, 0, par); SET_VECTOR_ELT(out, 1, val); SET_VECTOR_ELT(out, 2, its);
  UNPROTECT(4);
  if (rc) Rf_error("dk_polish: %s", dk_last_error());
  return out;
}

/* sampleFromEI's draw on the device (R/sampleFromEI.R:109-135): 1-based row indices of the n drawn points */
SEXP dkR_sample_from_ei(SEXP h, SEXP x, SEXP T, SEXP minimization, SEXP n, SEXP seed) {
  dk_model* m = handle_of(h);
  int nn = Rf_asInteger(n);
  SEXP idx = PROTECT(Rf_allocVector(REALSXP, nn));
  int64_t* tmp = (int64_t*)R_alloc(nn, sizeof(int64_t));
  unsigned long long s = ((unsigned long long)(unsigned)INTEGER(seed)[1] << 32) | (unsigned)INTEGER(seed)[0];
  int rc = dk_sample_from_ei(m, REAL(x), (int64_t)Rf_nrows(x), Rf_asReal(T), Rf_asLogical(minimization), nn, s, tmp);
  for (int i = 0; i < nn; i++) REAL(idx)[i] = (double)(tmp[i] + 1);
  UNPROTECT(1);
  if (rc) Rf_error("dk_sample_from_ei: %s", dk_last_error());
  return idx;
}

/* One more observation in place (the refresh of the constant-liar loop, R/max_qEI.R:305-316), on every replica */
SEXP dkR_model_append(SEXP h, SEXP x_new, SEXP y_new, SEXP new_noise_var) {
  if (dk_multi_append(multi_of(h), REAL(x_new), Rf_asReal(y_new), Rf_asReal(new_noise_var)))
    Rf_error("dk_model_append: %s", dk_last_error());
  return R_NilValue;
}
#endif /* DK_HAVE_R */
