/* Registration: the reference's .C table (/root/reference/src/packagename_init.c:12-21) extended with the
 * .Call table of the GPU engine.  Compiled only with R headers (DK_HAVE_R). */
#ifdef DK_HAVE_R
#include <stdlib.h>
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>

extern void davies(double* lb1, double* nc1, int* n1, int* r1, double* sigma, double* c1, int* lq, int* lim1,
                   double* acc, double* trace, int* ifault, double* res);
extern SEXP dkR_model_create(SEXP, SEXP);
extern SEXP dkR_ei(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_ei_grad(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_quantile(SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_akg(SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_qei(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_screen_ei(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_sample_weights(SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_check_predict(SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_feasible_improvement(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_crit_sur_cst(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_release(SEXP);
extern SEXP dkR_model_key(SEXP);
extern SEXP dkR_predict(SEXP, SEXP, SEXP);
extern SEXP dkR_akg_grad(SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_aei(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_eqi(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_screen_generated(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_polish(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_sample_from_ei(SEXP, SEXP, SEXP, SEXP, SEXP, SEXP);
extern SEXP dkR_model_append(SEXP, SEXP, SEXP, SEXP);

static const R_CMethodDef CEntries[] = {{"davies", (DL_FUNC)&davies, 13}, {NULL, NULL, 0}};
static const R_CallMethodDef CallEntries[] = {
    {"dkR_model_create", (DL_FUNC)&dkR_model_create, 2}, {"dkR_ei", (DL_FUNC)&dkR_ei, 7},
    {"dkR_ei_grad", (DL_FUNC)&dkR_ei_grad, 6},           {"dkR_quantile", (DL_FUNC)&dkR_quantile, 5},
    {"dkR_akg", (DL_FUNC)&dkR_akg, 4},                   {"dkR_qei", (DL_FUNC)&dkR_qei, 7},
    {"dkR_screen_ei", (DL_FUNC)&dkR_screen_ei, 7},       {"dkR_sample_weights", (DL_FUNC)&dkR_sample_weights, 4},
    {"dkR_check_predict", (DL_FUNC)&dkR_check_predict, 4},
    {"dkR_feasible_improvement", (DL_FUNC)&dkR_feasible_improvement, 8},
    {"dkR_crit_sur_cst", (DL_FUNC)&dkR_crit_sur_cst, 9},
    {"dkR_release", (DL_FUNC)&dkR_release, 1},           {"dkR_model_key", (DL_FUNC)&dkR_model_key, 1},
    {"dkR_predict", (DL_FUNC)&dkR_predict, 3},           {"dkR_akg_grad", (DL_FUNC)&dkR_akg_grad, 4},
    {"dkR_aei", (DL_FUNC)&dkR_aei, 6},                   {"dkR_eqi", (DL_FUNC)&dkR_eqi, 7},
    {"dkR_screen_generated", (DL_FUNC)&dkR_screen_generated, 7}, {"dkR_polish", (DL_FUNC)&dkR_polish, 7},
    {"dkR_sample_from_ei", (DL_FUNC)&dkR_sample_from_ei, 6},     {"dkR_model_append", (DL_FUNC)&dkR_model_append, 4},
    {NULL, NULL, 0}};

void R_init_DiceOptim(DllInfo* dll) {
  R_registerRoutines(dll, CEntries, CallEntries, NULL, NULL);
  R_useDynamicSymbols(dll, FALSE);
}
#endif
