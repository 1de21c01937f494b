/* Minimal DECLARATIONS of the R C API used by the r-package/src sources, for a syntax/type check only
 * (tests/test_abi_symbols.py).  Not an R runtime: nothing here is linked or executed. */
#ifndef R_STUB_INTERNALS_H
#define R_STUB_INTERNALS_H
#include <stddef.h>
typedef struct SEXPREC* SEXP;
typedef ptrdiff_t R_xlen_t;
typedef enum { FALSE = 0, TRUE } Rboolean;
typedef void (*R_CFinalizer_t)(SEXP);
extern SEXP R_NilValue, R_NamesSymbol;
extern double R_NaReal;
extern int R_NaInt;
#define NA_REAL R_NaReal
#define NA_INTEGER R_NaInt
#define INTSXP 13
#define REALSXP 14
#define VECSXP 19
double* REAL(SEXP); int* INTEGER(SEXP); int* LOGICAL(SEXP); const char* CHAR(SEXP);
SEXP STRING_ELT(SEXP, R_xlen_t); SEXP VECTOR_ELT(SEXP, R_xlen_t); SEXP SET_VECTOR_ELT(SEXP, R_xlen_t, SEXP);
SEXP Rf_getAttrib(SEXP, SEXP); R_xlen_t Rf_xlength(SEXP); int Rf_nrows(SEXP); int Rf_ncols(SEXP);
int Rf_asInteger(SEXP); int Rf_asLogical(SEXP); double Rf_asReal(SEXP);
SEXP Rf_allocVector(unsigned int, R_xlen_t); SEXP Rf_allocMatrix(unsigned int, int, int);
SEXP Rf_protect(SEXP); void Rf_unprotect(int);
#define PROTECT(s) Rf_protect(s)
#define UNPROTECT(n) Rf_unprotect(n)
void Rf_error(const char*, ...);
void* R_ExternalPtrAddr(SEXP); void R_ClearExternalPtr(SEXP); SEXP R_MakeExternalPtr(void*, SEXP, SEXP);
void R_RegisterCFinalizerEx(SEXP, R_CFinalizer_t, Rboolean);
char* R_alloc(size_t, int);
SEXP Rf_mkString(const char*);
#endif
