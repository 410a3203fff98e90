/* TEST INFRASTRUCTURE: the few declarations of R's C API that combat-seq_b200/R/pkg/src/r_shim.c uses, so that the shim
 * can be syntax- and type-checked in an image without R (tests/test_abi_host.py).  Not R's header; never shipped. */
#ifndef CBS_R_STUB_H
#define CBS_R_STUB_H
typedef struct SEXPREC* SEXP;
extern SEXP R_DimSymbol;
#define REALSXP 14
SEXP Rf_getAttrib(SEXP, SEXP);
int* INTEGER(SEXP);
double* REAL(SEXP);
int Rf_asInteger(SEXP);
int Rf_asLogical(SEXP);
double Rf_asReal(SEXP);
SEXP Rf_allocMatrix(unsigned int, int, int);
SEXP Rf_protect(SEXP);
void Rf_unprotect(int);
void Rf_error(const char*, ...);
#define PROTECT(s) Rf_protect(s)
#define UNPROTECT(n) Rf_unprotect(n)
#endif
