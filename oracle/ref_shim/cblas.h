/* Test infrastructure only (oracle/_ref build).
 * Minimal CBLAS declaration so the UNMODIFIED reference TU src/data/grm/grm.cpp
 * (which does `#include <cblas.h>`, /root/reference/src/data/grm/grm.cpp:22-27) compiles
 * in an image that has no system cblas.h.  The symbol is provided by the OpenBLAS that
 * ships inside the scipy wheel (exported there as scipy_cblas_dsyrk). */
#ifndef GGRM_ORACLE_REF_SHIM_CBLAS_H_
#define GGRM_ORACLE_REF_SHIM_CBLAS_H_
#ifdef __cplusplus
extern "C" {
#endif
typedef enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef enum CBLAS_UPLO { CblasUpper = 121, CblasLower = 122 } CBLAS_UPLO;
typedef enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;
#ifdef GGRM_REF_SCIPY_OPENBLAS
#define cblas_dsyrk scipy_cblas_dsyrk
#endif
void cblas_dsyrk(CBLAS_ORDER order, CBLAS_UPLO uplo, CBLAS_TRANSPOSE trans, int n, int k,
                 double alpha, const double* a, int lda, double beta, double* c, int ldc);
#ifdef __cplusplus
}
#endif
#endif
