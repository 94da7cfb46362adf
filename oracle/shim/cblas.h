/* TEST INFRASTRUCTURE ONLY (oracle build).
 *
 * Minimal cblas.h so that the reference CPU backend (/root/reference/src/cpu_backend.cpp,
 * which does `#include <cblas.h>` at :10) compiles in an image that has no system
 * OpenBLAS headers.  The only CBLAS entry it uses is cblas_sgemm (:182-256); the LAPACK
 * Fortran symbols it declares itself (:12-22).  The OpenBLAS bundled with the scipy wheel
 * prefixes every export with `scipy_`, hence the renames below (LP64 build: 32-bit ints,
 * matching the reference's prototypes).
 */
#ifndef DR4G_ORACLE_CBLAS_SHIM_H
#define DR4G_ORACLE_CBLAS_SHIM_H

#define cblas_sgemm scipy_cblas_sgemm
#define sgesdd_ scipy_sgesdd_
#define sgesvd_ scipy_sgesvd_
#define sgeqrf_ scipy_sgeqrf_
#define sorgqr_ scipy_sorgqr_

#ifdef __cplusplus
extern "C" {
#endif

typedef enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;

void cblas_sgemm(CBLAS_ORDER order, CBLAS_TRANSPOSE ta, CBLAS_TRANSPOSE tb, int m, int n, int k,
                 float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                 float* c, int ldc);

#ifdef __cplusplus
}
#endif
#endif
