/* TEST INFRASTRUCTURE ONLY (oracle/): minimal column-major CBLAS used to compile the
 * UNMODIFIED reference sources under /root/reference into oracle/_ref/.
 * The reference links a system BLAS that is absent from this image; the eight routines
 * it calls (bbfmm/blas_wrapper.cpp:17-138) are restated here as plain loops so that the
 * oracle is deterministic.  Only the call patterns used by the reference are supported
 * (CblasColMajor, unit strides are NOT assumed, dsbmv only with K=0). */
#ifndef ANISO_ORACLE_CBLAS_SHIM_H
#define ANISO_ORACLE_CBLAS_SHIM_H
#ifdef __cplusplus
extern "C" {
#endif
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
enum CBLAS_UPLO { CblasUpper = 121, CblasLower = 122 };
void cblas_dscal(int n, double alpha, double *x, int incx);
void cblas_daxpy(int n, double alpha, const double *x, int incx, double *y, int incy);
double cblas_ddot(int n, const double *x, int incx, const double *y, int incy);
double cblas_dnrm2(int n, const double *x, int incx);
void cblas_dgemv(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE trans, int m, int n, double alpha,
                 const double *a, int lda, const double *x, int incx, double beta, double *y, int incy);
void cblas_dgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, int m, int n,
                 int k, double alpha, const double *a, int lda, const double *b, int ldb, double beta,
                 double *c, int ldc);
void cblas_dger(enum CBLAS_ORDER order, int m, int n, double alpha, const double *x, int incx,
                const double *y, int incy, double *a, int lda);
void cblas_dsbmv(enum CBLAS_ORDER order, enum CBLAS_UPLO uplo, int n, int k, double alpha,
                 const double *a, int lda, const double *x, int incx, double beta, double *y, int incy);
#ifdef __cplusplus
}
#endif
#endif
