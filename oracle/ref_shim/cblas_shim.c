/* TEST INFRASTRUCTURE ONLY (oracle/): loop CBLAS for the compiled reference, see cblas.h. */
#include "cblas.h"
#include <math.h>

void cblas_dscal(int n, double alpha, double *x, int incx) {
    for (int i = 0; i < n; ++i) x[i * incx] *= alpha;
}
void cblas_daxpy(int n, double alpha, const double *x, int incx, double *y, int incy) {
    for (int i = 0; i < n; ++i) y[i * incy] += alpha * x[i * incx];
}
double cblas_ddot(int n, const double *x, int incx, const double *y, int incy) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += x[i * incx] * y[i * incy];
    return s;
}
double cblas_dnrm2(int n, const double *x, int incx) {
    /* scaled accumulation as in the netlib reference dnrm2 */
    double scale = 0.0, ssq = 1.0;
    for (int i = 0; i < n; ++i) {
        double v = x[i * incx];
        if (v != 0.0) {
            double a = fabs(v);
            if (scale < a) { ssq = 1.0 + ssq * (scale / a) * (scale / a); scale = a; }
            else ssq += (a / scale) * (a / scale);
        }
    }
    return scale * sqrt(ssq);
}
void cblas_dgemv(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE trans, int m, int n, double alpha,
                 const double *a, int lda, const double *x, int incx, double beta, double *y, int incy) {
    (void)order;
    if (trans == CblasNoTrans) {
        for (int i = 0; i < m; ++i) y[i * incy] = (beta == 0.0) ? 0.0 : beta * y[i * incy];
        for (int j = 0; j < n; ++j) {
            double t = alpha * x[j * incx];
            for (int i = 0; i < m; ++i) y[i * incy] += t * a[i + (long)j * lda];
        }
    } else {
        for (int j = 0; j < n; ++j) {
            double s = 0.0;
            for (int i = 0; i < m; ++i) s += a[i + (long)j * lda] * x[i * incx];
            y[j * incy] = ((beta == 0.0) ? 0.0 : beta * y[j * incy]) + alpha * s;
        }
    }
}
void cblas_dgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, int m, int n,
                 int k, double alpha, const double *a, int lda, const double *b, int ldb, double beta,
                 double *c, int ldc) {
    (void)order;
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i) {
            double s = 0.0;
            for (int l = 0; l < k; ++l) {
                double av = (ta == CblasNoTrans) ? a[i + (long)l * lda] : a[l + (long)i * lda];
                double bv = (tb == CblasNoTrans) ? b[l + (long)j * ldb] : b[j + (long)l * ldb];
                s += av * bv;
            }
            c[i + (long)j * ldc] = ((beta == 0.0) ? 0.0 : beta * c[i + (long)j * ldc]) + alpha * s;
        }
}
void cblas_dger(enum CBLAS_ORDER order, int m, int n, double alpha, const double *x, int incx,
                const double *y, int incy, double *a, int lda) {
    (void)order;
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i) a[i + (long)j * lda] += alpha * x[i * incx] * y[j * incy];
}
void cblas_dsbmv(enum CBLAS_ORDER order, enum CBLAS_UPLO uplo, int n, int k, double alpha,
                 const double *a, int lda, const double *x, int incx, double beta, double *y, int incy) {
    (void)order; (void)uplo; (void)k; /* K = 0: diagonal matrix stored with stride lda */
    for (int i = 0; i < n; ++i)
        y[i * incy] = ((beta == 0.0) ? 0.0 : beta * y[i * incy]) + alpha * a[(long)i * lda] * x[i * incx];
}
