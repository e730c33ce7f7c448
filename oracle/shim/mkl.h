/* TEST INFRASTRUCTURE (oracle).  Not shipped, not on the product path.
 *
 * Stand-in for the "mkl.h" that the reference includes at src/util/shared.h:24.
 * It only declares the CBLAS / LAPACKE entry points the reference's engines call
 * (src/blas/interface.hpp:54,74,92; src/lapack/interface.hpp:39,54,69,84) so the
 * unmodified reference headers can be compiled against OpenBLAS 0.3.15.
 */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
typedef enum { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef enum { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;
typedef enum { CblasUpper = 121, CblasLower = 122 } CBLAS_UPLO;
typedef enum { CblasNonUnit = 131, CblasUnit = 132 } CBLAS_DIAG;
typedef enum { CblasLeft = 141, CblasRight = 142 } CBLAS_SIDE;

void cblas_dgemm(CBLAS_ORDER order, CBLAS_TRANSPOSE ta, CBLAS_TRANSPOSE tb, int m, int n, int k,
                 double alpha, const double* a, int lda, const double* b, int ldb,
                 double beta, double* c, int ldc);
void cblas_dtrmm(CBLAS_ORDER order, CBLAS_SIDE side, CBLAS_UPLO uplo, CBLAS_TRANSPOSE ta,
                 CBLAS_DIAG diag, int m, int n, double alpha, const double* a, int lda,
                 double* b, int ldb);
void cblas_dsyrk(CBLAS_ORDER order, CBLAS_UPLO uplo, CBLAS_TRANSPOSE trans, int n, int k,
                 double alpha, const double* a, int lda, double beta, double* c, int ldc);

#define LAPACK_ROW_MAJOR 101
#define LAPACK_COL_MAJOR 102
int LAPACKE_dpotrf(int layout, char uplo, int n, double* a, int lda);
int LAPACKE_dtrtri(int layout, char uplo, char diag, int n, double* a, int lda);
int LAPACKE_dgeqrf(int layout, int m, int n, double* a, int lda, double* tau);
int LAPACKE_dorgqr(int layout, int m, int n, int k, double* a, int lda, const double* tau);

void openblas_set_num_threads(int n);
#ifdef __cplusplus
}
#endif
