/* Stand-in for "mkl.h", only so that the CBLAS / LAPACKE enum types named in the reference's src/blas/interface.h:35-42 and
 * src/lapack/interface.hpp:4-28 exist in the engine check: no CBLAS or LAPACKE function is declared, let alone linked --
 * with integration/{blas,lapack}/interface.hpp in place the reference's engines do not call any. */
#pragma once
typedef enum { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef enum { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;
typedef enum { CblasUpper = 121, CblasLower = 122 } CBLAS_UPLO;
typedef enum { CblasNonUnit = 131, CblasUnit = 132 } CBLAS_DIAG;
typedef enum { CblasLeft = 141, CblasRight = 142 } CBLAS_SIDE;
#define LAPACK_ROW_MAJOR 101
#define LAPACK_COL_MAJOR 102
