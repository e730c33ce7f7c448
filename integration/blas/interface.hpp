// capital-b200 binding of the reference's BLAS engine: drop this file in place of src/blas/interface.hpp
// (huttered40/capital).  src/blas/interface.h:68 includes "interface.hpp" -- with this one there instead, the three
// double specialisations the algorithms call (blas::engine::_gemm/_trmm/_syrk, declared src/blas/interface.h:57-66,
// defined over CBLAS in src/blas/interface.hpp:43-97) run on the B200 through the C ABI of include/capital_b200.h.
// Pointers are DEVICE pointers (cap_malloc); only column-major operands exist on the path.
// Compiled against the UNMODIFIED src/blas/engine.h + interface.h by integration/Makefile (tests/test_gpu_grid.py runs it).

#include "capital_b200.h"

// one handle per process (= per GPU), created by the application: cap_create(&capital_b200_handle, local_rank)
extern cap_handle_t capital_b200_handle;

namespace blas{
namespace capital_b200_detail{
inline char tr(Transpose x){ return x == Transpose::AblasTrans ? 'T' : 'N'; }
inline char ul(UpLo x){ return x == UpLo::AblasUpper ? 'U' : 'L'; }
inline void must(int rc, const char* what){
  if (rc != CAP_OK){ fprintf(stderr, "capital_b200: %s failed: %s\n", what, cap_last_error()); abort(); }
}
}

// the helper's CBLAS translations stay defined (they are declared in interface.h:35-42); nothing here uses them
template<typename T>
void helper::setInfoParameters_gemm(const ArgPack_gemm<T>& srcPackage, CBLAS_ORDER& destArg1, CBLAS_TRANSPOSE& destArg2, CBLAS_TRANSPOSE& destArg3){
  destArg1 = (srcPackage.order == Order::AblasRowMajor ? CblasRowMajor : CblasColMajor);
  destArg2 = (srcPackage.transposeA == Transpose::AblasTrans ? CblasTrans : CblasNoTrans);
  destArg3 = (srcPackage.transposeB == Transpose::AblasTrans ? CblasTrans : CblasNoTrans);
}
template<typename T>
void helper::setInfoParameters_trmm(const ArgPack_trmm<T>& srcPackage, CBLAS_ORDER& destArg1, CBLAS_SIDE& destArg2, CBLAS_UPLO& destArg3, CBLAS_TRANSPOSE& destArg4, CBLAS_DIAG& destArg5){
  destArg1 = (srcPackage.order == Order::AblasRowMajor ? CblasRowMajor : CblasColMajor);
  destArg2 = (srcPackage.side == Side::AblasLeft ? CblasLeft : CblasRight);
  destArg3 = (srcPackage.uplo == UpLo::AblasLower ? CblasLower : CblasUpper);
  destArg4 = (srcPackage.transposeA == Transpose::AblasTrans ? CblasTrans : CblasNoTrans);
  destArg5 = (srcPackage.diag == Diag::AblasUnit ? CblasUnit : CblasNonUnit);
}
template<typename T>
void helper::setInfoParameters_syrk(const ArgPack_syrk<T>& srcPackage, CBLAS_ORDER& destArg1, CBLAS_UPLO& destArg2, CBLAS_TRANSPOSE& destArg3){
  destArg1 = (srcPackage.order == Order::AblasRowMajor ? CblasRowMajor : CblasColMajor);
  destArg2 = (srcPackage.uplo == UpLo::AblasLower ? CblasLower : CblasUpper);
  destArg3 = (srcPackage.transposeA == Transpose::AblasTrans ? CblasTrans : CblasNoTrans);
}

// replaces cblas_dgemm (src/blas/interface.hpp:54)
template<>
inline void engine::_gemm(double* matrixA, double* matrixB, double* matrixC, int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb, int64_t ldc, const ArgPack_gemm<double>& srcPackage){
  assert(srcPackage.order == Order::AblasColumnMajor);
  capital_b200_detail::must(cap_dgemm(capital_b200_handle, capital_b200_detail::tr(srcPackage.transposeA), capital_b200_detail::tr(srcPackage.transposeB), m, n, k,
                                      srcPackage.alpha, matrixA, lda, matrixB, ldb, srcPackage.beta, matrixC, ldc), "cap_dgemm");
}

// replaces cblas_dtrmm (src/blas/interface.hpp:74): B <- alpha op(A) B or alpha B op(A), in place
template<>
inline void engine::_trmm(double* matrixA, double* matrixB, int64_t m, int64_t n, int64_t lda, int64_t ldb, const ArgPack_trmm<double>& srcPackage){
  assert(srcPackage.order == Order::AblasColumnMajor);
  capital_b200_detail::must(cap_dtrmm(capital_b200_handle, srcPackage.side == Side::AblasLeft ? 'L' : 'R', capital_b200_detail::ul(srcPackage.uplo),
                                      capital_b200_detail::tr(srcPackage.transposeA), srcPackage.diag == Diag::AblasUnit ? 'U' : 'N', m, n, srcPackage.alpha,
                                      matrixA, lda, matrixB, ldb), "cap_dtrmm");
}

// replaces cblas_dsyrk (src/blas/interface.hpp:92)
template<>
inline void engine::_syrk(double* matrixA, double* matrixC, int64_t n, int64_t k, int64_t lda, int64_t ldc, const ArgPack_syrk<double>& srcPackage){
  assert(srcPackage.order == Order::AblasColumnMajor);
  capital_b200_detail::must(cap_dsyrk(capital_b200_handle, capital_b200_detail::ul(srcPackage.uplo), capital_b200_detail::tr(srcPackage.transposeA), n, k,
                                      srcPackage.alpha, matrixA, lda, srcPackage.beta, matrixC, ldc), "cap_dsyrk");
}
}
