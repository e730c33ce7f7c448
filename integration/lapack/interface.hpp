// capital-b200 binding of the reference's LAPACK engine: drop this file in place of src/lapack/interface.hpp
// (huttered40/capital).  lapack::engine::_potrf/_trtri (declared src/lapack/interface.h:49-53, defined over LAPACKE in
// src/lapack/interface.hpp:31-58) run on the B200 through include/capital_b200.h; device pointers, column-major.
// LAPACK's `info` stays on the device (cap_get_info); the reference discards it (interface.hpp:39,54).
// _geqrf / _orgqr (Householder QR) are not on the CholeskyQR / Cholesky path: they stop with a message.

#include "capital_b200.h"

extern cap_handle_t capital_b200_handle;

namespace lapack{
inline void helper::setInfoParameters_potrf(const ArgPack_potrf& srcPackage, int& destArg1, char& destArg2){
  destArg1 = (srcPackage.order == Order::AlapackRowMajor ? LAPACK_ROW_MAJOR : LAPACK_COL_MAJOR);
  destArg2 = (srcPackage.uplo == UpLo::AlapackUpper ? 'U' : 'L');
}
inline void helper::setInfoParameters_trtri(const ArgPack_trtri& srcPackage, int& destArg1, char& destArg2, char& destArg3){
  destArg1 = (srcPackage.order == Order::AlapackRowMajor ? LAPACK_ROW_MAJOR : LAPACK_COL_MAJOR);
  destArg2 = (srcPackage.uplo == UpLo::AlapackUpper ? 'U' : 'L');
  destArg3 = (srcPackage.diag == Diag::AlapackUnit ? 'U' : 'N');
}
inline void helper::setInfoParameters_geqrf(const ArgPack_geqrf& srcPackage, int& destArg1){
  destArg1 = (srcPackage.order == Order::AlapackRowMajor ? LAPACK_ROW_MAJOR : LAPACK_COL_MAJOR);
}
inline void helper::setInfoParameters_orgqr(const ArgPack_orgqr& srcPackage, int& destArg1){
  destArg1 = (srcPackage.order == Order::AlapackRowMajor ? LAPACK_ROW_MAJOR : LAPACK_COL_MAJOR);
}

// replaces LAPACKE_dpotrf (src/lapack/interface.hpp:39): A = R^T R in place (upper), N^3/3 look-ahead schedule
template<>
inline void engine::_potrf(double* matrixA, int n, int lda, const ArgPack_potrf& srcPackage){
  assert(srcPackage.order == Order::AlapackColumnMajor);
  if (cap_dpotrf(capital_b200_handle, srcPackage.uplo == UpLo::AlapackUpper ? 'U' : 'L', n, matrixA, lda) != CAP_OK){
    fprintf(stderr, "capital_b200: cap_dpotrf failed: %s\n", cap_last_error()); abort();
  }
}

// replaces LAPACKE_dtrtri (src/lapack/interface.hpp:54)
template<>
inline void engine::_trtri(double* matrixA, int n, int lda, const ArgPack_trtri& srcPackage){
  assert(srcPackage.order == Order::AlapackColumnMajor);
  if (cap_dtrtri(capital_b200_handle, srcPackage.uplo == UpLo::AlapackUpper ? 'U' : 'L', srcPackage.diag == Diag::AlapackUnit ? 'U' : 'N', n, matrixA, lda) != CAP_OK){
    fprintf(stderr, "capital_b200: cap_dtrtri failed: %s\n", cap_last_error()); abort();
  }
}

template<>
inline void engine::_geqrf(double*, double*, int, int, int, const ArgPack_geqrf&){
  fprintf(stderr, "capital_b200: lapack::engine::_geqrf is not on the CholeskyQR / Cholesky path (no device implementation)\n"); abort();
}
template<>
inline void engine::_orgqr(double*, double*, int, int, int, int, const ArgPack_orgqr&){
  fprintf(stderr, "capital_b200: lapack::engine::_orgqr is not on the CholeskyQR / Cholesky path (no device implementation)\n"); abort();
}
}
