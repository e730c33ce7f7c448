// Engine check: the reference's UNMODIFIED src/blas/engine.h + interface.h and src/lapack/engine.h + interface.h, compiled
// with integration/{blas,lapack}/interface.hpp standing where the reference keeps its CBLAS / LAPACKE specialisations
// (integration/Makefile builds that view with symbolic links; nothing of the reference is copied), driving device memory.
//
//   engine_check <dir>     reads <dir>/cases.txt, one case per line:
//     gemm  <ta> <tb> m n k alpha beta  A.bin B.bin C.bin out.bin
//     syrk  <uplo> <trans> n k alpha beta  A.bin C.bin out.bin
//     trmm  <side> <uplo> <trans> <diag> m n alpha  T.bin B.bin out.bin
//     potrf <uplo> n  A.bin out.bin
//     trtri <uplo> <diag> n  A.bin out.bin
//   operands are raw column-major doubles (leading dimension = rows); tests/test_gpu_grid.py writes them from
//   tests/golden/engines.npz (outputs of the reference's own engines on the CPU) and compares what comes back.
#include "src/blas/engine.h"
#include "src/lapack/engine.h"

#include <fstream>
#include <sstream>
#include <string>
#include <vector>

cap_handle_t capital_b200_handle = nullptr;

static std::vector<double> slurp(const std::string& path, size_t count) {
  std::vector<double> v(count);
  FILE* f = fopen(path.c_str(), "rb");
  if (!f || fread(v.data(), sizeof(double), count, f) != count) { fprintf(stderr, "engine_check: cannot read %zu doubles from %s\n", count, path.c_str()); exit(2); }
  fclose(f);
  return v;
}
static void spit(const std::string& path, const std::vector<double>& v) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f || fwrite(v.data(), sizeof(double), v.size(), f) != v.size()) { fprintf(stderr, "engine_check: cannot write %s\n", path.c_str()); exit(2); }
  fclose(f);
}
struct dev {
  double* p = nullptr; size_t n;
  explicit dev(const std::vector<double>& h) : n(h.size()) {
    if (cap_malloc((void**)&p, n * sizeof(double)) != CAP_OK || cap_h2d_async(capital_b200_handle, p, h.data(), n * sizeof(double)) != CAP_OK) { fprintf(stderr, "engine_check: upload failed: %s\n", cap_last_error()); exit(2); }
    cap_synchronize(capital_b200_handle);
  }
  std::vector<double> host() const {
    std::vector<double> h(n);
    cap_d2h_async(capital_b200_handle, h.data(), p, n * sizeof(double));
    if (cap_synchronize(capital_b200_handle) != CAP_OK) { fprintf(stderr, "engine_check: %s\n", cap_last_error()); exit(2); }
    return h;
  }
  ~dev() { cap_free(p); }
};

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: engine_check <dir>\n"); return 2; }
  const std::string dir = std::string(argv[1]) + "/";
  if (cap_create(&capital_b200_handle, 0) != CAP_OK) { fprintf(stderr, "engine_check: cap_create: %s\n", cap_last_error()); return 3; }
  std::ifstream in(dir + "cases.txt");
  std::string line;
  int done = 0;
  while (std::getline(in, line)) {
    std::istringstream ss(line);
    std::string op;
    if (!(ss >> op) || op[0] == '#') continue;
    if (op == "gemm") {
      char ta, tb; int64_t m, n, k; double alpha, beta; std::string fa, fb, fc, fo;
      ss >> ta >> tb >> m >> n >> k >> alpha >> beta >> fa >> fb >> fc >> fo;
      const int64_t ar = ta == 'T' ? k : m, ac = ta == 'T' ? m : k, br = tb == 'T' ? n : k, bc = tb == 'T' ? k : n;
      dev A(slurp(dir + fa, ar * ac)), B(slurp(dir + fb, br * bc)), C(slurp(dir + fc, m * n));
      blas::ArgPack_gemm<double> pack(blas::Order::AblasColumnMajor, ta == 'T' ? blas::Transpose::AblasTrans : blas::Transpose::AblasNoTrans,
                                      tb == 'T' ? blas::Transpose::AblasTrans : blas::Transpose::AblasNoTrans, alpha, beta);
      blas::engine::_gemm(A.p, B.p, C.p, m, n, k, ar, br, m, pack);
      spit(dir + fo, C.host());
    } else if (op == "syrk") {
      char uplo, trans; int64_t n, k; double alpha, beta; std::string fa, fc, fo;
      ss >> uplo >> trans >> n >> k >> alpha >> beta >> fa >> fc >> fo;
      const int64_t ar = trans == 'T' ? k : n, ac = trans == 'T' ? n : k;
      dev A(slurp(dir + fa, ar * ac)), C(slurp(dir + fc, n * n));
      blas::ArgPack_syrk<double> pack(blas::Order::AblasColumnMajor, uplo == 'U' ? blas::UpLo::AblasUpper : blas::UpLo::AblasLower,
                                      trans == 'T' ? blas::Transpose::AblasTrans : blas::Transpose::AblasNoTrans, alpha, beta);
      blas::engine::_syrk(A.p, C.p, n, k, ar, n, pack);
      spit(dir + fo, C.host());
    } else if (op == "trmm") {
      char side, uplo, trans, diag; int64_t m, n; double alpha; std::string ft, fb, fo;
      ss >> side >> uplo >> trans >> diag >> m >> n >> alpha >> ft >> fb >> fo;
      const int64_t nt = side == 'L' ? m : n;
      dev T(slurp(dir + ft, nt * nt)), B(slurp(dir + fb, m * n));
      blas::ArgPack_trmm<double> pack(blas::Order::AblasColumnMajor, side == 'L' ? blas::Side::AblasLeft : blas::Side::AblasRight,
                                      uplo == 'U' ? blas::UpLo::AblasUpper : blas::UpLo::AblasLower,
                                      trans == 'T' ? blas::Transpose::AblasTrans : blas::Transpose::AblasNoTrans,
                                      diag == 'U' ? blas::Diag::AblasUnit : blas::Diag::AblasNonUnit, alpha);
      blas::engine::_trmm(T.p, B.p, m, n, nt, m, pack);
      spit(dir + fo, B.host());
    } else if (op == "potrf") {
      char uplo; int n; std::string fa, fo;
      ss >> uplo >> n >> fa >> fo;
      dev A(slurp(dir + fa, (size_t)n * n));
      lapack::ArgPack_potrf pack(lapack::Order::AlapackColumnMajor, uplo == 'U' ? lapack::UpLo::AlapackUpper : lapack::UpLo::AlapackLower);
      lapack::engine::_potrf(A.p, n, n, pack);
      spit(dir + fo, A.host());
    } else if (op == "trtri") {
      char uplo, diag; int n; std::string fa, fo;
      ss >> uplo >> diag >> n >> fa >> fo;
      dev A(slurp(dir + fa, (size_t)n * n));
      lapack::ArgPack_trtri pack(lapack::Order::AlapackColumnMajor, uplo == 'U' ? lapack::UpLo::AlapackUpper : lapack::UpLo::AlapackLower,
                                 diag == 'U' ? lapack::Diag::AlapackUnit : lapack::Diag::AlapackNonUnit);
      lapack::engine::_trtri(A.p, n, n, pack);
      spit(dir + fo, A.host());
    } else { fprintf(stderr, "engine_check: unknown op %s\n", op.c_str()); return 2; }
    ++done;
  }
  int info = -1;
  cap_get_info(capital_b200_handle, &info);
  printf("engine_check: %d cases through blas::engine / lapack::engine of the reference headers, lapack info %d, %lld kernel launches\n",
         done, info, cap_kernel_launches(capital_b200_handle));
  cap_destroy(capital_b200_handle);
  return done > 0 ? 0 : 1;
}
