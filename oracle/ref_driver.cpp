/* TEST INFRASTRUCTURE (oracle).  Not shipped, not on the product path.
 *
 * ref_driver: runs the UNMODIFIED reference (headers included from /root/reference where they
 * lie; nothing is copied) on its own CPU engines (OpenBLAS behind the mkl.h shim) under minimpi,
 * and dumps inputs/outputs as raw little-endian float64 files so that tests can compare the CUDA
 * path and the numpy restatement (oracle/capital_oracle.py) with the real thing.
 *
 * Output files: <prefix>.r<rank>.<name>.f64 ; rank 0 also prints one "key=value ..." line.
 *
 *   ref_driver cacqr   <prefix> variant m n rep complete_inv split bc layout chunks policy validate
 *        -> qr::cacqr<Serialize,SaveIntermediates>::factor   (src/alg/qr/cacqr/cacqr.hpp:219)
 *   ref_driver cholinv <prefix> N rep_div complete_inv split bc layout chunks policy validate
 *        -> cholesky::cholinv<Serialize,SaveIntermediates,BP>::factor (src/alg/cholesky/cholinv/cholinv.hpp:6)
 *   ref_driver summa   <prefix> M N K rep_div layout chunks
 *        -> matmult::summa::invoke gemm overload (src/alg/matmult/summa/summa.hpp:7)
 *   ref_driver engine  <prefix> op ...   (single rank; reads <prefix>.in.{A,B,C}.f64)
 *        -> blas::engine / lapack::engine (src/blas/interface.hpp:43-97, src/lapack/interface.hpp:31-58)
 *   ref_driver gen     <prefix> kind X Y        (kind: random | symmetric | lowertri; grid = ranks)
 *        -> src/matrix/structure.hpp:69-129,248-287
 *
 * policy: 0 NoReplication (bench default, incorrect on d>1 grids under recursion), 1 ReplicateCommComp,
 *         2 ReplicateComp (the trustworthy multi-rank oracle), 3 NoReplicationOverlap.
 */
#include <array>
#include "src/alg/qr/cacqr/cacqr.h"
#include "test/qr/validate.h"
#include "test/cholesky/validate.h"

#include <string>
#include <unistd.h>

using T = double;
using U = int64_t;
namespace cp = cholesky::policy::cholinv;

static std::string g_prefix;
static int g_rank = 0, g_size = 1;

static void dump(const char* name, const T* p, size_t n) {
  if (getenv("REF_NO_DUMP")) return;   // timing-only runs (bench.py cpu_baseline / --impl reference)
  std::string f = g_prefix + ".r" + std::to_string(g_rank) + "." + name + ".f64";
  FILE* fp = fopen(f.c_str(), "wb");
  if (!fp || fwrite(p, sizeof(T), n, fp) != n) { perror(f.c_str()); _exit(4); }
  fclose(fp);
}

static std::vector<T> slurp(const char* name, size_t n) {
  std::string f = g_prefix + ".in." + name + ".f64";
  std::vector<T> v(n);
  FILE* fp = fopen(f.c_str(), "rb");
  if (!fp || fread(v.data(), sizeof(T), n, fp) != n) { perror(f.c_str()); _exit(4); }
  fclose(fp);
  return v;
}

static double max_over_ranks(double t) { PMPI_Allreduce(MPI_IN_PLACE, &t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD); return t; }

template <class BP>
static int run_cacqr(char** a) {
  size_t variant = atoi(a[0]); U m = atol(a[1]), n = atol(a[2]); int rep = atoi(a[3]);
  bool complete_inv = atoi(a[4]); U split = atoi(a[5]); U bc = atoi(a[6]);
  size_t layout = atoi(a[7]), chunks = atoi(a[8]); int do_validate = atoi(a[10]);
  using qr_type = qr::cacqr<qr::policy::cacqr::Serialize, qr::policy::cacqr::SaveIntermediates>;
  using ci_type = cholesky::cholinv<cp::Serialize, cp::SaveIntermediates, BP>;
  auto RectTopo = topo::rect(MPI_COMM_WORLD, rep, layout, chunks);
  matrix<T, U, rect> A(n, m, RectTopo.c, RectTopo.d);
  A.distribute_random(RectTopo.x, RectTopo.y, RectTopo.c, RectTopo.d, g_rank / RectTopo.c);
  typename ci_type::template info<T, U> ci_pack(complete_inv, split, bc, 'U');
  typename qr_type::template info<T, U, ci_type> pack(variant, ci_pack);
  qr_type::factor(A, pack, RectTopo);  // warm-up, as bench/qr/cacqr.cpp:42-46
  PMPI_Barrier(MPI_COMM_WORLD);
  double t0 = MPI_Wtime();
  qr_type::factor(A, pack, RectTopo);
  double t = max_over_ranks(MPI_Wtime() - t0);
  dump("A", A.data(), (size_t)A.num_rows_local() * A.num_columns_local());
  dump("Q", pack.Q.data(), (size_t)pack.Q.num_rows_local() * pack.Q.num_columns_local());
  auto R = qr_type::construct_R(pack, RectTopo);
  dump("R", R.data(), (size_t)R.num_rows_local() * R.num_columns_local());
  double res = -1, orth = -1;
  if (do_validate) {
    res = max_over_ranks(qr::validate<qr_type>::residual(A, pack, RectTopo));
    orth = max_over_ranks(qr::validate<qr_type>::orthogonality(A, pack, RectTopo));
  }
  if (g_rank == 0)
    printf("alg=cacqr m=%ld n=%ld ranks=%d c=%zu d=%zu mloc=%ld nloc=%ld seconds=%.6f residual=%.3e orthogonality=%.3e\n", (long)m, (long)n, g_size,
           RectTopo.c, RectTopo.d, (long)A.num_rows_local(), (long)A.num_columns_local(), t, res, orth);
  return 0;
}

template <class BP>
static int run_cholinv(char** a) {
  U N = atol(a[0]); U rep_div = atoi(a[1]); bool complete_inv = atoi(a[2]); U split = atoi(a[3]); U bc = atoi(a[4]);
  size_t layout = atoi(a[5]), chunks = atoi(a[6]); int do_validate = atoi(a[8]);
  using ci_type = cholesky::cholinv<cp::Serialize, cp::SaveIntermediates, BP>;
  size_t cube = std::nearbyint(std::ceil(pow(g_size, 1. / 3.)));
  auto SquareTopo = topo::square(MPI_COMM_WORLD, cube / rep_div, layout, chunks);
  matrix<T, U, rect> A(N, N, SquareTopo.d, SquareTopo.d);
  A.distribute_symmetric(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, g_rank / SquareTopo.c, true);
  typename ci_type::template info<T, U> pack(complete_inv, split, bc, 'U');
  ci_type::factor(A, pack, SquareTopo);  // warm-up, as bench/cholesky/cholinv.cpp:44
  MPI_Barrier(MPI_COMM_WORLD);
  double t0 = MPI_Wtime();
  ci_type::factor(A, pack, SquareTopo);
  double t = max_over_ranks(MPI_Wtime() - t0);
  dump("A", A.data(), (size_t)A.num_rows_local() * A.num_columns_local());
  auto R = ci_type::construct_R(pack, SquareTopo);
  auto Ri = ci_type::construct_Rinv(pack, SquareTopo);
  dump("R", R.data(), (size_t)R.num_rows_local() * R.num_columns_local());
  dump("Rinv", Ri.data(), (size_t)Ri.num_rows_local() * Ri.num_columns_local());
  double res = -1;
  if (do_validate) res = max_over_ranks(cholesky::validate<ci_type>::residual(A, pack, SquareTopo));
  if (g_rank == 0)
    printf("alg=cholinv N=%ld ranks=%d c=%zu d=%zu lloc=%ld bcdim=%ld seconds=%.6f residual=%.3e\n", (long)N, g_size, SquareTopo.c, SquareTopo.d,
           (long)A.num_rows_local(), (long)pack.bcDimension, t, res);
  return 0;
}

static int run_summa(char** a) {
  U M = atol(a[0]), N = atol(a[1]), K = atol(a[2]); U rep_div = atoi(a[3]); size_t layout = atoi(a[4]), chunks = atoi(a[5]);
  size_t cube = std::nearbyint(std::ceil(pow(g_size, 1. / 3.)));
  auto SquareTopo = topo::square(MPI_COMM_WORLD, cube / rep_div, layout, chunks);
  // same construction and keys as bench/matmult/summa_gemm.cpp:32-38
  matrix<T, U, rect> matA(K, M, SquareTopo.d, SquareTopo.d), matB(N, K, SquareTopo.d, SquareTopo.d), matC(N, M, SquareTopo.d, SquareTopo.d);
  blas::ArgPack_gemm<T> blasArgs(blas::Order::AblasColumnMajor, blas::Transpose::AblasNoTrans, blas::Transpose::AblasNoTrans, 1., 0.);
  matA.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, g_rank / SquareTopo.c);
  matB.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, g_rank / SquareTopo.c * (-1));
  matC.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, g_rank / SquareTopo.c * (-1));
  dump("A", matA.data(), (size_t)matA.num_elems());
  dump("B", matB.data(), (size_t)matB.num_elems());
  MPI_Barrier(MPI_COMM_WORLD);
  double t0 = MPI_Wtime();
  matmult::summa::invoke(matA, matB, matC, SquareTopo, blasArgs);
  double t = max_over_ranks(MPI_Wtime() - t0);
  dump("C", matC.data(), (size_t)matC.num_elems());
  if (g_rank == 0) printf("alg=summa M=%ld N=%ld K=%ld ranks=%d c=%zu d=%zu seconds=%.6f\n", (long)M, (long)N, (long)K, g_size, SquareTopo.c, SquareTopo.d, t);
  return 0;
}

static int run_engine(char** a, int n_args) {
  std::string op = a[0];
  auto order = blas::Order::AblasColumnMajor;
  auto tr = [](const char* s) { return s[0] == 'T' ? blas::Transpose::AblasTrans : blas::Transpose::AblasNoTrans; };
  if (op == "gemm") {  // gemm ta tb m n k lda ldb ldc alpha beta
    U m = atol(a[3]), n = atol(a[4]), k = atol(a[5]), lda = atol(a[6]), ldb = atol(a[7]), ldc = atol(a[8]);
    double alpha = atof(a[9]), beta = atof(a[10]);
    bool ta = a[1][0] == 'T', tb = a[2][0] == 'T';
    auto A = slurp("A", (size_t)lda * (ta ? m : k)), B = slurp("B", (size_t)ldb * (tb ? k : n)), C = slurp("C", (size_t)ldc * n);
    blas::ArgPack_gemm<T> pk(order, tr(a[1]), tr(a[2]), alpha, beta);
    blas::engine::_gemm(A.data(), B.data(), C.data(), m, n, k, lda, ldb, ldc, pk);
    dump("C", C.data(), C.size());
  } else if (op == "syrk") {  // syrk uplo trans n k lda ldc alpha beta
    U n = atol(a[3]), k = atol(a[4]), lda = atol(a[5]), ldc = atol(a[6]); double alpha = atof(a[7]), beta = atof(a[8]);
    bool t = a[2][0] == 'T';
    auto A = slurp("A", (size_t)lda * (t ? n : k)), C = slurp("C", (size_t)ldc * n);
    blas::ArgPack_syrk<T> pk(order, a[1][0] == 'U' ? blas::UpLo::AblasUpper : blas::UpLo::AblasLower, tr(a[2]), alpha, beta);
    blas::engine::_syrk(A.data(), C.data(), n, k, lda, ldc, pk);
    dump("C", C.data(), C.size());
  } else if (op == "trmm") {  // trmm side uplo trans diag m n lda ldb alpha
    U m = atol(a[5]), n = atol(a[6]), lda = atol(a[7]), ldb = atol(a[8]); double alpha = atof(a[9]);
    bool left = a[1][0] == 'L';
    auto A = slurp("A", (size_t)lda * (left ? m : n)), B = slurp("B", (size_t)ldb * n);
    blas::ArgPack_trmm<T> pk(order, left ? blas::Side::AblasLeft : blas::Side::AblasRight, a[2][0] == 'U' ? blas::UpLo::AblasUpper : blas::UpLo::AblasLower,
                             tr(a[3]), a[4][0] == 'U' ? blas::Diag::AblasUnit : blas::Diag::AblasNonUnit, alpha);
    blas::engine::_trmm(A.data(), B.data(), m, n, lda, ldb, pk);
    dump("B", B.data(), B.size());
  } else if (op == "potrf") {  // potrf uplo n lda
    int n = atoi(a[2]), lda = atoi(a[3]);
    auto A = slurp("A", (size_t)lda * n);
    lapack::ArgPack_potrf pk(lapack::Order::AlapackColumnMajor, a[1][0] == 'U' ? lapack::UpLo::AlapackUpper : lapack::UpLo::AlapackLower);
    lapack::engine::_potrf(A.data(), n, lda, pk);
    dump("A", A.data(), A.size());
  } else if (op == "trtri") {  // trtri uplo diag n lda
    int n = atoi(a[3]), lda = atoi(a[4]);
    auto A = slurp("A", (size_t)lda * n);
    lapack::ArgPack_trtri pk(lapack::Order::AlapackColumnMajor, a[1][0] == 'U' ? lapack::UpLo::AlapackUpper : lapack::UpLo::AlapackLower,
                             a[2][0] == 'U' ? lapack::Diag::AlapackUnit : lapack::Diag::AlapackNonUnit);
    lapack::engine::_trtri(A.data(), n, lda, pk);
    dump("A", A.data(), A.size());
  } else { fprintf(stderr, "engine: unknown op %s\n", op.c_str()); return 2; }
  (void)n_args;
  return 0;
}

static int run_gen(char** a) {
  std::string kind = a[0]; U X = atol(a[1]), Y = atol(a[2]);
  if (kind == "random") {  // 1D row-cyclic like bench/qr/cacqr.cpp:32-34: grid 1 x ranks, key = rank
    matrix<T, U, rect> A(X, Y, 1, g_size);
    A.distribute_random(0, g_rank, 1, g_size, g_rank);
    dump("A", A.data(), (size_t)A.num_elems());
  } else {
    size_t d = std::nearbyint(std::sqrt((double)g_size));
    size_t x = g_rank / d, y = g_rank % d;
    if (kind == "random2d") {
      matrix<T, U, rect> A(X, Y, d, d);
      A.distribute_random(x, y, d, d, -(int64_t)g_rank);
      dump("A", A.data(), (size_t)A.num_elems());
    } else if (kind == "symmetric") {
      matrix<T, U, rect> A(X, Y, d, d);
      A.distribute_symmetric(x, y, d, d, g_rank, true);
      dump("A", A.data(), (size_t)A.num_elems());
    } else if (kind == "lowertri") {
      matrix<T, U, lowertri> A(X, Y, d, d);
      A.distribute_random(x, y, d, d, g_rank);
      dump("A", A.data(), (size_t)A.num_elems());
    } else { fprintf(stderr, "gen: unknown kind\n"); return 2; }
  }
  return 0;
}

int main(int argc, char** argv) {
  int provided;
  MPI_Init_thread(&argc, &argv, MPI_THREAD_SINGLE, &provided);
  MPI_Comm_rank(MPI_COMM_WORLD, &g_rank); MPI_Comm_size(MPI_COMM_WORLD, &g_size);
  if (argc < 3) { fprintf(stderr, "usage: ref_driver <alg> <prefix> args...\n"); return 2; }
  std::string alg = argv[1]; g_prefix = argv[2];
  char** a = argv + 3; int rc = 2;
  if (alg == "cacqr") {
    switch (atoi(a[9])) {
      case 0: rc = run_cacqr<cp::NoReplication>(a); break;
      case 1: rc = run_cacqr<cp::ReplicateCommComp>(a); break;
      case 2: rc = run_cacqr<cp::ReplicateComp>(a); break;
      default: rc = run_cacqr<cp::NoReplicationOverlap>(a); break;
    }
  } else if (alg == "cholinv") {
    switch (atoi(a[7])) {
      case 0: rc = run_cholinv<cp::NoReplication>(a); break;
      case 1: rc = run_cholinv<cp::ReplicateCommComp>(a); break;
      case 2: rc = run_cholinv<cp::ReplicateComp>(a); break;
      default: rc = run_cholinv<cp::NoReplicationOverlap>(a); break;
    }
  } else if (alg == "summa") rc = run_summa(a);
  else if (alg == "engine") rc = run_engine(a, argc - 3);
  else if (alg == "gen") rc = run_gen(a);
  MPI_Finalize();
  return rc;
}
