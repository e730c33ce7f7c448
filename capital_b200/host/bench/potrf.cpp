// Cholesky driver for the look-ahead schedule (cap_potrf_1d); no counterpart among the reference's bench mains
// (bench/cholesky/cholinv.cpp times the recursion that also forms R^-1).  Same timing protocol as the other drivers:
// warm-up, barrier, time one call, max over ranks.
//   potrf num_rows block_width num_iter
// Input: the reference's SPD generator for this rank's block columns; restored from a device copy before every call.
#include "common.hpp"

int main(int argc, char** argv) {
  if (argc < 4) { std::cerr << "usage: potrf num_rows block_width num_iter\n"; return 2; }
  auto& rt = capital::runtime::get();
  const int64_t N = atoll(argv[1]); const int nb = atoi(argv[2]); const int num_iter = atoi(argv[3]);
  const auto L = cholesky::potrf1d::make_layout(N, nb);
  const size_t bytes = sizeof(double) * (size_t)N * (size_t)(L.ncols_local > 0 ? L.ncols_local : 1);
  double *A0 = nullptr, *A = nullptr;
  capital::check(cap_malloc((void**)&A0, bytes), "cap_malloc");
  capital::check(cap_malloc((void**)&A, bytes), "cap_malloc");
  cholesky::potrf1d::distribute_symmetric(A0, N, L);
  auto restore = [&] { capital::check(cap_copy2d(rt.handle(), A, N, A0, N, N, L.ncols_local, 0), "cap_copy2d"); };
  restore(); cholesky::potrf1d::factor(A, N, L);             // warm-up (allocates the library workspace)
  for (int i = 0; i < num_iter; ++i) {
    restore();
    const double sec = benchutil::timed([&] { cholesky::potrf1d::factor(A, N, L); });
    const double n = (double)N;
    if (rt.rank() == 0) std::cout << "total time - " << sec << "   # " << n * n * n / 3. / sec * 1e-12 << " TFLOP/s (N^3/3) on " << rt.size()
                                  << " GPU(s), block columns of " << nb << " dealt 1-D" << std::endl;
  }
  if (getenv("CAPITAL_B200_VALIDATE") && rt.size() == 1) {   // ||R^T R - A||_F / ||A||_F over the upper triangle (test/cholesky/validate.hpp:7-30)
    double* Rm = nullptr; double res = -1;
    capital::check(cap_malloc((void**)&Rm, bytes), "cap_malloc");
    capital::check(cap_copy2d(rt.handle(), Rm, N, A, N, N, N, 3), "cap_copy2d");          // R with an explicit zero lower triangle
    topo::square One(MPI_COMM_WORLD, 1);
    capital::check(cap_validate_cholinv(rt.handle(), One.grid, N, A0, N, Rm, N, &res), "cap_validate_cholinv");
    std::cout << "residual ||R^T R - A||_F / ||A||_F (upper) - " << res << std::endl;
    cap_free(Rm);
  }
  int info = 0; cap_get_info(rt.handle(), &info);
  cap_free(A0); cap_free(A);
  if (info != 0) { std::cerr << "potrf info = " << info << std::endl; return 1; }
  return 0;
}
