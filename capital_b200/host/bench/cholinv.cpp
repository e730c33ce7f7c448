// Recursive Cholesky-inverse driver, argv-compatible with the reference's bench/cholesky/cholinv.cpp:15-22:
//   cholinv num_rows rep_div complete_inv split bcMultiplier layout num_chunks num_iter
// Prints "total time - seconds" per iteration like the reference (:59) plus the N^3/3 rate.
#include "common.hpp"

int main(int argc, char** argv) {
  using T = double; using U = int64_t; using MatrixType = matrix<T, U, rect>;
  if (argc < 9) { std::cerr << "usage: cholinv num_rows rep_div complete_inv split bcMultiplier layout num_chunks num_iter\n"; return 2; }
  auto& rt = capital::runtime::get();
  const U num_rows = atoll(argv[1]), rep_div = atoll(argv[2]); const bool complete_inv = atoi(argv[3]); const U split = atoll(argv[4]);
  const long bcMultiplier = atol(argv[5]); const size_t layout = atoi(argv[6]), num_chunks = atoi(argv[7]), num_iter = atoi(argv[8]);
  using cholesky_type = cholesky::cholinv<cholesky::policy::cholinv::Serialize, cholesky::policy::cholinv::SaveIntermediates, cholesky::policy::cholinv::ReplicateComp>;
  topo::square SquareTopo(MPI_COMM_WORLD, benchutil::grid_depth(rt.size(), rep_div), layout, num_chunks);
  MatrixType A(num_rows, num_rows, SquareTopo.d, SquareTopo.d);
  A.distribute_symmetric(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, rt.rank() / SquareTopo.c, true);
  cholesky_type::info<T, U> pack(complete_inv, split, bcMultiplier, 'U');
  cholesky_type::factor(A, pack, SquareTopo);   // warm-up (allocates the library workspace)
  for (size_t i = 0; i < num_iter; ++i) {
    const double sec = benchutil::timed([&] { cholesky_type::factor(A, pack, SquareTopo); });
    const double n = (double)num_rows;
    if (rt.rank() == 0) std::cout << "total time - " << sec << "   # " << n * n * n / 3. / sec * 1e-12 << " TFLOP/s (N^3/3) on " << rt.size()
                                  << " GPU(s), grid " << SquareTopo.c << "x" << SquareTopo.d << "x" << SquareTopo.d << std::endl;
  }
  if (getenv("CAPITAL_B200_VALIDATE")) {   // the reference's test/cholesky/cholinv.cpp prints the same metric
    const double res = cholesky::validate<cholesky_type>::residual(A, pack, SquareTopo);
    if (rt.rank() == 0) std::cout << "residual ||R^T R - A||_F / ||A||_F (upper) - " << res << std::endl;
  }
  int info = 0; cap_get_info(rt.handle(), &info);
  if (info != 0) { std::cerr << "potrf info = " << info << std::endl; return 1; }
  return 0;
}
