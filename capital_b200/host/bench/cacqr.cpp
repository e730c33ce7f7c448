// CholeskyQR / CholeskyQR2 driver, argv-compatible with the reference's bench/qr/cacqr.cpp:14-25:
//   cacqr variant num_rows num_columns rep_start rep_end complete_inv split bc_start bc_end layout num_chunks num_iter
// Prints "num_rows num_columns rep bc seconds" like the reference (bench/qr/cacqr.cpp:53) plus the FP64 rate.
#include "common.hpp"

int main(int argc, char** argv) {
  using T = double; using U = int64_t; using MatrixType = matrix<T, U, rect>;
  if (argc < 13) { std::cerr << "usage: cacqr variant rows cols rep_start rep_end complete_inv split bc_start bc_end layout num_chunks num_iter\n"; return 2; }
  auto& rt = capital::runtime::get();
  const size_t variant = atoi(argv[1]); const U num_rows = atoll(argv[2]), num_columns = atoll(argv[3]);
  const U rep_start = atoll(argv[4]), rep_end = atoll(argv[5]); const bool complete_inv = atoi(argv[6]); const U split = atoll(argv[7]);
  const U bc_start = atoll(argv[8]), bc_end = atoll(argv[9]); const size_t layout = atoi(argv[10]), num_chunks = atoi(argv[11]), num_iter = atoi(argv[12]);
  using qr_type = qr::cacqr<qr::policy::cacqr::Serialize, qr::policy::cacqr::SaveIntermediates>;
  using ci_type = cholesky::cholinv<cholesky::policy::cholinv::Serialize, cholesky::policy::cholinv::SaveIntermediates, cholesky::policy::cholinv::ReplicateComp>;
  for (U i = rep_start; i <= rep_end; ++i) {
    topo::rect RectTopo(MPI_COMM_WORLD, i, layout, num_chunks);
    MatrixType A(num_columns, num_rows, RectTopo.c, RectTopo.d);
    for (U j = bc_start; j <= bc_end; ++j) {
      ci_type::info<T, U> ci_pack(complete_inv, split, j, 'U');
      qr_type::info<T, U, ci_type> pack(variant, ci_pack);
      for (size_t k = 0; k < num_iter; ++k) {
        A.distribute_random(RectTopo.x, RectTopo.y, RectTopo.c, RectTopo.d, rt.rank() / RectTopo.c);
        rt.barrier();
        qr_type::factor(A, pack, RectTopo);
      }
      A.distribute_random(RectTopo.x, RectTopo.y, RectTopo.c, RectTopo.d, rt.rank() / RectTopo.c);
      const double sec = benchutil::timed([&] { qr_type::factor(A, pack, RectTopo); });
      const double m = (double)num_rows, n = (double)num_columns;
      const double flops = (variant == 2 ? 4. : 2.) * m * n * n + (variant == 2 ? 5. / 3. : 2. / 3.) * n * n * n;
      if (getenv("CAPITAL_B200_VALIDATE")) {   // the reference's test/qr/cacqr.cpp prints the same two metrics
        double orth, resid;
        qr::validate<qr_type>::both(A, pack, RectTopo, &orth, &resid);
        if (rt.rank() == 0) std::cout << "orthogonality ||Q^T Q - I||_F / n - " << orth << "   residual ||Q R - A||_F / ||A||_F - " << resid << std::endl;
      }
      if (rt.rank() == 0) std::cout << num_rows << " " << num_columns << " " << i << " " << j << " " << sec << "   # " << flops / sec * 1e-12 << " TFLOP/s on " << rt.size() << " GPU(s)" << std::endl;
    }
  }
  return 0;
}
