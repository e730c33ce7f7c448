// 3D SUMMA gemm driver, argv-compatible with the reference's bench/matmult/summa_gemm.cpp:19-25:
//   summa_gemm M N K pGridDimensionC layout num_chunks numIterations
#include "common.hpp"

int main(int argc, char** argv) {
  using T = double; using U = int64_t; using MatrixTypeR = matrix<T, U, rect>;
  if (argc < 8) { std::cerr << "usage: summa_gemm M N K pGridDimensionC layout num_chunks numIterations\n"; return 2; }
  auto& rt = capital::runtime::get();
  const U M = atoll(argv[1]), N = atoll(argv[2]), K = atoll(argv[3]), cdiv = atoll(argv[4]);
  const size_t layout = atoi(argv[5]), num_chunks = atoi(argv[6]), iters = atoi(argv[7]);
  topo::square SquareTopo(MPI_COMM_WORLD, benchutil::grid_depth(rt.size(), cdiv), layout, num_chunks);
  MatrixTypeR matA(K, M, SquareTopo.d, SquareTopo.d), matB(N, K, SquareTopo.d, SquareTopo.d), matC(N, M, SquareTopo.d, SquareTopo.d);
  blas::ArgPack_gemm<T> blasArgs(blas::Order::AblasColumnMajor, blas::Transpose::AblasNoTrans, blas::Transpose::AblasNoTrans, 1., 0.);
  matA.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, rt.rank() / SquareTopo.c);
  matB.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, rt.rank() / SquareTopo.c * (-1));
  matC.distribute_random(SquareTopo.x, SquareTopo.y, SquareTopo.d, SquareTopo.d, rt.rank() / SquareTopo.c * (-1));
  matmult::summa::invoke(matA, matB, matC, SquareTopo, blasArgs);   // warm-up
  for (size_t i = 0; i < iters; ++i) {
    const double sec = benchutil::timed([&] { matmult::summa::invoke(matA, matB, matC, SquareTopo, blasArgs); });
    if (rt.rank() == 0) std::cout << M << " " << N << " " << K << " " << sec << "   # " << 2. * M * N * K / sec * 1e-12 << " TFLOP/s on " << rt.size()
                                  << " GPU(s), grid " << SquareTopo.c << "x" << SquareTopo.d << "x" << SquareTopo.d << std::endl;
  }
  return 0;
}
