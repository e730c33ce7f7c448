// Shared by the bench mains: argv layout and timing protocol of the reference's bench/ drivers
// (warm-up call, barrier, time one call, max over ranks -- bench/qr/cacqr.cpp:42-52).
#pragma once
#include <cmath>
#include <iostream>

#include "../capital.hpp"

namespace benchutil {
inline double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
// depth of the c x d x d grid: the reference's cube rule (bench/cholesky/cholinv.cpp:36-37) where it yields a valid
// grid, else the 1D/2D grids this build adds for 2 and 4 GPUs
inline size_t grid_depth(int size, long rep_div) {
  size_t cube = (size_t)std::nearbyint(std::ceil(std::pow((double)size, 1. / 3.)));
  size_t c = rep_div > 0 ? cube / rep_div : cube;
  if (c < 1) c = 1;
  size_t d = 1; while (c * (d + 1) * (d + 1) <= (size_t)size) ++d;
  if (c * d * d == (size_t)size && (d % c == 0 || d == 1)) return c;
  return size == 2 ? 2 : 1;
}
template <class F>
double timed(F&& f) {
  auto& rt = capital::runtime::get();
  rt.barrier();
  const double t0 = now();
  f();
  capital::check(cap_synchronize(rt.handle()), "cap_synchronize");
  return rt.max_over_ranks(now() - t0);
}
}  // namespace benchutil
