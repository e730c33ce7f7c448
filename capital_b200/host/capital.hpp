// C++ host of capital-b200: the reference's engine / matrix / topology / algorithm entry points
// (huttered40/capital, src/) re-stated over the C ABI of include/capital_b200.h, so that code written
// against the reference -- e.g. its bench mains -- drives device-resident data on B200s.
//
//   blas::engine::_gemm/_trmm/_syrk      src/blas/interface.h:57-66, arg packs src/blas/engine.h:23-130
//   lapack::engine::_potrf/_trtri        src/lapack/interface.h:49-53, arg packs src/lapack/engine.h:23-84
//   matrix<T,U,Structure>                src/matrix/matrix.h:9-97 (local block of an element-cyclic matrix)
//   topo::square / topo::rect            src/util/topology.h:16-143
//   matmult::summa::invoke               src/alg/matmult/summa/summa.hpp:7-161
//   cholesky::cholinv<>::factor          src/alg/cholesky/cholinv/cholinv.hpp:6-47
//   cholesky::potrf1d, blas::engine::_trsm   additions without a reference counterpart (cap_potrf_1d, cap_dtrsm)
//   qr::cacqr<>::factor                  src/alg/qr/cacqr/cacqr.hpp:219-270
//
// Differences that a device back end forces: all matrix storage is DEVICE memory;
// triangular matrices are kept in full n x n storage with an explicit zero triangle (packed storage only
// through pack()/unpack()); one process per GPU, ranks come from torchrun-style environment variables and
// the NCCL id is exchanged through a file under /tmp (one node).  Header-only; link libcapital_b200.so.
#pragma once
#include <fcntl.h>
#include <unistd.h>

#include <ctime>

#include <cassert>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/capital_b200.h"

namespace capital {

inline void check(int rc, const char* what) {
  if (rc != CAP_OK) throw std::runtime_error(std::string(what) + ": " + cap_last_error());
}

// One process = one GPU.  RANK / WORLD_SIZE / LOCAL_RANK / MASTER_PORT as set by torch.distributed.run
// (or any launcher that exports them); a single process needs nothing.
class runtime {
public:
  static runtime& get() { static runtime r; return r; }
  cap_handle_t handle() const { return h_; }
  cap_comm_t world() const { return world_; }
  int rank() const { return rank_; }
  int size() const { return size_; }
  void barrier() {   // stream drain + a 1-element allreduce: the role of PMPI_Barrier in the bench mains
    if (size_ > 1) check(cap_allreduce_sum(h_, world_, token_, 1), "barrier");
    check(cap_synchronize(h_), "cap_synchronize");
  }
  double max_over_ranks(double v) {   // PMPI_Allreduce(MAX) of a host scalar, through the device
    if (size_ == 1) return v;
    std::vector<double> all(size_, 0.0);
    all[rank_] = v;
    check(cap_h2d_async(h_, gather_, all.data(), sizeof(double) * size_), "h2d");
    check(cap_allreduce_sum(h_, world_, gather_, size_), "allreduce");
    check(cap_d2h_async(h_, all.data(), gather_, sizeof(double) * size_), "d2h");
    check(cap_synchronize(h_), "sync");
    double m = 0; for (double x : all) m = x > m ? x : m;
    return m;
  }
  ~runtime() {
    if (token_) cap_free(token_);
    if (gather_) cap_free(gather_);
    if (world_) cap_comm_destroy(world_);
    if (h_) cap_destroy(h_);
  }

private:
  runtime() {
    auto env = [](const char* k, int dflt) { const char* v = getenv(k); return v ? atoi(v) : dflt; };
    rank_ = env("RANK", 0); size_ = env("WORLD_SIZE", 1);
    const int local = env("LOCAL_RANK", rank_);
    check(cap_create(&h_, local), "cap_create");
    check(cap_malloc((void**)&token_, 8), "cap_malloc");
    check(cap_memset_async(h_, token_, 0, 8), "memset");
    check(cap_malloc((void**)&gather_, 8 * (size_t)size_), "cap_malloc");
    if (size_ > 1) {
      char id[128];
      // One file per LAUNCH: the name carries the launcher's pid (the same torchrun agent forks every rank) next to the
      // port, rank 0 creates it with O_EXCL and mode 0600 after removing any left-over, writes a header with its own pid
      // and start time, and the other ranks accept only a complete file whose header is younger than their own start.
      const char* port = getenv("MASTER_PORT");
      const std::string path = "/tmp/capital_b200_ncclid_" + std::string(port ? port : "0") + "_" + std::to_string((long)getppid()) + "_" +
                               std::to_string((long)getuid());
      struct { long long magic, stamp; char id[128]; } rec;
      const long long now_s = (long long)time(nullptr);
      if (rank_ == 0) {
        check(cap_comm_unique_id(id), "cap_comm_unique_id");
        unlink(path.c_str());
        const std::string tmp = path + ".tmp";
        unlink(tmp.c_str());
        const int fd = open(tmp.c_str(), O_WRONLY | O_CREAT | O_EXCL, 0600);
        if (fd < 0) throw std::runtime_error("cannot create " + tmp);
        rec.magic = 0x43415042323030LL; rec.stamp = now_s; memcpy(rec.id, id, 128);
        const bool ok = write(fd, &rec, sizeof rec) == (ssize_t)sizeof rec;
        close(fd);
        if (!ok || rename(tmp.c_str(), path.c_str()) != 0) throw std::runtime_error("cannot write " + path);
      } else {
        bool got = false;
        for (int tries = 0; tries < 12000 && !got; ++tries) {          // two minutes, then an error instead of a hang
          FILE* f = fopen(path.c_str(), "rb");
          if (f) {
            got = fread(&rec, 1, sizeof rec, f) == sizeof rec && rec.magic == 0x43415042323030LL && rec.stamp >= now_s - 600;
            fclose(f);
          }
          if (!got) std::this_thread::sleep_for(std::chrono::milliseconds(10));
        }
        if (!got) throw std::runtime_error("no NCCL id from rank 0 within 120 s (" + path + ")");
        memcpy(id, rec.id, 128);
      }
      check(cap_comm_init_rank(&world_, id, size_, rank_), "cap_comm_init_rank");
      barrier();
      if (rank_ == 0) unlink(path.c_str());
    }
  }
  cap_handle_t h_ = nullptr;
  cap_comm_t world_ = nullptr;
  double *token_ = nullptr, *gather_ = nullptr;
  int rank_ = 0, size_ = 1;
};

inline cap_handle_t handle() { return runtime::get().handle(); }

// padXlen / padYlen of the reference's generators (src/matrix/structure.hpp:42-43,72-73,110-111): number of generated rows /
// columns of a local block of `dim` when the grid does not divide the global dimension.  Pure arithmetic: tests/test_host_cpp.py
// compiles it on the CPU and compares it with the oracle.
inline int64_t pad_len(int64_t dim, int64_t global, int64_t local_p, int64_t grid_p) {
  return ((global % grid_p != 0) && ((dim - 1) * grid_p + local_p >= global)) ? dim - 1 : dim;
}
inline int64_t local_dim(int64_t global, int64_t grid_p) { return (global + grid_p - 1) / grid_p; }   // matrix.hpp:142-150

// What stands where the reference passes an MPI communicator (topology.h:18,69): the ranks of the runtime.
struct world_t {};
constexpr world_t WORLD{};

}  // namespace capital

// Code written against the reference says topo::square(MPI_COMM_WORLD, c, layout, num_chunks): when no MPI header is in
// sight these two names make that line compile unchanged (a build with a real <mpi.h> passes capital::WORLD instead).
#if !defined(MPI_COMM_WORLD) && !defined(CAPITAL_B200_NO_MPI_NAMES)
using MPI_Comm = capital::world_t;
constexpr MPI_Comm MPI_COMM_WORLD{};
#endif

// ------------------------------------------------------------------------------------------ blas::engine
namespace blas {
enum class Order : unsigned char { AblasRowMajor = 0x0, AblasColumnMajor = 0x1 };
enum class Transpose : unsigned char { AblasNoTrans = 0x0, AblasTrans = 0x1 };
enum class Side : unsigned char { AblasLeft = 0x0, AblasRight = 0x1 };
enum class UpLo : unsigned char { AblasLower = 0x0, AblasUpper = 0x1 };
enum class Diag : unsigned char { AblasNonUnit = 0x0, AblasUnit = 0x1 };

template <typename T> struct ArgPack_gemm {
  ArgPack_gemm(Order o, Transpose ta, Transpose tb, T a, T b) : order(o), transposeA(ta), transposeB(tb), alpha(a), beta(b) {}
  Order order; Transpose transposeA, transposeB; T alpha, beta;
};
template <typename T> struct ArgPack_trmm {
  ArgPack_trmm(Order o, Side s, UpLo u, Transpose t, Diag d, T a) : order(o), side(s), uplo(u), transposeA(t), diag(d), alpha(a) {}
  Order order; Side side; UpLo uplo; Transpose transposeA; Diag diag; T alpha;
};
template <typename T> struct ArgPack_syrk {
  ArgPack_syrk(Order o, UpLo u, Transpose t, T a, T b) : order(o), uplo(u), transposeA(t), alpha(a), beta(b) {}
  Order order; UpLo uplo; Transpose transposeA; T alpha, beta;
};
inline char chr(Transpose t) { return t == Transpose::AblasTrans ? 'T' : 'N'; }
inline char chr(Side s) { return s == Side::AblasLeft ? 'L' : 'R'; }
inline char chr(UpLo u) { return u == UpLo::AblasUpper ? 'U' : 'L'; }
inline char chr(Diag d) { return d == Diag::AblasUnit ? 'U' : 'N'; }

// Same argument order as the reference's static engine; pointers are device pointers.  Only the column-major
// double case exists on the path (the reference specialises for double only, interface.hpp:43,62,82).
class engine {
public:
  static void _gemm(double* A, double* B, double* C, int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb, int64_t ldc,
                    const ArgPack_gemm<double>& p) {
    assert(p.order == Order::AblasColumnMajor);
    capital::check(cap_dgemm(capital::handle(), chr(p.transposeA), chr(p.transposeB), m, n, k, p.alpha, A, lda, B, ldb, p.beta, C, ldc), "cap_dgemm");
  }
  static void _trmm(double* A, double* B, int64_t m, int64_t n, int64_t lda, int64_t ldb, const ArgPack_trmm<double>& p) {
    assert(p.order == Order::AblasColumnMajor);
    capital::check(cap_dtrmm(capital::handle(), chr(p.side), chr(p.uplo), chr(p.transposeA), chr(p.diag), m, n, p.alpha, A, lda, B, ldb), "cap_dtrmm");
  }
  static void _syrk(double* A, double* C, int64_t n, int64_t k, int64_t lda, int64_t ldc, const ArgPack_syrk<double>& p) {
    assert(p.order == Order::AblasColumnMajor);
    capital::check(cap_dsyrk(capital::handle(), chr(p.uplo), chr(p.transposeA), n, k, p.alpha, A, lda, p.beta, C, ldc), "cap_dsyrk");
  }
  // Not a member of the reference's engine (src/blas/interface.h:57-66 has no _trsm; its schedules multiply by explicit
  // inverses, cholinv.hpp:113-124): op(A) X = alpha B (Left) or X op(A) = alpha B (Right), in place on B; takes the same
  // argument pack as _trmm.  Non-unit diagonal only.
  static void _trsm(double* A, double* B, int64_t m, int64_t n, int64_t lda, int64_t ldb, const ArgPack_trmm<double>& p) {
    assert(p.order == Order::AblasColumnMajor);
    capital::check(cap_dtrsm(capital::handle(), chr(p.side), chr(p.uplo), chr(p.transposeA), chr(p.diag), m, n, p.alpha, A, lda, B, ldb), "cap_dtrsm");
  }
};
}  // namespace blas

// ------------------------------------------------------------------------------------------ lapack::engine
namespace lapack {
enum class Order : unsigned char { AlapackRowMajor = 0x0, AlapackColumnMajor = 0x1 };
enum class UpLo : unsigned char { AlapackLower = 0x0, AlapackUpper = 0x1 };
enum class Diag : unsigned char { AlapackNonUnit = 0x0, AlapackUnit = 0x1 };
struct ArgPack_potrf { ArgPack_potrf(Order o, UpLo u) : order(o), uplo(u) {} Order order; UpLo uplo; };
struct ArgPack_trtri { ArgPack_trtri(Order o, UpLo u, Diag d) : order(o), uplo(u), diag(d) {} Order order; UpLo uplo; Diag diag; };
class engine {
public:
  // LAPACK `info` stays on the device (cap_get_info); the reference discards it (lapack/interface.hpp:39,54)
  static void _potrf(double* A, int n, int lda, const ArgPack_potrf& p) {
    assert(p.order == Order::AlapackColumnMajor);
    capital::check(cap_dpotrf(capital::handle(), p.uplo == UpLo::AlapackUpper ? 'U' : 'L', n, A, lda), "cap_dpotrf");
  }
  static void _trtri(double* A, int n, int lda, const ArgPack_trtri& p) {
    assert(p.order == Order::AlapackColumnMajor);
    capital::check(cap_dtrtri(capital::handle(), p.uplo == UpLo::AlapackUpper ? 'U' : 'L', p.diag == Diag::AlapackUnit ? 'U' : 'N', n, A, lda), "cap_dtrtri");
  }
};
}  // namespace lapack

// ------------------------------------------------------------------------------------------ topology
namespace topo {
// c x d x d grid (topology.h:67-143, layout 0); P = c*d*d with d % c == 0 or d == 1
class square {
public:
  // topology.h:69: square(MPI_Comm comm, size_t c, size_t layout = 0, size_t num_chunks = 0).  layout 0 / 1 are the
  // reference's two rank orders; num_chunks is the number of pieces per panel of the pipelined SUMMA (0 = automatic).
  square(capital::world_t, size_t c_, size_t layout_ = 0, size_t num_chunks_ = 0) : layout(layout_), num_chunks(num_chunks_) {
    auto& rt = capital::runtime::get();
    rank = rt.rank(); size = rt.size();
    capital::check(cap_grid_create2(rt.world(), (int)c_, (int)layout_, (int)num_chunks_, &grid), "cap_grid_create");
    int ci, di, xi, yi, zi;
    cap_grid_info(grid, &ci, &di, &xi, &yi, &zi);
    c = ci; d = di; x = xi; y = yi; z = zi;
  }
  explicit square(size_t c_, size_t layout_ = 0, size_t num_chunks_ = 0) : square(capital::WORLD, c_, layout_, num_chunks_) {}
  ~square() { cap_grid_destroy(grid); }
  square(const square&) = delete;
  cap_grid_t grid = nullptr;
  int rank, size;
  size_t c, d, x, y, z, layout, num_chunks;
};
// Tall-skinny grid of the CholeskyQR path (topology.h:16-65).  c == 1: the 1D grid (d = P, cacqr.hpp:229);
// c > 1 with c^3 == P: the cubic grid of the 3D algorithm (c == d, cacqr.hpp:232), backed by the same cap_grid_t as
// topo::square.  Other c x d x c shapes of the reference's tunable variant: use topo::square(c), whose c x d x d
// grids (d % c == 0 or d == 1) are what cap_cacqr_factor_3d generalises to.
class rect {
public:
  // topology.h:18: rect(MPI_Comm comm, size_t c, size_t layout = 0, size_t num_chunks = 0)
  rect(capital::world_t, size_t c_, size_t layout_ = 0, size_t num_chunks_ = 0) : layout(layout_), num_chunks(num_chunks_) {
    auto& rt = capital::runtime::get();
    rank = rt.rank(); size = rt.size();
    if (c_ <= 1) { c = 1; d = size; x = 0; y = rank; z = 0; return; }
    if (c_ * c_ * c_ != (size_t)size) throw std::runtime_error("topo::rect: c > 1 needs c^3 ranks here; use topo::square(c) for c x d x d grids");
    capital::check(cap_grid_create2(rt.world(), (int)c_, (int)layout_, (int)num_chunks_, &grid), "cap_grid_create");
    int ci, di, xi, yi, zi;
    cap_grid_info(grid, &ci, &di, &xi, &yi, &zi);
    c = ci; d = di; x = xi; y = yi; z = zi;
  }
  explicit rect(size_t c_, size_t layout_ = 0, size_t num_chunks_ = 0) : rect(capital::WORLD, c_, layout_, num_chunks_) {}
  ~rect() { if (grid) cap_grid_destroy(grid); }
  rect(const rect&) = delete;
  cap_grid_t grid = nullptr;
  int rank, size;
  size_t c, d, x, y, z, layout, num_chunks;
};
}  // namespace topo

// ------------------------------------------------------------------------------------------ matrix
struct rect {};
struct uppertri {};
struct lowertri {};

// Local block of an element-cyclic distributed matrix (matrix.h:9-97) in DEVICE memory, column-major,
// leading dimension = local rows.  X = columns, Y = rows, as everywhere in the reference.  Triangular
// structures keep full storage (zeros in the other triangle); num_elems() is the full local count.
template <typename ScalarT = double, typename DimensionT = int64_t, typename StructurePolicy = rect>
class matrix {
  static_assert(sizeof(ScalarT) == sizeof(double), "FP64 only: the reference specialises its engines for double only");
public:
  using ScalarType = ScalarT; using DimensionType = DimensionT; using StructureType = StructurePolicy;
  matrix() = default;
  matrix(DimensionT globalDimensionX, DimensionT globalDimensionY, int64_t globalPgridX, int64_t globalPgridY) {
    _register_(globalDimensionX, globalDimensionY, globalPgridX, globalPgridY);
  }
  // injection constructor (matrix.hpp:53-74): wrap a caller-owned device pointer
  matrix(ScalarT* data, DimensionT dimensionX, DimensionT dimensionY, DimensionT globalPgridX, DimensionT globalPgridY)
      : data_(data), owned_(false), dimX_(dimensionX), dimY_(dimensionY), gX_(dimensionX * globalPgridX), gY_(dimensionY * globalPgridY), ld_(dimensionY) {}
  matrix(const matrix& rhs) { copy_from(rhs); }
  matrix(matrix&& rhs) noexcept { steal(rhs); }
  matrix& operator=(const matrix& rhs) { if (this != &rhs) { _destroy_(); copy_from(rhs); } return *this; }
  matrix& operator=(matrix&& rhs) noexcept { if (this != &rhs) { _destroy_(); steal(rhs); } return *this; }
  ~matrix() { _destroy_(); }

  void _register_(DimensionT globalDimensionX, DimensionT globalDimensionY, int64_t globalPgridX, int64_t globalPgridY) {
    _destroy_();
    gX_ = globalDimensionX; gY_ = globalDimensionY;
    dimX_ = (globalDimensionX + globalPgridX - 1) / globalPgridX;     // ceil: zero padding past the edge (matrix.hpp:142-150)
    dimY_ = (globalDimensionY + globalPgridY - 1) / globalPgridY;
    ld_ = dimY_;
    _fill_();
  }
  // matrix.hpp:127-139: second half of the two-part construction -- allocate (zero-filled) unless a pointer was injected
  void _fill_() {
    if (data_ || dimX_ * dimY_ == 0) return;
    capital::check(cap_malloc((void**)&data_, sizeof(double) * (size_t)dimX_ * dimY_), "cap_malloc");
    capital::check(cap_memset_async(capital::handle(), data_, 0, sizeof(double) * (size_t)dimX_ * dimY_), "cap_memset_async");
    owned_ = true;
  }
  void _destroy_() {
    if (saved_) _derestrict_();
    if (owned_ && data_) cap_free(data_);
    if (scratch_) cap_free(scratch_);
    if (pad_) cap_free(pad_);
    data_ = scratch_ = pad_ = nullptr; owned_ = false;
  }
  // matrix.hpp:172-180: work on the sub-block [startX, endX) x [startY, endY) of the local block.  Here a restricted
  // matrix is a (pointer, leading dimension) view: data() moves, ld() stays, the local dimensions shrink.
  void _restrict_(DimensionT startX, DimensionT endX, DimensionT startY, DimensionT endY) {
    assert(!saved_ && startX >= 0 && endX <= dimX_ && startY >= 0 && endY <= dimY_);
    sv_data_ = data_; sv_scratch_ = scratch_; sv_dimX_ = dimX_; sv_dimY_ = dimY_; saved_ = true;
    data_ += startY + startX * ld_;
    if (scratch_) scratch_ += startY + startX * ld_;
    dimX_ = endX - startX; dimY_ = endY - startY;
  }
  void _derestrict_() {
    if (!saved_) return;
    data_ = sv_data_; scratch_ = sv_scratch_; dimX_ = sv_dimX_; dimY_ = sv_dimY_; saved_ = false;
  }

  ScalarT*& data() { return data_; }
  ScalarT* data() const { return data_; }
  // matrix.h:41-44: second (and, for triangular structures, third) buffer of the same size for one-to-all / all-to-one
  // exchanges; device memory, allocated on first use
  ScalarT*& scratch() { if (!scratch_) alloc_like(&scratch_); return scratch_; }
  ScalarT* scratch() const { return scratch_; }
  ScalarT*& pad() { if (!pad_) alloc_like(&pad_); return pad_; }
  ScalarT* pad() const { return pad_; }
  void swap() { ScalarT* t = data(); data() = scratch(); scratch() = t; }          // matrix.h:55
  void swap_pad() { ScalarT* t = scratch(); scratch() = pad(); pad() = t; }        // matrix.h:56
  DimensionT ld() const { return ld_; }                                            // leading dimension of data() (column-major)
  DimensionT num_elems() const { return dimX_ * dimY_; }
  DimensionT num_elems(DimensionT rangeX, DimensionT rangeY) const { return rangeX * rangeY; }
  DimensionT num_rows_local() const { return dimY_; }
  DimensionT num_columns_local() const { return dimX_; }
  DimensionT num_rows_global() const { return gY_; }
  DimensionT num_columns_global() const { return gX_; }
  DimensionT offset_local(DimensionT coordX, DimensionT coordY, size_t = 0) const { return coordY + coordX * ld_; }
  void set_num_rows_local(DimensionT v) { dimY_ = v; }
  void set_num_columns_local(DimensionT v) { dimX_ = v; }
  void set_num_rows_global(DimensionT v) { gY_ = v; }
  void set_num_columns_global(DimensionT v) { gX_ = v; }

  // Edge padding as in the reference (structure.hpp:42-43,72-73,110-111): when the grid does not divide a global dimension,
  // the ranks past the remainder hold one zero row / column at the end of their block and draw one value fewer per column.
  DimensionT pad_len(DimensionT dim, DimensionT global, int64_t local_p, int64_t grid_p) const {
    return (DimensionT)capital::pad_len((int64_t)dim, (int64_t)global, local_p, grid_p);
  }
  // structure.hpp:106-129: element t of the UNPADDED local block (column by column) = t-th drand48 value after srand48(key)
  void distribute_random(int64_t localPgridX, int64_t localPgridY, int64_t globalPgridX, int64_t globalPgridY, int64_t key) {
    const DimensionT px = pad_len(dimX_, gX_, localPgridX, globalPgridX), py = pad_len(dimY_, gY_, localPgridY, globalPgridY);
    zero_all();
    capital::check(cap_fill_random(capital::handle(), data_, ld_, py, px, key), "cap_fill_random");
  }
  // structure.hpp:69-103: a(gy,gx) seeded by max + N*min, + N on the diagonal when diagonallyDominant
  void distribute_symmetric(int64_t localPgridX, int64_t localPgridY, int64_t globalPgridX, int64_t globalPgridY, int64_t, bool diagonallyDominant) {
    const DimensionT px = pad_len(dimX_, gX_, localPgridX, globalPgridX), py = pad_len(dimY_, gY_, localPgridY, globalPgridY);
    zero_all();
    capital::check(cap_fill_symmetric(capital::handle(), data_, ld_, py, px, localPgridX, localPgridY, globalPgridX, globalPgridY,
                                      gY_, diagonallyDominant ? 1 : 0), "cap_fill_symmetric");
  }
  // structure.hpp:37-66: val where local row == local column on the processes of the grid diagonal, zero elsewhere
  void distribute_identity(int64_t localPgridX, int64_t localPgridY, int64_t globalPgridX, int64_t globalPgridY, ScalarT val = 1.) {
    const DimensionT px = pad_len(dimX_, gX_, localPgridX, globalPgridX), py = pad_len(dimY_, gY_, localPgridY, globalPgridY);
    zero_all();
    if (localPgridX == localPgridY) capital::check(cap_fill_diagonal(capital::handle(), data_, ld_, py < px ? py : px, val), "cap_fill_diagonal");
  }
  std::vector<double> to_host() const {
    std::vector<double> out((size_t)dimX_ * dimY_);
    if (ld_ == dimY_) capital::check(cap_d2h_async(capital::handle(), out.data(), data_, sizeof(double) * out.size()), "cap_d2h_async");
    else for (DimensionT c = 0; c < dimX_; ++c)
      capital::check(cap_d2h_async(capital::handle(), out.data() + (size_t)c * dimY_, data_ + (size_t)c * ld_, sizeof(double) * dimY_), "cap_d2h_async");
    capital::check(cap_synchronize(capital::handle()), "cap_synchronize");
    return out;
  }

private:
  void alloc_like(ScalarT** p) {
    const size_t n = (size_t)(saved_ ? sv_dimX_ : dimX_) * ld_;
    capital::check(cap_malloc((void**)p, sizeof(double) * n), "cap_malloc");
    capital::check(cap_memset_async(capital::handle(), *p, 0, sizeof(double) * n), "cap_memset_async");
  }
  void zero_all() {
    if (ld_ == dimY_) capital::check(cap_memset_async(capital::handle(), data_, 0, sizeof(double) * (size_t)dimX_ * dimY_), "cap_memset_async");
    else capital::check(cap_axpby2d(capital::handle(), data_, ld_, data_, ld_, dimY_, dimX_, 0.0, 0.0), "cap_axpby2d");
  }
  void copy_from(const matrix& rhs) {
    dimX_ = rhs.dimX_; dimY_ = rhs.dimY_; gX_ = rhs.gX_; gY_ = rhs.gY_; ld_ = rhs.dimY_; owned_ = false; data_ = scratch_ = pad_ = nullptr; saved_ = false;
    if (!rhs.data_) return;
    capital::check(cap_malloc((void**)&data_, sizeof(double) * (size_t)dimX_ * dimY_), "cap_malloc");
    owned_ = true;
    capital::check(cap_copy2d(capital::handle(), data_, ld_, rhs.data_, rhs.ld_, dimY_, dimX_, 0), "cap_copy2d");
  }
  void steal(matrix& rhs) {
    data_ = rhs.data_; scratch_ = rhs.scratch_; pad_ = rhs.pad_; owned_ = rhs.owned_; dimX_ = rhs.dimX_; dimY_ = rhs.dimY_; gX_ = rhs.gX_; gY_ = rhs.gY_; ld_ = rhs.ld_;
    saved_ = rhs.saved_; sv_data_ = rhs.sv_data_; sv_scratch_ = rhs.sv_scratch_; sv_dimX_ = rhs.sv_dimX_; sv_dimY_ = rhs.sv_dimY_;
    rhs.data_ = rhs.scratch_ = rhs.pad_ = nullptr; rhs.owned_ = false; rhs.saved_ = false;
  }
  ScalarT *data_ = nullptr, *scratch_ = nullptr, *pad_ = nullptr;
  bool owned_ = false;
  DimensionT dimX_ = 0, dimY_ = 0, gX_ = 0, gY_ = 0, ld_ = 0;
  bool saved_ = false;                       // _restrict_ state
  ScalarT *sv_data_ = nullptr, *sv_scratch_ = nullptr;
  DimensionT sv_dimX_ = 0, sv_dimY_ = 0;
};

// ------------------------------------------------------------------------------------------ matmult::summa
namespace matmult {
class summa {
public:
  // gemm (summa.hpp:7-44)
  template <typename MA, typename MB, typename MC, typename CommType>
  static void invoke(MA& A, MB& B, MC& C, CommType&& ci, blas::ArgPack_gemm<double>& p) {
    const bool ta = p.transposeA == blas::Transpose::AblasTrans, tb = p.transposeB == blas::Transpose::AblasTrans;
    const int64_t m = ta ? A.num_columns_local() : A.num_rows_local(), n = tb ? B.num_rows_local() : B.num_columns_local();
    const int64_t k = ta ? A.num_rows_local() : A.num_columns_local();
    capital::check(cap_summa_gemm(capital::handle(), ci.grid, blas::chr(p.transposeA), blas::chr(p.transposeB), m, n, k, p.alpha,
                                  A.data(), A.ld(), B.data(), B.ld(), p.beta, C.data(), C.ld()), "cap_summa_gemm");
  }
  // trmm (summa.hpp:47-83): A triangular, B rectangular and overwritten
  template <typename MA, typename MB, typename CommType>
  static void invoke(MA& A, MB& B, CommType&& ci, blas::ArgPack_trmm<double>& p) {
    capital::check(cap_summa_trmm(capital::handle(), ci.grid, blas::chr(p.side), blas::chr(p.uplo), blas::chr(p.transposeA), blas::chr(p.diag),
                                  B.num_rows_local(), B.num_columns_local(), p.alpha, A.data(), A.ld(), B.data(), B.ld()), "cap_summa_trmm");
  }
  // syrk (summa.hpp:86-96): the grid-transposed copy is made inside
  template <typename MS, typename MD, typename CommType>
  static void invoke(MS& A, MD& C, CommType&& ci, blas::ArgPack_syrk<double>& p) {
    const bool tr = p.transposeA == blas::Transpose::AblasTrans;
    capital::check(cap_summa_syrk(capital::handle(), ci.grid, blas::chr(p.uplo), blas::chr(p.transposeA), C.num_columns_local(),
                                  tr ? A.num_rows_local() : A.num_columns_local(), p.alpha, A.data(), A.ld(), nullptr, 0, p.beta,
                                  C.data(), C.ld()), "cap_summa_syrk");
  }
};
}  // namespace matmult

// ------------------------------------------------------------------------------------------ cholesky::potrf1d
// Cholesky (R only, A = R^T R) over the ranks of the runtime with the look-ahead schedule of cap_potrf_1d.  NOT a class
// of the reference: it is the B200-native route to the Cholesky target of BASELINE.json (north_star: "on a 2x2x2 (or
// 1D/2D) process grid") when the inverse factor that cholesky::cholinv also produces is not wanted.  The matrix lives in
// its own distribution: block columns of width nb, block j on rank (j / P even ? j % P : P-1 - j % P), stored side by
// side as an N x ncols_local column-major array.
namespace cholesky {
class potrf1d {
public:
  struct layout {
    int64_t N = 0; int nb = 0, ranks = 1, rank = 0;
    int64_t ncols_local = 0;
    std::vector<int64_t> blocks;                            // global block-column index of every local block, in local order
    int64_t width(int64_t j) const { const int64_t w = N - j * nb; return w < nb ? w : nb; }
  };
  static layout make_layout(int64_t N, int nb) {
    auto& rt = capital::runtime::get();
    layout L; L.N = N; L.nb = nb; L.ranks = rt.size(); L.rank = rt.rank();
    int nblocks = 0;
    capital::check(cap_potrf_1d_layout(N, nb, L.ranks, L.rank, &L.ncols_local, &nblocks), "cap_potrf_1d_layout");
    const int64_t nblk = (N + nb - 1) / nb;
    for (int64_t j = 0; j < nblk; ++j) {
      int owner = 0, local = 0;
      capital::check(cap_potrf_1d_block(j, L.ranks, &owner, &local), "cap_potrf_1d_block");
      if (owner == L.rank) L.blocks.push_back(j);
    }
    assert((int)L.blocks.size() == nblocks);
    return L;
  }
  // the reference's SPD generator (rect::_distribute_symmetric, src/matrix/structure.hpp:69-103) for this rank's columns.
  // The generator adds N to the diagonal only where LOCAL row == LOCAL column (structure.hpp:87-89), so a shifted block
  // gets its diagonal square generated by a second call whose local origin is the global diagonal.
  static void distribute_symmetric(double* A_loc, int64_t lda, const layout& L) {
    for (size_t r = 0; r < L.blocks.size(); ++r) {
      const int64_t j = L.blocks[r], w = L.width(j), c0 = (int64_t)r * L.nb;
      double* col = A_loc + c0 * lda;
      capital::check(cap_fill_symmetric(capital::handle(), col, lda, L.N, w, j * L.nb, 0, 1, 1, L.N, 1), "cap_fill_symmetric");
      if (j > 0) capital::check(cap_fill_symmetric(capital::handle(), col + j * L.nb, lda, w, w, j * L.nb, j * L.nb, 1, 1, L.N, 1), "cap_fill_symmetric");
    }
  }
  // in place: the upper triangle of the global matrix becomes R; nothing below the diagonal is written
  static void factor(double* A_loc, int64_t lda, const layout& L) {
    auto& rt = capital::runtime::get();
    capital::check(cap_reset_info(rt.handle()), "cap_reset_info");
    capital::check(cap_potrf_1d(rt.handle(), rt.size() > 1 ? rt.world() : nullptr, L.N, L.nb, A_loc, lda), "cap_potrf_1d");
  }
};
}  // namespace cholesky

// ------------------------------------------------------------------------------------------ cholesky::cholinv
namespace cholesky {
namespace policy { namespace cholinv {
// Policy tags of the reference (cholinv/policy.h).  Storage and buffering choices are the library's here;
// every base-case tag maps to the replicated base case (ReplicateComp, policy.h:226-305) -- the reference's
// NoReplication variants return wrong factors on d > 1 grids (header).
struct Serialize { using structure = uppertri; }; struct NoSerialize { using structure = rect; };
struct SaveIntermediates {}; struct FlushIntermediates {};
struct ReplicateCommComp {}; struct ReplicateComp {}; struct NoReplication {}; struct NoReplicationOverlap {};
}}  // namespace policy::cholinv

template <class SerializePolicy = policy::cholinv::Serialize, class IntermediatesPolicy = policy::cholinv::SaveIntermediates,
          class BaseCasePolicy = policy::cholinv::ReplicateComp>
class cholinv {
public:
  template <typename ScalarT, typename DimensionT>
  class info {
  public:
    using ScalarType = ScalarT; using DimensionType = DimensionT;
    using alg_type = cholinv<SerializePolicy, IntermediatesPolicy, BaseCasePolicy>;
    info(bool complete_inv_, DimensionT split_, int64_t bc_mult_dim_, char dir_) : complete_inv(complete_inv_), split(split_), bc_mult_dim(bc_mult_dim_), dir(dir_) {}
    bool complete_inv; DimensionT split; int64_t bc_mult_dim; char dir;
    matrix<ScalarT, DimensionT, uppertri> R, Rinv;      // full local storage, zeros outside the triangle
    int lapack_info = 0;
  };

  // cholinv.hpp:6-29.  Base-case size follows cholinv.hpp:15-18: bcDimLocal = c*d scaled by 2^-bc_mult_dim
  // (negative bc_mult_dim = smaller base case), clamped to [1, l]; recursion stops at global size d*l/bcDimLocal.
  template <typename MatrixType, typename ArgType, typename CommType>
  static void factor(const MatrixType& A, ArgType& args, CommType&& ci) {
    assert(args.split > 0); assert(args.dir == 'U');
    const int64_t l = A.num_rows_local();
    int64_t bc = (int64_t)(ci.c * ci.d), mult = args.bc_mult_dim;
    if (mult < 0) { for (int64_t i = 0; i < -mult; ++i) bc *= 2; } else { for (int64_t i = 0; i < mult; ++i) bc /= 2; }
    bc = bc < 1 ? 1 : bc; bc = bc > l ? l : bc;
    const int64_t bc_global = (int64_t)ci.d * (l / bc);
    args.R = matrix<double, typename ArgType::DimensionType, uppertri>(A.num_columns_global(), A.num_rows_global(), ci.d, ci.d);
    args.Rinv = matrix<double, typename ArgType::DimensionType, uppertri>(A.num_columns_global(), A.num_rows_global(), ci.d, ci.d);
    capital::check(cap_copy2d(capital::handle(), args.R.data(), l, A.data(), l, l, l, 0), "cap_copy2d");
    capital::check(cap_reset_info(capital::handle()), "cap_reset_info");
    capital::check(cap_cholinv_factor(capital::handle(), ci.grid, l, args.R.data(), l, args.Rinv.data(), l, args.complete_inv ? 1 : 0,
                                      (int)args.split, bc_global), "cap_cholinv_factor");
  }
  template <typename ArgType, typename CommType>
  static matrix<double, typename ArgType::DimensionType, rect> construct_R(ArgType& args, CommType&&) { return to_rect(args.R); }
  template <typename ArgType, typename CommType>
  static matrix<double, typename ArgType::DimensionType, rect> construct_Rinv(ArgType& args, CommType&&) { return to_rect(args.Rinv); }

private:
  template <typename M>
  static matrix<double, typename M::DimensionType, rect> to_rect(const M& src) {
    matrix<double, typename M::DimensionType, rect> out(src.num_columns_local(), src.num_rows_local(), 1, 1);
    capital::check(cap_copy2d(capital::handle(), out.data(), src.num_rows_local(), src.data(), src.num_rows_local(), src.num_rows_local(),
                              src.num_columns_local(), 0), "cap_copy2d");
    return out;
  }
};

// cholesky::validate (test/cholesky/validate.hpp:7-30): ||R^T R - A||_F / ||A||_F over the upper triangle, on the device
template <class AlgType>
struct validate {
  template <typename MatrixType, typename ArgType, typename CommType>
  static double residual(const MatrixType& A, ArgType& args, CommType&& ci) {
    double r = 0;
    capital::check(cap_validate_cholinv(capital::handle(), ci.grid, A.num_rows_local(), A.data(), A.ld(),
                                        args.R.data(), args.R.num_rows_local(), &r), "cap_validate_cholinv");
    return r;
  }
};
}  // namespace cholesky

// ------------------------------------------------------------------------------------------ qr::cacqr
namespace qr {
namespace policy { namespace cacqr {
struct Serialize { using structure = uppertri; }; struct NoSerialize { using structure = rect; };   // packed vs full Gram allreduce (policy.h:9-85)
struct SaveIntermediates {}; struct FlushIntermediates {};
}}  // namespace policy::cacqr

template <class SerializePolicy = policy::cacqr::Serialize, class IntermediatesPolicy = policy::cacqr::SaveIntermediates>
class cacqr {
public:
  template <typename ScalarT, typename DimensionT, typename CholeskyInversionType>
  class info {
  public:
    using ScalarType = ScalarT; using DimensionType = DimensionT;
    using alg_type = cacqr<SerializePolicy, IntermediatesPolicy>;
    using cholesky_inverse_type = CholeskyInversionType;
    template <typename CholeskyInversionArgType>
    info(size_t num_iter_, CholeskyInversionArgType&& ci_args) : num_iter(num_iter_), cholesky_inverse_args(ci_args) {}
    const size_t num_iter;                                   // 1 = CholeskyQR, 2 = CholeskyQR2
    typename CholeskyInversionType::template info<ScalarT, DimensionT> cholesky_inverse_args;
    matrix<ScalarT, DimensionT, rect> Q;
    matrix<ScalarT, DimensionT, uppertri> R;                 // n x n, full storage
  };

  // cacqr.hpp:219-248: invoke_1d (:174-193) on the 1D grid, invoke_3d (:197-216) on a process grid (topo::rect with
  // c > 1 or any topo::square).  A is untouched, Q and R are (re)allocated on first use.  3D: A is the local block of
  // matrix(n, m, d, d) -- columns dealt over x, rows over y -- and R the local n/d x n/d block of the factor.
  template <typename MatrixType, typename ArgType, typename CommType>
  static void factor(const MatrixType& A, ArgType& args, CommType&& ci) {
    const int64_t m = A.num_rows_local(), n = A.num_columns_local();
    auto& rt = capital::runtime::get();
    if (ci.grid == nullptr) {
      if (!args.Q.data() || args.Q.num_rows_local() != m || args.Q.num_columns_local() != n)
        args.Q = matrix<double, typename ArgType::DimensionType, rect>(A.num_columns_global(), A.num_rows_global(), ci.c, ci.d);
      if (!args.R.data() || args.R.num_rows_local() != n)
        args.R = matrix<double, typename ArgType::DimensionType, uppertri>(n, n, 1, 1);
      capital::check(cap_cacqr_factor_1d(rt.handle(), rt.world(), m, n, A.data(), m, args.Q.data(), m, args.R.data(), n, (int)args.num_iter), "cap_cacqr_factor_1d");
      return;
    }
    if (!args.Q.data() || args.Q.num_rows_local() != m || args.Q.num_columns_local() != n)
      args.Q = matrix<double, typename ArgType::DimensionType, rect>(A.num_columns_global(), A.num_rows_global(), ci.d, ci.d);
    if (!args.R.data() || args.R.num_rows_local() != n)
      args.R = matrix<double, typename ArgType::DimensionType, uppertri>(A.num_columns_global(), A.num_columns_global(), ci.d, ci.d);
    auto& ca = args.cholesky_inverse_args;
    int64_t bc = (int64_t)(ci.c * ci.d), mult = ca.bc_mult_dim;          // base-case rule of cholinv.hpp:15-18
    if (mult < 0) { for (int64_t i = 0; i < -mult; ++i) bc *= 2; } else { for (int64_t i = 0; i < mult; ++i) bc /= 2; }
    bc = bc < 1 ? 1 : bc; bc = bc > n ? n : bc;
    const int64_t bc_global = (int64_t)ci.d * (n / bc);
    capital::check(cap_reset_info(rt.handle()), "cap_reset_info");
    capital::check(cap_set_cacqr_complete_inv(rt.handle(), ca.complete_inv ? 1 : 0), "cap_set_cacqr_complete_inv");   // 0: cacqr::solve (cacqr.hpp:46-73)
    capital::check(cap_cacqr_factor_3d(rt.handle(), ci.grid, m, n, A.data(), m, args.Q.data(), m, args.R.data(), n, nullptr, 0,
                                       (int)args.num_iter, (int)ca.split, bc_global), "cap_cacqr_factor_3d");
  }
  template <typename ArgType, typename CommType>
  static matrix<double, typename ArgType::DimensionType, rect> construct_Q(ArgType& args, CommType&&) { return args.Q; }
  template <typename ArgType, typename CommType>
  static matrix<double, typename ArgType::DimensionType, rect> construct_R(ArgType& args, CommType&&) {
    matrix<double, typename ArgType::DimensionType, rect> out(args.R.num_columns_local(), args.R.num_rows_local(), 1, 1);
    capital::check(cap_copy2d(capital::handle(), out.data(), args.R.num_rows_local(), args.R.data(), args.R.num_rows_local(),
                              args.R.num_rows_local(), args.R.num_columns_local(), 0), "cap_copy2d");
    return out;
  }
};

// qr::validate (test/qr/validate.hpp:7-62): ||Q^T Q - I||_F / n and ||Q R - A||_F / ||A||_F, on the device
template <class AlgType>
struct validate {
  template <typename MatrixType, typename ArgType, typename CommType>
  static void both(const MatrixType& A, ArgType& args, CommType&& ci, double* orth, double* resid) {
    auto& rt = capital::runtime::get();
    const int64_t m = A.num_rows_local(), n = A.num_columns_local();
    if (ci.grid == nullptr)
      capital::check(cap_validate_cacqr_1d(rt.handle(), rt.world(), m, n, A.data(), m, args.Q.data(), m, args.R.data(), n, orth, resid), "cap_validate_cacqr_1d");
    else
      capital::check(cap_validate_cacqr_3d(rt.handle(), ci.grid, m, n, A.data(), m, args.Q.data(), m, args.R.data(), n, orth, resid), "cap_validate_cacqr_3d");
  }
  template <typename MatrixType, typename ArgType, typename CommType>
  static double orthogonality(const MatrixType& A, ArgType& args, CommType&& ci) { double o, r; both(A, args, ci, &o, &r); return o; }
  template <typename MatrixType, typename ArgType, typename CommType>
  static double residual(const MatrixType& A, ArgType& args, CommType&& ci) { double o, r; both(A, args, ci, &o, &r); return r; }
};
}  // namespace qr
