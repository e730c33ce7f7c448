// Process-grid algorithms: topo::square, util::transpose, matmult::summa::invoke (gemm / trmm / syrk) and the
// distributed cholesky::cholinv recursion, with NCCL (NVLink 5 / NVSwitch) in place of the reference's MPI.
//   topo::square          reference src/util/topology.h:67-143
//   util::transpose       src/util/util.hpp:233-247
//   summa::invoke x4      src/alg/matmult/summa/summa.hpp:7-161, distribute :164-221, collect :224-253
//   cholinv::invoke       src/alg/cholesky/cholinv/cholinv.hpp:87-165, base case policy.h:226-305 (ReplicateComp)
// Included at the end of capi.cu (one translation unit: uses its static helpers).
//
// Distribution (reference src/matrix/structure.hpp:74-98): a global matrix is dealt element-cyclically over the
// d x d face, process (x, y) holds A_loc[i, j] = A[y + d*i, x + d*j]; every depth layer z holds a replica.
// A product C = A*B then is C_loc(x,y) = sum_kp A_loc(kp, y) * B_loc(x, kp): layer z takes the panels
// kp = z, z+c, ... (d/c of them; exactly one when c == d, the reference's cube), receives A_loc(kp, y) along its
// row and B_loc(x, kp) along its column, multiplies locally, and the layers are summed over depth.
// Grids accepted: c x d x d with d % c == 0 (1x1x1, 1x2x2 = 2D SUMMA on 4 GPUs, 2x2x2 = the reference's cube on 8)
// and c x 1 x 1 (2 GPUs: everything replicated, the k range of every product cut in c slices).
#pragma once

struct cap_grid_s {
  cap_comm_t world = nullptr, row = nullptr, column = nullptr, depth = nullptr, slice = nullptr;
  int c = 1, d = 1, x = 0, y = 0, z = 0, rank = 0, size = 1;
  int num_chunks = 0;          // pieces per panel of the pipelined SUMMA (0 = chosen from the contraction length); topo's num_chunks
  int layout = 0;              // 0: z fastest, then x, then y; 1: y fastest, then x, then z (topology.h:80-110)
  int rank_of(int xx, int yy, int zz) const { return layout == 0 ? zz + c * (xx + d * yy) : yy + d * (xx + d * zz); }
};

int cap_grid_create2(cap_comm_t world, int c, int layout, int num_chunks, cap_grid_t* out) {
  if (layout != 0 && layout != 1) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_grid_create", "layout must be 0 or 1 (topology.h:80-110)");
  if (num_chunks < 0) return cap_fail(CAP_ERR_ARG, "cap_grid_create", "num_chunks must be >= 0");
  cap_grid_s* g = new cap_grid_s();
  g->world = world; g->c = c; g->layout = layout; g->num_chunks = num_chunks;
  if (world) { cap_comm_rank(world, &g->rank); cap_comm_size(world, &g->size); }
  int d = 1;
  while ((int64_t)c * (d + 1) * (d + 1) <= g->size) ++d;
  if (c < 1 || (int64_t)c * d * d != g->size || !(d % c == 0 || d == 1)) {
    delete g;
    return cap_fail(CAP_ERR_ARG, "cap_grid_create", "need size == c*d*d with d % c == 0 (or d == 1)");
  }
  g->d = d;
  if (layout == 0) {   // topology.h:80-95: z fastest, then x, then y
    g->z = g->rank % c; g->y = g->rank / (d * c); g->x = (g->rank % (d * c)) / c;
  } else {             // topology.h:96-110: y fastest, then x, then z
    g->y = g->rank % d; g->x = (g->rank % (d * d)) / d; g->z = g->rank / (d * d);
  }
  if (g->size > 1) {
    CAP_TRY(cap_comm_split(world, g->x + d * g->y, g->z, &g->depth));       // same (x, y): rank in comm = z
    CAP_TRY(cap_comm_split(world, g->z, g->x + d * g->y, &g->slice));       // same z: rank in comm = x + d*y
    CAP_TRY(cap_comm_split(g->slice, g->y, g->x, &g->row));                 // same (y, z): rank in comm = x
    CAP_TRY(cap_comm_split(g->slice, g->x, g->y, &g->column));              // same (x, z): rank in comm = y
  }
  *out = g;
  return CAP_OK;
}
int cap_grid_create(cap_comm_t world, int c, cap_grid_t* out) { return cap_grid_create2(world, c, 0, 0, out); }
int cap_grid_set_chunks(cap_grid_t g, int num_chunks) {
  if (num_chunks < 0) return cap_fail(CAP_ERR_ARG, "cap_grid_set_chunks", "num_chunks must be >= 0");
  g->num_chunks = num_chunks;
  return CAP_OK;
}
int cap_grid_destroy(cap_grid_t g) {
  if (!g) return CAP_OK;
  cap_comm_destroy(g->row); cap_comm_destroy(g->column); cap_comm_destroy(g->slice); cap_comm_destroy(g->depth);
  delete g;
  return CAP_OK;
}
int cap_grid_info(cap_grid_t g, int* c, int* d, int* x, int* y, int* z) {
  if (c) *c = g->c; if (d) *d = g->d; if (x) *x = g->x; if (y) *y = g->y; if (z) *z = g->z;
  return CAP_OK;
}

// util::transpose (util.hpp:233-247): exchange the local block with the process at (y, x) of the same layer.
int cap_grid_transpose(cap_handle_t h, cap_grid_t g, const double* src, double* dst, int64_t count) {
  const int partner = g->rank_of(g->y, g->x, g->z);
  return cap_sendrecv(h, g->size > 1 ? g->world : nullptr, src, dst, count, partner == g->rank ? g->rank : partner);
}

__global__ void axpby2d_kernel(double* __restrict__ y, int64_t ldy, const double* __restrict__ x, int64_t ldx, int64_t m, int64_t n,
                               double alpha, double beta, int tri) {
  for (int64_t c = blockIdx.y; c < n; c += gridDim.y)
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x) {
    if (tri == CTRI_UPPER && r > c) continue;
    if (tri == CTRI_LOWER && r < c) continue;
    const double v = alpha * x[r + c * ldx];
    y[r + c * ldy] = (beta == 0.0) ? v : v + beta * y[r + c * ldy];
  }
}
static int axpby2d(cap_handle_t h, double* y, int64_t ldy, const double* x, int64_t ldx, int64_t m, int64_t n, double alpha, double beta, int tri) {
  if (m <= 0 || n <= 0) return CAP_OK;
  dim3 gr((unsigned)std::min<int64_t>((m + 255) / 256, 65535), (unsigned)std::min<int64_t>(n, 65535));
  axpby2d_kernel<<<gr, 256, 0, h->stream>>>(y, ldy, x, ldx, m, n, alpha, beta, tri);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// Contiguous copy of a (ptr, ld) view into workspace unless it already is contiguous.
static int contiguous(cap_handle_t h, int slot, const double* src, int64_t ld, int64_t rows, int64_t cols, const double** out) {
  if (ld == rows || cols <= 1) { *out = src; return CAP_OK; }
  double* w;
  CAP_TRY(cap_workspace(h, slot, (size_t)rows * cols * sizeof(double), &w));
  CAP_TRY(copy2d(h, w, rows, src, ld, rows, cols, 0));
  *out = w;
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ SUMMA core
enum { SUMMA_GEMM = 0, SUMMA_TRMM_L = 1, SUMMA_TRMM_R = 2 };
struct SummaOp {
  int kind = SUMMA_GEMM;
  char ta = 'N', tb = 'N';                 // gemm: local op flags (the caller grid-transposes, like the reference)
  char uplo = 'U', trans = 'N', diag = 'N';  // trmm: the triangular operand
  bool t_clean = false;                    // trmm: the unreferenced triangle of every local T block holds zeros
  int c_tri = CTRI_FULL;                   // gemm: store (and compute) only one triangle of the local C block
  double alpha = 1.0;
};

// One local multiply-accumulate: target = alpha * op(L) * op(R) + beta * target, restricted to the k-slice
// [k0, k1) of the local contraction range when the grid has d == 1.
static int summa_local(cap_handle_t h, const SummaOp& op, int64_t m, int64_t n, int64_t k, int64_t k0, int64_t k1,
                       const double* L, int64_t ldl, const double* R, int64_t ldr, double beta, double* target, int64_t ldt) {
  const bool whole = (k0 == 0 && k1 == k);
  if (op.kind == SUMMA_GEMM || !whole) {
    GemmParams p{};
    bool ta, tb;
    if (op.kind == SUMMA_GEMM) {
      ta = is_t(op.ta); tb = is_t(op.tb);
      p.A = ta ? L + k0 : L + k0 * ldl; p.B = tb ? R + k0 * ldr : R + k0;
    } else if (op.kind == SUMMA_TRMM_L) {     // op(T)[:, ks] * B[ks, :]
      ta = is_t(op.trans); tb = false;
      p.A = ta ? L + k0 : L + k0 * ldl; p.B = R + k0;
    } else {                                  // B[:, ks] * op(T)[ks, :]
      ta = false; tb = is_t(op.trans);
      p.A = L + k0 * ldl; p.B = tb ? R + k0 * ldr : R + k0;
    }
    p.lda = ldl; p.ldb = ldr; p.C = target; p.Cin = target; p.ldc = ldt; p.ldcin = ldt;
    p.M = (int)m; p.N = (int)n; p.K = (int)(k1 - k0); p.alpha = op.alpha; p.beta = beta; p.c_tri = op.c_tri;
    if (p.K <= 0) return (beta == 1.0) ? CAP_OK : axpby2d(h, target, ldt, target, ldt, m, n, 0.0, beta, CTRI_FULL);
    return gemm_launch(h, p, ta, tb, -1);
  }
  if (op.kind == SUMMA_TRMM_L)
    return trmm_impl(h, 'L', op.uplo, op.trans, op.diag, m, n, op.alpha, L, ldl, R, ldr, target, ldt, op.t_clean, false, beta);
  return trmm_impl(h, 'R', op.uplo, op.trans, op.diag, m, n, op.alpha, R, ldr, L, ldl, target, ldt, op.t_clean, false, beta);
}

__global__ void mask_cyclic_upper_kernel(double* __restrict__ A, int64_t lda, int len, int d, int x, int y);

// d > 1: the panel loop of summa::distribute / collect with communication and computation overlapped -- the job of the
// reference's num_chunks pipeline (summa.hpp:196-215 chunked MPI_Ibcast, :238-249 chunked MPI_Iallreduce).  Both operands
// of a panel are cut along the CONTRACTION index into num_chunks pieces; piece q+1 travels (row broadcast of the left
// piece, column broadcast of the right piece, on the handle's communication stream) while piece q is multiplied on the
// compute stream.  With depth c > 1 the columns of C are done in two halves so that the depth all-reduce of the first
// half runs beside the products of the second.  The root packs each piece into the landing buffer (one device copy) and
// broadcasts in place, so pieces are dense whatever the leading dimension or orientation of the operand.
// Triangular operands: a piece only touches the rows (left) / columns (right) of C on its side of the diagonal, so the
// product is restricted to that range -- the structural zeros of a piece cost at most 1 / (2 num_chunks) of the flops.
static int summa_streams(cap_handle_t h) {
  if (h->summa_stream) return CAP_OK;
  int lo = 0, hi = 0;
  CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CUDA_TRY(cudaStreamCreateWithPriority(&h->summa_stream, cudaStreamNonBlocking, hi));
  for (cudaEvent_t* e : {&h->ev_sready[0], &h->ev_sready[1], &h->ev_sfree[0], &h->ev_sfree[1], &h->ev_shalf, &h->ev_sred[0], &h->ev_sred[1], &h->ev_sstart})
    CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  return CAP_OK;
}

static int summa_pipelined(cap_handle_t h, cap_grid_t g, const SummaOp& op, int64_t m, int64_t n, int64_t k,
                           const double* L, int64_t ldl, int64_t lr, int64_t lc, const double* R, int64_t ldr, int64_t rr, int64_t rc,
                           double beta, double* C, int64_t ldc, double* S, int64_t lds, bool direct) {
  const int c = g->c, d = g->d, steps = d / c;
  CAP_TRY(summa_streams(h));
  cudaStream_t ms = h->stream, cs = h->summa_stream;
  // how each stored operand is cut along the contraction index: by columns (dense, ld = rows) or by rows (packed, ld = piece)
  bool l_by_col, r_by_col;
  if (op.kind == SUMMA_GEMM) { l_by_col = !is_t(op.ta); r_by_col = is_t(op.tb); }
  else if (op.kind == SUMMA_TRMM_L) { l_by_col = !is_t(op.trans); r_by_col = false; }
  else { l_by_col = true; r_by_col = is_t(op.trans); }
  const bool tri = op.kind != SUMMA_GEMM;
  if (tri && (!op.t_clean || is_u(op.diag))) return cap_fail(CAP_ERR_UNSUPPORTED, "summa trmm", "pipelined path needs a clean non-unit triangular operand");
  int nq = g->num_chunks > 0 ? g->num_chunks : (int)std::max<int64_t>(1, std::min<int64_t>(8, k / 2048));
  nq = (int)std::min<int64_t>(nq, std::max<int64_t>(1, k / 128));
  const int64_t kc_max = ((k + nq - 1) / nq + 127) / 128 * 128;
  nq = (int)((k + kc_max - 1) / kc_max);
  const int nh = (c > 1 && n >= 256 && h->summa_halves) ? 2 : 1;
  const int64_t nh0 = nh == 2 ? ((n / 2 + 127) / 128) * 128 : n;       // columns of the first half
  double *Lbuf, *Rbuf;
  // several panels: the left landing buffer alternates between two halves, so a panel's pieces never land on the pieces a
  // product of the previous panel still reads (the panel before that is done: its products precede, on the compute
  // stream, the ones whose completion the right-operand buffers already wait for)
  CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_A, (size_t)(steps > 1 ? 2 : 1) * lr * lc * sizeof(double), &Lbuf));
  const int64_t r_other = r_by_col ? rr : rc;                              // extent of R across the contraction index
  CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_B, (size_t)2 * kc_max * r_other * sizeof(double), &Rbuf));
  const bool zero_first = tri || !direct;                                  // restricted products never write the whole of S
  if (zero_first && !direct) CAP_TRY(memset2d(h, S, lds, m, n));
  // the communication stream starts behind everything enqueued so far (the landing buffers may still be read)
  CUDA_TRY(cudaEventRecord(h->ev_sstart, ms));
  CUDA_TRY(cudaStreamWaitEvent(cs, h->ev_sstart, 0));
  int seq = 0;                                                             // running piece number -> R buffer seq & 1
  bool used[2] = {false, false};
  for (int hf = 0; hf < nh; ++hf) {
    const int64_t n0 = hf == 0 ? 0 : nh0, n1 = (hf == 0 && nh == 2) ? nh0 : n;
    for (int s = 0; s < steps; ++s) {
      const int kp = g->z + s * c;
      for (int q = 0; q < nq; ++q, ++seq) {
        const int64_t k0 = (int64_t)q * kc_max, k1 = std::min<int64_t>(k, k0 + kc_max), kc = k1 - k0;
        const int bufi = seq & 1;
        double* Lq = Lbuf + (size_t)((hf * steps + s) & 1) * (steps > 1 ? lr * lc : 0) + (l_by_col ? k0 * lr : k0 * lc);   // piece q of the left operand: lr x kc (ld lr) or kc x lc (ld kc)
        const int64_t ldlq = l_by_col ? lr : kc;
        double* Rq = Rbuf + (size_t)bufi * kc_max * r_other;              // piece of the right operand restricted to the half's columns
        const int64_t rcols = n1 - n0;
        const int64_t ldrq = r_by_col ? rcols : kc;                       // by column: (n1-n0) x kc stored; by row: kc x (n1-n0)
        {
          StreamScope scope(h, cs, 0);
          if (used[bufi]) CUDA_TRY(cudaStreamWaitEvent(cs, h->ev_sfree[bufi], 0));
          if (hf == 0 || steps > 1) {      // the left piece travels once per panel; with several panels its landing buffer is reused
            if (g->x == kp) {
              if (l_by_col) CAP_TRY(copy2d(h, Lq, lr, L + k0 * ldl, ldl, lr, kc, 0));
              else CAP_TRY(copy2d(h, Lq, kc, L + k0, ldl, kc, lc, 0));
            }
            CAP_TRY(cap_bcast(h, g->row, Lq, (l_by_col ? lr : lc) * kc, kp));
          }
          if (g->y == kp) {
            if (r_by_col) CAP_TRY(copy2d(h, Rq, rcols, R + n0 + k0 * ldr, ldr, rcols, kc, 0));
            else CAP_TRY(copy2d(h, Rq, kc, R + k0 + n0 * ldr, ldr, kc, rcols, 0));
          }
          CAP_TRY(cap_bcast(h, g->column, Rq, rcols * kc, kp));
          CUDA_TRY(cudaEventRecord(h->ev_sready[bufi], cs));
        }
        CUDA_TRY(cudaStreamWaitEvent(ms, h->ev_sready[bufi], 0));
        // ---- the product of this piece on the compute stream: S[mlo:mhi, n0+nlo : n0+nhi] += alpha * op(Lq) * op(Rq)
        int64_t mlo = 0, mhi = m, nlo = n0, nhi = n1;
        if (op.kind == SUMMA_TRMM_L) {
          const bool opupper = is_u(op.uplo) != is_t(op.trans);            // op(T) upper: rows <= contraction index
          if (opupper) mhi = std::min<int64_t>(m, k1); else mlo = k0;
        } else if (op.kind == SUMMA_TRMM_R) {
          const bool opupper = is_u(op.uplo) != is_t(op.trans);            // op(T) upper: columns >= contraction index
          if (opupper) nlo = std::max<int64_t>(n0, k0); else nhi = std::min<int64_t>(n1, k1);
        }
        const bool first = (s == 0 && q == 0);
        const double bq = (first && !zero_first) ? beta : 1.0;
        auto product = [&](int64_t r0, int64_t r1, int64_t c0, int64_t c1, int ctri) -> int {
          if (r1 <= r0 || c1 <= c0) return CAP_OK;
          SmCapScope cap(h, h->summa_reserve_sms > 0 ? h->num_sms - h->summa_reserve_sms : 0);   // room for the NCCL kernels of the next piece
          GemmParams p{};
          p.A = l_by_col ? Lq + r0 : Lq + r0 * ldlq; p.lda = ldlq;
          p.B = r_by_col ? Rq + (c0 - n0) : Rq + (c0 - n0) * ldrq; p.ldb = ldrq;
          p.C = S + r0 + c0 * lds; p.Cin = p.C; p.ldc = lds; p.ldcin = lds;
          p.M = (int)(r1 - r0); p.N = (int)(c1 - c0); p.K = (int)kc; p.alpha = op.alpha; p.beta = bq; p.c_tri = ctri;
          return gemm_launch(h, p, !l_by_col, r_by_col, -1);
        };
        if (op.c_tri == CTRI_UPPER) {          // only rows <= columns of the local block: a rectangle above the half's square + the square
          CAP_TRY(product(0, std::min<int64_t>(m, nlo), nlo, nhi, CTRI_FULL));
          CAP_TRY(product(nlo, std::min<int64_t>(m, nhi), nlo, nhi, CTRI_UPPER));
        } else if (op.c_tri == CTRI_LOWER) {
          CAP_TRY(product(nlo, std::min<int64_t>(m, nhi), nlo, nhi, CTRI_LOWER));
          CAP_TRY(product(std::min<int64_t>(m, nhi), m, nlo, nhi, CTRI_FULL));
        } else {
          CAP_TRY(product(mlo, mhi, nlo, nhi, CTRI_FULL));
        }
        CUDA_TRY(cudaEventRecord(h->ev_sfree[bufi], ms));
        used[bufi] = true;
      }
    }
    if (!direct && c > 1) {                    // collect (summa.hpp:236) of this half beside the products of the next one
      CUDA_TRY(cudaEventRecord(h->ev_shalf, ms));
      StreamScope scope(h, cs, 0);
      CUDA_TRY(cudaStreamWaitEvent(cs, h->ev_shalf, 0));
      if (lds == m) { CAP_TRY(cap_allreduce_sum(h, g->depth, S + n0 * lds, m * (n1 - n0))); }
      else for (int64_t col = n0; col < n1; ++col) CAP_TRY(cap_allreduce_sum(h, g->depth, S + col * lds, m));
      CUDA_TRY(cudaEventRecord(h->ev_sred[hf], cs));
    }
  }
  if (!direct) {
    if (c > 1) for (int hf = 0; hf < nh; ++hf) CUDA_TRY(cudaStreamWaitEvent(ms, h->ev_sred[hf], 0));
    CAP_TRY(axpby2d(h, C, ldc, S, lds, m, n, 1.0, beta, op.c_tri));          // C = beta*C + S (summa.hpp:32-35,153)
  }
  return CAP_OK;
}

// C_loc = alpha * sum over panels (left operand received along the row) x (right operand received along the
// column) + beta * C_loc.  L is lr x lc, R is rr x rc as stored (any leading dimension); k = local contraction length.
static int summa_core(cap_handle_t h, cap_grid_t g, SummaOp op, int64_t m, int64_t n, int64_t k,
                      const double* L, int64_t ldl, int64_t lr, int64_t lc, const double* R, int64_t ldr, int64_t rr, int64_t rc,
                      double beta, double* C, int64_t ldc) {
  if (m <= 0 || n <= 0) return CAP_OK;
  const int c = g->c, d = g->d;
  double* S = C; int64_t lds = ldc;
  if (c > 1) { CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_C, (size_t)m * n * sizeof(double), &S)); lds = m; }
  const bool in_place_hazard = (c == 1) && (C == L || C == R);   // trmm writes its own operand: go through the partial buffer
  if (in_place_hazard) { CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_C, (size_t)m * n * sizeof(double), &S)); lds = m; }
  const bool direct = (S == C);
  // a triangular product writes one triangle of the partial buffer only, and the depth all-reduce sums all of it: the other
  // triangle must not carry whatever the workspace held (uninitialised, possibly NaN, data through the collective)
  if (!direct && op.c_tri != CTRI_FULL) CAP_TRY(memset2d(h, S, lds, m, n));
  if (d == 1) {
    if (op.kind != SUMMA_GEMM && c > 1 && (!op.t_clean || is_u(op.diag))) {
      // k-slices of a triangular operand are multiplied as general panels: materialise the zero triangle once
      const bool left = op.kind == SUMMA_TRMM_L;
      const int64_t nt = left ? m : n;
      const double* T = left ? L : R; const int64_t ldt = left ? ldl : ldr;
      if (is_u(op.diag)) return cap_fail(CAP_ERR_UNSUPPORTED, "summa trmm", "unit diagonal on a c x 1 x 1 grid");
      double* tc;
      CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_T, (size_t)nt * nt * sizeof(double), &tc));
      CAP_TRY(copy2d(h, tc, nt, T, ldt, nt, nt, is_u(op.uplo) ? 3 : 4));
      if (left) { L = tc; ldl = nt; } else { R = tc; ldr = nt; }
      op.t_clean = true;
    }
    const int64_t kz = (k + c - 1) / c;
    const int64_t k0 = std::min<int64_t>(k, g->z * kz), k1 = std::min<int64_t>(k, k0 + kz);
    CAP_TRY(summa_local(h, op, m, n, k, k0, k1, L, ldl, R, ldr, direct ? beta : 0.0, S, lds));
  } else {
    if (op.kind != SUMMA_GEMM && !op.t_clean) {
      // the pipelined products multiply pieces of the triangular operand as general panels: materialise its zero triangle
      // once (global row <= global column for upper; the local triangle for lower)
      const bool left = op.kind == SUMMA_TRMM_L;
      const int64_t nt = left ? m : n;
      const double* T = left ? L : R; const int64_t ldt = left ? ldl : ldr;
      double* tc;
      CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_T, (size_t)nt * nt * sizeof(double), &tc));
      if (is_u(op.uplo)) {
        CAP_TRY(copy2d(h, tc, nt, T, ldt, nt, nt, 0));
        dim3 gl((unsigned)((nt + 255) / 256), (unsigned)std::min<int64_t>(nt, 65535));
        // op(T) = T^T: the caller hands in the grid-transposed block (util::transpose, as cholinv.hpp:116-118 does), i.e. the
        // block of the process at (y, x)
        const bool tr = is_t(op.trans);
        mask_cyclic_upper_kernel<<<gl, 256, 0, h->stream>>>(tc, nt, (int)nt, g->d, tr ? g->y : g->x, tr ? g->x : g->y);
        LAUNCH_CHECK(h);
      } else {
        CAP_TRY(copy2d(h, tc, nt, T, ldt, nt, nt, 4));
      }
      if (left) { L = tc; ldl = nt; } else { R = tc; ldr = nt; }
      op.t_clean = true;
    }
    if (h->summa_pipeline == 1 || (h->summa_pipeline < 0 && g->num_chunks > 0)) return summa_pipelined(h, g, op, m, n, k, L, ldl, lr, lc, R, ldr, rr, rc, beta, C, ldc, S, lds, direct);
    // serial panel loop (round 1): bcast -> bcast -> product, every collective on the compute stream
    const int steps = d / c;
    for (int s = 0; s < steps; ++s) {
      const int kp = g->z + s * c;
      const double *Lb, *Rb;
      if (g->x == kp) { CAP_TRY(contiguous(h, CAP_WS_SUMMA_A, L, ldl, lr, lc, &Lb)); }
      else { double* w; CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_A, (size_t)lr * lc * sizeof(double), &w)); Lb = w; }
      CAP_TRY(cap_bcast(h, g->row, const_cast<double*>(Lb), lr * lc, kp));
      if (g->y == kp) { CAP_TRY(contiguous(h, CAP_WS_SUMMA_B, R, ldr, rr, rc, &Rb)); }
      else { double* w; CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_B, (size_t)rr * rc * sizeof(double), &w)); Rb = w; }
      CAP_TRY(cap_bcast(h, g->column, const_cast<double*>(Rb), rr * rc, kp));
      const int64_t ldlb = (Lb == L) ? ldl : lr, ldrb = (Rb == R) ? ldr : rr;
      const double b = (s == 0) ? (direct ? beta : 0.0) : 1.0;
      CAP_TRY(summa_local(h, op, m, n, k, 0, k, Lb, ldlb, Rb, ldrb, b, S, lds));
    }
  }
  if (!direct) {
    if (c > 1) { CAP_TRY(cap_allreduce_sum(h, g->depth, S, m * n)); }       // collect (summa.hpp:236)
    CAP_TRY(axpby2d(h, C, ldc, S, lds, m, n, 1.0, beta, op.c_tri));          // C = beta*C + S (summa.hpp:32-35,153)
  }
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ summa::invoke
// gemm (summa.hpp:7-44).  m, n, k are LOCAL dimensions of op(A) (m x k), op(B) (k x n), C (m x n).
int cap_summa_gemm(cap_handle_t h, cap_grid_t g, char transa, char transb, int64_t m, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc) {
  if (m < 0 || n < 0 || k < 0 || m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_summa_gemm", "bad dimension");
  SummaOp op; op.kind = SUMMA_GEMM; op.ta = transa; op.tb = transb; op.alpha = alpha;
  const bool ta = is_t(transa), tb = is_t(transb);
  return summa_core(h, g, op, m, n, k, A, lda, ta ? k : m, ta ? m : k, B, ldb, tb ? n : k, tb ? k : n, beta, C, ldc);
}

// trmm (summa.hpp:47-83): B <- alpha*op(T)*B (side 'L') or alpha*B*op(T) (side 'R'), B is m x n local.
static int summa_trmm_impl(cap_handle_t h, cap_grid_t g, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                           const double* T, int64_t ldt, double* B, int64_t ldb, bool t_clean) {
  if (m < 0 || n < 0 || m > INT32_MAX || n > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_summa_trmm", "bad dimension");
  SummaOp op; op.uplo = uplo; op.trans = trans; op.diag = diag; op.alpha = alpha; op.t_clean = t_clean;
  if (is_l(side)) {
    op.kind = SUMMA_TRMM_L;
    return summa_core(h, g, op, m, n, m, T, ldt, m, m, B, ldb, m, n, 0.0, B, ldb);
  }
  op.kind = SUMMA_TRMM_R;
  return summa_core(h, g, op, m, n, n, B, ldb, m, n, T, ldt, n, n, 0.0, B, ldb);
}
int cap_summa_trmm(cap_handle_t h, cap_grid_t g, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                   const double* T, int64_t ldt, double* B, int64_t ldb) {
  return summa_trmm_impl(h, g, side, uplo, trans, diag, m, n, alpha, T, ldt, B, ldb, false);
}

// syrk (summa.hpp:86-161), trans = 'T': C(uplo) = alpha * A^T A + beta * C with A k x n local.  The reference
// runs a full gemm on both triangles and packs one (summa.hpp:115,139-148); here only the wanted local
// triangle is computed.  At: the grid-transposed copy of A (util::transpose); NULL = made here.
int cap_summa_syrk(cap_handle_t h, cap_grid_t g, char uplo, char trans, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* At, int64_t ldat, double beta, double* C, int64_t ldc) {
  if (n < 0 || k < 0 || n > INT32_MAX || k > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_summa_syrk", "bad dimension");
  const bool tr = is_t(trans);
  const int64_t ar = tr ? k : n, ac = tr ? n : k;          // A as stored
  if (!At) {
    const double* Ac;
    CAP_TRY(contiguous(h, CAP_WS_SUMMA_T, A, lda, ar, ac, &Ac));
    double* t;
    if (Ac == A) { CAP_TRY(cap_workspace(h, CAP_WS_SUMMA_T, (size_t)ar * ac * sizeof(double), &t)); }
    else { CAP_TRY(cap_workspace(h, CAP_WS_CI_T2, (size_t)ar * ac * sizeof(double), &t)); }
    CAP_TRY(cap_grid_transpose(h, g, Ac, t, ar * ac));
    At = t; ldat = ar;
  }
  SummaOp op; op.kind = SUMMA_GEMM; op.alpha = alpha; op.c_tri = is_u(uplo) ? CTRI_UPPER : CTRI_LOWER;
  if (tr) {   // C = At^T * A: the transposed copy travels along rows, A along columns (summa.hpp:129-131,143-145)
    op.ta = 'T'; op.tb = 'N';
    return summa_core(h, g, op, n, n, k, At, ldat, k, n, A, lda, k, n, beta, C, ldc);
  }
  op.ta = 'N'; op.tb = 'T';   // C = A * At^T (summa.hpp:126-128,138-140)
  return summa_core(h, g, op, n, n, k, A, lda, n, k, At, ldat, n, k, beta, C, ldc);
}

// ------------------------------------------------------------------------------------------ distributed cholinv
// util::block_to_cyclic_rect (util.hpp:99-133): the d*d gathered local blocks (slice-rank order = x + d*y) of
// len x len doubles each -> one dense (d*len) x (d*len) column-major matrix, G[yy + d*r, xx + d*c] = block(xx,yy)[r, c].
__global__ void blocks_to_cyclic_kernel(double* __restrict__ G, const double* __restrict__ blocks, int len, int d) {
  const int64_t N = (int64_t)len * d;
  const int64_t gc = blockIdx.y;
  for (int64_t gr = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gr < N; gr += (int64_t)gridDim.x * blockDim.x) {
    const int xx = (int)(gc % d), yy = (int)(gr % d);
    G[gr + gc * N] = blocks[((int64_t)(xx + d * yy) * len + gc / d) * len + gr / d];
  }
}
// util::cyclic_to_local (util.hpp:136-164): own block back out of the dense matrix, upper triangle (global
// row <= global column) only, zeros elsewhere.
__global__ void cyclic_to_local_kernel(double* __restrict__ dst, int64_t ldd, const double* __restrict__ G, int len, int d, int x, int y) {
  const int64_t N = (int64_t)len * d;
  const int c = blockIdx.y;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < len; r += gridDim.x * blockDim.x) {
    const int64_t gr = y + (int64_t)d * r, gc = x + (int64_t)d * c;
    dst[r + (int64_t)c * ldd] = (gr <= gc) ? G[gr + gc * N] : 0.0;
  }
}

__global__ void mask_cyclic_upper_kernel(double* __restrict__ A, int64_t lda, int len, int d, int x, int y) {
  for (int c = blockIdx.y; c < len; c += gridDim.y)
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < len; r += gridDim.x * blockDim.x)
    if (y + (int64_t)d * r > x + (int64_t)d * c) A[r + (int64_t)c * lda] = 0.0;
}

// Optional phase profile of the distributed recursion (CAPITAL_B200_PROFILE=1): synchronises around every phase and
// accumulates wall time per category; printed by rank 0 at the end of cap_cholinv_factor.  Off by default.
struct PhaseProf {
  bool on = false; double t[8] = {0, 0, 0, 0, 0, 0, 0, 0}; const char* names[8] = {"base", "trsm", "syrk", "inv", "copy", "transpose", "", ""};
  std::chrono::steady_clock::time_point t0;
  void begin(cap_handle_t h) { if (on) { cudaStreamSynchronize(h->stream); t0 = std::chrono::steady_clock::now(); } }
  void end(cap_handle_t h, int cat) { if (on) { cudaStreamSynchronize(h->stream); t[cat] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); } }
};
static PhaseProf g_prof;

struct CholinvCtx {
  cap_handle_t h; cap_grid_t g;
  double* A; int64_t lda; double* X; int64_t ldx;
  int complete_inv, split; int64_t bc_global;
};

// Base case = the reference's ReplicateComp policy (policy.h:226-305): gather the diagonal block of the layer,
// factor + invert it redundantly on every rank, keep the own cyclic part.
static int cholinv_grid_base(const CholinvCtx& cx, int off, int len) {
  cap_handle_t h = cx.h; cap_grid_t g = cx.g;
  const int d = g->d;
  double* A11 = cx.A + off + (int64_t)off * cx.lda;
  double* X11 = cx.X + off + (int64_t)off * cx.ldx;
  if (d == 1) {
    const int64_t n = len;
    CAP_TRY(cap_dcholinv(h, (int)n, A11, cx.lda, X11, cx.ldx, 1));
    return copy2d(h, A11, cx.lda, A11, cx.lda, n, n, 3);      // R only: zero the strict lower part
  }
  const int64_t blk = (int64_t)len * len, N = (int64_t)len * d;
  double *mine, *all, *G, *GX;
  CAP_TRY(cap_workspace(h, CAP_WS_CI_BC, (size_t)(blk * (d * d + 1)) * sizeof(double), &mine));
  all = mine + blk;
  CAP_TRY(cap_workspace(h, CAP_WS_CI_BC2, (size_t)(2 * N * N) * sizeof(double), &G));
  GX = G + N * N;
  CAP_TRY(copy2d(h, mine, len, A11, cx.lda, len, len, 0));
  CAP_TRY(cap_allgather(h, g->slice, mine, all, blk));                              // policy.h:240
  dim3 gg((unsigned)std::min<int64_t>((N + 255) / 256, 65535), (unsigned)N);
  blocks_to_cyclic_kernel<<<gg, 256, 0, h->stream>>>(G, all, len, d);
  LAUNCH_CHECK(h);
  CAP_TRY(cap_dcholinv(h, (int)N, G, N, GX, N, 1));                                 // policy.h:265-267
  dim3 gl((unsigned)((len + 255) / 256), (unsigned)len);
  cyclic_to_local_kernel<<<gl, 256, 0, h->stream>>>(A11, cx.lda, G, len, d, g->x, g->y);
  LAUNCH_CHECK(h);
  cyclic_to_local_kernel<<<gl, 256, 0, h->stream>>>(X11, cx.ldx, GX, len, d, g->x, g->y);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// cholinv::invoke (cholinv.hpp:87-165) on local index ranges: the local block [off, off+len)^2 of every process
// is the cyclic image of one global diagonal block.
static int cholinv_grid_rec(const CholinvCtx& cx, int off, int len, bool top) {
  cap_handle_t h = cx.h; cap_grid_t g = cx.g;
  const int l1 = len >> cx.split;
  if ((int64_t)len * g->d <= cx.bc_global || l1 < cx.split || l1 < 1) {   // cholinv.hpp:93
    g_prof.begin(h);
    const int rc = cholinv_grid_base(cx, off, len);
    g_prof.end(h, 0);
    return rc;
  }
  const int l2 = len - l1;
  double* A12 = cx.A + off + (int64_t)(off + l1) * cx.lda;
  double* A22 = A12 + l1;
  double* X11 = cx.X + off + (int64_t)off * cx.ldx;
  double* X12 = X11 + (int64_t)l1 * cx.ldx;
  double* X22 = X12 + l1;
  CAP_TRY(cholinv_grid_rec(cx, off, l1, false));
  // trsm (cholinv.hpp:113-124): R12 = R11^{-T} A12 = trmm(Left, Upper, Trans) with the grid-transposed inverse
  double *T1, *T1t, *W, *Wt;
  CAP_TRY(cap_workspace(h, CAP_WS_CI_T1, (size_t)2 * l1 * l1 * sizeof(double), &T1));
  T1t = T1 + (int64_t)l1 * l1;
  CAP_TRY(cap_workspace(h, CAP_WS_CI_W, (size_t)2 * l1 * l2 * sizeof(double), &W));
  Wt = W + (int64_t)l1 * l2;
  g_prof.begin(h);
  CAP_TRY(copy2d(h, T1, l1, X11, cx.ldx, l1, l1, 0));
  g_prof.end(h, 4); g_prof.begin(h);
  CAP_TRY(cap_grid_transpose(h, g, T1, T1t, (int64_t)l1 * l1));
  g_prof.end(h, 5); g_prof.begin(h);
  CAP_TRY(copy2d(h, W, l1, A12, cx.lda, l1, l2, 0));
  g_prof.end(h, 4); g_prof.begin(h);
  CAP_TRY(summa_trmm_impl(h, g, 'L', 'U', 'T', 'N', l1, l2, 1.0, T1t, l1, W, l1, true));
  g_prof.end(h, 1); g_prof.begin(h);
  CAP_TRY(copy2d(h, A12, cx.lda, W, l1, l1, l2, 0));
  g_prof.end(h, 4);
  // trailing matrix update (cholinv.hpp:129-134): A22 -= R12^T R12, upper triangle of the local block only
  g_prof.begin(h);
  CAP_TRY(cap_grid_transpose(h, g, W, Wt, (int64_t)l1 * l2));
  g_prof.end(h, 5); g_prof.begin(h);
  CAP_TRY(cap_summa_syrk(h, g, 'U', 'T', l2, l1, -1.0, W, l1, Wt, l1, 1.0, A22, cx.lda));
  g_prof.end(h, 2);
  CAP_TRY(cholinv_grid_rec(cx, off + l1, l2, false));
  if (!(top && !cx.complete_inv)) {
    // inverse completion (cholinv.hpp:147-156): Rinv12 = -(Rinv11 R12) Rinv22
    // (the right recursion reused the workspace slots: re-stage Rinv11 and R12 from where they are stored)
    double* T2;
    CAP_TRY(cap_workspace(h, CAP_WS_CI_T1, (size_t)2 * l1 * l1 * sizeof(double), &T1));
    CAP_TRY(cap_workspace(h, CAP_WS_CI_W, (size_t)2 * l1 * l2 * sizeof(double), &W));
    CAP_TRY(cap_workspace(h, CAP_WS_CI_T2, (size_t)l2 * l2 * sizeof(double), &T2));
    g_prof.begin(h);
    CAP_TRY(copy2d(h, T1, l1, X11, cx.ldx, l1, l1, 0));
    CAP_TRY(copy2d(h, W, l1, A12, cx.lda, l1, l2, 0));
    g_prof.end(h, 4); g_prof.begin(h);
    CAP_TRY(summa_trmm_impl(h, g, 'L', 'U', 'N', 'N', l1, l2, 1.0, T1, l1, W, l1, true));
    g_prof.end(h, 3); g_prof.begin(h);
    CAP_TRY(copy2d(h, T2, l2, X22, cx.ldx, l2, l2, 0));
    g_prof.end(h, 4); g_prof.begin(h);
    CAP_TRY(summa_trmm_impl(h, g, 'R', 'U', 'N', 'N', l1, l2, -1.0, T2, l2, W, l1, true));
    g_prof.end(h, 3); g_prof.begin(h);
    CAP_TRY(copy2d(h, X12, cx.ldx, W, l1, l1, l2, 0));
    g_prof.end(h, 4);
  }
  return CAP_OK;
}

// cholesky::cholinv::factor (cholinv.hpp:6-29).  A: the local n_loc x n_loc block of the SPD matrix (element-cyclic
// over the d x d face, replicated over depth); on exit its cyclic part of R (global row <= column, zeros
// elsewhere).  Rinv likewise.  bc_global: global size at or below which a diagonal block is gathered and
// factored redundantly (the reference's bcDimension, cholinv.hpp:15-18).
int cap_cholinv_factor(cap_handle_t h, cap_grid_t g, int64_t n_loc, double* A, int64_t lda, double* Rinv, int64_t ldri,
                       int complete_inv, int split, int64_t bc_global) {
  if (n_loc <= 0 || n_loc * g->d > INT32_MAX || split < 1) return cap_fail(CAP_ERR_ARG, "cap_cholinv_factor", "bad argument");
  CAP_TRY(memset2d(h, Rinv, ldri, n_loc, n_loc));
  CholinvCtx cx{h, g, A, lda, Rinv, ldri, complete_inv, split, bc_global};
  g_prof.on = getenv("CAPITAL_B200_PROFILE") != nullptr;
  for (double& v : g_prof.t) v = 0;
  CAP_TRY(cholinv_grid_rec(cx, 0, (int)n_loc, true));
  if (g_prof.on && g->rank == 0) {
    fprintf(stderr, "[capital_b200 profile] cholinv n_loc=%lld grid %dx%dx%d:", (long long)n_loc, g->c, g->d, g->d);
    for (int i = 0; i < 6; ++i) fprintf(stderr, " %s %.2f ms", g_prof.names[i], g_prof.t[i] * 1e3);
    fprintf(stderr, "\n");
  }
  // the strictly-lower local part of A still holds input values: R is upper (remove_triangle, util.hpp:267-293)
  dim3 gl((unsigned)((n_loc + 255) / 256), (unsigned)std::min<int64_t>(n_loc, 65535));
  mask_cyclic_upper_kernel<<<gl, 256, 0, h->stream>>>(A, lda, (int)n_loc, g->d, g->x, g->y);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ 3D CholeskyQR2
// qr::cacqr::factor on a process grid (cacqr.hpp:219-248 -> invoke_3d :197-216 -> sweep_3d :77-122).
// A, Q: local m_loc x n_loc blocks, element-cyclic over the d x d face (rows over y, columns over x), replicated
// over depth; R (and Rinv, optional) local n_loc x n_loc blocks of the upper factor in the same distribution.
// One sweep:  G = Q^T Q by SUMMA-syrk (the reference's row-broadcast / column-reduce / depth-broadcast of
// cacqr.hpp:86-98 is the c == d special case of the same contraction; here the grid-transposed copy of Q travels
// along rows and Q along columns, so every c x d x d grid works), R = chol(G) and R^-1 by the distributed cholinv,
// Q <- Q R^-1 by SUMMA-trmm (cacqr.hpp:104-108).  Two sweeps and R = R2 R1 (cacqr.hpp:203-211) for CholeskyQR2.
// The block back-substitution variant (`solve`, cacqr.hpp:46-73, complete_inv == 0) produces the same Q without the
// top-level inverse completion; selected by cap_set_cacqr_complete_inv(h, 0) (Rinv then lacks its top-level off-diagonal block).
int cap_cacqr_factor_3d(cap_handle_t h, cap_grid_t g, int64_t m_loc, int64_t n_loc, const double* A, int64_t lda,
                        double* Q, int64_t ldq, double* R, int64_t ldr, double* Rinv, int64_t ldri,
                        int num_iter, int split, int64_t bc_global) {
  if (m_loc <= 0 || n_loc <= 0 || m_loc > INT32_MAX || n_loc * g->d > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_3d", "bad dimension");
  if (num_iter < 1 || num_iter > 2) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_3d", "num_iter must be 1 or 2");
  const int64_t nn = n_loc * n_loc;
  double* ws;
  CAP_TRY(cap_workspace(h, CAP_WS_CQR3, (size_t)(2 * nn) * sizeof(double), &ws));
  double* R2 = ws;
  double* X = Rinv ? Rinv : ws + nn;
  const int64_t ldx = Rinv ? ldri : n_loc;
  if (Q != A) CAP_TRY(copy2d(h, Q, ldq, A, lda, m_loc, n_loc, 0));               // cacqr.hpp:226
  for (int it = 0; it < num_iter; ++it) {
    double* G = it == 0 ? R : R2;
    const int64_t ldg = it == 0 ? ldr : n_loc;
    CAP_TRY(memset2d(h, G, ldg, n_loc, n_loc));
    CAP_TRY(cap_summa_syrk(h, g, 'U', 'T', n_loc, m_loc, 1.0, Q, ldq, nullptr, 0, 0.0, G, ldg));
    const int64_t s1 = n_loc >> split, s2 = n_loc - s1;
    if (h->cacqr_complete_inv || s1 < 1 || s2 < 1) {
      CAP_TRY(cap_cholinv_factor(h, g, n_loc, G, ldg, X, ldx, 1, split, bc_global));
      CAP_TRY(cap_summa_trmm(h, g, 'R', 'U', 'N', 'N', m_loc, n_loc, 1.0, X, ldx, Q, ldq));
    } else {
      // cacqr::solve (cacqr.hpp:46-73): the inverse is left without its top-level off-diagonal block (cholinv.hpp:147) and
      // Q = Q R^-1 is finished by block back-substitution:  Q1 <- Q1 Rinv11;  Q2 <- (Q2 - Q1 R12) Rinv22
      CAP_TRY(cap_cholinv_factor(h, g, n_loc, G, ldg, X, ldx, 0, split, bc_global));
      double* Q2 = Q + s1 * ldq;
      CAP_TRY(summa_trmm_impl(h, g, 'R', 'U', 'N', 'N', m_loc, s1, 1.0, X, ldx, Q, ldq, true));
      CAP_TRY(cap_summa_gemm(h, g, 'N', 'N', m_loc, s2, s1, -1.0, Q, ldq, G + s1 * ldg, ldg, 1.0, Q2, ldq));
      CAP_TRY(summa_trmm_impl(h, g, 'R', 'U', 'N', 'N', m_loc, s2, 1.0, X + s1 + s1 * ldx, ldx, Q2, ldq, true));
    }
  }
  if (num_iter == 2) {
    CAP_TRY(cap_summa_trmm(h, g, 'L', 'U', 'N', 'N', n_loc, n_loc, 1.0, R2, n_loc, R, ldr));   // R = R2 R1
  }
  return CAP_OK;
}


// ------------------------------------------------------------------------------------------ validators on the grid
// cholesky::validate::residual (test/cholesky/validate.hpp:7-30): ||R^T R - A||_F / ||A||_F over the upper triangle.
// A: the ORIGINAL local block (the caller keeps a copy; cap_cholinv_factor overwrites its argument), R: local block of the factor.
int cap_validate_cholinv(cap_handle_t h, cap_grid_t g, int64_t n_loc, const double* A, int64_t lda, const double* R, int64_t ldr, double* resid) {
  if (n_loc <= 0 || n_loc > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_validate_cholinv", "bad dimension");
  double* T;
  CAP_TRY(cap_workspace(h, CAP_WS_STAGE0, (size_t)n_loc * n_loc * sizeof(double), &T));
  CAP_TRY(memset2d(h, T, n_loc, n_loc, n_loc));
  CAP_TRY(cap_summa_syrk(h, g, 'U', 'T', n_loc, n_loc, 1.0, R, ldr, nullptr, 0, 0.0, T, n_loc));
  CAP_TRY(cap_residual_begin(h));
  CAP_TRY(cap_residual_accumulate(h, n_loc, n_loc, T, n_loc, A, lda, g->d, g->x, g->y, 1, 1.0));
  return cap_residual_finish(h, g->slice, resid);
}

// qr::validate::orthogonality / residual (test/qr/validate.hpp:7-62) for the grid algorithm:
// orth = ||Q^T Q - I||_F / n, resid = ||Q R - A||_F / ||A||_F, both reduced over one layer of the grid.
int cap_validate_cacqr_3d(cap_handle_t h, cap_grid_t g, int64_t m_loc, int64_t n_loc, const double* A, int64_t lda,
                          const double* Q, int64_t ldq, const double* R, int64_t ldr, double* orth, double* resid) {
  if (m_loc <= 0 || n_loc <= 0 || m_loc > INT32_MAX || n_loc > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_validate_cacqr_3d", "bad dimension");
  double* T;
  CAP_TRY(cap_workspace(h, CAP_WS_STAGE0, (size_t)std::max(m_loc, n_loc) * n_loc * sizeof(double), &T));
  CAP_TRY(memset2d(h, T, n_loc, n_loc, n_loc));
  CAP_TRY(cap_summa_syrk(h, g, 'U', 'T', n_loc, m_loc, 1.0, Q, ldq, nullptr, 0, 0.0, T, n_loc));
  CAP_TRY(cap_residual_begin(h));
  CAP_TRY(cap_residual_accumulate(h, n_loc, n_loc, T, n_loc, nullptr, 0, g->d, g->x, g->y, 1, 2.0));
  CAP_TRY(cap_residual_finish(h, g->slice, orth));
  CAP_TRY(copy2d(h, T, m_loc, Q, ldq, m_loc, n_loc, 0));
  CAP_TRY(cap_summa_trmm(h, g, 'R', 'U', 'N', 'N', m_loc, n_loc, 1.0, R, ldr, T, m_loc));
  CAP_TRY(cap_residual_begin(h));
  CAP_TRY(cap_residual_accumulate(h, m_loc, n_loc, T, m_loc, A, lda, g->d, g->x, g->y, 0, 1.0));
  return cap_residual_finish(h, g->slice, resid);
}
