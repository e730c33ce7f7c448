// Right-looking blocked Cholesky with one-panel look-ahead, on one GPU or on P GPUs with the block columns dealt
// out 1-D (BASELINE.json north_star: "the recursive Cholesky ... on a 2x2x2 (or 1D/2D) process grid"; SURVEY.md 8(e)
// "1D/2D block-cyclic right-looking variant").  A = R^T R, upper, N^3/3 flops; what lapack::engine::_potrf
// (reference src/lapack/interface.hpp:31-43) computes, scheduled for a machine where the serial part (the diagonal
// blocks) must hide behind the tensor-pipe part (the trailing update).  Included at the end of capi.cu.
//
// Layout: block columns of width nb; block j belongs to round r = j / P and goes to rank (r even ? j % P : P-1 - j % P)
// (boustrophedon: in the upper triangle a later column is a taller column; alternating the direction evens the work out),
// where it is local block r.  A rank stores its blocks side by side as one N x (sum of widths) column-major array.
//
// Per panel j (two streams per rank: M = the handle's stream, P = a high-priority panel stream):
//   panel(j)  on P   owner: R_jj and X_j = R_jj^-1 (cholinv recursion on the nb x nb block)  -> broadcast X_j
//                    every rank: R_j,loc = X_j^T A_j,loc for its columns right of j (triangular multiply on the tile engine)
//                    all-gather of the row panel, put in global column order (Rrow)
//   U1(j)     on M   block row j+1 of the trailing matrix -= R_j,j+1^T R_j,:      (all that panel(j+1) needs)
//   U2(j)     on M   the rest of the trailing matrix, per run of locally owned consecutive block columns:
//                    a plain gemm for the rows above the run's diagonal square + a triangular syrk on the square
// panel(j+1) is enqueued behind U1(j) and runs WHILE U2(j) runs: U2 launches its persistent tile-engine CTAs on
// num_sms - reserve SMs so that the panel's one-CTA leaf kernels, small products and NCCL kernels have SMs to run on.
// Buffers of consecutive panels alternate (U2(j) still reads panel j's row while panel(j+1) writes its own).
//
// Pipelined variant (cap_set_lookahead's `pipeline` = 1; the default on ONE rank only, see cap_potrf_1d): the all-gather is taken
// off the critical chain.  The owner of block j+1 needs nothing from the other ranks to update and factor its
// diagonal block (R_j,j+1 is its own column), so
//   P: leaf(j) -> bcast X_j -> [wait U1rest(j-1)] R_j,loc -> [owner(j+1): wait U2(j-1); diagonal block -= R_j,j+1^T R_j,j+1] -> leaf(j+1) ...
//   C: [wait R_j,loc] all-gather(j) on a duplicate communicator -> ordered row panel
//   M: [wait gather(j)] U1rest(j) = block row j+1 right of the diagonal block -> U2(j)
// and the chain per panel shrinks from leaf + bcast + solve + all-gather + U1 to leaf + bcast + solve + one block syrk.
#pragma once

struct Layout1D {
  int P, nb; int64_t N; int nblk;
  int owner(int j) const { const int r = j / P, pos = j % P; return (r & 1) ? P - 1 - pos : pos; }
  int global_block(int r, int p) const { return r * P + ((r & 1) ? P - 1 - p : p); }
  int width(int j) const { return (int)std::min<int64_t>(nb, N - (int64_t)j * nb); }
  int nloc(int p) const { int n = 0; while (global_block(n, p) < nblk) ++n; return n; }
  // number of p's local blocks with global index <= j
  int upto(int p, int j) const { const int rj = j / P, pos = j % P; return rj + ((((rj & 1) ? P - 1 - p : p) <= pos) ? 1 : 0); }
  int64_t cols(int p) const { const int n = nloc(p); return n == 0 ? 0 : (int64_t)(n - 1) * nb + width(global_block(n - 1, p)); }
  // local columns of p strictly right of block j
  int64_t cols_right(int p, int j) const { return std::max<int64_t>(0, cols(p) - (int64_t)std::min(upto(p, j), nloc(p)) * nb); }
};

int cap_potrf_1d_layout(int64_t N, int nb, int nranks, int rank, int64_t* ncols_loc, int* nblocks_loc) {
  if (N < 0 || nb < 1 || nranks < 1 || rank < 0 || rank >= nranks) return cap_fail(CAP_ERR_ARG, "cap_potrf_1d_layout", "bad argument");
  Layout1D L{nranks, nb, N, (int)((N + nb - 1) / nb)};
  if (ncols_loc) *ncols_loc = L.cols(rank);
  if (nblocks_loc) *nblocks_loc = L.nloc(rank);
  return CAP_OK;
}
int cap_potrf_1d_block(int64_t j, int nranks, int* owner, int* local_block) {
  if (j < 0 || nranks < 1) return cap_fail(CAP_ERR_ARG, "cap_potrf_1d_block", "bad argument");
  Layout1D L{nranks, 1, 0, 0};
  if (owner) *owner = L.owner((int)j);
  if (local_block) *local_block = (int)(j / nranks);
  return CAP_OK;
}

// gathered row panel (rank-major, nb x maxcl per rank, every block nb x nb contiguous) -> global column order
__global__ void assemble_row_panel_kernel(double* __restrict__ Rrow, const double* __restrict__ gath, int j, int P, int nb, int64_t per_rank, int64_t N) {
  const int l = j + 1 + blockIdx.y;
  const int r = l / P, pos = l % P, own = (r & 1) ? P - 1 - pos : pos;
  const int rj = j / P, posj = j % P;
  const int upto = rj + ((((rj & 1) ? P - 1 - own : own) <= posj) ? 1 : 0);
  const int64_t w = min((int64_t)nb, N - (int64_t)l * nb);
  const double2* src = reinterpret_cast<const double2*>(gath + own * per_rank + (int64_t)(r - upto) * nb * nb);
  double2* dst = reinterpret_cast<double2*>(Rrow + (int64_t)blockIdx.y * nb * nb);
  const int64_t n2 = w * nb / 2;     // nb is a multiple of 128
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

struct StreamScope {   // every helper in this file launches on h->stream: switch it (and the split-K workspace bank) for a scope
  cap_handle_t h; cudaStream_t saved; int saved_bank;
  StreamScope(cap_handle_t h_, cudaStream_t s, int bank) : h(h_), saved(h_->stream), saved_bank(h_->ws_bank) { h->stream = s; h->ws_bank = bank; }
  ~StreamScope() { h->stream = saved; h->ws_bank = saved_bank; }
};
struct SmCapScope {
  cap_handle_t h; int saved;
  SmCapScope(cap_handle_t h_, int cap) : h(h_), saved(h_->num_sms) { if (cap > 0 && cap < saved) h->num_sms = cap; }
  ~SmCapScope() { h->num_sms = saved; }
};

// Optional timeline of one factorization (CAPITAL_B200_PROFILE=1: rank 0 prints; =2: every rank): timing events recorded
// on the three streams at the phase boundaries of every panel, read back after the call has been synchronised.  This is
// the stand-in for an nsys timeline (not in this image): it shows, per panel, when the diagonal chain and the trailing
// update started and ended on this rank.  Off by default (no events, no synchronisation).
struct Timeline1D {
  struct Stamp { cudaEvent_t ev; int panel; const char* what; };
  bool on = false;
  std::vector<Stamp> stamps;
  void mark(cudaStream_t s, int panel, const char* what) {
    if (!on) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) { on = false; return; }
    cudaEventRecord(e, s);
    stamps.push_back({e, panel, what});
  }
  void report(cudaStream_t main_stream, int rank, const Layout1D& L) {
    if (!on || stamps.empty()) return;
    cudaStreamSynchronize(main_stream);
    fprintf(stderr, "[capital_b200 timeline] rank %d  potrf_1d N=%lld nb=%d P=%d  (ms since the first stamp; begin..end per phase)\n", rank, (long long)L.N, L.nb, L.P);
    int cur = -1;
    for (size_t i = 0; i + 1 < stamps.size(); i += 2) {       // stamps come in begin/end pairs
      float t0 = 0, t1 = 0;
      cudaEventElapsedTime(&t0, stamps[0].ev, stamps[i].ev);
      cudaEventElapsedTime(&t1, stamps[0].ev, stamps[i + 1].ev);
      if (stamps[i].panel != cur) { cur = stamps[i].panel; fprintf(stderr, "%s  panel %3d (owner %d):", i ? "\n" : "", cur, L.owner(cur)); }
      fprintf(stderr, "  %s %.3f..%.3f", stamps[i].what, t0, t1);
    }
    fprintf(stderr, "\n");
    for (auto& st : stamps) cudaEventDestroy(st.ev);
    stamps.clear();
  }
};

struct Potrf1D {
  Timeline1D tl;
  cap_handle_t h; cap_comm_t comm; Layout1D L; int me;
  double* A; int64_t lda;
  double *X[2], *S[2], *G[2], *R[2], *Wleaf;   // inverse of the diagonal block, local row panel, gathered, ordered
  int64_t maxcl0;
  cudaStream_t sm, sp, sc;
  cudaEvent_t ev_panel[2], ev_u1;
  cap_comm_t comm2;                             // pipelined variant: duplicate communicator of the all-gathers
  cudaEvent_t ev_trmm, ev_gath[2], ev_u2;

  double* blockptr(int row_block, int local_block) const { return A + (int64_t)row_block * L.nb + (int64_t)local_block * L.nb * lda; }

  // ---- panel(j) on the panel stream --------------------------------------------------------------------------
  int panel(int j) {
    const int nb = L.nb, b = j & 1, own = L.owner(j), w = L.width(j);
    StreamScope scope(h, sp, 1);
    const bool last = (j == L.nblk - 1);
    if (me == own) {
      tl.mark(sp, j, "leaf");
      CUDA_TRY(cudaMemsetAsync(X[b], 0, (size_t)nb * nb * sizeof(double), h->stream));
      CAP_TRY(cholinv_rec(h, blockptr(j, j / L.P), lda, X[b], nb, 0, w, true, last ? 0 : 1, Wleaf, j * nb));
      tl.mark(sp, j, "leaf");
    }
    if (last) return CAP_OK;                             // nothing to the right of the last block
    tl.mark(sp, j, "bcast");
    CAP_TRY(cap_bcast(h, comm, X[b], (int64_t)nb * nb, own));
    tl.mark(sp, j, "bcast");
    const int r0 = std::min(L.upto(me, j), L.nloc(me));
    const int64_t ncl = L.cols_right(me, j);
    if (ncl > 0) {                                      // R_j,loc = X_j^T A_j,loc, kept in S (the all-gather payload) and in A
      double* Bp = blockptr(j, r0);
      tl.mark(sp, j, "solve");
      CAP_TRY(trmm_impl(h, 'L', 'U', 'T', 'N', nb, ncl, 1.0, X[b], nb, Bp, lda, S[b], nb, true, false));
      CAP_TRY(copy2d(h, Bp, lda, S[b], nb, nb, ncl, 0));
      tl.mark(sp, j, "solve");
    }
    if (L.P > 1) {
      int64_t maxcl = 0;
      for (int p = 0; p < L.P; ++p) maxcl = std::max(maxcl, L.cols_right(p, j));
      tl.mark(sp, j, "gather");
      CAP_TRY(cap_allgather(h, comm, S[b], G[b], (int64_t)nb * maxcl));
      const int nright = L.nblk - 1 - j;
      assemble_row_panel_kernel<<<dim3(16, nright), 256, 0, h->stream>>>(R[b], G[b], j, L.P, nb, (int64_t)nb * maxcl, L.N);
      LAUNCH_CHECK(h);
      tl.mark(sp, j, "gather");
    }
    return CAP_OK;
  }

  // ---- U1(j): block row j+1 of the trailing matrix (on M) ------------------------------------------------------
  int update_next_row(int j) {
    const int nb = L.nb, b = j & 1, k = j + 1, wk = L.width(k);
    const int r0 = std::min(L.upto(me, j), L.nloc(me));          // first local block right of j
    int rr = r0;
    if (me == L.owner(k)) {                                       // diagonal block (k, k): upper triangle only
      const double* Sk = S[b] + (int64_t)(k / L.P - r0) * nb * nb;
      CAP_TRY(cap_dsyrk(h, 'U', 'T', wk, nb, -1.0, Sk, nb, 1.0, blockptr(k, k / L.P), lda));
      rr = k / L.P + 1;
    }
    const int64_t ncl = L.cols(me) - (int64_t)std::min(rr, L.nloc(me)) * nb;
    if (ncl > 0)
      CAP_TRY(cap_dgemm(h, 'T', 'N', wk, ncl, nb, -1.0, R[b], nb, S[b] + (int64_t)(rr - r0) * nb * nb, nb, 1.0, blockptr(k, rr), lda));
    return CAP_OK;
  }

  // ---- U2(j): rows >= j+2 of the trailing matrix (on M, SM-capped while a panel runs beside it) -----------------------
  int update_rest(int j) {
    const int nb = L.nb, b = j & 1, n = L.nloc(me);
    const int r0 = std::min(L.upto(me, j), n);
    int r = r0;
    while (r < n && L.global_block(r, me) < j + 2) ++r;
    while (r < n) {
      int r_end = r + 1;                                          // run of consecutive global blocks [k0, k1]
      while (r_end < n && L.global_block(r_end, me) == L.global_block(r_end - 1, me) + 1) ++r_end;
      const int k0 = L.global_block(r, me);
      const int64_t wrun = (int64_t)(r_end - 1 - r) * nb + L.width(L.global_block(r_end - 1, me));
      const double* Srun = S[b] + (int64_t)(r - r0) * nb * nb;
      const int64_t hrect = (int64_t)(k0 - j - 2) * nb;
      if (hrect > 0)
        CAP_TRY(cap_dgemm(h, 'T', 'N', hrect, wrun, nb, -1.0, R[b] + (int64_t)nb * nb, nb, Srun, nb, 1.0, blockptr(j + 2, r), lda));
      CAP_TRY(cap_dsyrk(h, 'U', 'T', wrun, nb, -1.0, Srun, nb, 1.0, blockptr(k0, r), lda));
      r = r_end;
    }
    return CAP_OK;
  }

  // ---- pipelined variant: the pieces of panel() and update_next_row() on their own streams ---------------------------
  int leaf(int j) {                                      // on P, owner only
    const int nb = L.nb, b = j & 1, w = L.width(j);
    CUDA_TRY(cudaMemsetAsync(X[b], 0, (size_t)nb * nb * sizeof(double), h->stream));
    return cholinv_rec(h, blockptr(j, j / L.P), lda, X[b], nb, 0, w, true, j == L.nblk - 1 ? 0 : 1, Wleaf, j * nb);
  }
  int row_solve(int j) {                                 // on P: R_j,loc = X_j^T A_j,loc
    const int nb = L.nb, b = j & 1;
    const int r0 = std::min(L.upto(me, j), L.nloc(me));
    const int64_t ncl = L.cols_right(me, j);
    if (ncl <= 0) return CAP_OK;
    double* Bp = blockptr(j, r0);
    CAP_TRY(trmm_impl(h, 'L', 'U', 'T', 'N', nb, ncl, 1.0, X[b], nb, Bp, lda, S[b], nb, true, false));
    return copy2d(h, Bp, lda, S[b], nb, nb, ncl, 0);
  }
  int gather(int j) {                                    // on C
    const int nb = L.nb, b = j & 1;
    int64_t maxcl = 0;
    for (int p = 0; p < L.P; ++p) maxcl = std::max(maxcl, L.cols_right(p, j));
    CAP_TRY(cap_allgather(h, comm2, S[b], G[b], (int64_t)nb * maxcl));
    assemble_row_panel_kernel<<<dim3(16, L.nblk - 1 - j), 256, 0, h->stream>>>(R[b], G[b], j, L.P, nb, (int64_t)nb * maxcl, L.N);
    LAUNCH_CHECK(h);
    return CAP_OK;
  }
  int update_next_diag(int j) {                          // on P, owner(j+1) only: A_kk -= R_jk^T R_jk from its own row-panel block
    const int nb = L.nb, b = j & 1, k = j + 1;
    const int r0 = std::min(L.upto(me, j), L.nloc(me));
    return cap_dsyrk(h, 'U', 'T', L.width(k), nb, -1.0, S[b] + (int64_t)(k / L.P - r0) * nb * nb, nb, 1.0, blockptr(k, k / L.P), lda);
  }
  int update_next_row_rest(int j) {                      // on M: block row j+1, columns right of block j+1
    const int nb = L.nb, b = j & 1, k = j + 1, wk = L.width(k);
    const int r0 = std::min(L.upto(me, j), L.nloc(me));
    const int rr = (me == L.owner(k)) ? k / L.P + 1 : r0;
    const int64_t ncl = L.cols(me) - (int64_t)std::min(rr, L.nloc(me)) * nb;
    if (ncl <= 0) return CAP_OK;
    return cap_dgemm(h, 'T', 'N', wk, ncl, nb, -1.0, R[b], nb, S[b] + (int64_t)(rr - r0) * nb * nb, nb, 1.0, blockptr(k, rr), lda);
  }

  int run_pipelined() {
    const int nblk = L.nblk;
    auto panel_front = [&](int j) -> int {               // leaf (owner) -> bcast -> [row j complete] solve -> [owner(j+1)] next diagonal block
      StreamScope scope(h, sp, 1);
      if (me == L.owner(j)) { tl.mark(sp, j, "leaf"); CAP_TRY(leaf(j)); tl.mark(sp, j, "leaf"); }
      if (j == nblk - 1) return CAP_OK;
      tl.mark(sp, j, "bcast");
      CAP_TRY(cap_bcast(h, comm, X[j & 1], (int64_t)L.nb * L.nb, L.owner(j)));
      tl.mark(sp, j, "bcast");
      if (j > 0) CUDA_TRY(cudaStreamWaitEvent(sp, ev_u1, 0));          // U1rest(j-1) and every U2 before it
      tl.mark(sp, j, "solve");
      CAP_TRY(row_solve(j));
      tl.mark(sp, j, "solve");
      CUDA_TRY(cudaEventRecord(ev_trmm, sp));
      if (me == L.owner(j + 1)) {
        if (j > 0) CUDA_TRY(cudaStreamWaitEvent(sp, ev_u2, 0));        // U2(j-1): the last earlier update of block (j+1, j+1)
        tl.mark(sp, j, "U1diag");
        CAP_TRY(update_next_diag(j));
        tl.mark(sp, j, "U1diag");
      }
      return CAP_OK;
    };
    CUDA_TRY(cudaEventRecord(ev_u2, sm));                 // so that a wait on it is never a wait on a stale record of an earlier call
    CAP_TRY(panel_front(0));
    for (int j = 0; j + 1 < nblk; ++j) {
      const int b = j & 1;
      if (L.P > 1) {
        StreamScope scope(h, sc, 1);
        CUDA_TRY(cudaStreamWaitEvent(sc, ev_trmm, 0));
        tl.mark(sc, j, "gather");
        CAP_TRY(gather(j));
        tl.mark(sc, j, "gather");
        CUDA_TRY(cudaEventRecord(ev_gath[b], sc));
        CUDA_TRY(cudaStreamWaitEvent(sm, ev_gath[b], 0));
      } else {
        CUDA_TRY(cudaStreamWaitEvent(sm, ev_trmm, 0));
      }
      tl.mark(sm, j, "U1rest");
      CAP_TRY(update_next_row_rest(j));
      tl.mark(sm, j, "U1rest");
      CUDA_TRY(cudaEventRecord(ev_u1, sm));
      if (j + 2 < nblk) {
        SmCapScope cap(h, h->lookahead_reserve_sms > 0 ? h->num_sms - h->lookahead_reserve_sms : 0);
        tl.mark(sm, j, "U2");
        CAP_TRY(update_rest(j));
        tl.mark(sm, j, "U2");
        CUDA_TRY(cudaEventRecord(ev_u2, sm));
      }
      CAP_TRY(panel_front(j + 1));
    }
    // join: everything on P (the last leaf) and C is ordered before whatever follows on the caller's stream
    CUDA_TRY(cudaEventRecord(ev_panel[0], sp));
    CUDA_TRY(cudaStreamWaitEvent(sm, ev_panel[0], 0));
    if (L.P > 1) { CUDA_TRY(cudaEventRecord(ev_panel[1], sc)); CUDA_TRY(cudaStreamWaitEvent(sm, ev_panel[1], 0)); }
    return CAP_OK;
  }

  int run() {
    const int nblk = L.nblk;
    CAP_TRY(panel(0));
    CUDA_TRY(cudaEventRecord(ev_panel[0], sp));
    for (int j = 0; j < nblk; ++j) {
      CUDA_TRY(cudaStreamWaitEvent(sm, ev_panel[j & 1], 0));
      if (j + 1 >= nblk) break;
      tl.mark(sm, j, "U1");
      CAP_TRY(update_next_row(j));
      tl.mark(sm, j, "U1");
      CUDA_TRY(cudaEventRecord(ev_u1, sm));
      CUDA_TRY(cudaStreamWaitEvent(sp, ev_u1, 0));
      CAP_TRY(panel(j + 1));
      CUDA_TRY(cudaEventRecord(ev_panel[(j + 1) & 1], sp));
      if (j + 2 < nblk) {
        SmCapScope cap(h, h->lookahead_reserve_sms > 0 ? h->num_sms - h->lookahead_reserve_sms : 0);
        tl.mark(sm, j, "U2");
        CAP_TRY(update_rest(j));
        tl.mark(sm, j, "U2");
      }
    }
    return CAP_OK;
  }
};

int cap_set_lookahead(cap_handle_t h, int reserve_sms, int pipeline) {
  if (reserve_sms < 0 || reserve_sms >= h->num_sms) return cap_fail(CAP_ERR_ARG, "cap_set_lookahead", "reserve must be in [0, #SMs)");
  if (pipeline < -1 || pipeline > 1) return cap_fail(CAP_ERR_ARG, "cap_set_lookahead", "pipeline: -1 = by rank count, 0 = two streams, 1 = all-gather on its own stream and communicator");
  h->lookahead_reserve_sms = reserve_sms;
  h->lookahead_pipeline = pipeline;
  return CAP_OK;
}

int cap_potrf_1d(cap_handle_t h, cap_comm_t comm, int64_t N, int nb, double* A_loc, int64_t lda) {
  if (N < 0 || N > INT32_MAX || nb < BASE_NB || nb % BASE_NB || lda < std::max<int64_t>(1, N)) return cap_fail(CAP_ERR_ARG, "cap_potrf_1d", "bad argument (nb: multiple of 128; lda >= N)");
  if (N == 0) return CAP_OK;
  int P = 1, me = 0;
  if (comm) { cap_comm_size(comm, &P); cap_comm_rank(comm, &me); }
  Potrf1D s{};
  s.h = h; s.comm = P > 1 ? comm : nullptr; s.me = me; s.A = A_loc; s.lda = lda;
  s.L = Layout1D{P, nb, N, (int)((N + nb - 1) / nb)};
  int64_t maxcl0 = 0;
  for (int p = 0; p < P; ++p) maxcl0 = std::max(maxcl0, s.L.cols_right(p, 0));
  maxcl0 = std::max<int64_t>(maxcl0, 1);
  s.maxcl0 = maxcl0;
  const size_t xs = (size_t)nb * nb, ss = (size_t)nb * maxcl0, gs = P > 1 ? (size_t)P * ss : 0, rs = P > 1 ? (size_t)nb * std::max<int64_t>(N - nb, 1) : 0;
  const size_t leaf = cholinv_scratch_doubles(nb);
  double* ws;
  CAP_TRY(cap_workspace(h, CAP_WS_POTRF1D, (2 * (xs + ss + gs + rs) + leaf) * sizeof(double), &ws));
  for (int b = 0; b < 2; ++b) {
    s.X[b] = ws; ws += xs;
    s.S[b] = ws; ws += ss;
    s.G[b] = P > 1 ? ws : nullptr; ws += gs;
    s.R[b] = P > 1 ? ws : s.S[b]; ws += rs;      // one rank: the local row panel IS the whole row panel
  }
  s.Wleaf = ws;
  if (!h->panel_stream) {
    int lo = 0, hi = 0;
    CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CUDA_TRY(cudaStreamCreateWithPriority(&h->panel_stream, cudaStreamNonBlocking, hi));
    CUDA_TRY(cudaStreamCreateWithPriority(&h->gather_stream, cudaStreamNonBlocking, hi));
    for (cudaEvent_t* e : {&h->ev_panel[0], &h->ev_panel[1], &h->ev_u1, &h->ev_trmm, &h->ev_gath[0], &h->ev_gath[1], &h->ev_u2})
      CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  }
  // measured (profiles/r01_potrf1d_*.jsonl): the pipelined order wins on one rank (N = 16384: 53.1 vs 54.0 ms), the
  // two-stream order on 2 and 8 ranks (N = 32768 on 8: 86.0 vs 97.6 ms)
  const bool pipelined = h->lookahead_pipeline == 1 || (h->lookahead_pipeline < 0 && P == 1);
  if (pipelined && P > 1 && h->dup_parent != comm) {     // duplicate communicator for the all-gathers, made once per parent
    if (h->dup_comm) { cap_comm_destroy(h->dup_comm); h->dup_comm = nullptr; h->dup_parent = nullptr; }
    CAP_TRY(cap_comm_split(comm, 0, me, &h->dup_comm));
    h->dup_parent = comm;
  }
  s.sm = h->stream; s.sp = h->panel_stream; s.sc = h->gather_stream; s.comm2 = h->dup_comm;
  s.ev_panel[0] = h->ev_panel[0]; s.ev_panel[1] = h->ev_panel[1]; s.ev_u1 = h->ev_u1;
  s.ev_trmm = h->ev_trmm; s.ev_gath[0] = h->ev_gath[0]; s.ev_gath[1] = h->ev_gath[1]; s.ev_u2 = h->ev_u2;
  // the panel stream starts behind everything already enqueued on the caller's stream
  CUDA_TRY(cudaEventRecord(s.ev_u1, s.sm));
  CUDA_TRY(cudaStreamWaitEvent(s.sp, s.ev_u1, 0));
  CUDA_TRY(cudaStreamWaitEvent(s.sc, s.ev_u1, 0));
  const char* prof = getenv("CAPITAL_B200_PROFILE");
  s.tl.on = prof && (atoi(prof) >= 2 || (atoi(prof) == 1 && me == 0));
  const int rc = pipelined ? s.run_pipelined() : s.run();
  s.tl.report(s.sm, me, s.L);
  return rc;
}
