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

struct GridMultScope {
  cap_handle_t h; int saved;
  GridMultScope(cap_handle_t h_, int mult) : h(h_), saved(h_->grid_mult) { if (mult > 0) h->grid_mult = mult; }
  ~GridMultScope() { h->grid_mult = saved; }
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
    cudaDeviceSynchronize();
    std::string out;
    char line[256];
    snprintf(line, sizeof line, "[capital_b200 timeline] rank %d  potrf_1d N=%lld nb=%d P=%d  (ms since the first stamp; begin..end per phase)\n", rank, (long long)L.N, L.nb, L.P);
    out += line;
    int cur = -1;
    for (size_t i = 0; i + 1 < stamps.size(); i += 2) {       // stamps come in begin/end pairs
      float t0 = 0, t1 = 0;
      cudaEventElapsedTime(&t0, stamps[0].ev, stamps[i].ev);
      cudaEventElapsedTime(&t1, stamps[0].ev, stamps[i + 1].ev);
      if (stamps[i].panel != cur) { cur = stamps[i].panel; snprintf(line, sizeof line, "%s  panel %3d (owner %d):", i ? "\n" : "", cur, L.owner(cur)); out += line; }
      snprintf(line, sizeof line, "  %s %.3f..%.3f", stamps[i].what, t0, t1);
      out += line;
    }
    out += "\n";
    // CAPITAL_B200_PROFILE_FILE=<prefix>: one file per rank (ranks print at the same time; a shared stderr interleaves them)
    const char* prefix = getenv("CAPITAL_B200_PROFILE_FILE");
    FILE* f = nullptr;
    if (prefix) { snprintf(line, sizeof line, "%s.rank%d.txt", prefix, rank); f = fopen(line, "w"); }
    fwrite(out.data(), 1, out.size(), f ? f : stderr);
    if (f) fclose(f);
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

// ------------------------------------------------------------------------------------------ asynchronous schedule
// Round 2.  Same arithmetic and the same 1-D layout as above; what changes is WHO waits for WHAT.
//
// The serial chain of a right-looking Cholesky is   leaf(k) -> X_k -> R_k,k+1 = X_k^T A_k,k+1 -> A_k+1,k+1 -= R_k,k+1^T R_k,k+1 -> leaf(k+1):
// four small operations per panel that involve ONE block column, owned by ONE rank, and no collective.  Everything else --
// solving the rest of row k, shipping it to the other ranks, the trailing update -- is throughput work that only has to
// stay ahead of the chain by one panel.  So each rank runs four streams:
//   P (high priority)  the chain, only on the owner of the block column concerned:
//                      [X_k-1 here] [U1rest(k-2) done] solve block (k-1,k) -> [U2a(k-2) done] diagonal update -> leaf(k) -> (X) push X_k
//   M (the caller's)   [row k data here] U1rest(k): block row k+1 -> U2a(k): diagonal block (k+2,k+2) -> solve the rest of row k+1 ->
//                      (C) ship row k+1 -> U2b(k): the rest of the trailing update, SM-capped only on the rank whose chain runs beside it
//   C                  panel exchange: every solved nb x nb block is written straight into its place in every peer's ordered
//                      row-panel buffer by the COPY ENGINES (cudaMemcpyAsync into CUDA-IPC-mapped peer memory over NVSwitch)
//   X                  X_k pushes of the owner, and the acknowledgements that release landing buffers
// The chain of panel k+2 needs U1rest(k) and U2a(k) only, both at the START of panel k's trailing work, so it runs two panels
// ahead of the trailing update instead of one, and no NCCL kernel ever sits (spinning on a peer) on an SM beside the factorization.
// Ordering between ranks is by 32-bit flags in the peer-mapped buffer: the producer's stream writes the panel's sequence number
// after its copies (cuStreamWriteValue32 into a local word + a 4-byte peer copy behind the payload, same stream); the consumer's
// stream blocks in cuStreamWaitValue32(>=) -- no SM, no host involvement.  Landing buffers form a ring of RING panels; a peer's
// buffer is overwritten only after that peer acknowledged (same mechanism, opposite direction) the panel that last used it.
typedef CUresult (*stream_value32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

struct Async1D {
  static constexpr int RING = cap_handle_s::RING;
  Timeline1D tl;
  cap_handle_t h; cap_peer_s* peer; Layout1D L; int me;
  double* A; int64_t lda;
  double *X[RING], *S[RING], *R[RING], *Wleaf;
  size_t off_X[RING], off_R[RING], off_flag_x, off_flag_r, off_ack;   // byte offsets inside the symmetric buffer
  uint32_t base;
  cudaStream_t sm, sr, sp, sc, sx;     // bulk trailing update (the caller's stream), row work, chain, exchange, signals

  double* blockptr(int row_block, int local_block) const { return A + (int64_t)row_block * L.nb + (int64_t)local_block * L.nb * lda; }
  uint32_t seq(int j) const { return base + (uint32_t)j + 1u; }
  uint32_t ack_needed(int j) const { return j < RING ? base : base + (uint32_t)(j - RING) + 1u; }   // the peer released the panel that last used ring slot j % RING
  int r0_of(int j) const { return std::min(L.upto(me, j), L.nloc(me)); }                             // first local block right of panel j

  int wait_flag(cudaStream_t s, size_t off, uint32_t value) {
    const CUresult r = ((stream_value32_fn)h->fn_wait32)((CUstream)s, (CUdeviceptr)(peer->ptr[me] + off), value, CU_STREAM_WAIT_VALUE_GEQ);
    return r == CUDA_SUCCESS ? CAP_OK : cap_fail(CAP_ERR_CUDA, "cuStreamWaitValue32", "failed");
  }
  int signal(cudaStream_t s, int slot, int q, size_t off, uint32_t value) {
    const CUresult r = ((stream_value32_fn)h->fn_write32)((CUstream)s, (CUdeviceptr)(h->sig_scratch + slot), value, CU_STREAM_WRITE_VALUE_DEFAULT);
    if (r != CUDA_SUCCESS) return cap_fail(CAP_ERR_CUDA, "cuStreamWriteValue32", "failed");
    CUDA_TRY(cudaMemcpyAsync(peer->ptr[q] + off, h->sig_scratch + slot, sizeof(uint32_t), cudaMemcpyDefault, s));
    return CAP_OK;
  }

  // X_j is usable on stream s
  int get_x(cudaStream_t s, int j) {
    const int b = j % RING;
    if (me == L.owner(j)) { CUDA_TRY(cudaStreamWaitEvent(s, h->ev_leaf[b], 0)); return CAP_OK; }
    return wait_flag(s, off_flag_x + (size_t)b * 4, seq(j));
  }
  // R_j,[local blocks r_first, r_first + nblocks) = X_j^T A_j,.  ->  S (the exchange payload) and A; on the handle's current stream
  int solve_cols(int j, int r_first, int64_t ncols) {
    if (ncols <= 0) return CAP_OK;
    const int nb = L.nb, b = j % RING;
    double* Bp = blockptr(j, r_first);
    double* Sp = S[b] + (int64_t)(r_first - r0_of(j)) * nb * nb;
    CAP_TRY(trmm_impl(h, 'L', 'U', 'T', 'N', nb, ncols, 1.0, X[b], nb, Bp, lda, Sp, nb, true, false));
    return copy2d(h, Bp, lda, Sp, nb, nb, ncols, 0);
  }
  int64_t cols_from(int r_first) const { return std::max<int64_t>(0, L.cols(me) - (int64_t)std::min(r_first, L.nloc(me)) * L.nb); }
  // own solved blocks [r_first, r_end) of panel j -> their places in the own ordered row panel (several ranks only)
  int self_copy(int j, int r_first, int r_end) {
    if (!peer) return CAP_OK;
    const int nb = L.nb, b = j % RING, r0 = r0_of(j);
    for (int r = r_first; r < r_end; ++r) {
      const int l = L.global_block(r, me);
      CUDA_TRY(cudaMemcpyAsync(R[b] + (int64_t)(l - j - 1) * nb * nb, S[b] + (int64_t)(r - r0) * nb * nb, (size_t)nb * L.width(l) * 8, cudaMemcpyDeviceToDevice, h->stream));
    }
    return CAP_OK;
  }

  // ---- chain -----------------------------------------------------------------------------------------------------------
  int leaf_and_push(int k) {                       // on P, owner(k) only
    const int nb = L.nb, b = k % RING, w = L.width(k);
    const bool last = (k == L.nblk - 1);
    if (k >= RING && peer) CUDA_TRY(cudaStreamWaitEvent(sp, h->ev_xpush[b], 0));      // X[b] is still the source of panel k-RING's pushes
    tl.mark(sp, k, "leaf");
    CUDA_TRY(cudaMemsetAsync(X[b], 0, (size_t)nb * nb * sizeof(double), sp));
    CAP_TRY(cholinv_rec(h, blockptr(k, k / L.P), lda, X[b], nb, 0, w, true, last ? 0 : 1, Wleaf, k * nb));
    tl.mark(sp, k, "leaf");
    CUDA_TRY(cudaEventRecord(h->ev_leaf[b], sp));
    if (last || !peer) return CAP_OK;
    CUDA_TRY(cudaStreamWaitEvent(sx, h->ev_leaf[b], 0));
    const int first = L.owner(k + 1);                // the next link of the chain gets X_k first
    for (int i = 0; i < L.P; ++i) {
      const int q = (first + i) % L.P;
      if (q == me) continue;
      CAP_TRY(wait_flag(sx, off_ack + (size_t)q * 4, ack_needed(k)));
      CUDA_TRY(cudaMemcpyAsync(peer->ptr[q] + off_X[b], X[b], (size_t)nb * nb * 8, cudaMemcpyDefault, sx));
      CAP_TRY(signal(sx, 1, q, off_flag_x + (size_t)b * 4, seq(k)));
    }
    CUDA_TRY(cudaEventRecord(h->ev_xpush[b], sx));
    return CAP_OK;
  }
  int chain_front(int k) {                          // k >= 1, on P, owner(k) only: everything between X_k-1 and X_k
    if (me != L.owner(k)) return CAP_OK;
    StreamScope scope(h, sp, 1);
    const int j = k - 1, b = j % RING, nb = L.nb;
    CAP_TRY(get_x(sp, j));
    if (k >= 2) CUDA_TRY(cudaStreamWaitEvent(sp, h->ev_r1[(k - 2) % RING], 0));       // U1rest(k-2): block row k-1 is final
    if (j >= RING && peer) CUDA_TRY(cudaStreamWaitEvent(sp, h->ev_xchg[b], 0));        // S[b] still read by the exchange of panel j-RING
    tl.mark(sp, k, "csolve");
    CAP_TRY(solve_cols(j, k / L.P, L.width(k)));
    tl.mark(sp, k, "csolve");
    CUDA_TRY(cudaEventRecord(h->ev_csolve[b], sp));
    if (k >= 2) CUDA_TRY(cudaStreamWaitEvent(sp, h->ev_r2a[(k - 2) % RING], 0));      // U2a(k-2): the last earlier update of block (k,k)
    tl.mark(sp, k, "cdiag");
    CAP_TRY(cap_dsyrk(h, 'U', 'T', L.width(k), nb, -1.0, S[b] + (int64_t)(k / L.P - r0_of(j)) * nb * nb, nb, 1.0, blockptr(k, k / L.P), lda));
    tl.mark(sp, k, "cdiag");
    return leaf_and_push(k);
  }

  // ---- trailing work on M ------------------------------------------------------------------------------------------------
  int solve_rest(int j) {                            // row j right of block j, minus the block the chain solves; then ship it
    const int b = j % RING, nb = L.nb;
    if (j >= L.nblk - 1) return CAP_OK;
    CAP_TRY(get_x(sr, j));
    if (j >= RING && peer) CUDA_TRY(cudaStreamWaitEvent(sr, h->ev_xchg[b], 0));
    int r_first = r0_of(j);
    const bool chain_block = (me == L.owner(j + 1));
    if (chain_block) ++r_first;
    tl.mark(sr, j, "solve");
    CAP_TRY(solve_cols(j, r_first, cols_from(r_first)));
    CAP_TRY(self_copy(j, r_first, L.nloc(me)));
    tl.mark(sr, j, "solve");
    CUDA_TRY(cudaEventRecord(h->ev_solve[b], sr));
    if (!peer) return CAP_OK;
    // exchange(j) on C
    CUDA_TRY(cudaStreamWaitEvent(sc, h->ev_solve[b], 0));
    if (chain_block) CUDA_TRY(cudaStreamWaitEvent(sc, h->ev_csolve[b], 0));
    const int r0 = r0_of(j), n = L.nloc(me);
    tl.mark(sc, j, "xchg");
    if (r0 < n) {
      for (int i = 1; i < L.P; ++i) {
        const int q = (me + i) % L.P;
        CAP_TRY(wait_flag(sc, off_ack + (size_t)q * 4, ack_needed(j)));
        for (int r = r0; r < n; ++r) {
          const int l = L.global_block(r, me);
          CUDA_TRY(cudaMemcpyAsync(peer->ptr[q] + off_R[b] + (size_t)(l - j - 1) * nb * nb * 8, S[b] + (int64_t)(r - r0) * nb * nb,
                                   (size_t)nb * L.width(l) * 8, cudaMemcpyDefault, sc));
        }
        CAP_TRY(signal(sc, 0, q, off_flag_r + ((size_t)me * RING + b) * 4, seq(j)));
      }
    }
    tl.mark(sc, j, "xchg");
    CUDA_TRY(cudaEventRecord(h->ev_xchg[b], sc));
    return CAP_OK;
  }
  int row_data_ready(int j) {                        // every block of the ordered row panel j is in R[b]
    const int b = j % RING;
    if (me == L.owner(j + 1)) {
      CUDA_TRY(cudaStreamWaitEvent(sr, h->ev_csolve[b], 0));
      CAP_TRY(self_copy(j, r0_of(j), r0_of(j) + 1));
    }
    if (!peer) return CAP_OK;
    for (int p = 0; p < L.P; ++p)
      if (p != me && L.cols_right(p, j) > 0) CAP_TRY(wait_flag(sr, off_flag_r + ((size_t)p * RING + b) * 4, seq(j)));
    return CAP_OK;
  }
  int update_row(int j) {                            // U1rest(j): block row j+1, columns right of block j+1
    const int nb = L.nb, b = j % RING, k = j + 1, wk = L.width(k);
    const int r0 = r0_of(j);
    const int rr = (me == L.owner(k)) ? k / L.P + 1 : r0;
    const int64_t ncl = cols_from(rr);
    if (ncl <= 0) return CAP_OK;
    return cap_dgemm(h, 'T', 'N', wk, ncl, nb, -1.0, R[b], nb, S[b] + (int64_t)(rr - r0) * nb * nb, nb, 1.0, blockptr(k, rr), lda);
  }
  int update_diag2(int j) {                          // U2a(j): diagonal block (j+2, j+2), all the chain of panel j+2 still needs from panel j
    const int nb = L.nb, b = j % RING, k = j + 2;
    if (k >= L.nblk || me != L.owner(k)) return CAP_OK;
    return cap_dsyrk(h, 'U', 'T', L.width(k), nb, -1.0, S[b] + (int64_t)(k / L.P - r0_of(j)) * nb * nb, nb, 1.0, blockptr(k, k / L.P), lda);
  }
  int update_rest(int j) {                           // U2b(j): rows >= j+2 of the block columns >= j+3
    const int nb = L.nb, b = j % RING, n = L.nloc(me);
    const int r0 = r0_of(j);
    int r = r0;
    while (r < n && L.global_block(r, me) < j + 3) ++r;
    if (r >= n) return CAP_OK;
    if (h->lookahead_stair) {
      // ONE launch: every tile of this rank's block columns on or above the global diagonal (no per-run tails, no 36-tile syrk launches)
      const int g_last = L.global_block(n - 1, me);
      const int64_t rows = std::min<int64_t>(L.N, (int64_t)(g_last + 1) * nb) - (int64_t)(j + 2) * nb;
      const int rc = stair_launch(h, R[b] + (int64_t)nb * nb, nb, S[b] + (int64_t)(r - r0) * nb * nb, nb, blockptr(j + 2, r), lda,
                                  rows, cols_from(r), nb, -1.0, nb, L.P, me, r, j + 2);
      if (rc != CAP_ERR_UNSUPPORTED) return rc;
    }
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

  // Host order = the order in which every event is recorded before anything waits on it.
  //   R (row stream, above the caller's priority): [row j here] [bulk(j-1) done] U1rest(j) -> U2a(j) -> solve rest of row j+1 -> (C) ship it
  //   M (the caller's stream):                      [row j here] U2b(j), the bulk of the trailing update, as ONE staircase launch
  // U1rest(j) touches block row j+1 only and U2b(j) rows >= j+2, so they run side by side: the row work of a panel (two
  // partial waves of the tile engine, a triangular solve, the copy-engine exchange) hides inside the bulk of the same panel
  // instead of delaying it.  bulk(j) -> U1rest(j+1) is ordered by an event (both update block row j+2).
  int run() {
    const int nblk = L.nblk;
    {
      StreamScope scope(h, sp, 1);
      if (me == L.owner(0)) CAP_TRY(leaf_and_push(0));
    }
    if (nblk > 1) CAP_TRY(chain_front(1));            // before solve_rest(0): its events are waited on by the exchange of panel 0
    { StreamScope scope(h, sr, 2); CAP_TRY(solve_rest(0)); }
    // SMs left to the chain: only while a chain runs on THIS rank beside the work being launched
    auto reserve_for = [&](int k) { return (k < nblk && (L.P == 1 || me == L.owner(k)) && h->lookahead_reserve_sms > 0) ? h->num_sms - h->lookahead_reserve_sms : 0; };
    for (int j = 0; j + 1 < nblk; ++j) {
      const int b = j % RING;
      {
        StreamScope scope(h, sr, 2);
        SmCapScope cap(h, reserve_for(j + 1));         // leaf(j+1) may still be running here
        CAP_TRY(row_data_ready(j));
        CUDA_TRY(cudaEventRecord(h->ev_rowready[b], sr));
        if (j > 0) CUDA_TRY(cudaStreamWaitEvent(sr, h->ev_r2[(j - 1) % RING], 0));     // bulk(j-1) also updated block row j+1
        tl.mark(sr, j, "U1");
        CAP_TRY(update_row(j));
        CUDA_TRY(cudaEventRecord(h->ev_r1[b], sr));
        CAP_TRY(update_diag2(j));
        tl.mark(sr, j, "U1");
        CUDA_TRY(cudaEventRecord(h->ev_r2a[b], sr));
      }
      if (j + 2 < nblk) CAP_TRY(chain_front(j + 2));
      {
        StreamScope scope(h, sr, 2);
        SmCapScope cap(h, reserve_for(j + 2));         // the chain of panel j+2 runs beside the rest of this iteration
        CAP_TRY(solve_rest(j + 1));
      }
      {
        StreamScope scope(h, sm, 0);
        SmCapScope cap(h, reserve_for(j + 2));
        CUDA_TRY(cudaStreamWaitEvent(sm, h->ev_rowready[b], 0));
        if (j + 2 < nblk) {
          GridMultScope gm(h, h->lookahead_grid_mult);
          tl.mark(sm, j, "U2");
          CAP_TRY(update_rest(j));
          tl.mark(sm, j, "U2");
        }
        CUDA_TRY(cudaEventRecord(h->ev_r2[b], sm));
      }
      if (peer) {                                     // panel j's landing buffers here are free again (bulk AND row work read them): tell every peer
        CUDA_TRY(cudaStreamWaitEvent(sx, h->ev_r2[b], 0));
        CUDA_TRY(cudaStreamWaitEvent(sx, h->ev_r2a[b], 0));
        for (int i = 1; i < L.P; ++i) CAP_TRY(signal(sx, 1, (me + i) % L.P, off_ack + (size_t)me * 4, seq(j)));
      }
    }
    // join: the last leaf (P) and the last row work (R) are ordered before whatever follows on the caller's stream; then the final acknowledgement
    CUDA_TRY(cudaEventRecord(h->ev_panel[0], sp));
    CUDA_TRY(cudaStreamWaitEvent(sm, h->ev_panel[0], 0));
    CUDA_TRY(cudaEventRecord(h->ev_panel[0], sr));
    CUDA_TRY(cudaStreamWaitEvent(sm, h->ev_panel[0], 0));
    if (peer) {
      CUDA_TRY(cudaEventRecord(h->ev_panel[1], sm));
      CUDA_TRY(cudaStreamWaitEvent(sx, h->ev_panel[1], 0));
      for (int i = 1; i < L.P; ++i) CAP_TRY(signal(sx, 1, (me + i) % L.P, off_ack + (size_t)me * 4, base + (uint32_t)nblk));
      peer->seq = base + (uint32_t)nblk;
    }
    return CAP_OK;
  }
};

static int async_streams(cap_handle_t h) {
  if (h->xchg_stream) return CAP_OK;
  int lo = 0, hi = 0;
  CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CUDA_TRY(cudaStreamCreateWithPriority(&h->xchg_stream, cudaStreamNonBlocking, hi));
  CUDA_TRY(cudaStreamCreateWithPriority(&h->sig_stream, cudaStreamNonBlocking, hi));
  CUDA_TRY(cudaStreamCreateWithPriority(&h->row_stream, cudaStreamNonBlocking, hi < lo - 1 ? hi + 1 : hi));   // between the chain and the caller's stream
  for (int i = 0; i < cap_handle_s::RING; ++i)
    for (cudaEvent_t* e : {&h->ev_leaf[i], &h->ev_csolve[i], &h->ev_solve[i], &h->ev_xchg[i], &h->ev_xpush[i], &h->ev_rowready[i], &h->ev_r1[i], &h->ev_r2a[i], &h->ev_r2[i]})
      CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  CUDA_TRY(cudaMalloc((void**)&h->sig_scratch, 64));
  for (const char* name : {"cuStreamWaitValue32", "cuStreamWriteValue32"}) {
    cudaDriverEntryPointQueryResult qres;
    void* fn = nullptr;
    if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) { cudaGetLastError(); fn = nullptr; }
    (name[9] == 'a' ? h->fn_wait32 : h->fn_write32) = fn;      // cuStreamWa.. / cuStreamWr..
  }
  return CAP_OK;
}

// Landing buffers of the exchange: one symmetric allocation per handle, re-made (collectively) when a call needs more.
static int async_peer(cap_handle_t h, cap_comm_t comm, size_t bytes, cap_peer_s** out) {
  *out = nullptr;
  if (h->peer_exchange == 0) return CAP_OK;
  if (!h->fn_wait32 || !h->fn_write32)
    return h->peer_exchange == 1 ? cap_fail(CAP_ERR_UNSUPPORTED, "cap_potrf_1d", "cuStreamWaitValue32 / cuStreamWriteValue32 are not available from this driver") : CAP_OK;
  if (h->peer1d && (h->peer1d->comm != comm || h->peer1d->bytes < bytes)) { CAP_TRY(cap_peer_free_collective(h, h->peer1d)); h->peer1d = nullptr; }
  if (!h->peer1d) {
    if (h->peer_failed_comm == comm) return CAP_OK;                 // tried before on this communicator: CUDA IPC is not available
    const int rc = cap_peer_alloc(h, comm, bytes, &h->peer1d);
    if (rc != CAP_OK) { h->peer1d = nullptr; h->peer_failed_comm = comm; return h->peer_exchange == 1 ? rc : CAP_OK; }
  }
  *out = h->peer1d;
  return CAP_OK;
}

int cap_set_lookahead(cap_handle_t h, int reserve_sms, int pipeline) {
  if (reserve_sms < 0 || reserve_sms >= h->num_sms) return cap_fail(CAP_ERR_ARG, "cap_set_lookahead", "reserve must be in [0, #SMs)");
  if (pipeline < -1 || pipeline > 1) return cap_fail(CAP_ERR_ARG, "cap_set_lookahead", "pipeline: -1 = asynchronous schedule (peer copy engines), 0 = two streams + NCCL, 1 = NCCL all-gather on its own stream and communicator");
  h->lookahead_reserve_sms = reserve_sms;
  h->lookahead_pipeline = pipeline;
  return CAP_OK;
}

int cap_potrf_1d(cap_handle_t h, cap_comm_t comm, int64_t N, int nb, double* A_loc, int64_t lda) {
  if (N < 0 || N > INT32_MAX || nb < BASE_NB || nb % BASE_NB || lda < std::max<int64_t>(1, N)) return cap_fail(CAP_ERR_ARG, "cap_potrf_1d", "bad argument (nb: multiple of 128; lda >= N)");
  if (N == 0) return CAP_OK;
  int P = 1, me = 0;
  if (comm) { cap_comm_size(comm, &P); cap_comm_rank(comm, &me); }
  // ---- round 2: the asynchronous schedule (one rank, or several ranks with CUDA-IPC peer buffers) -------------------------
  if (h->lookahead_pipeline < 0) {
    CAP_TRY(async_streams(h));
    Layout1D L{P, nb, N, (int)((N + nb - 1) / nb)};
    constexpr int RING = cap_handle_s::RING;
    const size_t xs = (size_t)nb * nb, rs = (size_t)nb * std::max<int64_t>(N - nb, 1);
    const size_t flag_bytes = 4096;
    cap_peer_s* peer = nullptr;
    if (P > 1) CAP_TRY(async_peer(h, comm, flag_bytes + RING * (xs + rs) * sizeof(double), &peer));
    if (P == 1 || peer) {
      if (!h->panel_stream) {
        int lo = 0, hi = 0;
        CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CUDA_TRY(cudaStreamCreateWithPriority(&h->panel_stream, cudaStreamNonBlocking, hi));
        CUDA_TRY(cudaStreamCreateWithPriority(&h->gather_stream, cudaStreamNonBlocking, hi));
        for (cudaEvent_t* e : {&h->ev_panel[0], &h->ev_panel[1], &h->ev_u1, &h->ev_trmm, &h->ev_gath[0], &h->ev_gath[1], &h->ev_u2})
          CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
      }
      Async1D a{};
      a.h = h; a.peer = peer; a.L = L; a.me = me; a.A = A_loc; a.lda = lda;
      int64_t maxcl0 = 1;
      for (int p = 0; p < P; ++p) maxcl0 = std::max(maxcl0, L.cols_right(p, 0));
      const size_t ss = (size_t)nb * maxcl0, leaf = cholinv_scratch_doubles(nb);
      double* ws;
      CAP_TRY(cap_workspace(h, CAP_WS_POTRF1D, (RING * (ss + (peer ? 0 : xs)) + leaf) * sizeof(double), &ws));
      for (int b = 0; b < RING; ++b) { a.S[b] = ws; ws += ss; }
      if (!peer) for (int b = 0; b < RING; ++b) { a.X[b] = ws; ws += xs; a.R[b] = a.S[b]; }
      a.Wleaf = ws;
      if (peer) {
        a.off_flag_x = 0; a.off_flag_r = 64; a.off_ack = 64 + (size_t)CAP_PEER_MAX * RING * 4;
        size_t off = flag_bytes;
        for (int b = 0; b < RING; ++b) { a.off_X[b] = off; a.X[b] = (double*)(peer->ptr[me] + off); off += xs * sizeof(double); }
        for (int b = 0; b < RING; ++b) { a.off_R[b] = off; a.R[b] = (double*)(peer->ptr[me] + off); off += rs * sizeof(double); }
        a.base = peer->seq;
      }
      a.sm = h->stream; a.sr = h->row_stream; a.sp = h->panel_stream; a.sc = h->xchg_stream; a.sx = h->sig_stream;
      // the side streams start behind everything already enqueued on the caller's stream
      CUDA_TRY(cudaEventRecord(h->ev_u1, a.sm));
      for (cudaStream_t s : {a.sr, a.sp, a.sc, a.sx}) CUDA_TRY(cudaStreamWaitEvent(s, h->ev_u1, 0));
      const char* prof = getenv("CAPITAL_B200_PROFILE");
      a.tl.on = prof && (atoi(prof) >= 2 || (atoi(prof) == 1 && me == 0));
      const int rc = a.run();
      a.tl.report(a.sm, me, a.L);
      return rc;
    }
  }
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
  const bool pipelined = h->lookahead_pipeline == 1;     // (< 0 only reaches here when CUDA IPC is unavailable: two-stream NCCL order)
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
