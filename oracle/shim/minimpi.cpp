/* TEST INFRASTRUCTURE (oracle).  Not shipped, not on the product path.
 *
 * minimpi: a fork + shared-memory miniature MPI, just large enough to run the reference's
 * own bench mains and headers unmodified inside one container without an MPI install.
 *
 *   - ranks are PROCESSES (the reference's generators use the process-global drand48 state,
 *     src/matrix/structure.hpp:72-115), forked by the real main() below from MINIMPI_NP;
 *   - every collective is: copy my contribution to my shared staging buffer, barrier on the
 *     communicator, read what I need from the peers' staging buffers, barrier;
 *   - reductions are summed in communicator-rank order, so the oracle is deterministic.
 *
 * Environment: MINIMPI_NP (default 1), MINIMPI_STAGE_MB (per-rank staging, default 1024).
 */
#define MINIMPI_IMPLEMENTATION
#include "mpi.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <sched.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

namespace {

constexpr int kMaxRanks = 64;
constexpr int kMaxBarriers = 8192;

struct Barrier {
  std::atomic<int> count;
  std::atomic<int> generation;
};

struct SplitSlot {
  long color, key;
  int barrier_id;
};

struct Control {
  Barrier barriers[kMaxBarriers];
  std::atomic<int> next_barrier;
  SplitSlot split[kMaxRanks];
  std::atomic<long> p2p_post[kMaxRanks][kMaxRanks];
  std::atomic<long> p2p_done[kMaxRanks][kMaxRanks];
};

struct Comm {
  std::vector<int> members;  // world ranks in communicator-rank order
  int my_index;
  int barrier_id;
};

Control* g_ctl = nullptr;
char* g_stage = nullptr;  // np * stage_bytes
size_t g_stage_bytes = 0;
int g_np = 1, g_me = 0;
std::vector<Comm> g_comms;
long g_posted[kMaxRanks], g_finished[kMaxRanks];

inline char* stage_of(int world_rank) { return g_stage + (size_t)world_rank * g_stage_bytes; }

void spin_pause(int& spins) {
  if (++spins > 64) { sched_yield(); spins = 0; }
}

void barrier_on(const Comm& c) {
  const int n = (int)c.members.size();
  if (n == 1) return;
  Barrier& b = g_ctl->barriers[c.barrier_id];
  const int gen = b.generation.load(std::memory_order_acquire);
  if (b.count.fetch_add(1, std::memory_order_acq_rel) == n - 1) {
    b.count.store(0, std::memory_order_relaxed);
    b.generation.fetch_add(1, std::memory_order_acq_rel);
  } else {
    int spins = 0;
    while (b.generation.load(std::memory_order_acquire) == gen) spin_pause(spins);
  }
}

void publish(const void* src, size_t bytes) {
  if (bytes > g_stage_bytes) {
    fprintf(stderr, "minimpi: message of %zu bytes exceeds staging (%zu); raise MINIMPI_STAGE_MB\n", bytes, g_stage_bytes);
    _exit(3);
  }
  memcpy(stage_of(g_me), src, bytes);
}

template <typename T>
void combine(T* dst, const T* src, long n, MPI_Op op) {
  if (op == MPI_SUM) for (long i = 0; i < n; ++i) dst[i] += src[i];
  else for (long i = 0; i < n; ++i) dst[i] = std::max(dst[i], src[i]);
}

void reduce_from_stage(const Comm& c, void* out, long count, MPI_Datatype t, MPI_Op op) {
  memcpy(out, stage_of(c.members[0]), (size_t)count * t);
  for (size_t r = 1; r < c.members.size(); ++r) {
    if (t == MPI_DOUBLE) combine((double*)out, (const double*)stage_of(c.members[r]), count, op);
    else combine((float*)out, (const float*)stage_of(c.members[r]), count, op);
  }
}

}  // namespace

extern "C" {

int MPI_Init_thread(int*, char***, int, int* provided) { if (provided) *provided = MPI_THREAD_SINGLE; return 0; }
int MPI_Finalize(void) { return 0; }
int MPI_Comm_rank(MPI_Comm c, int* rank) { *rank = g_comms[c].my_index; return 0; }
int MPI_Comm_size(MPI_Comm c, int* size) { *size = (int)g_comms[c].members.size(); return 0; }
int MPI_Comm_free(MPI_Comm*) { return 0; }
int MPI_Barrier(MPI_Comm c) { barrier_on(g_comms[c]); return 0; }
int PMPI_Barrier(MPI_Comm c) { return MPI_Barrier(c); }
double MPI_Wtime(void) { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
int MPI_Wait(MPI_Request*, MPI_Status*) { return 0; }

int MPI_Comm_split(MPI_Comm ci, int color, int key, MPI_Comm* out) {
  const Comm parent = g_comms[ci];
  g_ctl->split[g_me].color = color; g_ctl->split[g_me].key = key; g_ctl->split[g_me].barrier_id = -1;
  barrier_on(parent);
  std::vector<std::pair<long, int>> group;  // (key, world rank) of same-colour members
  for (int w : parent.members) if (g_ctl->split[w].color == color) group.emplace_back(g_ctl->split[w].key, w);
  std::sort(group.begin(), group.end());
  Comm child;
  for (auto& kv : group) child.members.push_back(kv.second);
  child.my_index = (int)(std::find(child.members.begin(), child.members.end(), g_me) - child.members.begin());
  const int leader = child.members[0];
  if (leader == g_me) {
    int id = g_ctl->next_barrier.fetch_add(1);
    if (id >= kMaxBarriers) { fprintf(stderr, "minimpi: out of barrier slots\n"); _exit(3); }
    g_ctl->split[g_me].barrier_id = id;
  }
  barrier_on(parent);
  child.barrier_id = g_ctl->split[leader].barrier_id;
  barrier_on(parent);
  g_comms.push_back(child);
  *out = (int)g_comms.size() - 1;
  return 0;
}

int MPI_Comm_dup(MPI_Comm c, MPI_Comm* out) { return MPI_Comm_split(c, 0, g_comms[c].my_index, out); }

int MPI_Bcast(void* buf, long count, MPI_Datatype t, int root, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  if (c.members.size() == 1 || count == 0) return 0;
  if (c.my_index == root) publish(buf, (size_t)count * t);
  barrier_on(c);
  if (c.my_index != root) memcpy(buf, stage_of(c.members[root]), (size_t)count * t);
  barrier_on(c);
  return 0;
}
int MPI_Ibcast(void* b, long n, MPI_Datatype t, int root, MPI_Comm c, MPI_Request* r) { *r = 0; return MPI_Bcast(b, n, t, root, c); }

int MPI_Allreduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  const void* src = (s == MPI_IN_PLACE) ? r : s;
  if (c.members.size() == 1) { if (src != r) memcpy(r, src, (size_t)count * t); return 0; }
  publish(src, (size_t)count * t);
  barrier_on(c);
  reduce_from_stage(c, r, count, t, op);
  barrier_on(c);
  return 0;
}
int PMPI_Allreduce(const void* s, void* r, long n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { return MPI_Allreduce(s, r, n, t, op, c); }
int MPI_Iallreduce(const void* s, void* r, long n, MPI_Datatype t, MPI_Op op, MPI_Comm c, MPI_Request* q) { *q = 0; return MPI_Allreduce(s, r, n, t, op, c); }

int MPI_Reduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  const void* src = (s == MPI_IN_PLACE) ? r : s;
  if (c.members.size() == 1) { if (src != r) memcpy(r, src, (size_t)count * t); return 0; }
  publish(src, (size_t)count * t);
  barrier_on(c);
  if (c.my_index == root) reduce_from_stage(c, r, count, t, op);
  barrier_on(c);
  return 0;
}

int MPI_Allgather(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  publish(s, (size_t)sc * st);
  barrier_on(c);
  for (size_t i = 0; i < c.members.size(); ++i) memcpy((char*)r + i * (size_t)rc * rt, stage_of(c.members[i]), (size_t)rc * rt);
  barrier_on(c);
  return 0;
}

int MPI_Gather(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  publish(s, (size_t)sc * st);
  barrier_on(c);
  if (c.my_index == root)
    for (size_t i = 0; i < c.members.size(); ++i) memcpy((char*)r + i * (size_t)rc * rt, stage_of(c.members[i]), (size_t)rc * rt);
  barrier_on(c);
  return 0;
}

int MPI_Scatter(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  if (c.my_index == root) publish(s, (size_t)sc * st * c.members.size());
  barrier_on(c);
  memcpy(r, stage_of(c.members[root]) + (size_t)c.my_index * rc * rt, (size_t)rc * rt);
  barrier_on(c);
  return 0;
}
int MPI_Iscatter(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request* q) {
  *q = 0; return MPI_Scatter(s, sc, st, r, rc, rt, root, c);
}

int MPI_Alltoall(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, MPI_Comm ci) {
  const Comm& c = g_comms[ci];
  publish(s, (size_t)sc * st * c.members.size());
  barrier_on(c);
  for (size_t i = 0; i < c.members.size(); ++i)
    memcpy((char*)r + i * (size_t)rc * rt, stage_of(c.members[i]) + (size_t)c.my_index * sc * st, (size_t)rc * rt);
  barrier_on(c);
  return 0;
}

int MPI_Sendrecv_replace(void* buf, long count, MPI_Datatype t, int dest, int, int src, int, MPI_Comm ci, MPI_Status*) {
  const Comm& c = g_comms[ci];
  if (dest != src) { fprintf(stderr, "minimpi: Sendrecv_replace only supports a symmetric partner\n"); _exit(3); }
  const int partner = c.members[dest];
  if (partner == g_me) return 0;
  publish(buf, (size_t)count * t);
  int spins = 0;
  const long post = ++g_posted[partner];
  g_ctl->p2p_post[g_me][partner].store(post, std::memory_order_release);
  while (g_ctl->p2p_post[partner][g_me].load(std::memory_order_acquire) < post) spin_pause(spins);
  memcpy(buf, stage_of(partner), (size_t)count * t);
  const long fin = ++g_finished[partner];
  g_ctl->p2p_done[g_me][partner].store(fin, std::memory_order_release);
  while (g_ctl->p2p_done[partner][g_me].load(std::memory_order_acquire) < fin) spin_pause(spins);
  return 0;
}

}  // extern "C"

int main(int argc, char** argv) {
  const char* e = getenv("MINIMPI_NP");
  g_np = e ? atoi(e) : 1;
  if (g_np < 1 || g_np > kMaxRanks) { fprintf(stderr, "minimpi: bad MINIMPI_NP\n"); return 2; }
  e = getenv("MINIMPI_STAGE_MB");
  g_stage_bytes = (size_t)(e ? atol(e) : 1024) << 20;
  g_ctl = (Control*)mmap(nullptr, sizeof(Control), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  g_stage = (char*)mmap(nullptr, g_stage_bytes * g_np, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
  if (g_ctl == MAP_FAILED || g_stage == MAP_FAILED) { perror("minimpi: mmap"); return 2; }
  memset((void*)g_ctl, 0, sizeof(Control));
  g_ctl->next_barrier.store(1);
  Comm world;
  for (int r = 0; r < g_np; ++r) world.members.push_back(r);
  world.barrier_id = 0;

  if (g_np == 1) {
    world.my_index = 0; g_me = 0; g_comms.push_back(world);
    return minimpi_rank_main(argc, argv);
  }
  std::vector<pid_t> kids;
  for (int r = 0; r < g_np; ++r) {
    pid_t pid = fork();
    if (pid < 0) { perror("minimpi: fork"); return 2; }
    if (pid == 0) {
      g_me = r; world.my_index = r; g_comms.push_back(world);
      int rc = minimpi_rank_main(argc, argv);
      fflush(stdout); fflush(stderr);
      _exit(rc);
    }
    kids.push_back(pid);
  }
  int failed = 0;
  for (pid_t pid : kids) {
    int status = 0;
    waitpid(pid, &status, 0);
    if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) failed = 1;
  }
  return failed;
}
