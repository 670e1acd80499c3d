// b2_slab.cu -- native driver of the slab-decomposed substep: the whole sequence
//   clear, P2G | X: ghost layer -> r+1 and owned layer -> r-1 (both ranks end up with the sums) | grid-op, G2P | C: migrants <-> r+-1
// is enqueued from C++: kernels on the runtime's stream, the three exchanges (ncclSend/ncclRecv over NVLink/NVSwitch
// inside ncclGroupStart/End) on a second stream, ordered by events.  No Python and no host synchronisation in the loop.  (New: the reference has no multi-GPU path, cuda_context.cpp:22.)
// NCCL is bound at run time (dlopen of the libnccl.so.2 that torch already loaded), so the
// library has no link-time dependency on it and single-GPU users never touch it.
#include <dlfcn.h>

#include <algorithm>
#include <cstring>

#include "b2_host.h"

typedef struct ncclComm *ncclComm_t;
typedef struct {
  char internal[128];
} ncclUniqueId;
enum { NCCL_UINT8 = 1, NCCL_UINT64 = 5, NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2 };

struct NcclApi {
  int (*GetUniqueId)(ncclUniqueId *);
  int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
  int (*CommDestroy)(ncclComm_t);
  int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t);
  int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t);
  int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t);
  int (*GroupStart)();
  int (*GroupEnd)();
  const char *(*GetErrorString)(int);
  bool ok = false;
};
static NcclApi &nccl() {
  static NcclApi a;
  if (a.ok) return a;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) throw B2Error(TI_ERROR_NOT_SUPPORTED, std::string("libnccl.so.2 not found: ") + dlerror());
#define B2_SYM(field, name)                                                    \
  a.field = (decltype(a.field))dlsym(h, name);                                 \
  if (!a.field) throw B2Error(TI_ERROR_NOT_SUPPORTED, std::string("NCCL symbol missing: ") + name)
  B2_SYM(GetUniqueId, "ncclGetUniqueId");
  B2_SYM(CommInitRank, "ncclCommInitRank");
  B2_SYM(CommDestroy, "ncclCommDestroy");
  B2_SYM(Send, "ncclSend");
  B2_SYM(Recv, "ncclRecv");
  B2_SYM(AllReduce, "ncclAllReduce");
  B2_SYM(GroupStart, "ncclGroupStart");
  B2_SYM(GroupEnd, "ncclGroupEnd");
  B2_SYM(GetErrorString, "ncclGetErrorString");
#undef B2_SYM
  a.ok = true;
  return a;
}
#define B2_NCCL(expr)                                                                                     \
  do {                                                                                                    \
    int _r = (expr);                                                                                      \
    if (_r != 0) throw B2Error(TI_ERROR_INVALID_STATE, std::string(#expr) + ": " + nccl().GetErrorString(_r)); \
  } while (0)

struct B2Slab_t;
struct B2Comm_t {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  // loopback transport: all slabs of the domain live in THIS process on one device and one stream; a "send" is a
  // device-to-device copy into the peer's receive buffer (same call order and message layout as the NCCL path), so
  // the native driver can be tested at world > 1 on a single GPU.
  bool loopback = false;
  std::vector<B2Slab_t *> peers;
};

static const int64_t BLOCK_REC = 1040, PART_REC = 80, HEADER = 16;

struct B2Slab_t {
  B2Mpm mpm;
  B2Comm_t *comm;
  TiRuntime_t *rt;
  int rank, world, up, down, lo, hi;
  int64_t cap_blocks, cap_particles;
  // bytes actually put on the wire per message (<= capacity; see b2_slab_tune)
  int64_t sendA_bytes, recvA_bytes, sendB_bytes, recvB_bytes, sendL_bytes, recvL_bytes, sendR_bytes, recvR_bytes;
  uint8_t *sendA, *recvA, *sendB, *recvB, *sendL, *sendR, *recvL, *recvR;
  bool tuned = false;   // message sizes only grow after the first tune
  // The exchanges run on a second stream: ghost layers and migrants travel while the compute stream works on blocks / bins
  // that do not depend on them.  ready[k]: the messages of exchange k are packed (compute stream); arrived_*[k]: they are in MY
  // receive buffers (recorded by my own NCCL group, or -- loopback -- by the peer whose copy filled them).
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ready[2], arrived_lo[2], arrived_hi[2], sent;  // arrived_lo / _hi: the message from the lower / upper neighbour is in my buffers;
                                                            // sent: my own sends / copies of the substep have left their buffers
  bool overlap = true;
  // B2_SLAB_PROF=1: CUDA events on the compute stream around every part of the substep; averages printed at destroy
  bool prof = false;
  long long prof_every = 0;  // B2_SLAB_PROF = n > 1: print and reset after every n substeps
  std::vector<cudaEvent_t> pev;
  size_t pev_n = 0;
  double psum[12] = {0};
  long long pcount = 0;
  void mark(cudaStream_t st) {
    if (!prof) return;
    if (pev_n == pev.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      pev.push_back(e);
    }
    cudaEventRecord(pev[pev_n++], st);
  }
  uint32_t *sizes_dev;  // tune: 4 counts out, 4 counts in
  double *diag_dev;     // diagnostics: 8 doubles, reduced over the slabs in place
  std::vector<void *> allocs;
  uint8_t *buf(int64_t bytes) {
    void *p = nullptr;
    B2_CUDA(cudaMalloc(&p, bytes));
    B2_CUDA(cudaMemsetAsync(p, 0, bytes, b2_current_stream()));
    allocs.push_back(p);
    return (uint8_t *)p;
  }
};

static void require(bool ok, TiError code, const char *msg) {
  if (!ok) throw B2Error(code, msg);
}
static void ok0(int32_t rc) {
  if (rc != 0) {
    uint64_t n = 0;
    ti_get_last_error(&n, nullptr);
    std::string m(n ? n : 1, 0);
    ti_get_last_error(&n, &m[0]);
    throw B2Error(TI_ERROR_INVALID_STATE, m);
  }
}

extern "C" {
static void slab_prof_print(B2Slab_t *s);
int32_t b2_comm_unique_id(uint8_t *out128) {
  B2_API_BEGIN
  require(out128 != nullptr, TI_ERROR_ARGUMENT_NULL, "out");
  ncclUniqueId id;
  B2_NCCL(nccl().GetUniqueId(&id));
  memcpy(out128, id.internal, 128);
  return 0;
  B2_API_END(-1)
}
B2Comm b2_comm_init(TiRuntime rt, const uint8_t *id128, int32_t rank, int32_t world) {
  B2_API_BEGIN
  require(rt != nullptr && id128 != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime/id");
  require(world >= 1 && rank >= 0 && rank < world, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "rank/world");
  B2_CUDA(cudaSetDevice(rt->device));
  B2Comm_t *c = new B2Comm_t();
  c->rank = rank, c->world = world;
  ncclUniqueId id;
  memcpy(id.internal, id128, 128);
  B2_NCCL(nccl().CommInitRank(&c->comm, world, id, rank));
  return c;
  B2_API_END(nullptr)
}
B2Comm b2_comm_loopback(TiRuntime rt, int32_t world) {
  B2_API_BEGIN
  require(rt != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(world >= 1, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "world");
  B2Comm_t *c = new B2Comm_t();
  c->world = world, c->loopback = true;
  c->peers.assign((size_t)world, nullptr);
  return c;
  B2_API_END(nullptr)
}
void b2_comm_destroy(B2Comm c) {
  if (!c) return;
  if (c->comm) nccl().CommDestroy(c->comm);
  delete c;
}

B2Slab b2_slab_create(B2Mpm mpm, B2Comm comm, int32_t rank, int32_t world, int32_t lo, int32_t hi, int64_t cap_blocks, int64_t cap_particles) {
  B2_API_BEGIN
  require(mpm != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  require(world == 1 || comm != nullptr, TI_ERROR_ARGUMENT_NULL, "comm");
  require(cap_blocks > 0 && cap_particles > 0, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "capacities");
  B2Slab_t *s = new B2Slab_t();
  s->mpm = mpm, s->comm = comm, s->rank = rank, s->world = world, s->lo = lo, s->hi = hi;
  s->rt = (TiRuntime_t *)b2_mpm_runtime(mpm);
  s->up = rank + 1 < world ? rank + 1 : -1;
  s->down = rank > 0 ? rank - 1 : -1;
  s->cap_blocks = cap_blocks, s->cap_particles = cap_particles;
  const int64_t bb = HEADER + cap_blocks * BLOCK_REC, pb = HEADER + cap_particles * PART_REC;
  s->sendA = s->buf(bb), s->recvA = s->buf(bb), s->sendB = s->buf(bb), s->recvB = s->buf(bb);
  s->sendL = s->buf(pb), s->sendR = s->buf(pb), s->recvL = s->buf(pb), s->recvR = s->buf(pb);
  s->sizes_dev = (uint32_t *)s->buf(64);
  s->diag_dev = (double *)s->buf(64);
  s->sendA_bytes = s->recvA_bytes = s->sendB_bytes = s->recvB_bytes = bb;
  s->sendL_bytes = s->recvL_bytes = s->sendR_bytes = s->recvR_bytes = pb;
  if (comm && comm->loopback) {
    require(world == comm->world && rank >= 0 && rank < world, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "rank / world of the loopback communicator");
    comm->peers[(size_t)rank] = s;
  }
  if (const char *e = getenv("B2_SLAB_OVERLAP")) s->overlap = atoi(e) != 0;
  if (const char *e = getenv("B2_SLAB_PROF")) s->prof = atoi(e) != 0, s->prof_every = atoi(e) > 1 ? atoi(e) : 0;
  B2_CUDA(cudaStreamCreateWithFlags(&s->comm_stream, cudaStreamNonBlocking));
  for (int k = 0; k < 2; k++) {
    B2_CUDA(cudaEventCreateWithFlags(&s->ready[k], cudaEventDisableTiming));
    B2_CUDA(cudaEventCreateWithFlags(&s->arrived_lo[k], cudaEventDisableTiming));
    B2_CUDA(cudaEventCreateWithFlags(&s->arrived_hi[k], cudaEventDisableTiming));
  }
  B2_CUDA(cudaEventCreateWithFlags(&s->sent, cudaEventDisableTiming));
  return s;
  B2_API_END(nullptr)
}
void b2_slab_destroy(B2Slab s) {
  if (!s) return;
  cudaStreamSynchronize(b2_current_stream());
  if (s->prof) slab_prof_print(s);
  for (cudaEvent_t e : s->pev) cudaEventDestroy(e);
  if (s->comm_stream) {
    cudaStreamSynchronize(s->comm_stream);
    for (int k = 0; k < 2; k++) cudaEventDestroy(s->ready[k]), cudaEventDestroy(s->arrived_lo[k]), cudaEventDestroy(s->arrived_hi[k]);
    cudaEventDestroy(s->sent);
    cudaStreamDestroy(s->comm_stream);
  }
  for (void *p : s->allocs) cudaFree(p);
  delete s;
}

// The decomposed substep: three compute phases on the runtime's stream, two exchanges on the slab's comm stream.
//   phase 0: clear, P2G, pack my ghost layer hi (for rank + 1) and my owned layer lo (for rank - 1)        -> ready[0]
//   X      : both layers travel in ONE group: every node of a shared layer ends up with BOTH ranks' contributions on BOTH
//            ranks (a + b == b + a bit for bit), so the owner's post-grid-op values never have to travel back
//   phase 1: wait for X; add what arrived (from below: into my layer lo, activating; from above: into my ghost layer, existing
//            blocks only); grid-op on every block incl. the ghost layer (pointwise: identical on both ranks); G2P; extract
//            migrants                                                                                        -> ready[1]
//   C      : migrants <-> rank +- 1
//   phase 2: wait for C; insert arrivals
// r1 sent the ghost layer up (A), ran the grid-op, and sent the finished layer back down (B): two dependent exchanges per
// substep.  Measured on 2 B200 (B2_SLAB_PROF): running interior blocks / bins under A and B hid 40-60 us of waiting but the
// boundary-only G2P launch that followed cost 56 us at 11 % machine fill; the symmetric exchange removes B altogether.
static void slab_phase(B2Slab_t *s, int ph) {
  B2Mpm m = s->mpm;
  cudaStream_t st = s->rt->stream();
  const int64_t capA = (s->sendA_bytes - HEADER) / BLOCK_REC, capB = (s->sendB_bytes - HEADER) / BLOCK_REC;
  const int64_t capL = (s->sendL_bytes - HEADER) / PART_REC, capR = (s->sendR_bytes - HEADER) / PART_REC;
  const bool multi = s->world > 1;
  switch (ph) {
    case 0:
      s->mark(st);  // 0
      ok0(b2_mpm_begin_substep(m));
      ok0(b2_mpm_clear_grid(m));
      if (multi) ok0(b2_mpm_zero_headers(m, s->up >= 0 ? s->sendA : nullptr, s->down >= 0 ? s->sendB : nullptr, s->down >= 0 ? s->sendL : nullptr,
                                         s->up >= 0 ? s->sendR : nullptr));
      ok0(b2_mpm_p2g(m));
      s->mark(st);  // 1: clear + P2G
      if (s->up >= 0) ok0(b2_mpm_pack_blocks(m, s->hi, s->sendA, capA));
      if (s->down >= 0) ok0(b2_mpm_pack_blocks(m, s->lo, s->sendB, capB));
      if (multi) B2_CUDA(cudaEventRecord(s->ready[0], st));
      s->mark(st);  // 2: pack both layers
      break;
    case 1:
      if (multi) {
        B2_CUDA(cudaStreamWaitEvent(st, s->arrived_lo[0], 0));
        B2_CUDA(cudaStreamWaitEvent(st, s->arrived_hi[0], 0));
      }
      s->mark(st);  // 3: wait for X
      if (s->down >= 0) ok0(b2_mpm_unpack_blocks(m, s->recvA, (s->recvA_bytes - HEADER) / BLOCK_REC, 0));
      if (s->up >= 0) ok0(b2_mpm_unpack_blocks(m, s->recvB, (s->recvB_bytes - HEADER) / BLOCK_REC, 2));
      s->mark(st);  // 4: add arrivals
      ok0(b2_mpm_grid_op(m));
      s->mark(st);  // 5: grid-op
      ok0(b2_mpm_g2p(m));
      s->mark(st);  // 6: G2P
      if (multi) ok0(b2_mpm_extract_migrants2(m, s->down >= 0 ? s->sendL : nullptr, capL, s->up >= 0 ? s->sendR : nullptr, capR));
      if (multi) B2_CUDA(cudaEventRecord(s->ready[1], st));
      s->mark(st);  // 7: extract migrants
      break;
    default:
      if (multi) {
        B2_CUDA(cudaStreamWaitEvent(st, s->arrived_lo[1], 0));
        B2_CUDA(cudaStreamWaitEvent(st, s->arrived_hi[1], 0));
        B2_CUDA(cudaStreamWaitEvent(st, s->sent, 0));  // the send buffers are packed again by the next substep
      }
      s->mark(st);  // 8: wait for C
      if (multi)
        ok0(b2_mpm_insert_records2(m, s->up >= 0 ? s->recvR : nullptr, (s->recvR_bytes - HEADER) / PART_REC, s->down >= 0 ? s->recvL : nullptr,
                                   (s->recvL_bytes - HEADER) / PART_REC));
      s->mark(st);  // 9: insert arrivals
      ok0(b2_mpm_end_substep(m));
  }
}
// exchange ex (0: ghost / owned layers, 1: migrants) of this rank, enqueued on its comm stream behind ready[ex]
static void slab_exchange(B2Slab_t *s, int ex) {
  if (s->world == 1) return;
  cudaStream_t sc = s->overlap ? s->comm_stream : s->rt->stream();
  B2_CUDA(cudaStreamWaitEvent(sc, s->ready[ex], 0));
  const uint8_t *to_up = ex == 0 ? s->sendA : s->sendR, *to_down = ex == 0 ? s->sendB : s->sendL;
  const int64_t up_bytes = ex == 0 ? s->sendA_bytes : s->sendR_bytes, down_bytes = ex == 0 ? s->sendB_bytes : s->sendL_bytes;
  uint8_t *from_down = ex == 0 ? s->recvA : s->recvL, *from_up = ex == 0 ? s->recvB : s->recvR;
  const int64_t from_down_bytes = ex == 0 ? s->recvA_bytes : s->recvL_bytes, from_up_bytes = ex == 0 ? s->recvB_bytes : s->recvR_bytes;
  if (s->comm->loopback) {
    // a "send" is a device-to-device copy into the peer's receive buffer; the peer's arrived event follows the copy
    if (s->up >= 0) {
      B2Slab_t *p = s->comm->peers[s->up];
      B2_CUDA(cudaMemcpyAsync(ex == 0 ? p->recvA : p->recvL, to_up, (size_t)up_bytes, cudaMemcpyDeviceToDevice, sc));
      B2_CUDA(cudaEventRecord(p->arrived_lo[ex], sc));
    } else {
      B2_CUDA(cudaEventRecord(s->arrived_hi[ex], sc));  // nothing will arrive from above: the wait must not dangle
    }
    if (s->down >= 0) {
      B2Slab_t *p = s->comm->peers[s->down];
      B2_CUDA(cudaMemcpyAsync(ex == 0 ? p->recvB : p->recvR, to_down, (size_t)down_bytes, cudaMemcpyDeviceToDevice, sc));
      B2_CUDA(cudaEventRecord(p->arrived_hi[ex], sc));
    } else {
      B2_CUDA(cudaEventRecord(s->arrived_lo[ex], sc));
    }
    if (ex == 1) B2_CUDA(cudaEventRecord(s->sent, sc));
    return;
  }
  NcclApi &n = nccl();
  B2_NCCL(n.GroupStart());
  if (s->down >= 0) {
    B2_NCCL(n.Send(to_down, (size_t)down_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
    B2_NCCL(n.Recv(from_down, (size_t)from_down_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
  }
  if (s->up >= 0) {
    B2_NCCL(n.Send(to_up, (size_t)up_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
    B2_NCCL(n.Recv(from_up, (size_t)from_up_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
  }
  B2_NCCL(n.GroupEnd());
  B2_CUDA(cudaEventRecord(s->arrived_lo[ex], sc));
  B2_CUDA(cudaEventRecord(s->arrived_hi[ex], sc));
  if (ex == 1) B2_CUDA(cudaEventRecord(s->sent, sc));
}

static void slab_prof_print(B2Slab_t *s) {
  static const char *names[9] = {"clear+p2g", "pack", "wait_X", "unpack", "grid_op", "g2p", "extract", "wait_C", "insert"};
  if (!s->pcount) return;
  fprintf(stderr, "[b2_slab rank %d/%d] per substep over %lld substeps (ms):", s->rank, s->world, s->pcount);
  double tot = 0;
  for (int k = 0; k < 9; k++) fprintf(stderr, " %s %.4f", names[k], s->psum[k] / s->pcount), tot += s->psum[k] / s->pcount;
  fprintf(stderr, " | total %.4f\n", tot);
  for (int k = 0; k < 12; k++) s->psum[k] = 0;
  s->pcount = 0;
}
static void slab_prof_collect(B2Slab_t *s) {
  if (!s->prof || s->pev_n == 0) return;
  cudaEventSynchronize(s->pev[s->pev_n - 1]);
  for (size_t b = 0; b + 10 <= s->pev_n; b += 10) {
    for (int k = 0; k < 9; k++) {
      float ms = 0;
      cudaEventElapsedTime(&ms, s->pev[b + k], s->pev[b + k + 1]);
      s->psum[k] += ms;
    }
    s->pcount++;
  }
  s->pev_n = 0;
  if (s->prof_every > 0 && s->pcount >= s->prof_every) slab_prof_print(s);
}
int32_t b2_slab_substep(B2Slab s, int32_t nsteps) {
  B2_API_BEGIN
  require(s != nullptr, TI_ERROR_ARGUMENT_NULL, "slab");
  require(s->world == 1 || !s->comm->loopback, TI_ERROR_INVALID_STATE, "loopback slabs advance together: b2_slab_substep_lockstep");
  for (int it = 0; it < nsteps; it++)
    for (int ph = 0; ph < 3; ph++) {
      slab_phase(s, ph);
      if (ph < 2) slab_exchange(s, ph);
    }
  slab_prof_collect(s);
  return 0;
  B2_API_END(-1)
}
// All slabs of a loopback communicator, phase by phase (what the ranks of an NCCL job do concurrently).
int32_t b2_slab_substep_lockstep(B2Slab *slabs, int32_t world, int32_t nsteps) {
  B2_API_BEGIN
  require(slabs != nullptr, TI_ERROR_ARGUMENT_NULL, "slabs");
  for (int r = 0; r < world; r++)
    require(slabs[r] != nullptr && slabs[r]->world == world && slabs[r]->rank == r && (world == 1 || slabs[r]->comm->loopback), TI_ERROR_INVALID_ARGUMENT,
            "slabs[r] must be rank r of one loopback communicator");
  for (int it = 0; it < nsteps; it++)
    for (int ph = 0; ph < 3; ph++) {
      for (int r = 0; r < world; r++) slab_phase(slabs[r], ph);
      if (ph < 2)
        for (int r = 0; r < world; r++) slab_exchange(slabs[r], ph);
    }
  return 0;
  B2_API_END(-1)
}

// Whole-domain particle diagnostics: every slab reduces its own particles (k_mpm_diag), then
// ncclAllReduce over the slabs -- sums for out[0..5], maxima (as ordered bit patterns) for out[6..7] -- on the
// same stream, one group.  Every rank gets the result.  Synchronises.
int32_t b2_slab_diagnostics(B2Slab s, double *out8) {
  B2_API_BEGIN
  require(s != nullptr && out8 != nullptr, TI_ERROR_ARGUMENT_NULL, "slab/out");
  cudaStream_t st = s->rt->stream();
  if (s->world > 1 && s->comm->loopback) {  // peers live in this process: reduce their vectors on the host
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (B2Slab_t *p : s->comm->peers) {
      double v[8];
      ok0(b2_mpm_diagnostics(p->mpm, v));
      for (int k = 0; k < 6; k++) acc[k] += v[k];
      for (int k = 6; k < 8; k++) acc[k] = std::max(acc[k], v[k]);
    }
    memcpy(out8, acc, sizeof(acc));
    return 0;
  }
  ok0(b2_mpm_diagnostics_device(s->mpm, s->diag_dev));
  if (s->world > 1) {
    NcclApi &n = nccl();
    B2_NCCL(n.GroupStart());
    B2_NCCL(n.AllReduce(s->diag_dev, s->diag_dev, 6, NCCL_FLOAT64, NCCL_SUM, s->comm->comm, st));
    B2_NCCL(n.AllReduce(s->diag_dev + 6, s->diag_dev + 6, 2, NCCL_UINT64, NCCL_MAX, s->comm->comm, st));
    B2_NCCL(n.GroupEnd());
  }
  B2_CUDA(cudaMemcpyAsync(out8, s->diag_dev, 64, cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  return 0;
  B2_API_END(-1)
}
// Shrinks the bytes put on the wire to 2x the record counts of the last substep (+256 records,
// never above the capacity); neighbours learn each other's sizes through one tiny exchange.
// Synchronises.  An overflow after tuning raises the device-side error flag (bin overflow).
int32_t b2_slab_tune(B2Slab s) {
  B2_API_BEGIN
  require(s != nullptr, TI_ERROR_ARGUMENT_NULL, "slab");
  if (s->world == 1) return 0;
  cudaStream_t st = s->rt->stream();
  uint32_t cnt[4] = {0, 0, 0, 0};  // A (to up), B (to down), L (to down), R (to up)
  const uint8_t *src[4] = {s->sendA, s->sendB, s->sendL, s->sendR};
  for (int k = 0; k < 4; k++) B2_CUDA(cudaMemcpyAsync(&cnt[k], src[k], 4, cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  auto want = [&](uint32_t c, int64_t cap, int64_t rec) { return HEADER + std::min<int64_t>(cap, 2 * (int64_t)c + 256) * rec; };
  int64_t sb[4] = {want(cnt[0], s->cap_blocks, BLOCK_REC), want(cnt[1], s->cap_blocks, BLOCK_REC), want(cnt[2], s->cap_particles, PART_REC),
                   want(cnt[3], s->cap_particles, PART_REC)};
  if (s->tuned) {  // ghost layers and migration grow as material reaches a boundary: after the first tune a message never shrinks
    const int64_t prev[4] = {s->sendA_bytes, s->sendB_bytes, s->sendL_bytes, s->sendR_bytes};
    for (int k = 0; k < 4; k++) sb[k] = std::max(sb[k], prev[k]);
  }
  s->tuned = true;
  uint32_t out[4];
  for (int k = 0; k < 4; k++) out[k] = (uint32_t)((sb[k] - HEADER) / (k < 2 ? BLOCK_REC : PART_REC));
  if (s->comm->loopback) {  // the peers are in this process: set their receive sizes directly
    s->sendA_bytes = sb[0], s->sendB_bytes = sb[1], s->sendL_bytes = sb[2], s->sendR_bytes = sb[3];
    if (s->up >= 0) s->comm->peers[s->up]->recvA_bytes = sb[0], s->comm->peers[s->up]->recvL_bytes = sb[3];
    if (s->down >= 0) s->comm->peers[s->down]->recvB_bytes = sb[1], s->comm->peers[s->down]->recvR_bytes = sb[2];
    return 0;
  }
  // to up: {A, R}; to down: {B, L}.  from down: its {A, R}; from up: its {B, L}
  uint32_t to_up[2] = {out[0], out[3]}, to_down[2] = {out[1], out[2]}, from_down[2] = {0, 0}, from_up[2] = {0, 0};
  B2_CUDA(cudaMemcpyAsync(s->sizes_dev, to_up, 8, cudaMemcpyHostToDevice, st));
  B2_CUDA(cudaMemcpyAsync(s->sizes_dev + 2, to_down, 8, cudaMemcpyHostToDevice, st));
  NcclApi &n = nccl();
  B2_NCCL(n.GroupStart());
  if (s->up >= 0) {
    B2_NCCL(n.Send(s->sizes_dev, 8, NCCL_UINT8, s->up, s->comm->comm, st));
    B2_NCCL(n.Recv(s->sizes_dev + 4, 8, NCCL_UINT8, s->up, s->comm->comm, st));
  }
  if (s->down >= 0) {
    B2_NCCL(n.Send(s->sizes_dev + 2, 8, NCCL_UINT8, s->down, s->comm->comm, st));
    B2_NCCL(n.Recv(s->sizes_dev + 6, 8, NCCL_UINT8, s->down, s->comm->comm, st));
  }
  B2_NCCL(n.GroupEnd());
  B2_CUDA(cudaMemcpyAsync(from_up, s->sizes_dev + 4, 8, cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaMemcpyAsync(from_down, s->sizes_dev + 6, 8, cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  s->sendA_bytes = sb[0], s->sendB_bytes = sb[1], s->sendL_bytes = sb[2], s->sendR_bytes = sb[3];
  if (s->down >= 0) {  // down's {A, R} land in my recvA / recvL
    s->recvA_bytes = HEADER + (int64_t)from_down[0] * BLOCK_REC;
    s->recvL_bytes = HEADER + (int64_t)from_down[1] * PART_REC;
  }
  if (s->up >= 0) {  // up's {B, L} land in my recvB / recvR
    s->recvB_bytes = HEADER + (int64_t)from_up[0] * BLOCK_REC;
    s->recvR_bytes = HEADER + (int64_t)from_up[1] * PART_REC;
  }
  return 0;
  B2_API_END(-1)
}
}  // extern "C"
