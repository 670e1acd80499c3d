// b2_slab.cu -- native driver of the slab-decomposed substep: the whole sequence
//   clear, P2G | A: ghost layer -> r+1 | grid-op | B: owned layer -> r-1 | G2P | C: migrants <-> r+-1
// is enqueued from C++: kernels on the runtime's stream, the three exchanges (ncclSend/ncclRecv over NVLink/NVSwitch
// inside ncclGroupStart/End) on a second stream, ordered by events.  While A travels the grid-op runs on every block
// outside the receiving layer; while B travels G2P runs on every bin that does not read the ghost layer (SURVEY 8e:
// "interior blocks overlap").  No Python and no host synchronisation in the loop.  (New: the reference has no multi-GPU path, cuda_context.cpp:22.)
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
  // that do not depend on them.  ready[k]: message k packed (compute stream); arrived[k]: message k is in MY receive buffers
  // (recorded by my own NCCL group, or -- loopback -- by the peer whose copy filled them).
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ready[3], arrived[3], arrived2, sent;  // sent: my own sends / copies of the substep have left their buffers  // (exchange C has two senders in loopback mode: arrived[2] from below, arrived2 from above)
  bool overlap = true;
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

static uint8_t *recv_buffer(B2Slab_t *s, int which) { return which == 0 ? s->recvA : which == 1 ? s->recvB : which == 2 ? s->recvL : s->recvR; }
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

static uint8_t *recv_buffer(B2Slab_t *s, int which);
extern "C" {
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
  B2_CUDA(cudaStreamCreateWithFlags(&s->comm_stream, cudaStreamNonBlocking));
  for (int k = 0; k < 3; k++) {
    B2_CUDA(cudaEventCreateWithFlags(&s->ready[k], cudaEventDisableTiming));
    B2_CUDA(cudaEventCreateWithFlags(&s->arrived[k], cudaEventDisableTiming));
  }
  B2_CUDA(cudaEventCreateWithFlags(&s->arrived2, cudaEventDisableTiming));
  B2_CUDA(cudaEventCreateWithFlags(&s->sent, cudaEventDisableTiming));
  return s;
  B2_API_END(nullptr)
}
void b2_slab_destroy(B2Slab s) {
  if (!s) return;
  cudaStreamSynchronize(b2_current_stream());
  if (s->comm_stream) {
    cudaStreamSynchronize(s->comm_stream);
    for (int k = 0; k < 3; k++) cudaEventDestroy(s->ready[k]), cudaEventDestroy(s->arrived[k]);
    cudaEventDestroy(s->arrived2), cudaEventDestroy(s->sent);
    cudaStreamDestroy(s->comm_stream);
  }
  for (void *p : s->allocs) cudaFree(p);
  delete s;
}

// The decomposed substep: four compute phases on the runtime's stream, three exchanges on the slab's comm stream.
//   phase 0: clear, P2G, pack ghost layer hi                                  -> ready[0]
//   A      : ghost layer -> rank + 1                                            (waits ready[0]; -> arrived[0] on the receiver)
//   phase 1: grid-op on every block outside layer lo  ||  A in flight
//            wait arrived[0]: add ghosts into layer lo, grid-op on layer lo, pack layer lo   -> ready[1]
//   B      : owned layer lo -> rank - 1 (becomes its ghost layer)               (waits ready[1]; -> arrived[1])
//   phase 2: G2P on every bin outside layer hi - 1  ||  B in flight
//            wait arrived[1]: overwrite ghost layer, G2P on the bins of layer hi - 1, extract migrants  -> ready[2]
//   C      : migrants <-> rank +- 1                                             (waits ready[2]; -> arrived[2])
//   phase 3: wait arrived[2]: insert arrivals
// With overlap off (B2_SLAB_OVERLAP=0) the same calls run unsplit and the comm stream is the compute stream.
static void slab_phase(B2Slab_t *s, int ph) {
  B2Mpm m = s->mpm;
  cudaStream_t st = s->rt->stream();
  const int64_t capA = (s->sendA_bytes - HEADER) / BLOCK_REC, capB = (s->sendB_bytes - HEADER) / BLOCK_REC;
  const int64_t capL = (s->sendL_bytes - HEADER) / PART_REC, capR = (s->sendR_bytes - HEADER) / PART_REC;
  const bool multi = s->world > 1, split = multi && s->overlap;
  switch (ph) {
    case 0:
      ok0(b2_mpm_begin_substep(m));
      ok0(b2_mpm_clear_grid(m));
      if (multi) ok0(b2_mpm_zero_headers(m, s->up >= 0 ? s->sendA : nullptr, s->down >= 0 ? s->sendB : nullptr, s->down >= 0 ? s->sendL : nullptr,
                                         s->up >= 0 ? s->sendR : nullptr));
      ok0(b2_mpm_p2g(m));
      if (s->up >= 0) ok0(b2_mpm_pack_blocks(m, s->hi, s->sendA, capA));
      if (multi) B2_CUDA(cudaEventRecord(s->ready[0], st));
      break;
    case 1:
      if (split && s->down >= 0) ok0(b2_mpm_grid_op_layer(m, 2, s->lo));  // everything but the layer that receives ghosts
      if (multi) B2_CUDA(cudaStreamWaitEvent(st, s->arrived[0], 0));
      if (s->down >= 0) ok0(b2_mpm_unpack_blocks(m, s->recvA, (s->recvA_bytes - HEADER) / BLOCK_REC, 0));
      if (split && s->down >= 0)
        ok0(b2_mpm_grid_op_layer(m, 1, s->lo));
      else
        ok0(b2_mpm_grid_op(m));
      if (s->down >= 0) ok0(b2_mpm_pack_blocks(m, s->lo, s->sendB, capB));
      if (multi) B2_CUDA(cudaEventRecord(s->ready[1], st));
      break;
    case 2:
      if (split && s->up >= 0) ok0(b2_mpm_g2p_layer(m, 2, s->hi - 1, 0));  // bins that never read the ghost layer
      if (multi) B2_CUDA(cudaStreamWaitEvent(st, s->arrived[1], 0));
      if (s->up >= 0) ok0(b2_mpm_unpack_blocks(m, s->recvB, (s->recvB_bytes - HEADER) / BLOCK_REC, 1));
      if (split && s->up >= 0)
        ok0(b2_mpm_g2p_layer(m, 1, s->hi - 1, 1));
      else
        ok0(b2_mpm_g2p(m));
      if (multi) ok0(b2_mpm_extract_migrants2(m, s->down >= 0 ? s->sendL : nullptr, capL, s->up >= 0 ? s->sendR : nullptr, capR));
      if (multi) B2_CUDA(cudaEventRecord(s->ready[2], st));
      break;
    default:
      if (multi) {
        B2_CUDA(cudaStreamWaitEvent(st, s->arrived[2], 0));
        if (s->comm->loopback) B2_CUDA(cudaStreamWaitEvent(st, s->arrived2, 0));
        B2_CUDA(cudaStreamWaitEvent(st, s->sent, 0));  // the send buffers are packed again by the next substep
        ok0(b2_mpm_insert_records2(m, s->up >= 0 ? s->recvR : nullptr, (s->recvR_bytes - HEADER) / PART_REC, s->down >= 0 ? s->recvL : nullptr,
                                   (s->recvL_bytes - HEADER) / PART_REC));
      }
      ok0(b2_mpm_end_substep(m));
  }
}
// exchange ex of this rank, enqueued on its comm stream behind ready[ex]
static void slab_exchange(B2Slab_t *s, int ex) {
  if (s->world == 1) return;
  cudaStream_t sc = s->overlap ? s->comm_stream : s->rt->stream();
  B2_CUDA(cudaStreamWaitEvent(sc, s->ready[ex], 0));
  if (s->comm->loopback) {
    // a "send" is a device-to-device copy into the peer's receive buffer; the peer's arrived event follows the copy
    auto put = [&](const uint8_t *src, int64_t bytes, int dst, int which, bool second) {
      B2Slab_t *p = s->comm->peers[dst];
      B2_CUDA(cudaMemcpyAsync(recv_buffer(p, which), src, (size_t)bytes, cudaMemcpyDeviceToDevice, sc));
      B2_CUDA(cudaEventRecord(second ? p->arrived2 : p->arrived[ex], sc));
    };
    if (ex == 0 && s->up >= 0) put(s->sendA, s->sendA_bytes, s->up, 0, false);
    if (ex == 1 && s->down >= 0) put(s->sendB, s->sendB_bytes, s->down, 1, false);
    if (ex == 2) {
      if (s->down >= 0) put(s->sendL, s->sendL_bytes, s->down, 3, true);   // my left migrants arrive at the lower rank from above
      if (s->up >= 0) put(s->sendR, s->sendR_bytes, s->up, 2, false);      // my right migrants arrive at the upper rank from below
    }
    // ranks at the ends of the chain receive nothing in some exchanges: their wait must not dangle
    if (ex == 0 && s->down < 0) B2_CUDA(cudaEventRecord(s->arrived[0], sc));
    if (ex == 1 && s->up < 0) B2_CUDA(cudaEventRecord(s->arrived[1], sc));
    if (ex == 2 && s->down < 0) B2_CUDA(cudaEventRecord(s->arrived[2], sc));
    if (ex == 2 && s->up < 0) B2_CUDA(cudaEventRecord(s->arrived2, sc));
    if (ex == 2) B2_CUDA(cudaEventRecord(s->sent, sc));
    return;
  }
  NcclApi &n = nccl();
  B2_NCCL(n.GroupStart());
  if (ex == 0) {
    if (s->up >= 0) B2_NCCL(n.Send(s->sendA, (size_t)s->sendA_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
    if (s->down >= 0) B2_NCCL(n.Recv(s->recvA, (size_t)s->recvA_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
  } else if (ex == 1) {
    if (s->down >= 0) B2_NCCL(n.Send(s->sendB, (size_t)s->sendB_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
    if (s->up >= 0) B2_NCCL(n.Recv(s->recvB, (size_t)s->recvB_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
  } else {
    if (s->down >= 0) {
      B2_NCCL(n.Send(s->sendL, (size_t)s->sendL_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
      B2_NCCL(n.Recv(s->recvL, (size_t)s->recvL_bytes, NCCL_UINT8, s->down, s->comm->comm, sc));
    }
    if (s->up >= 0) {
      B2_NCCL(n.Send(s->sendR, (size_t)s->sendR_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
      B2_NCCL(n.Recv(s->recvR, (size_t)s->recvR_bytes, NCCL_UINT8, s->up, s->comm->comm, sc));
    }
  }
  B2_NCCL(n.GroupEnd());
  B2_CUDA(cudaEventRecord(s->arrived[ex], sc));
  if (ex == 2) B2_CUDA(cudaEventRecord(s->sent, sc));
}

int32_t b2_slab_substep(B2Slab s, int32_t nsteps) {
  B2_API_BEGIN
  require(s != nullptr, TI_ERROR_ARGUMENT_NULL, "slab");
  require(s->world == 1 || !s->comm->loopback, TI_ERROR_INVALID_STATE, "loopback slabs advance together: b2_slab_substep_lockstep");
  for (int it = 0; it < nsteps; it++)
    for (int ph = 0; ph < 4; ph++) {
      slab_phase(s, ph);
      if (ph < 3) slab_exchange(s, ph);
    }
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
    for (int ph = 0; ph < 4; ph++) {
      for (int r = 0; r < world; r++) slab_phase(slabs[r], ph);
      if (ph < 3)
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
