// b2_runtime.cu -- sparse-SNode runtime for B200 (sm_100a): device pool, lock-free NodeManager,
// activation, warp-scan element-list generation, fused GC, struct-for built-in bodies.
//
// Reference behaviour restated (not translated):
//   element_listgen_root / _nonroot   runtime_module/runtime.cpp:1282-1383
//   clear_list                        runtime.cpp:1270-1273   (folded into the root listgen)
//   parallel_struct_for               runtime.cpp:1422-1464   (k_struct_for<Body>)
//   gc_parallel_0/1/2                 runtime.cpp:1695-1785   (one kernel: k_gc)
//   reduce_* / warp_reduce_*          runtime.cpp:1204-1266   (k_reduce)
//   runtime_initialize*, snodes       runtime.cpp:911-1040    (host side, no 1-thread kernels)
#include <algorithm>
#include <cstdio>
#include <cstring>

#include "b2_host.h"

// ------------------------------------------------------------------------------------------
// errors / stream (thread-local last error: taichi_core_impl.cpp:63-70)
// ------------------------------------------------------------------------------------------
static thread_local TiError g_last_error = TI_ERROR_SUCCESS;
static thread_local std::string g_last_msg;
static cudaStream_t g_stream = nullptr;
static bool g_stream_user = false;
static uint64_t g_pool_size = 1ull << 30;  // device_memory_GB default = 1 (compile_config.cpp:63)

void b2_record_error(TiError code, const std::string &msg) {
  g_last_error = code;
  g_last_msg = msg;
}
cudaStream_t b2_current_stream() { return g_stream; }

extern "C" {
TiError ti_get_last_error(uint64_t *message_size, char *message) {
  if (message_size) {
    uint64_t cap = *message_size;
    uint64_t need = g_last_msg.size() + 1;
    if (message && cap > 0) {
      uint64_t n = std::min(cap - 1, (uint64_t)g_last_msg.size());
      memcpy(message, g_last_msg.data(), n);
      message[n] = 0;
    }
    *message_size = need;
  }
  return g_last_error;
}
void ti_set_last_error(TiError error, const char *message) {
  g_last_error = error;
  g_last_msg = (error != TI_ERROR_SUCCESS && message) ? message : "";
}
void ti_set_cuda_stream(void *stream) {
  g_stream = (cudaStream_t)stream;
  g_stream_user = true;
}
void ti_get_cuda_stream(void **stream) {
  if (stream) *stream = (void *)g_stream;
}
void b2_set_pool_size(uint64_t bytes) { g_pool_size = bytes; }
static int32_t g_random_seed = 0;  // ti.init(random_seed=), compile_config.cpp
void b2_set_random_seed(int32_t seed) { g_random_seed = seed; }
}

// ------------------------------------------------------------------------------------------
// runtime
// ------------------------------------------------------------------------------------------
static void b2_noalloc_setup(int device);  // address of the static scratch of b2_reduce_noalloc (below)
TiRuntime_t::TiRuntime_t(int device_index) : device(device_index) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    throw B2Error(TI_ERROR_NOT_SUPPORTED, "no CUDA device: libb200mpm has no CPU fallback");
  if (device_index < 0 || device_index >= count)
    throw B2Error(TI_ERROR_ARGUMENT_OUT_OF_RANGE, "device_index out of range");
  B2_CUDA(cudaSetDevice(device_index));
  cudaDeviceProp prop;
  B2_CUDA(cudaGetDeviceProperties(&prop, device_index));
  if (prop.major < 10)
    throw B2Error(TI_ERROR_NOT_SUPPORTED, "libb200mpm is built for sm_100a only");
  num_sms = prop.multiProcessorCount;
  {  // host<->device staging buffers come from the stream-ordered pool: keep them cached across
     // synchronisations (the default threshold of 0 hands the memory back at every sync)
    cudaMemPool_t mp;
    if (cudaDeviceGetDefaultMemPool(&mp, device_index) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  pool_size = g_pool_size;
  B2_CUDA(cudaMalloc(&pool_mem, pool_size));
  B2_CUDA(cudaMemsetAsync(pool_mem, 0, pool_size, stream()));
  B2_CUDA(cudaMalloc((void **)&pool_dev, sizeof(B2Pool)));
  B2Pool h;
  memset(&h, 0, sizeof(h));
  h.base = (u64)pool_mem;
  h.size = pool_size;
  B2_CUDA(cudaMemcpyAsync(pool_dev, &h, sizeof(h), cudaMemcpyHostToDevice, stream()));
  B2_CUDA(cudaStreamSynchronize(stream()));
  b2_noalloc_setup(device_index);
}
// initialize_rand_state (runtime.cpp:535-541): state i of the run = starting_rand_state + i with
// starting_rand_state = random_seed * 1048391 (llvm_runtime_executor.cpp:639)
__global__ void k_init_rand_states(unsigned *st, int n, unsigned start) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned k = start + (unsigned)i;
    st[4 * i] = 123456789u * k * 1000000007u, st[4 * i + 1] = 362436069u, st[4 * i + 2] = 521288629u, st[4 * i + 3] = 88675123u;
  }
}
unsigned *TiRuntime_t::ensure_rand_states() {
  if (rand_states) return rand_states;
  num_rand_states = num_sms * 2048;  // the widest launch of a run-time body: 8 CTAs x 256 or 16 CTAs x 128 threads per SM
  B2_CUDA(cudaMalloc((void **)&rand_states, (size_t)num_rand_states * 16));
  k_init_rand_states<<<num_sms * 4, 256, 0, stream()>>>(rand_states, num_rand_states, (unsigned)g_random_seed * 1048391u);
  count_launch();
  B2_CUDA(cudaGetLastError());
  return rand_states;
}
TiRuntime_t::~TiRuntime_t() {
  cudaSetDevice(device);
  cudaDeviceSynchronize();
  if (rand_states) cudaFree(rand_states);
  for (auto *m : memories) {
    if (m && !m->imported && m->dev) {
      if (m->host_access)
        cudaFreeHost(m->dev);
      else
        cudaFree(m->dev);
    }
    if (m && m->staging) cudaFreeHost(m->staging);
    delete m;
  }
  cudaFree(pool_mem);
  cudaFree(pool_dev);
}
void TiRuntime_t::sync() { B2_CUDA(cudaStreamSynchronize(stream())); }
int TiRuntime_t::grid_for(int64_t work_items, int block, int per_sm) const {
  int64_t need = (work_items + block - 1) / block;
  int64_t cap = (int64_t)num_sms * per_sm;
  return (int)std::max<int64_t>(1, std::min(need, cap));
}
int TiRuntime_t::check_error(std::string *msg) {
  B2Pool h;
  B2_CUDA(cudaMemcpyAsync(&h, pool_dev, sizeof(h), cudaMemcpyDeviceToHost, stream()));
  sync();
  if (h.error == 0) return 0;
  static const char *names[] = {"", "device memory pool exhausted (raise device_memory_GB)", "element list overflow",
                                "NodeManager chunk table overflow",
                                "particle bin overflow (more than 8192 particles in one 4^3 block, more occupied blocks than B2MpmParams.max_bins, or a full "
                                "slab message)",
                                "assertion failed"};
  if (msg) *msg = std::string(names[h.error < 6 ? h.error : 5]) + " [arg " + std::to_string(h.error_arg) + "]";
  int code = h.error;
  int zero[2] = {0, 0};
  B2_CUDA(cudaMemcpyAsync(&pool_dev->error, zero, 8, cudaMemcpyHostToDevice, stream()));
  return code;
}

bool b2_is_device_pointer(const void *p) {
  if (!p) return false;
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
void B2Staged::init(const void *any, size_t nbytes, bool in, bool out, cudaStream_t s) {
  st = s;
  bytes = nbytes;
  if (!any || nbytes == 0) return;
  if (b2_is_device_pointer(any)) {
    dev = (void *)any;
    return;
  }
  host = (void *)any;
  owned = true;
  writeback = out;
  B2_CUDA(cudaMallocAsync(&dev, nbytes, s));
  if (in)
    B2_CUDA(cudaMemcpyAsync(dev, any, nbytes, cudaMemcpyHostToDevice, s));
  else
    B2_CUDA(cudaMemsetAsync(dev, 0, nbytes, s));
}
void B2Staged::finish() {
  if (!owned) return;
  if (writeback) {
    B2_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
  }
  cudaFreeAsync(dev, st);
  owned = false;
}
B2Staged::~B2Staged() {
  if (owned) {
    cudaStreamSynchronize(st);
    cudaFreeAsync(dev, st);
  }
}

// ------------------------------------------------------------------------------------------
// tree geometry + layout (host)
// ------------------------------------------------------------------------------------------
static int64_t roundup(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

void B2Tree_t::layout(int id) {
  B2SNodeHost &n = s[id];
  int64_t off = 0;
  int align = 1;
  for (int c : n.ch) {
    layout(c);
    B2SNodeHost &ch = s[c];
    off = roundup(off, ch.node_align);
    ch.offset_in_parent_cell = off;
    off += ch.node_size;
    align = std::max(align, ch.node_align);
  }
  n.cell_size = roundup(off, align);
  n.cell_align = align;
  switch (n.type) {
    case B2T_PLACE:
      n.node_size = n.dtype_bytes;
      n.node_align = n.dtype_bytes;
      n.cell_size = n.dtype_bytes;
      break;
    case B2T_ROOT:
      n.node_size = n.cell_size;
      n.node_align = std::max(align, 16);
      break;
    case B2T_DENSE:
      n.node_size = n.cell_size * n.num_cells;
      n.node_align = align;
      break;
    case B2T_BITMASKED: {
      int64_t data = n.cell_size * n.num_cells;
      n.aux_offset = roundup(data, 4);
      int a = std::max(align, 4);
      n.node_size = roundup(n.aux_offset + 4 * ((n.num_cells + 31) / 32), a);
      n.node_align = a;
      break;
    }
    case B2T_POINTER:
      n.aux_offset = 8 * n.num_cells;
      n.node_size = roundup(12 * n.num_cells, 8);
      n.node_align = 8;
      n.alloc_size = roundup(std::max<int64_t>(n.cell_size, 1), 128);
      break;
    case B2T_DYNAMIC:
      n.node_size = 16;
      n.node_align = 8;
      n.alloc_size = roundup(16 + n.cell_size * n.chunk_size, 128);
      break;
  }
}

std::vector<int> B2Tree_t::path_ids(int target) const {
  std::vector<int> p;
  for (int i = target; i >= 0; i = s[i].parent) p.push_back(i);
  std::reverse(p.begin(), p.end());
  return p;
}
B2Path B2Tree_t::path_to(int target) const {
  B2Path P;
  memset(&P, 0, sizeof(P));
  auto ids = path_ids(target);
  if ((int)ids.size() > B2_MAX_LEVELS) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "SNode tree too deep");
  P.n = (int)ids.size();
  int total[B2_MAX_AXES];
  for (int a = 0; a < B2_MAX_AXES; a++) {
    total[a] = 1;
    for (int id : ids) total[a] *= s[id].shape[a];
  }
  P.pot = 1;
  auto ispot = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
  for (int l = 0; l < P.n; l++) {
    P.id[l] = ids[l];
    for (int a = 0; a < B2_MAX_AXES; a++) {
      total[a] /= s[ids[l]].shape[a];
      P.div[l][a] = total[a];
      if (!ispot(total[a]) || !ispot(s[ids[l]].shape[a])) P.pot = 0;
      int sh = 0;
      while ((1 << sh) < total[a]) sh++;
      P.shift[l][a] = sh;
    }
  }
  return P;
}
void B2Tree_t::check_snode(int id, const char *what) const {
  if (id < 0 || id >= (int)s.size()) throw B2Error(TI_ERROR_ARGUMENT_OUT_OF_RANGE, std::string(what) + ": bad snode id");
}

B2Tree_t::B2Tree_t(TiRuntime_t *rt_, const B2SNodeDesc *desc, int n) : rt(rt_) {
  if (!desc) throw B2Error(TI_ERROR_ARGUMENT_NULL, "desc");
  if (n <= 0 || n > B2_MAX_SNODES || desc[0].type != B2_ROOT)
    throw B2Error(TI_ERROR_INVALID_ARGUMENT, "tree must have 1..1024 snodes (taichi_max_num_snodes) and start with the root");
  s.resize(n);
  for (int i = 0; i < n; i++) {
    B2SNodeHost &x = s[i];
    x.type = desc[i].type;
    x.parent = desc[i].parent;
    x.chunk_size = desc[i].chunk_size;
    x.dtype_bytes = desc[i].dtype_bytes;
    if (x.type < B2T_ROOT || x.type > B2T_PLACE || (i > 0 && x.type == B2T_ROOT))
      throw B2Error(TI_ERROR_INVALID_ARGUMENT, "bad snode type");
    if (x.type == B2T_PLACE && x.dtype_bytes != 4 && x.dtype_bytes != 8)
      throw B2Error(TI_ERROR_INVALID_ARGUMENT, "place: dtype_bytes must be 4 or 8");
    if (x.type == B2T_DYNAMIC && (x.chunk_size <= 0 || desc[i].num_axes != 1))
      throw B2Error(TI_ERROR_INVALID_ARGUMENT, "dynamic: one axis and chunk_size > 0 required");
    for (int a = 0; a < B2_MAX_AXES; a++) {
      x.shape[a] = 1;
      x.active[a] = 0;
      x.from_root[a] = i == 0 ? 1 : 0;
    }
    if (i > 0) {
      if (x.parent < 0 || x.parent >= i) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "parent must precede child");
      if (s[x.parent].type == B2T_PLACE) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "a place has no children");
      s[x.parent].ch.push_back(i);
      x.depth = s[x.parent].depth + 1;
      for (int a = 0; a < B2_MAX_AXES; a++) x.from_root[a] = s[x.parent].from_root[a];
    }
    if (desc[i].num_axes < 0 || desc[i].num_axes > B2_MAX_AXES) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "num_axes");
    for (int k = 0; k < desc[i].num_axes; k++) {
      int a = desc[i].axes[k];
      if (a < 0 || a >= B2_MAX_AXES || desc[i].sizes[k] <= 0) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "bad axis/size");
      x.active[a] = 1;
      x.shape[a] = desc[i].sizes[k];
      x.from_root[a] *= desc[i].sizes[k];
    }
    int64_t acc = 1;
    for (int a = B2_MAX_AXES - 1; a >= 0; a--) {
      x.acc[a] = (int)acc;
      acc *= x.shape[a];
    }
    if (acc > (1ll << 30)) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "container too large (indices are i32, runtime.cpp:424)");
    x.num_cells = acc;
  }
  layout(0);

  // device descriptor
  memset(&host_copy, 0, sizeof(host_copy));
  host_copy.pool = rt->pool_dev;
  host_copy.n = n;
  cudaStream_t st = rt->stream();
  int64_t root_bytes = roundup(std::max<int64_t>(s[0].node_size, 16), 4096);
  B2_CUDA(cudaMalloc(&root_mem, root_bytes));
  B2_CUDA(cudaMemsetAsync(root_mem, 0, root_bytes, st));
  host_copy.root = (u8 *)root_mem;
  B2_CUDA(cudaMalloc(&lists_mem, sizeof(B2CList) * n));
  B2_CUDA(cudaMemsetAsync(lists_mem, 0, sizeof(B2CList) * n, st));
  int nm = 0;
  for (int i = 0; i < n; i++)
    if (s[i].type == B2T_POINTER || s[i].type == B2T_DYNAMIC) nm++;
  if (nm) {
    B2_CUDA(cudaMalloc(&mgr_mem, sizeof(B2NodeMgr) * nm));
    B2_CUDA(cudaMemsetAsync(mgr_mem, 0, sizeof(B2NodeMgr) * nm, st));
  }
  B2_CUDA(cudaMalloc(&scratch, 4096));
  B2_CUDA(cudaMemsetAsync(scratch, 0, 4096, st));
  std::vector<B2CList> hl(n);
  std::vector<B2NodeMgr> *hm = new std::vector<B2NodeMgr>(nm);
  int im = 0;
  for (int i = 0; i < n; i++) {
    B2SNodeDev &d = host_copy.s[i];
    const B2SNodeHost &x = s[i];
    d.type = x.type;
    d.parent = x.parent;
    d.depth = x.depth;
    d.nch = (int)x.ch.size();
    d.nax = 0;
    for (int a = 0; a < B2_MAX_AXES; a++) {
      d.shape[a] = x.shape[a];
      d.acc[a] = x.acc[a];
      if (x.shape[a] > 1) d.ax[d.nax++] = a;
    }
    d.pot = 1;
    for (int a = 0; a < B2_MAX_AXES; a++) {
      auto is_pot = [](long long v) { return v > 0 && (v & (v - 1)) == 0; };
      if (!is_pot(x.shape[a]) || !is_pot(x.acc[a])) d.pot = 0;
      int sh = 0;
      while ((1ll << (sh + 1)) <= (long long)x.acc[a]) sh++;
      d.acc_shift[a] = sh;
    }
    d.num_cells = x.num_cells;
    d.cell_size = x.cell_size;
    d.node_size = x.node_size;
    d.offset_in_parent_cell = x.offset_in_parent_cell;
    d.aux_offset = x.aux_offset;
    d.alloc_size = x.alloc_size;
    d.chunk_size = x.chunk_size;
    d.dtype_bytes = x.dtype_bytes;
    memset(&hl[i], 0, sizeof(B2CList));
    hl[i].elem_size = sizeof(B2Element);
    hl[i].log2_epc = B2_ELEM_LOG2_EPC;
    host_copy.lists[i] = (B2CList *)lists_mem + i;
    if (x.type == B2T_POINTER || x.type == B2T_DYNAMIC) {
      B2NodeMgr &m = (*hm)[im];
      memset(&m, 0, sizeof(m));
      m.elem_size = (u64)x.alloc_size;
      int lg = 0;  // elements per data chunk: chunks of <= 64 MB, and <= 1/16 of the pool so that a small pool still holds several
      const u64 chunk_cap = std::min<u64>(64ull << 20, rt->pool_size / 16);
      while (lg < 14 && ((u64)x.alloc_size << (lg + 1)) <= chunk_cap) lg++;
      m.log2_epc = lg;
      for (int k = 0; k < 2; k++) {
        m.free_list[k].elem_size = 4;
        m.free_list[k].log2_epc = B2_IDX_LOG2_EPC;
      }
      m.recycled.elem_size = 4;
      m.recycled.log2_epc = B2_IDX_LOG2_EPC;
      void *amb = nullptr;  // runtime_allocate_ambient (runtime.cpp:1033-1040)
      B2_CUDA(cudaMalloc(&amb, x.alloc_size));
      B2_CUDA(cudaMemsetAsync(amb, 0, x.alloc_size, st));
      ambient.push_back(amb);
      m.ambient = (u64)amb;
      host_copy.mgr[i] = (B2NodeMgr *)mgr_mem + im;
      im++;
    }
  }
  // only the small headers need a copy; chunk tables are already zero
  for (int i = 0; i < n; i++)
    B2_CUDA(cudaMemcpyAsync((B2CList *)lists_mem + i, &hl[i], 16, cudaMemcpyHostToDevice, st));
  for (int k = 0; k < nm; k++) {
    B2NodeMgr &m = (*hm)[k];
    B2NodeMgr *dm = (B2NodeMgr *)mgr_mem + k;
    B2_CUDA(cudaMemcpyAsync(dm, &m, offsetof(B2NodeMgr, data_chunks), cudaMemcpyHostToDevice, st));
    for (int q = 0; q < 2; q++) B2_CUDA(cudaMemcpyAsync(&dm->free_list[q], &m.free_list[q], 16, cudaMemcpyHostToDevice, st));
    B2_CUDA(cudaMemcpyAsync(&dm->recycled, &m.recycled, 16, cudaMemcpyHostToDevice, st));
  }
  // runtime_initialize_snodes seeds the root list with one Element (runtime.cpp:1006-1016):
  // here the root list stays virtual -- k_listgen_root synthesises it.
  B2_CUDA(cudaMalloc((void **)&dev, sizeof(B2TreeDev)));
  B2_CUDA(cudaMemcpyAsync(dev, &host_copy, sizeof(B2TreeDev), cudaMemcpyHostToDevice, st));
  B2_CUDA(cudaStreamSynchronize(st));
  delete hm;
}
B2Tree_t::~B2Tree_t() {
  cudaStreamSynchronize(rt->stream());
  cudaFree(root_mem);
  cudaFree(lists_mem);
  cudaFree(mgr_mem);
  cudaFree(scratch);
  cudaFree(dev);
  for (void *a : ambient) cudaFree(a);
}

// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
// Exclusive scan of one int per thread over the CTA; *total gets the CTA sum.
// smem: 33 ints. All threads of the CTA must call it.
__device__ __forceinline__ int b2_block_scan(int v, int *total, int *smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  int inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += t;
  }
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < nw ? smem[lane] : 0;
    int winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += t;
    }
    if (lane < nw) smem[lane] = winc - w;
    if (lane == 31) smem[32] = winc;
  }
  __syncthreads();
  int res = smem[warp] + inc - v;
  *total = smem[32];
  __syncthreads();
  return res;
}

__device__ __forceinline__ void b2_store_element(B2Element *dst, u8 *node, int lo, int hi, const int *pc) {
  uint4 *d = (uint4 *)dst;
  u64 p = (u64)node;
  d[0] = make_uint4((u32)p, (u32)(p >> 32), (u32)lo, (u32)hi);
  d[1] = make_uint4(pc[0], pc[1], pc[2], pc[3]);
  d[2] = make_uint4(pc[4], pc[5], pc[6], pc[7]);
  d[3] = make_uint4(pc[8], pc[9], pc[10], pc[11]);
}

// ------------------------------------------------------------------------------------------
// listgen
// ------------------------------------------------------------------------------------------
// Root level: the single child container of root is split into <=1024-cell Elements
// (element_listgen_root).  Also zeroes the counters of the deeper lists on the path
// (clear_list tasks of offload.cpp:204-213 folded in).
__global__ void k_listgen_root(B2TreeDev *T, B2Path P) {
  const int cid = P.id[1];
  const B2SNodeDev &cs = T->s[cid];
  B2CList *CL = T->lists[cid];
  u8 *chn = T->root + cs.offset_in_parent_cell;
  const int n = b2_num_elements(cs, chn);
  const int cnt = (n + B2_LISTGEN_MAX_ELEMENT - 1) / B2_LISTGEN_MAX_ELEMENT;
  const int step = n < B2_LISTGEN_MAX_ELEMENT ? n : B2_LISTGEN_MAX_ELEMENT;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gtid == 0) {
    CL->count = cnt;
    for (int l = 2; l < P.n; l++) T->lists[P.id[l]]->count = 0;
  }
  int zero[B2_MAX_AXES];
#pragma unroll
  for (int a = 0; a < B2_MAX_AXES; a++) zero[a] = 0;
  for (int k = gtid; k < cnt; k += gridDim.x * blockDim.x) {
    B2Element *dst = (B2Element *)b2_list_slot(T->pool, CL, (u32)k);
    int lo = k * step;
    int hi = lo + step < n ? lo + step : n;
    if (dst) b2_store_element(dst, chn, lo, hi, zero);
  }
}

// One level below: for every parent Element (CTA-stride) and every cell in it (thread-stride):
// active? -> child container, its element count, ceil(n/1024) child Elements.  Output slots
// come from a CTA-wide prefix sum and ONE atomicAdd per CTA pass (the reference does one
// atomic per Element on a single address, runtime.cpp:451-456).
__global__ void __launch_bounds__(256) k_listgen_nonroot(B2TreeDev *T, int pid, int cid) {
  __shared__ int scan_smem[33];
  __shared__ u32 base_smem;
  const B2SNodeDev &ps = T->s[pid];
  const B2SNodeDev &cs = T->s[cid];
  B2CList *PL = T->lists[pid];
  B2CList *CL = T->lists[cid];
  const u32 np = PL->count;
  // work unit = blockDim consecutive cells of one parent Element (an Element spans <= 1024 cells), so a
  // short parent list (C4: 32 Elements of 1024 pointer cells) still spreads over 128 CTAs
  const u32 upe = (B2_LISTGEN_MAX_ELEMENT + blockDim.x - 1) / blockDim.x;
  for (u32 u = blockIdx.x; u < np * upe; u += gridDim.x) {
    const B2Element E = *(const B2Element *)b2_list_get(PL, u / upe);
    {
      const int j0 = E.lo + (int)((u % upe) * blockDim.x);
      if (j0 >= E.hi) continue;  // uniform
      const int j = j0 + threadIdx.x;
      int cnt = 0, n = 0;
      u8 *chn = nullptr;
      if (j < E.hi && b2_is_active(ps, E.node, j)) {
        chn = b2_lookup(T, pid, E.node, j) + cs.offset_in_parent_cell;
        n = b2_num_elements(cs, chn);
        cnt = (n + B2_LISTGEN_MAX_ELEMENT - 1) / B2_LISTGEN_MAX_ELEMENT;
      }
      int total;
      int off = b2_block_scan(cnt, &total, scan_smem);
      if (total == 0) continue;  // uniform
      if (threadIdx.x == 0) base_smem = atomicAdd(&CL->count, (u32)total);
      __syncthreads();
      const u32 base = base_smem;
      if (cnt) {
        int rc[B2_MAX_AXES];
        b2_refine(ps, E.pcoord, rc, j);
        const int step = n < B2_LISTGEN_MAX_ELEMENT ? n : B2_LISTGEN_MAX_ELEMENT;
        for (int k = 0; k < cnt; k++) {
          B2Element *dst = (B2Element *)b2_list_slot(T->pool, CL, base + off + k);
          int lo = k * step;
          int hi = lo + step < n ? lo + step : n;
          if (dst) b2_store_element(dst, chn, lo, hi, rc);
        }
      }
      __syncthreads();  // base_smem reuse
    }
  }
}

// Bitmasked parent whose child containers are single Elements (static child, <= 1024 cells): one
// WARP per parent Element.  Lane l owns mask word l: popc -> warp prefix sum -> one atomicAdd per
// warp -> every lane writes the Elements of its set bits (the "warp-ballot prefix-sum listgen";
// the reference takes one atomic per Element, runtime.cpp:451-456,1367-1379).
__global__ void __launch_bounds__(256) k_listgen_bitmasked(B2TreeDev *T, int pid, int cid) {
  const B2SNodeDev &ps = T->s[pid];
  const B2SNodeDev &cs = T->s[cid];
  B2CList *PL = T->lists[pid];
  B2CList *CL = T->lists[cid];
  const u32 np = PL->count;
  const int lane = threadIdx.x & 31;
  const u32 wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  const int ncc = (int)cs.num_cells;
  // work unit = (parent Element, run of S mask words); S shrinks when there are few parents so that
  // the whole chip stays busy (np = 200 containers -> one word per warp, np >= #warps -> 32 words)
  int S = 32;
  while (S > 1 && (unsigned long long)np * (32 / S) < 2ull * nw) S >>= 1;
  const u32 upe = 32 / S;  // units per Element (an Element spans <= 1024 cells = 32 words)
  for (u32 u = wid; u < np * upe; u += nw) {
    const u32 e = u / upe;
    const B2Element E = *(const B2Element *)b2_list_get(PL, e);
    const u32 *mask = b2_mask_words(ps, E.node);
    {
      const int w0 = (E.lo >> 5) + (int)(u % upe) * S;
      const int w = w0 + lane;
      u32 word = 0;
      if (lane < S && w * 32 < E.hi) {
        word = mask[w];
        const int first = w * 32;
        if (first < E.lo) word &= ~0u << (E.lo - first);
        if (first + 32 > E.hi) word &= ~0u >> (first + 32 - E.hi);
      }
      const int cnt = __popc(word);
      int inc = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int tv = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += tv;
      }
      const int total = __shfl_sync(0xffffffffu, inc, 31);
      if (total == 0) continue;
      u32 base = 0;
      if (lane == 31) base = atomicAdd(&CL->count, (u32)total);
      base = __shfl_sync(0xffffffffu, base, 31);
      // cooperative expansion: lane handles output ranks lane, lane+32, ... so that the 64-byte
      // Element stores of a warp are contiguous; rank k -> (word holding it, bit inside the word)
      for (int k0 = 0; k0 < total; k0 += 32) {
        const int k = k0 + lane;
        const int kk = k < total ? k : total - 1;
        int l = 0;  // smallest lane with inc > kk (binary search over the inclusive prefix)
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
          const int probe = __shfl_sync(0xffffffffu, inc, l + step - 1);
          if (probe <= kk) l += step;
        }
        const u32 wsel = __shfl_sync(0xffffffffu, word, l);
        const int excl = __shfl_sync(0xffffffffu, inc - cnt, l);
        if (k < total) {
          const int bit = (int)__fns(wsel, 0, kk - excl + 1);
          const int j = (w0 + l) * 32 + bit;
          int rc[B2_MAX_AXES];
          b2_refine(ps, E.pcoord, rc, j);
          B2Element *dst = (B2Element *)b2_list_slot(T->pool, CL, base + (u32)k);
          if (dst) b2_store_element(dst, E.node + ps.cell_size * j + cs.offset_in_parent_cell, 0, ncc, rc);
        }
      }
    }
  }
}

void B2Tree_t::listgen_async(int target) {
  check_snode(target, "listgen");
  if (s[target].type == B2T_PLACE) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "listgen: target is a place");
  B2Path P = path_to(target);
  if (P.n < 2) return;  // struct-for over root itself: nothing to generate
  cudaStream_t st = rt->stream();
  {
    int64_t cnt = (s[P.id[1]].num_cells + B2_LISTGEN_MAX_ELEMENT - 1) / B2_LISTGEN_MAX_ELEMENT;
    k_listgen_root<<<rt->grid_for(cnt, 128), 128, 0, st>>>(dev, P);
    rt->count_launch();
  }
  for (int l = 2; l < P.n; l++) {
    // parent elements in flight: one CTA each, capped at a few waves of the chip
    const B2SNodeHost &par = s[P.id[l - 1]], &ch = s[P.id[l]];
    if (par.type == B2T_BITMASKED && ch.type != B2T_DYNAMIC && ch.num_cells <= B2_LISTGEN_MAX_ELEMENT) {
      k_listgen_bitmasked<<<rt->num_sms * 8, 256, 0, st>>>(dev, P.id[l - 1], P.id[l]);
    } else {
      int block = (int)std::min<int64_t>(256, std::max<int64_t>(32, roundup(std::min<int64_t>(par.num_cells, 1024), 32)));
      k_listgen_nonroot<<<rt->num_sms * 16, block, 0, st>>>(dev, P.id[l - 1], P.id[l]);
    }
    rt->count_launch();
  }
  B2_CUDA(cudaGetLastError());
}
int64_t B2Tree_t::list_count(int target) {
  if (target == 0) return 1;
  u32 c = 0;
  B2_CUDA(cudaMemcpyAsync(&c, &host_copy.lists[target]->count, 4, cudaMemcpyDeviceToHost, rt->stream()));
  rt->sync();
  return c;
}

// ------------------------------------------------------------------------------------------
// struct-for driver with built-in bodies (parallel_struct_for: one CTA per Element,
// threads stride the cells, inactive leaf cells skipped: runtime.cpp:1422-1464,
// codegen_llvm.cpp:2222-2262)
// ------------------------------------------------------------------------------------------
template <typename Body>
__global__ void __launch_bounds__(256) k_struct_for(B2TreeDev *T, int sid, Body body) {
  const B2SNodeDev &ts = T->s[sid];
  B2CList *L = T->lists[sid];
  const u32 ne = L->count;
  for (u32 e = blockIdx.x; e < ne; e += gridDim.x) {
    const B2Element E = *(const B2Element *)b2_list_get(L, e);
    for (int j0 = E.lo; j0 < E.hi; j0 += blockDim.x) {  // uniform trip count per CTA
      const int j = j0 + threadIdx.x;
      const bool act = j < E.hi && b2_is_active(ts, E.node, j);
      body(T, ts, E, j, act);
    }
  }
  body.finish(T);
}

struct BodyCollect {  // visited coordinates -> out[k][12]
  int *out;
  long long cap;
  unsigned long long *counter;
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &ts, const B2Element &E, int j, bool act) {
    const unsigned m = __ballot_sync(0xffffffffu, act);
    if (!m) return;
    const int lane = threadIdx.x & 31;
    unsigned long long base = 0;
    if (lane == (__ffs(m) - 1)) base = atomicAdd(counter, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (act) {
      unsigned long long k = base + __popc(m & ((1u << lane) - 1));
      if (out && (long long)k < cap) {
        int rc[B2_MAX_AXES];
        b2_refine(ts, E.pcoord, rc, j);
        uint4 *d = (uint4 *)(out + k * B2_MAX_AXES);
        d[0] = make_uint4(rc[0], rc[1], rc[2], rc[3]);
        d[1] = make_uint4(rc[4], rc[5], rc[6], rc[7]);
        d[2] = make_uint4(rc[8], rc[9], rc[10], rc[11]);
      }
    }
  }
  __device__ void finish(B2TreeDev *) {}
};

struct BodyDeactivate {  // snode_deactivate (_kernels.py:264-267)
  int sid;
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &, const B2Element &E, int j, bool act) {
    if (act) b2_deactivate_cell(T, sid, E.node, j);
  }
  __device__ void finish(B2TreeDev *) {}
};

struct BodyDeactivateDynamic {  // snode_deactivate_dynamic: loop over the parent's cells
  int psid, dsid;
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &, const B2Element &E, int j, bool act) {
    if (act) {
      u8 *d = b2_lookup(T, psid, E.node, j) + T->s[dsid].offset_in_parent_cell;
      b2_deactivate_cell(T, dsid, d, 0);
    }
  }
  __device__ void finish(B2TreeDev *) {}
};

struct BodyFill {  // field.fill over active cells
  long long place_off;
  int bytes;
  u64 bits;  // the value's bit pattern, zero-extended to 64 bits
  int psid;
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &, const B2Element &E, int j, bool act) {
    if (act) {
      u8 *c = b2_lookup(T, psid, E.node, j) + place_off;
      if (bytes == 8)
        *(u64 *)c = bits;
      else if (bytes == 4)
        *(u32 *)c = (u32)bits;
      else if (bytes == 2)
        *(unsigned short *)c = (unsigned short)bits;
      else
        *c = (u8)bits;
    }
  }
  __device__ void finish(B2TreeDev *) {}
};

struct BodyToDense {  // ext_arr[I] = field[I] for active I
  long long place_off;
  int bytes, psid;
  u8 *out;
  int dims[B2_MAX_AXES];  // extents (1 for unused axes)
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &ts, const B2Element &E, int j, bool act) {
    if (act) {
      int rc[B2_MAX_AXES];
      b2_refine(ts, E.pcoord, rc, j);
      long long lin = 0;
#pragma unroll
      for (int a = 0; a < B2_MAX_AXES; a++) lin = lin * dims[a] + rc[a];
      u8 *c = b2_lookup(T, psid, E.node, j) + place_off;
      if (bytes == 4)
        ((u32 *)out)[lin] = *(u32 *)c;
      else
        ((u64 *)out)[lin] = *(u64 *)c;
    }
  }
  __device__ void finish(B2TreeDev *) {}
};

// Reduction with a thread-local accumulator, then warp shuffle tree, CTA shared memory, and one
// atomic per CTA (reference: one atomic per warp, runtime.cpp:1235-1266).
template <typename V>
__device__ __forceinline__ V b2_red_op(V a, V b, int op) {
  return op == 0 ? a + b : (op == 1 ? (a < b ? a : b) : (a > b ? a : b));
}
__device__ __forceinline__ void b2_atomic_red(float *dst, float v, int op) {
  if (op == 0) {
    atomicAdd(dst, v);
    return;
  }
  int *ip = (int *)dst;  // CAS loop for min/max (atomic.h:5-88)
  int old = *ip, assumed;
  do {
    assumed = old;
    float cur = __int_as_float(assumed);
    float neu = b2_red_op(cur, v, op);
    if (neu == cur) break;
    old = atomicCAS(ip, assumed, __float_as_int(neu));
  } while (old != assumed);
}
__device__ __forceinline__ void b2_atomic_red(int *dst, int v, int op) {
  if (op == 0)
    atomicAdd(dst, v);
  else if (op == 1)
    atomicMin(dst, v);
  else
    atomicMax(dst, v);
}
template <typename V>
__device__ __forceinline__ void b2_block_reduce_to(V acc, V ident, int op, V *dst) {
  __shared__ V part[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc = b2_red_op(acc, __shfl_down_sync(0xffffffffu, acc, d), op);
  if (lane == 0) part[warp] = acc;
  __syncthreads();
  if (warp == 0) {
    V v = lane < nw ? part[lane] : ident;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = b2_red_op(v, __shfl_down_sync(0xffffffffu, v, d), op);
    if (lane == 0) b2_atomic_red(dst, v, op);
  }
}
template <typename V>
struct BodyReduce {
  long long place_off;
  int psid, op;
  V ident;
  V *dst;
  V acc;
  __device__ void operator()(B2TreeDev *T, const B2SNodeDev &, const B2Element &E, int j, bool act) {
    if (act) acc = b2_red_op(acc, *(V *)(b2_lookup(T, psid, E.node, j) + place_off), op);
  }
  __device__ void finish(B2TreeDev *) { b2_block_reduce_to<V>(acc, ident, op, dst); }
};
// 128-bit loads, four of them in flight per thread (8M f32 = 32 MB: the kernel is a pure HBM stream); dst is initialised by
// CTA-independent means (k_set_u32) so the launch needs no host round trip
template <typename V>
__global__ void __launch_bounds__(256) k_reduce(const V *data, long long n, int op, V ident, V *dst) {
  V acc = ident;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  if ((((unsigned long long)data) & 15ull) == 0) {
    const long long n4 = n >> 2;
    const uint4 *d4 = (const uint4 *)data;
    long long i = tid;
    for (; i + 3 * nth < n4; i += 4 * nth) {
      const uint4 a = d4[i], b = d4[i + nth], c = d4[i + 2 * nth], d = d4[i + 3 * nth];
      const uint4 q[4] = {a, b, c, d};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        V v[4];
        memcpy(v, &q[k], 16);
        acc = b2_red_op(b2_red_op(acc, v[0], op), b2_red_op(v[1], b2_red_op(v[2], v[3], op), op), op);
      }
    }
    for (; i < n4; i += nth) {
      V v[4];
      const uint4 a = d4[i];
      memcpy(v, &a, 16);
      acc = b2_red_op(b2_red_op(acc, v[0], op), b2_red_op(v[1], b2_red_op(v[2], v[3], op), op), op);
    }
    for (long long j = (n4 << 2) + tid; j < n; j += nth) acc = b2_red_op(acc, data[j], op);
  } else {
    for (long long i = tid; i < n; i += nth) acc = b2_red_op(acc, data[i], op);
  }
  b2_block_reduce_to<V>(acc, ident, op, dst);
}
__global__ void k_set_u32(u32 *dst, u32 v) { *dst = v; }

// ---- allocation-free reduction (b2_reduce_noalloc) -------------------------------------------------------------
// The per-CTA partials and the arrival ticket live in a file-scope device array instead of a buffer from cudaMalloc: every
// GPU that loads the library holds B2_NOALLOC_SCRATCH_BYTES of it until the process exits.  Word 0 is the ticket, the
// partials start at word 64 (their own 256-byte line).  All calls share the array: two calls that overlap on the GPU race.
#define B2_NOALLOC_MAX_CTAS 2048
#define B2_NOALLOC_MAX_N ((long long)B2_NOALLOC_MAX_CTAS * 256 * 512) /* 2^28 elements = 1 GiB of f32 / i32 */
#define B2_NOALLOC_SCRATCH_WORDS (64 + B2_NOALLOC_MAX_CTAS)
static __device__ __align__(256) u32 b2_noalloc_scratch[B2_NOALLOC_SCRATCH_WORDS];
static u32 *g_noalloc_scratch[64];  // the array's DEVICE address per device (the name above is only a host-side stand-in)

// set-up, once per device, with that device current (runtime construction)
static void b2_noalloc_setup(int device) {
  if (device < 0 || device >= 64 || g_noalloc_scratch[device]) return;
  u32 *ptr = nullptr;
  B2_CUDA(cudaGetSymbolAddress((void **)&ptr, b2_noalloc_scratch));
  if (!ptr) throw B2Error(TI_ERROR_INVALID_STATE, "cudaGetSymbolAddress(b2_noalloc_scratch) returned null");
  g_noalloc_scratch[device] = ptr;
}
template <typename V>
__device__ __forceinline__ V b2_block_reduce_value(V acc, V ident, int op) {  // valid in thread 0
  __shared__ V part[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc = b2_red_op(acc, __shfl_down_sync(0xffffffffu, acc, d), op);
  __syncthreads();
  if (lane == 0) part[warp] = acc;
  __syncthreads();
  V v = (warp == 0 && lane < nw) ? part[lane] : ident;
  if (warp == 0) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = b2_red_op(v, __shfl_down_sync(0xffffffffu, v, d), op);
  }
  return v;
}
// two stages in one launch: every CTA leaves its partial in the scratch array, the CTA that draws the last ticket folds
// the partials in index order (the result does not depend on which CTA that is) and hands the ticket back as zero
template <typename V>
__global__ void __launch_bounds__(256) k_reduce_noalloc(const V *data, long long n, int op, V ident, u32 *scratch, V *dst) {
  __shared__ int is_last;
  V *partial = (V *)(scratch + 64);
  V acc = ident;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  if ((((unsigned long long)data) & 15ull) == 0) {
    const long long n4 = n >> 2;
    const uint4 *d4 = (const uint4 *)data;
    for (long long i = tid; i < n4; i += nth) {
      V v[4];
      const uint4 a = d4[i];
      memcpy(v, &a, 16);
      acc = b2_red_op(b2_red_op(acc, v[0], op), b2_red_op(v[1], b2_red_op(v[2], v[3], op), op), op);
    }
    for (long long j = (n4 << 2) + tid; j < n; j += nth) acc = b2_red_op(acc, data[j], op);
  } else {
    for (long long i = tid; i < n; i += nth) acc = b2_red_op(acc, data[i], op);
  }
  const V mine = b2_block_reduce_value<V>(acc, ident, op);
  if (threadIdx.x == 0) {
    __stcg(partial + blockIdx.x, mine);
    __threadfence();
    is_last = atomicAdd(scratch, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  V fold = ident;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) fold = b2_red_op(fold, __ldcg(partial + i), op);
  fold = b2_block_reduce_value<V>(fold, ident, op);
  if (threadIdx.x == 0) {
    *dst = fold;
    *scratch = 0u;
  }
}

// ------------------------------------------------------------------------------------------
// coordinate-list kernels (activate / deactivate / is_active / read / write / dynamic)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void b2_load_coord(const int *coords, long long k, int *I) {
  const uint4 *c = (const uint4 *)(coords + k * B2_MAX_AXES);
  uint4 a = c[0], b = c[1], d = c[2];
  I[0] = a.x, I[1] = a.y, I[2] = a.z, I[3] = a.w;
  I[4] = b.x, I[5] = b.y, I[6] = b.z, I[7] = b.w;
  I[8] = d.x, I[9] = d.y, I[10] = d.z, I[11] = d.w;
}
// mode 0 activate, 1 deactivate, 2 is_active
__global__ void k_coord_op(B2TreeDev *T, B2Path P, const int *coords, long long n, int mode, u8 *out) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    int I[B2_MAX_AXES], cell;
    b2_load_coord(coords, k, I);
    const int sid = P.id[P.n - 1];
    if (mode == 0) {
      bool dead;
      u8 *node = b2_walk<true>(T, P, I, &cell, &dead);
      if (!dead) b2_activate_cell(T, sid, node, cell);
    } else {
      bool dead;
      u8 *node = b2_walk<false>(T, P, I, &cell, &dead);
      if (mode == 1) {
        if (!dead) b2_deactivate_cell(T, sid, node, cell);
      } else {
        out[k] = (!dead && b2_is_active(T->s[sid], node, cell)) ? 1 : 0;
      }
    }
  }
}
// P = path to the place's parent. mode 0 write, 1 atomic add f32, 2 read
__global__ void k_access(B2TreeDev *T, B2Path P, long long place_off, int bytes, const int *coords, long long n, int mode, u8 *vals) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    int I[B2_MAX_AXES], cell;
    b2_load_coord(coords, k, I);
    const int sid = P.id[P.n - 1];
    if (mode < 2) {
      bool dead;
      u8 *node = b2_walk<true>(T, P, I, &cell, &dead);
      if (dead) continue;  // out of memory (flag raised): the ambient node is never written
      b2_activate_cell(T, sid, node, cell);
      u8 *c = b2_lookup(T, sid, node, cell) + place_off;
      if (mode == 1)
        atomicAdd((float *)c, ((const float *)vals)[k]);
      else if (bytes == 4)
        *(u32 *)c = ((const u32 *)vals)[k];
      else
        *(u64 *)c = ((const u64 *)vals)[k];
    } else {
      u8 *node = b2_walk<false>(T, P, I, &cell, nullptr);
      u8 *c = b2_lookup(T, sid, node, cell) + place_off;
      if (bytes == 4)
        ((u32 *)vals)[k] = *(u32 *)c;
      else
        ((u64 *)vals)[k] = *(u64 *)c;
    }
  }
}
// from_numpy: every index of the dense image is written (activating), ext_arr_to_tensor
// (_kernels.py:52-60)
__global__ void k_from_dense(B2TreeDev *T, B2Path P, long long place_off, int bytes, const u8 *in, long long total, B2Path dimsP) {
  // dimsP.div[0][a] holds the extents
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (long long)gridDim.x * blockDim.x) {
    int I[B2_MAX_AXES], cell;
    long long r = k;
#pragma unroll
    for (int a = B2_MAX_AXES - 1; a >= 0; a--) {
      int d = dimsP.div[0][a];
      I[a] = (int)(r % d);
      r /= d;
    }
    const int sid = P.id[P.n - 1];
    bool dead;
    u8 *node = b2_walk<true>(T, P, I, &cell, &dead);
    if (dead) continue;
    b2_activate_cell(T, sid, node, cell);
    u8 *c = b2_lookup(T, sid, node, cell) + place_off;
    if (bytes == 4)
      *(u32 *)c = ((const u32 *)in)[k];
    else
      *(u64 *)c = ((const u64 *)in)[k];
  }
}
// ti.append (Dynamic_allocate, node_dynamic.h:61-88): slot = atomicAdd(n), chunk chain made on demand
__global__ void k_dynamic_append(B2TreeDev *T, B2Path P, const int *coords, long long n, const int *vals) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    int I[B2_MAX_AXES], cell;
    b2_load_coord(coords, k, I);
    const int sid = P.id[P.n - 1];
    const B2SNodeDev &s = T->s[sid];
    bool dead;
    B2DynNode *d = (B2DynNode *)b2_walk<true>(T, P, I, &cell, &dead);
    if (dead) continue;
    u32 i = atomicAdd(&d->n, 1u);
    B2DynChunk *c = b2_dynamic_chunk(T, sid, d, (int)(i / (u32)s.chunk_size));
    if (c) *(int *)((u8 *)c + 16 + (long long)(i % (u32)s.chunk_size) * s.cell_size) = vals[k];
  }
}
__global__ void k_dynamic_length(B2TreeDev *T, B2Path P, const int *coords, long long n, int *out) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    int I[B2_MAX_AXES], cell;
    b2_load_coord(coords, k, I);
    B2DynNode *d = (B2DynNode *)b2_walk<false>(T, P, I, &cell, nullptr);
    out[k] = (int)d->n;
  }
}
__global__ void k_export_list(B2TreeDev *T, int sid, int *out, long long cap) {
  B2CList *L = T->lists[sid];
  const u32 ne = L->count;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < ne && k < cap; k += (long long)gridDim.x * blockDim.x) {
    const B2Element E = *(const B2Element *)b2_list_get(L, (u32)k);
    int *o = out + k * 14;
#pragma unroll
    for (int a = 0; a < B2_MAX_AXES; a++) o[a] = E.pcoord[a];
    o[12] = E.lo;
    o[13] = E.hi;
  }
}

// ------------------------------------------------------------------------------------------
// GC: free-list compaction + recycled-node zero-fill + list swap in ONE launch.
//   part A  unused tail of the live free list -> front of the spare list   (gc_parallel_0)
//   part B  recycled indices appended behind it                            (gc_parallel_2, thread 0)
//   part C  recycled nodes zero-filled with 128-bit stores, 32 KiB segments so that one
//           512 KiB node is spread over 16 CTAs                            (gc_parallel_2)
//   last CTA to finish swaps the lists and resets the counters             (gc_parallel_1)
// ------------------------------------------------------------------------------------------
#define B2_GC_SEG (32 * 1024)
__device__ __forceinline__ void b2_gc_body(B2TreeDev *T, int sid) {
  B2NodeMgr *m = T->mgr[sid];
  const u32 cur = m->cur;
  B2CList *F0 = &m->free_list[cur], *F1 = &m->free_list[cur ^ 1];
  const u32 fc = m->free_count;
  const u32 used = m->free_used < fc ? m->free_used : fc;
  const u32 unused = fc - used;
  const u32 nrec = m->recycled.count;
  const u32 gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
  for (u32 i = gtid; i < unused; i += gsize) {
    u32 *d = (u32 *)b2_list_slot(T->pool, F1, i);
    if (d) *d = *(u32 *)b2_list_get(F0, used + i);
  }
  for (u32 i = gtid; i < nrec; i += gsize) {
    u32 *d = (u32 *)b2_list_slot(T->pool, F1, unused + i);
    if (d) *d = *(u32 *)b2_list_get(&m->recycled, i);
  }
  const u64 esz = m->elem_size;
  const u32 segs = (u32)((esz + B2_GC_SEG - 1) / B2_GC_SEG);
  const u64 items = (u64)nrec * segs;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (u64 it = blockIdx.x; it < items; it += gridDim.x) {
    const u32 r = (u32)(it / segs), sg = (u32)(it % segs);
    const u32 idx = *(u32 *)b2_list_get(&m->recycled, r);
    u8 *p = (u8 *)m->data_chunks[idx >> m->log2_epc] + (u64)(idx & ((1u << m->log2_epc) - 1)) * esz + (u64)sg * B2_GC_SEG;
    const u64 len = (esz - (u64)sg * B2_GC_SEG) < B2_GC_SEG ? (esz - (u64)sg * B2_GC_SEG) : B2_GC_SEG;
    for (u64 o = (u64)threadIdx.x * 16; o < len; o += (u64)blockDim.x * 16) *(uint4 *)(p + o) = z;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 t = atomicAdd(&m->gc_ticket, 1u);
    if (t == gridDim.x - 1) {
      m->free_count = unused + nrec;
      m->free_used = 0;
      m->recycled.count = 0;
      F1->count = unused + nrec;
      F0->count = 0;
      m->cur = cur ^ 1;
      m->gc_ticket = 0;
      m->gc_bar = 0;
      __threadfence();
    }
  }
}
__global__ void __launch_bounds__(256) k_gc(B2TreeDev *T, int sid) { b2_gc_body(T, sid); }

// SNode.deactivate_all + GC of a pointer SNode whose ONE container sits directly under the root, in a single launch (the sparse
// life cycle of the MPM solver runs it every substep): every active cell is deactivated and its node recycled, a grid barrier
// (all CTAs are resident: the grid is sized from the occupancy), then the fused GC above.  Also empties up to two element
// lists and a bitmap that the caller derives from the SNode's cells.  Replaces listgen + deactivate body + GC + three memsets.
__global__ void __launch_bounds__(256) k_deactivate_all_gc(B2TreeDev *T, int sid, u8 *node, int ncells, u32 *count_a, u32 *count_b, uint4 *zero16,
                                                           long long n_zero16) {
  B2NodeMgr *m = T->mgr[sid];
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x, gsize = (long long)gridDim.x * blockDim.x;
  for (long long i = gtid; i < ncells; i += gsize) b2_deactivate_cell(T, sid, node, (int)i);
  for (long long i = gtid; i < n_zero16; i += gsize) zero16[i] = make_uint4(0, 0, 0, 0);
  if (gtid == 0) {
    if (count_a) *count_a = 0;
    if (count_b) *count_b = 0;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(&m->gc_bar, 1u);
    unsigned ns = 32, polls = 0;
    while (*(volatile u32 *)&m->gc_bar < gridDim.x) {
      __nanosleep(ns);
      if (ns < 256) ns *= 2;
      if (++polls > (1u << 23)) {  // (seconds: part of the grid never became resident; flag it -- nodes recycled after this
        b2_set_error(T->pool, B2E_ASSERT, -1004);  // point are collected by the next GC)
        break;
      }
    }
    __threadfence();
  }
  __syncthreads();
  b2_gc_body(T, sid);
}

void B2Tree_t::gc_async(int sid) {
  check_snode(sid, "gc");
  if (s[sid].type != B2T_POINTER && s[sid].type != B2T_DYNAMIC) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "gc: snode is not gc-able (snode_types.cpp:21-23)");
  k_gc<<<rt->num_sms * 4, 256, 0, rt->stream()>>>(dev, sid);
  rt->count_launch();
  B2_CUDA(cudaGetLastError());
}

__global__ void k_deactivate_root_dynamic(B2TreeDev *T, int dsid) {
  b2_deactivate_cell(T, dsid, T->root + T->s[dsid].offset_in_parent_cell, 0);
}
void B2Tree_t::deactivate_all_gc_async(int sid, u8 *container, int list_a, int list_b, void *zero, size_t zero_bytes) {
  check_snode(sid, "deactivate_all_gc");
  if (s[sid].type != B2T_POINTER) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "deactivate_all_gc: pointer SNode");
  if (zero_bytes % 16 != 0) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "deactivate_all_gc: bitmap size");
  static int per_sm = 0;  // CTAs of the kernel that fit one SM (the grid barrier needs all of them resident)
  if (per_sm == 0) {
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_deactivate_all_gc, 256, 0));
    per_sm = std::max(1, std::min(per_sm, 4));
  }
  k_deactivate_all_gc<<<rt->num_sms * per_sm, 256, 0, rt->stream()>>>(dev, sid, container, (int)s[sid].num_cells,
                                                                      list_a >= 0 ? &host_copy.lists[list_a]->count : nullptr,
                                                                      list_b >= 0 ? &host_copy.lists[list_b]->count : nullptr, (uint4 *)zero,
                                                                      (long long)(zero_bytes / 16));
  rt->count_launch();
  B2_CUDA(cudaGetLastError());
}
void B2Tree_t::reset_list_async(int sid) {
  check_snode(sid, "reset_list");
  B2_CUDA(cudaMemsetAsync(&host_copy.lists[sid]->count, 0, 4, rt->stream()));
}
void B2Tree_t::deactivate_level_async(int sid) {
  check_snode(sid, "deactivate");
  if (s[sid].type != B2T_POINTER && s[sid].type != B2T_BITMASKED) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "deactivate_level: pointer or bitmasked SNode");
  cudaStream_t st = rt->stream();
  listgen_async(sid);
  int block = (int)std::min<int64_t>(256, std::max<int64_t>(32, roundup(std::min<int64_t>(s[sid].num_cells, 1024), 32)));
  k_struct_for<<<rt->num_sms * 8, block, 0, st>>>(dev, sid, BodyDeactivate{sid});
  rt->count_launch();
  if (s[sid].type == B2T_POINTER) gc_async(sid);
  B2_CUDA(cudaGetLastError());
}
void B2Tree_t::deactivate_all(int sid) {
  check_snode(sid, "deactivate_all");
  for (int c : s[sid].ch)
    if (s[c].type != B2T_PLACE) deactivate_all(c);
  cudaStream_t st = rt->stream();
  const int grid = rt->num_sms * 8;
  if (s[sid].type == B2T_POINTER || s[sid].type == B2T_BITMASKED) {
    deactivate_level_async(sid);
  } else if (s[sid].type == B2T_DYNAMIC) {
    int par = s[sid].parent;
    if (par != 0) listgen_async(par);
    int block = (int)std::min<int64_t>(256, std::max<int64_t>(32, roundup(std::min<int64_t>(s[par].num_cells, 1024), 32)));
    if (par == 0) {
      // ti.root.dynamic(...): the root cell holds the one list (test_dynamic.py:13, test_gc.py:64)
      k_deactivate_root_dynamic<<<1, 1, 0, st>>>(dev, sid);
    } else {
      k_struct_for<<<grid, block, 0, st>>>(dev, par, BodyDeactivateDynamic{par, sid});
    }
    rt->count_launch();
    gc_async(sid);
  }
  B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------
// C ABI (Part 2, tree functions)
// ------------------------------------------------------------------------------------------
static int block_for(const B2Tree_t *t, int sid) {
  return (int)std::min<int64_t>(256, std::max<int64_t>(32, roundup(std::min<int64_t>(t->s[sid].num_cells, 1024), 32)));
}
static void require(bool ok, TiError code, const char *msg) {
  if (!ok) throw B2Error(code, msg);
}

extern "C" {
B2Tree b2_create_tree(TiRuntime runtime, const B2SNodeDesc *desc, int32_t n) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  B2_CUDA(cudaSetDevice(runtime->device));
  return new B2Tree_t(runtime, desc, n);
  B2_API_END(nullptr)
}
void b2_destroy_tree(B2Tree tree) { delete tree; }
int64_t b2_snode_num_cells(B2Tree t, int32_t id) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(id, "num_cells");
  return t->s[id].num_cells;
  B2_API_END(-1)
}
int64_t b2_snode_cell_size(B2Tree t, int32_t id) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(id, "cell_size");
  return t->s[id].cell_size;
  B2_API_END(-1)
}
int64_t b2_snode_node_size(B2Tree t, int32_t id) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(id, "node_size");
  return t->s[id].node_size;
  B2_API_END(-1)
}

static int32_t coord_op(B2Tree t, int32_t sid, const int32_t *coords, int64_t n, int mode, uint8_t *out) {
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(sid, "coord op");
  require(t->s[sid].type != B2T_PLACE, TI_ERROR_INVALID_ARGUMENT, "snode is a place");
  if (n <= 0) return 0;
  require(coords != nullptr, TI_ERROR_ARGUMENT_NULL, "coords");
  cudaStream_t st = t->rt->stream();
  B2Staged c, o;
  c.init(coords, (size_t)n * 48, true, false, st);
  if (out) o.init(out, (size_t)n, false, true, st);
  B2Path P = t->path_to(sid);
  k_coord_op<<<t->rt->grid_for(n, 128), 128, 0, st>>>(t->dev, P, c.as<int>(), n, mode, o.as<u8>());
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  if (mode == 1 && (t->s[sid].type == B2T_POINTER || t->s[sid].type == B2T_DYNAMIC)) t->gc_async(sid);
  o.finish();
  c.finish();
  return 0;
}
int32_t b2_activate(B2Tree t, int32_t sid, const int32_t *coords, int64_t n) {
  B2_API_BEGIN
  return coord_op(t, sid, coords, n, 0, nullptr);
  B2_API_END(-1)
}
int32_t b2_deactivate(B2Tree t, int32_t sid, const int32_t *coords, int64_t n) {
  B2_API_BEGIN
  return coord_op(t, sid, coords, n, 1, nullptr);
  B2_API_END(-1)
}
int32_t b2_is_active(B2Tree t, int32_t sid, const int32_t *coords, int64_t n, uint8_t *out) {
  B2_API_BEGIN
  require(out != nullptr || n == 0, TI_ERROR_ARGUMENT_NULL, "out");
  return coord_op(t, sid, coords, n, 2, out);
  B2_API_END(-1)
}
int64_t b2_listgen(B2Tree t, int32_t sid) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->listgen_async(sid);
  return t->list_count(sid);
  B2_API_END(-1)
}
int32_t b2_listgen_async(B2Tree t, int32_t sid) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->listgen_async(sid);
  return 0;
  B2_API_END(-1)
}
int64_t b2_get_list(B2Tree t, int32_t sid, int32_t *out, int64_t cap) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(sid, "get_list");
  int64_t cnt = t->list_count(sid);
  if (sid == 0 || cnt == 0 || cap <= 0 || !out) return cnt;
  int64_t m = std::min(cnt, cap);
  cudaStream_t st = t->rt->stream();
  B2Staged o;
  o.init(out, (size_t)m * 56, false, true, st);
  k_export_list<<<t->rt->grid_for(m, 128), 128, 0, st>>>(t->dev, sid, o.as<int>(), m);
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  o.finish();
  return cnt;
  B2_API_END(-1)
}
int64_t b2_struct_for_coords(B2Tree t, int32_t sid, int32_t *out, int64_t cap) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(sid, "struct_for");
  require(t->s[sid].type != B2T_PLACE && sid != 0, TI_ERROR_INVALID_ARGUMENT, "struct-for target must be a non-root, non-place snode");
  cudaStream_t st = t->rt->stream();
  t->listgen_async(sid);
  unsigned long long *counter = (unsigned long long *)t->scratch;
  B2_CUDA(cudaMemsetAsync(counter, 0, 8, st));
  B2Staged o;
  if (out && cap > 0) o.init(out, (size_t)cap * 48, false, true, st);
  k_struct_for<<<t->rt->num_sms * 8, block_for(t, sid), 0, st>>>(t->dev, sid, BodyCollect{o.as<int>(), cap, counter});
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  unsigned long long cnt = 0;
  B2_CUDA(cudaMemcpyAsync(&cnt, counter, 8, cudaMemcpyDeviceToHost, st));
  t->rt->sync();
  o.finish();
  return (int64_t)cnt;
  B2_API_END(-1)
}
int32_t b2_deactivate_all(B2Tree t, int32_t sid) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->deactivate_all(sid);
  return 0;
  B2_API_END(-1)
}
int32_t b2_gc(B2Tree t, int32_t sid) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->gc_async(sid);
  return 0;
  B2_API_END(-1)
}
int32_t b2_alloc_stats(B2Tree t, int32_t sid, int32_t *out) {
  B2_API_BEGIN
  require(t && out, TI_ERROR_ARGUMENT_NULL, "tree/out");
  t->check_snode(sid, "alloc_stats");
  require(t->host_copy.mgr[sid] != nullptr, TI_ERROR_INVALID_ARGUMENT, "snode has no NodeManager");
  B2NodeMgr h;
  B2_CUDA(cudaMemcpyAsync(&h, t->host_copy.mgr[sid], offsetof(B2NodeMgr, data_chunks), cudaMemcpyDeviceToHost, t->rt->stream()));
  u32 rec = 0;
  B2_CUDA(cudaMemcpyAsync(&rec, &t->host_copy.mgr[sid]->recycled.count, 4, cudaMemcpyDeviceToHost, t->rt->stream()));
  t->rt->sync();
  out[0] = (int)h.data_count;
  out[1] = (int)h.free_count;
  out[2] = (int)rec;
  out[3] = (int)h.free_used;
  return 0;
  B2_API_END(-1)
}
int64_t b2_num_dynamically_allocated(B2Tree t, int32_t sid) {
  // NodeManager data_list size (llvm_runtime_executor.cpp get_snode_num_dynamically_allocated)
  int32_t st[4];
  if (b2_alloc_stats(t, sid, st) != 0) return -1;
  return st[0];
}

static int32_t access(B2Tree t, int32_t place, const int32_t *coords, int64_t n, void *vals, int mode) {
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(place, "access");
  require(t->s[place].type == B2T_PLACE, TI_ERROR_INVALID_ARGUMENT, "not a place");
  if (n <= 0) return 0;
  require(coords && vals, TI_ERROR_ARGUMENT_NULL, "coords/vals");
  const B2SNodeHost &p = t->s[place];
  require(p.parent != 0 || true, TI_ERROR_INVALID_ARGUMENT, "");
  cudaStream_t st = t->rt->stream();
  B2Staged c, v;
  c.init(coords, (size_t)n * 48, true, false, st);
  v.init(vals, (size_t)n * p.dtype_bytes, mode < 2, mode == 2, st);
  B2Path P = t->path_to(p.parent);
  k_access<<<t->rt->grid_for(n, 128), 128, 0, st>>>(t->dev, P, p.offset_in_parent_cell, p.dtype_bytes, c.as<int>(), n, mode, v.as<u8>());
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  v.finish();
  c.finish();
  return 0;
}
int32_t b2_write(B2Tree t, int32_t place, const int32_t *coords, int64_t n, const void *vals, int32_t atomic_add_f32) {
  B2_API_BEGIN
  return access(t, place, coords, n, (void *)vals, atomic_add_f32 ? 1 : 0);
  B2_API_END(-1)
}
int32_t b2_read(B2Tree t, int32_t place, const int32_t *coords, int64_t n, void *vals) {
  B2_API_BEGIN
  return access(t, place, coords, n, vals, 2);
  B2_API_END(-1)
}
int32_t b2_dynamic_append_i32(B2Tree t, int32_t dyn, const int32_t *coords, int64_t n, const int32_t *vals) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(dyn, "append");
  require(t->s[dyn].type == B2T_DYNAMIC, TI_ERROR_INVALID_ARGUMENT, "not a dynamic snode");
  if (n <= 0) return 0;
  require(coords && vals, TI_ERROR_ARGUMENT_NULL, "coords/vals");
  cudaStream_t st = t->rt->stream();
  B2Staged c, v;
  c.init(coords, (size_t)n * 48, true, false, st);
  v.init(vals, (size_t)n * 4, true, false, st);
  k_dynamic_append<<<t->rt->grid_for(n, 128), 128, 0, st>>>(t->dev, t->path_to(dyn), c.as<int>(), n, v.as<int>());
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  v.finish();
  c.finish();
  return 0;
  B2_API_END(-1)
}
int32_t b2_dynamic_length(B2Tree t, int32_t dyn, const int32_t *coords, int64_t n, int32_t *out) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(dyn, "length");
  require(t->s[dyn].type == B2T_DYNAMIC, TI_ERROR_INVALID_ARGUMENT, "not a dynamic snode");
  if (n <= 0) return 0;
  require(coords && out, TI_ERROR_ARGUMENT_NULL, "coords/out");
  cudaStream_t st = t->rt->stream();
  B2Staged c, o;
  c.init(coords, (size_t)n * 48, true, false, st);
  o.init(out, (size_t)n * 4, false, true, st);
  k_dynamic_length<<<t->rt->grid_for(n, 128), 128, 0, st>>>(t->dev, t->path_to(dyn), c.as<int>(), n, o.as<int>());
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  o.finish();
  c.finish();
  return 0;
  B2_API_END(-1)
}

int32_t b2_field_fill(B2Tree t, int32_t place, uint64_t bits) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  t->check_snode(place, "fill");
  const B2SNodeHost &p = t->s[place];
  require(p.type == B2T_PLACE && p.parent > 0, TI_ERROR_INVALID_ARGUMENT, "fill needs a place below a non-root snode");
  t->listgen_async(p.parent);
  k_struct_for<<<t->rt->num_sms * 8, block_for(t, p.parent), 0, t->rt->stream()>>>(
      t->dev, p.parent, BodyFill{p.offset_in_parent_cell, p.dtype_bytes, bits, p.parent});
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
static void dense_dims(const B2Tree_t *t, int par, int *dims, int64_t *total) {
  *total = 1;
  for (int a = 0; a < B2_MAX_AXES; a++) {
    dims[a] = t->s[par].from_root[a];
    *total *= dims[a];
  }
}
int32_t b2_field_to_dense(B2Tree t, int32_t place, void *out) {
  B2_API_BEGIN
  require(t && out, TI_ERROR_ARGUMENT_NULL, "tree/out");
  t->check_snode(place, "to_dense");
  const B2SNodeHost &p = t->s[place];
  require(p.type == B2T_PLACE && p.parent > 0, TI_ERROR_INVALID_ARGUMENT, "to_dense needs a place below a non-root snode");
  BodyToDense body;
  int64_t total;
  dense_dims(t, p.parent, body.dims, &total);
  cudaStream_t st = t->rt->stream();
  B2Staged o;
  o.init(out, (size_t)total * p.dtype_bytes, false, true, st);
  if (!o.owned) B2_CUDA(cudaMemsetAsync(o.dev, 0, (size_t)total * p.dtype_bytes, st));
  body.place_off = p.offset_in_parent_cell;
  body.bytes = p.dtype_bytes;
  body.psid = p.parent;
  body.out = o.as<u8>();
  t->listgen_async(p.parent);
  k_struct_for<<<t->rt->num_sms * 8, block_for(t, p.parent), 0, st>>>(t->dev, p.parent, body);
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  o.finish();
  return 0;
  B2_API_END(-1)
}
int32_t b2_field_from_dense(B2Tree t, int32_t place, const void *in) {
  B2_API_BEGIN
  require(t && in, TI_ERROR_ARGUMENT_NULL, "tree/in");
  t->check_snode(place, "from_dense");
  const B2SNodeHost &p = t->s[place];
  require(p.type == B2T_PLACE && p.parent > 0, TI_ERROR_INVALID_ARGUMENT, "from_dense needs a place below a non-root snode");
  B2Path dimsP;
  memset(&dimsP, 0, sizeof(dimsP));
  int64_t total;
  dense_dims(t, p.parent, dimsP.div[0], &total);
  cudaStream_t st = t->rt->stream();
  B2Staged i;
  i.init(in, (size_t)total * p.dtype_bytes, true, false, st);
  k_from_dense<<<t->rt->grid_for(total, 128), 128, 0, st>>>(t->dev, t->path_to(p.parent), p.offset_in_parent_cell, p.dtype_bytes, i.as<u8>(), total, dimsP);
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  i.finish();
  return 0;
  B2_API_END(-1)
}

static u32 ident_bits(int op, int dtype) {
  if (dtype == 0) {
    float f = op == 0 ? 0.f : (op == 1 ? INFINITY : -INFINITY);
    u32 b;
    memcpy(&b, &f, 4);
    return b;
  }
  int v = op == 0 ? 0 : (op == 1 ? INT32_MAX : INT32_MIN);
  return (u32)v;
}
int32_t b2_field_reduce(B2Tree t, int32_t place, int32_t op, int32_t dtype, void *out_host) {
  B2_API_BEGIN
  require(t && out_host, TI_ERROR_ARGUMENT_NULL, "tree/out");
  t->check_snode(place, "reduce");
  const B2SNodeHost &p = t->s[place];
  require(p.type == B2T_PLACE && p.parent > 0 && p.dtype_bytes == 4, TI_ERROR_INVALID_ARGUMENT, "reduce needs a 4-byte place below a non-root snode");
  require(op >= 0 && op <= 2 && (dtype == 0 || dtype == 1), TI_ERROR_INVALID_ARGUMENT, "op/dtype");
  cudaStream_t st = t->rt->stream();
  u32 id = ident_bits(op, dtype);
  u32 *dst = (u32 *)t->scratch + 16;
  B2_CUDA(cudaMemcpyAsync(dst, &id, 4, cudaMemcpyHostToDevice, st));
  t->listgen_async(p.parent);
  const int grid = t->rt->num_sms * 8, block = 256;
  if (dtype == 0) {
    float idf;
    memcpy(&idf, &id, 4);
    k_struct_for<<<grid, block, 0, st>>>(t->dev, p.parent, BodyReduce<float>{p.offset_in_parent_cell, p.parent, op, idf, (float *)dst, idf});
  } else {
    k_struct_for<<<grid, block, 0, st>>>(t->dev, p.parent, BodyReduce<int>{p.offset_in_parent_cell, p.parent, op, (int)id, (int *)dst, (int)id});
  }
  t->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpyAsync(out_host, dst, 4, cudaMemcpyDeviceToHost, st));
  t->rt->sync();
  return 0;
  B2_API_END(-1)
}
// enqueue only: the 4-byte result lands in device memory
int32_t b2_reduce_device(TiRuntime rt, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_dev) {
  B2_API_BEGIN
  require(rt && out_dev, TI_ERROR_ARGUMENT_NULL, "runtime/out");
  require(op >= 0 && op <= 2 && (dtype == 0 || dtype == 1), TI_ERROR_INVALID_ARGUMENT, "op/dtype");
  require(n == 0 || b2_is_device_pointer(data_dev), TI_ERROR_INVALID_ARGUMENT, "data must be a device pointer");
  require(b2_is_device_pointer(out_dev), TI_ERROR_INVALID_ARGUMENT, "out must be a device pointer");
  cudaStream_t st = rt->stream();
  const u32 id = ident_bits(op, dtype);
  k_set_u32<<<1, 1, 0, st>>>((u32 *)out_dev, id);
  rt->count_launch();
  if (n > 0) {
    const int grid = rt->grid_for(n, 256 * 16, 8);
    if (dtype == 0) {
      float idf;
      memcpy(&idf, &id, 4);
      k_reduce<float><<<grid, 256, 0, st>>>((const float *)data_dev, n, op, idf, (float *)out_dev);
    } else {
      k_reduce<int><<<grid, 256, 0, st>>>((const int *)data_dev, n, op, (int)id, (int *)out_dev);
    }
    rt->count_launch();
  }
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
// Allocation-free variant of b2_reduce_device: no cudaMalloc / cudaFree / stream-ordered allocation on this path; the
// scratch is the static device array above.  n <= B2_NOALLOC_MAX_N.
int32_t b2_reduce_noalloc(TiRuntime rt, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_dev) {
  B2_API_BEGIN
  require(rt && out_dev, TI_ERROR_ARGUMENT_NULL, "runtime/out");
  require(op >= 0 && op <= 2 && (dtype == 0 || dtype == 1), TI_ERROR_INVALID_ARGUMENT, "op/dtype");
  require(n >= 0, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "b2_reduce_noalloc: n < 0");
  if (n > B2_NOALLOC_MAX_N)
    throw B2Error(TI_ERROR_ARGUMENT_OUT_OF_RANGE, "b2_reduce_noalloc: n = " + std::to_string((long long)n) + " exceeds the static scratch's maximum of " +
                                                      std::to_string(B2_NOALLOC_MAX_N) + " elements (use b2_reduce_device)");
  require(n == 0 || data_dev, TI_ERROR_ARGUMENT_NULL, "data");
  require(rt->device >= 0 && rt->device < 64 && g_noalloc_scratch[rt->device], TI_ERROR_INVALID_STATE,
          "b2_reduce_noalloc: the scratch array was not set up for this device");
  u32 *scratch = g_noalloc_scratch[rt->device];
  cudaStream_t st = rt->stream();
  const u32 id = ident_bits(op, dtype);
  B2_CUDA(cudaMemsetAsync(scratch, 0, 4, st));  // the ticket starts at zero whatever an earlier (aborted) call left
  int grid = (int)std::min<long long>(B2_NOALLOC_MAX_CTAS, std::max<long long>(1, rt->grid_for(std::max<int64_t>(n, 1), 256 * 16, 8)));
  if (dtype == 0) {
    float idf;
    memcpy(&idf, &id, 4);
    k_reduce_noalloc<float><<<grid, 256, 0, st>>>((const float *)data_dev, n, op, idf, scratch, (float *)out_dev);
  } else {
    k_reduce_noalloc<int><<<grid, 256, 0, st>>>((const int *)data_dev, n, op, (int)id, scratch, (int *)out_dev);
  }
  rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_reduce(TiRuntime rt, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_host) {
  B2_API_BEGIN
  require(rt && out_host, TI_ERROR_ARGUMENT_NULL, "runtime/out");
  cudaStream_t st = rt->stream();
  u32 *dst = nullptr;
  B2_CUDA(cudaMallocAsync((void **)&dst, 4, st));
  const int32_t rc = b2_reduce_device(rt, data_dev, n, op, dtype, dst);
  if (rc == 0) B2_CUDA(cudaMemcpyAsync(out_host, dst, 4, cudaMemcpyDeviceToHost, st));
  rt->sync();
  cudaFreeAsync(dst, st);
  return rc;
  B2_API_END(-1)
}

void b2_sync(TiRuntime rt) {
  B2_API_BEGIN
  if (rt) rt->sync();
  B2_API_END_VOID
}
int32_t b2_device_index(TiRuntime rt) { return rt ? rt->device : -1; }
int64_t b2_pool_bytes_used(TiRuntime rt) {
  B2_API_BEGIN
  require(rt, TI_ERROR_ARGUMENT_NULL, "runtime");
  B2Pool h;
  B2_CUDA(cudaMemcpyAsync(&h, rt->pool_dev, sizeof(h), cudaMemcpyDeviceToHost, rt->stream()));
  rt->sync();
  return (int64_t)std::min(h.head, h.size);
  B2_API_END(-1)
}
// print_memory_profiler_info (llvm_runtime_executor.cpp:314-384): the same walk over the tree, one block of lines per
// non-place SNode -- its element list, and for SNodes with a NodeManager the data list and the allocator counters --
// followed by the bytes taken from the pool.  Written into `out` (size query like ti_get_last_error); synchronises.
int64_t b2_memory_profiler_info(B2Tree t, char *out, int64_t cap) {
  B2_API_BEGIN
  require(t, TI_ERROR_ARGUMENT_NULL, "tree");
  static const char *names[] = {"root", "dense", "pointer", "bitmasked", "dynamic", "place"};
  cudaStream_t st = t->rt->stream();
  std::string r = "[Memory Profiler]\n";
  char line[256];
  for (int sid = 0; sid < (int)t->s.size(); sid++) {
    const B2SNodeHost &sn = t->s[sid];
    if (sn.type == B2T_PLACE) continue;
    snprintf(line, sizeof(line), "SNode S%d%s\n", sid, names[sn.type]);
    r += line;
    if (t->host_copy.lists[sid]) {
      u32 hdr[4] = {0, 0, 0, 0};  // count, elem_size, log2_epc
      B2_CUDA(cudaMemcpyAsync(hdr, t->host_copy.lists[sid], 16, cudaMemcpyDeviceToHost, st));
      B2_CUDA(cudaStreamSynchronize(st));
      const unsigned long long epc = 1ull << hdr[2];
      snprintf(line, sizeof(line), "  active element list: num_elements=%u; element_size=%u B; %llu elements per chunk\n", hdr[0], hdr[1], epc);
      r += line;
    }
    if (t->host_copy.mgr[sid]) {
      int32_t a[4];
      if (b2_alloc_stats(t, sid, a) != 0) return -1;
      snprintf(line, sizeof(line), "  data list:           num_elements=%d; element_size=%lld B\n", a[0], (long long)t->s[sid].alloc_size);
      r += line;
      snprintf(line, sizeof(line), "  Allocated elements=%d; free list length=%d; recycled list length=%d\n", a[3], a[1], a[2]);
      r += line;
    }
  }
  const int64_t used = b2_pool_bytes_used(t->rt);
  snprintf(line, sizeof(line), "Total requested dynamic memory: %lld B of a %llu B pool\n", (long long)used, (unsigned long long)t->rt->pool_size);
  r += line;
  if (out && cap > 0) {
    const size_t n = std::min<size_t>((size_t)cap - 1, r.size());
    memcpy(out, r.data(), n);
    out[n] = 0;
  }
  return (int64_t)r.size() + 1;
  B2_API_END(-1)
}
uint64_t b2_launch_count(TiRuntime rt) { return rt ? rt->launches : 0; }
int32_t b2_check_runtime_error(TiRuntime rt, char *msg, uint64_t cap) {
  B2_API_BEGIN
  require(rt, TI_ERROR_ARGUMENT_NULL, "runtime");
  std::string m;
  int code = rt->check_error(&m);
  if (code && msg && cap) {
    size_t k = std::min<size_t>(cap - 1, m.size());
    memcpy(msg, m.data(), k);
    msg[k] = 0;
  }
  return code;
  B2_API_END(-1)
}
}  // extern "C"
