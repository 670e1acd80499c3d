// b2_jit.cu -- run-time compiled struct-for bodies: host side.
//
// What it stands in for in the reference: the per-kernel LLVM codegen of a struct-for task
// (codegen_llvm.cpp:2089-2317), the CUDA JIT session (runtime/cuda/jit_cuda.cpp:8-240: PTX -> module ->
// function) and the launcher's per-task launch (runtime/cuda/kernel_launcher.cpp:15-213).  Here the body is
// CUDA C++ text (contract in b2_jit.cuh); it is compiled ONCE per source with NVRTC straight to an sm_100a
// cubin (no PTX JIT in the driver), loaded through the runtime API (cudaLibraryLoadData) and launched over
// the element list of the loop SNode, which is regenerated before the launch like the reference does
// (offload.cpp:174-264).  libnvrtc is bound at run time: the library loads without it, and the first
// b2_jit_compile fails loudly if it cannot be found.
#include <dlfcn.h>

#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "b2_host.h"
#include "b2_jit.cuh"
#include "b2_embed.h"        // b2_src_internal / b2_src_jit / b2_src_svd: the device headers as text

namespace {
typedef struct _nvrtcProgram *nvrtcProgram;
struct Nvrtc {
  void *h = nullptr;
  int (*CreateProgram)(nvrtcProgram *, const char *, const char *, int, const char *const *, const char *const *) = nullptr;
  int (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
  int (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
  int (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
  int (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
  int (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
  int (*DestroyProgram)(nvrtcProgram *) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
Nvrtc &nvrtc() {
  static Nvrtc n;
  static std::once_flag once;
  std::call_once(once, [] {
    const char *env = getenv("B2_NVRTC_LIB");
    const char *cands[] = {env, "libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so"};
    for (const char *c : cands) {
      if (!c || !*c) continue;
      n.h = dlopen(c, RTLD_NOW | RTLD_LOCAL);
      if (n.h) break;
    }
    if (!n.h) return;
#define B2_SYM(name) n.name = (decltype(n.name))dlsym(n.h, "nvrtc" #name)
    B2_SYM(CreateProgram), B2_SYM(CompileProgram), B2_SYM(GetProgramLogSize), B2_SYM(GetProgramLog);
    B2_SYM(GetCUBINSize), B2_SYM(GetCUBIN), B2_SYM(DestroyProgram), B2_SYM(GetErrorString);
#undef B2_SYM
    if (!n.CreateProgram || !n.CompileProgram || !n.GetCUBINSize || !n.GetCUBIN || !n.DestroyProgram) n.h = nullptr;
  });
  return n;
}
}  // namespace

struct B2JitKernel_t {
  TiRuntime_t *rt = nullptr;
  cudaLibrary_t lib = nullptr;
  cudaKernel_t fn = nullptr;        // b2_jit_struct_for (B2_BODY), if the source has one
  cudaKernel_t fn_range = nullptr;  // b2_jit_range_for (B2_RANGE_BODY), if the source has one
  B2JitTable *tab_dev = nullptr;  // launch table (fields + args), rewritten in stream order before every launch
  std::string log;
  size_t cubin_bytes = 0;
  size_t bls_smem = 0;  // dynamic shared memory of the next struct-for launch (block-local pads)
};

static void jit_require(bool ok, TiError code, const char *msg) {
  if (!ok) throw B2Error(code, msg);
}

// source -> sm_100a cubin (throws with the compiler log on errors); needs no GPU
static void jit_build(const char *source, const char *name, std::vector<char> *cubin, std::string *log) {
  jit_require(source != nullptr, TI_ERROR_ARGUMENT_NULL, "source");
  Nvrtc &N = nvrtc();
  jit_require(N.h != nullptr, TI_ERROR_NOT_SUPPORTED, "libnvrtc.so.12 not found (set B2_NVRTC_LIB): struct-for bodies cannot be compiled");
  std::string src;
  if (strstr(source, "B2_RANGE_BODY")) src += "#define B2_JIT_HAS_RANGE 1\n";
  if (strstr(source, "B2_BODY")) src += "#define B2_JIT_HAS_BODY 1\n";  // ("B2_RANGE_BODY" does not contain it)
  src += "#include \"b2_internal.cuh\"\n#include \"b2_svd.cuh\"\n#include \"b2_jit.cuh\"\n#line 1 \"";
  src += (name && *name) ? name : "b2_body";
  src += "\"\n";
  src += source;
  src += "\n";
  const char *headers[3] = {b2_src_internal, b2_src_jit, b2_src_svd};
  const char *hnames[3] = {"b2_internal.cuh", "b2_jit.cuh", "b2_svd.cuh"};
  nvrtcProgram prog = nullptr;
  int rc = N.CreateProgram(&prog, src.c_str(), "b2_jit_kernel.cu", 3, headers, hnames);
  jit_require(rc == 0 && prog != nullptr, TI_ERROR_INVALID_STATE, "nvrtcCreateProgram failed");
  const char *opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-lineinfo"};
  rc = N.CompileProgram(prog, 3, opts);
  size_t ln = 0;
  log->clear();
  if (N.GetProgramLogSize && N.GetProgramLog && N.GetProgramLogSize(prog, &ln) == 0 && ln > 1) {
    log->resize(ln);
    N.GetProgramLog(prog, &(*log)[0]);
    while (!log->empty() && (log->back() == 0 || log->back() == '\n')) log->pop_back();
  }
  if (rc != 0) {
    N.DestroyProgram(&prog);
    throw B2Error(TI_ERROR_INVALID_ARGUMENT, "struct-for body does not compile:\n" + *log);
  }
  size_t nb = 0;
  rc = N.GetCUBINSize(prog, &nb);
  cubin->resize(nb);
  if (rc == 0 && nb) rc = N.GetCUBIN(prog, cubin->data());
  N.DestroyProgram(&prog);
  jit_require(rc == 0 && nb > 0, TI_ERROR_INVALID_STATE, "nvrtcGetCUBIN failed");
}

extern "C" {
// Compile only (no device needed): size of the sm_100a cubin, or -1 with the compiler log as last error.
int64_t b2_jit_check(const char *source, const char *name) {
  B2_API_BEGIN
  std::vector<char> cubin;
  std::string log;
  jit_build(source, name, &cubin, &log);
  return (int64_t)cubin.size();
  B2_API_END(-1)
}
B2JitKernel b2_jit_compile(TiRuntime rt, const char *source, const char *name) {
  B2_API_BEGIN
  jit_require(rt != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  std::vector<char> cubin;
  std::string log;
  jit_build(source, name, &cubin, &log);
  const size_t nb = cubin.size();
  B2_CUDA(cudaSetDevice(rt->device));
  std::unique_ptr<B2JitKernel_t> k(new B2JitKernel_t);
  k->rt = rt;
  k->log = log;
  k->cubin_bytes = nb;
  B2_CUDA(cudaLibraryLoadData(&k->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0));
  if (cudaLibraryGetKernel(&k->fn, k->lib, "b2_jit_struct_for") != cudaSuccess) k->fn = nullptr;
  if (cudaLibraryGetKernel(&k->fn_range, k->lib, "b2_jit_range_for") != cudaSuccess) k->fn_range = nullptr;
  cudaGetLastError();
  jit_require(k->fn != nullptr || k->fn_range != nullptr, TI_ERROR_INVALID_ARGUMENT, "source defines neither B2_BODY(c) nor B2_RANGE_BODY(c)");
  B2_CUDA(cudaMalloc((void **)&k->tab_dev, sizeof(B2JitTable)));
  return k.release();
  B2_API_END(nullptr)
}

void b2_jit_destroy(B2JitKernel k) {
  if (!k) return;
  cudaStreamSynchronize(k->rt->stream());
  if (k->tab_dev) cudaFree(k->tab_dev);
  if (k->lib) cudaLibraryUnload(k->lib);
  delete k;
}

// Compiler diagnostics (warnings) of a successful compile; size query like ti_get_last_error.
int64_t b2_jit_log(B2JitKernel k, char *out, int64_t cap) {
  if (!k) return -1;
  if (out && cap > 0) {
    const size_t n = std::min<size_t>((size_t)cap - 1, k->log.size());
    memcpy(out, k->log.data(), n);
    out[n] = 0;
  }
  return (int64_t)k->log.size() + 1;
}

// fields + args -> the kernel's launch table, uploaded in stream order
static void jit_upload_table(B2JitKernel k, B2Tree t, const int32_t *field_places, int32_t n_fields, const uint64_t *args, int32_t n_args,
                             int32_t loop_snode = -1, const B2BlsSpec *bls = nullptr, int32_t n_bls = 0) {
  jit_require(k != nullptr && t != nullptr, TI_ERROR_ARGUMENT_NULL, "kernel/tree");
  jit_require(k->rt == t->rt, TI_ERROR_INVALID_ARGUMENT, "kernel and tree belong to different runtimes");
  jit_require(n_fields >= 0 && n_fields <= B2_JIT_MAX_FIELDS, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "at most 8 fields per body");
  jit_require(n_args >= 0 && n_args <= B2_JIT_MAX_ARGS, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "at most 16 arguments per body");
  jit_require(n_fields == 0 || field_places != nullptr, TI_ERROR_ARGUMENT_NULL, "field_places");
  jit_require(n_args == 0 || args != nullptr, TI_ERROR_ARGUMENT_NULL, "args");
  static thread_local B2JitTable tab;  // 10 KB: not on the stack
  memset(&tab, 0, sizeof(tab));
  tab.n_fields = n_fields, tab.n_args = n_args;
  tab.rand_states = (unsigned long long)t->rt->ensure_rand_states(), tab.num_rand = t->rt->num_rand_states;
  for (int i = 0; i < n_args; i++) tab.args[i] = args[i];
  for (int i = 0; i < n_fields; i++) {
    const int pid = field_places[i];
    t->check_snode(pid, "jit_launch field");
    const B2SNodeHost &p = t->s[pid];
    jit_require(p.type == B2T_PLACE, TI_ERROR_INVALID_ARGUMENT, "fields are place SNodes");
    tab.f[i].path = t->path_to(p.parent);
    tab.f[i].off = p.offset_in_parent_cell;
    tab.f[i].bytes = p.dtype_bytes;
    tab.f[i].holder = p.parent;
  }
  // block-local pads: geometry from the declared offsets (ScratchPad), staged per Element by the kernel
  size_t smem = 0;
  tab.n_bls = n_bls;
  for (int b = 0; b < n_bls; b++) {
    const B2BlsSpec &S = bls[b];
    B2JitBls &B = tab.bls[b];
    jit_require(S.field >= 0 && S.field < n_fields, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "block_local: field index");
    jit_require(S.ndim >= 1 && S.ndim <= 3, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "block_local: 1..3 pad axes");
    jit_require(S.mode >= 1 && S.mode <= 4, TI_ERROR_INVALID_ARGUMENT, "block_local: mode 1 read, 2 / 3 / 4 accumulate f32 / i32 / f64");
    const B2SNodeHost &place = t->s[field_places[S.field]];
    const B2SNodeHost &holder = t->s[place.parent];
    B.field = S.field, B.mode = S.mode, B.ndim = S.ndim, B.halo = S.halo, B.bytes = place.dtype_bytes;
    jit_require(B.bytes == 4 || B.bytes == 8, TI_ERROR_INVALID_ARGUMENT, "block_local: 4- or 8-byte fields");
    jit_require(S.mode == 1 || (S.mode == 4) == (B.bytes == 8), TI_ERROR_INVALID_ARGUMENT, "block_local: accumulate mode does not match the field's width");
    B.count = 1;
    for (int d = 0; d < S.ndim; d++) {
      const int a = S.axis[d];
      jit_require(a >= 0 && a < B2_MAX_AXES, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "block_local: axis");
      int ext = S.extent[d];
      if (ext <= 0) ext = t->s[loop_snode].shape[a];  // the loop SNode's own block (dense leaf loops)
      jit_require(ext >= 1 && S.lo[d] <= S.hi[d], TI_ERROR_INVALID_ARGUMENT, "block_local: extent / bounds");
      B.axis[d] = a, B.extent[d] = ext, B.lo[d] = S.lo[d], B.hi[d] = S.hi[d];
      B.size[d] = ext + S.hi[d] - S.lo[d];  // ScratchPad::pad_size = bounds.high - bounds.low
      B.field_extent[d] = holder.from_root[a];
      B.count *= B.size[d];
    }
    smem = (smem + 15) / 16 * 16;
    B.smem_off = (int)smem;
    smem += (size_t)B.count * B.bytes;
  }
  jit_require(smem <= 48 * 1024, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "block_local: the pads need more than 48 KB of shared memory");
  k->bls_smem = smem;
  B2_CUDA(cudaSetDevice(t->rt->device));
  B2_CUDA(cudaMemcpyAsync(k->tab_dev, &tab, sizeof(tab), cudaMemcpyHostToDevice, t->rt->stream()));  // pageable source: staged before return
}

// ScratchPad geometry (ir/scratch_pad.h:30-180, analysis/bls_analyzer.cpp:8-106): a block of block_size[d] loop cells whose body
// reaches the field at the given offsets needs the index range [low, high) = [min offset, block_size + max offset) per axis;
// pad_size = high - low; the linear pad index of (i0, i1, ..) is row-major over pad_size.  Returns the number of pad cells.
int64_t b2_bls_pad_geometry(int32_t ndim, const int32_t *block_size, int32_t n_access, const int32_t *offsets, int32_t *low, int32_t *high,
                            int32_t *pad_size) {
  B2_API_BEGIN
  jit_require(ndim >= 1 && ndim <= B2_MAX_AXES && block_size && offsets && n_access >= 1, TI_ERROR_INVALID_ARGUMENT, "pad geometry arguments");
  int64_t total = 1;
  for (int d = 0; d < ndim; d++) {
    int lo = offsets[d], hi = offsets[d];
    for (int a = 1; a < n_access; a++) lo = std::min(lo, offsets[a * ndim + d]), hi = std::max(hi, offsets[a * ndim + d]);
    const int l = lo, h = block_size[d] + hi;
    if (low) low[d] = l;
    if (high) high[d] = h;
    if (pad_size) pad_size[d] = h - l;
    total *= h - l;
  }
  return total;
  B2_API_END(-1)
}

static int32_t jit_launch_struct_for(B2JitKernel k, B2Tree t, int32_t loop_snode, const int32_t *field_places, int32_t n_fields, const uint64_t *args,
                                     int32_t n_args, const B2BlsSpec *bls, int32_t n_bls);
int32_t b2_jit_launch(B2JitKernel k, B2Tree t, int32_t loop_snode, const int32_t *field_places, int32_t n_fields, const uint64_t *args, int32_t n_args) {
  B2_API_BEGIN
  return jit_launch_struct_for(k, t, loop_snode, field_places, n_fields, args, n_args, nullptr, 0);
  B2_API_END(-1)
}
int32_t b2_jit_launch_bls(B2JitKernel k, B2Tree t, int32_t loop_snode, const int32_t *field_places, int32_t n_fields, const uint64_t *args, int32_t n_args,
                          const B2BlsSpec *bls, int32_t n_bls) {
  B2_API_BEGIN
  jit_require(n_bls >= 0 && n_bls <= B2_JIT_MAX_BLS, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "at most 4 block-local pads per body");
  jit_require(n_bls == 0 || bls != nullptr, TI_ERROR_ARGUMENT_NULL, "bls");
  return jit_launch_struct_for(k, t, loop_snode, field_places, n_fields, args, n_args, bls, n_bls);
  B2_API_END(-1)
}
static int32_t jit_launch_struct_for(B2JitKernel k, B2Tree t, int32_t loop_snode, const int32_t *field_places, int32_t n_fields, const uint64_t *args,
                                     int32_t n_args, const B2BlsSpec *bls, int32_t n_bls) {
  {
  jit_upload_table(k, t, field_places, n_fields, args, n_args, loop_snode, bls, n_bls);
  jit_require(k->fn != nullptr, TI_ERROR_INVALID_ARGUMENT, "this kernel has no B2_BODY (struct-for) entry");
  t->check_snode(loop_snode, "jit_launch");
  jit_require(t->s[loop_snode].type != B2T_PLACE && loop_snode != 0, TI_ERROR_INVALID_ARGUMENT, "loop SNode must be a non-root container SNode");
  cudaStream_t st = t->rt->stream();
  t->listgen_async(loop_snode);
  B2TreeDev *dev = t->dev;
  int sid = loop_snode;
  const B2JitTable *tp = k->tab_dev;
  void *params[3] = {&dev, &sid, &tp};
  const int64_t nc = std::min<int64_t>(t->s[loop_snode].num_cells, 1024);
  const int block = (int)std::min<int64_t>(256, std::max<int64_t>(32, (nc + 31) / 32 * 32));
  B2_CUDA(cudaLaunchKernel((const void *)k->fn, dim3(t->rt->num_sms * 8), dim3(block), params, k->bls_smem, st));
  t->rt->count_launch();
  return 0;
  }
}

// `for i in range(begin, end)`: grid-stride loop, block 128 (the reference's default block_dim), grid capped at a
// few waves (gpu_parallel_range_for, runtime.cpp:1541-1565; launch shape codegen_cuda.cpp:606-705)
int32_t b2_jit_launch_range(B2JitKernel k, B2Tree t, int64_t begin, int64_t end, const int32_t *field_places, int32_t n_fields, const uint64_t *args, int32_t n_args) {
  B2_API_BEGIN
  jit_upload_table(k, t, field_places, n_fields, args, n_args);
  jit_require(k->fn_range != nullptr, TI_ERROR_INVALID_ARGUMENT, "this kernel has no B2_RANGE_BODY (range-for) entry");
  if (end <= begin) return 0;
  B2TreeDev *dev = t->dev;
  long long b = begin, e = end;
  const B2JitTable *tp = k->tab_dev;
  void *params[4] = {&dev, &b, &e, &tp};
  const int block = 128;
  const int grid = (int)std::min<int64_t>((end - begin + block - 1) / block, (int64_t)t->rt->num_sms * 16);
  B2_CUDA(cudaLaunchKernel((const void *)k->fn_range, dim3(grid), dim3(block), params, 0, t->rt->stream()));
  t->rt->count_launch();
  return 0;
  B2_API_END(-1)
}
}  // extern "C"
