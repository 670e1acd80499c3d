// b2_capi.cu -- the Taichi C-API subset (include/b200mpm.h Part 1) on top of the B200 runtime.
// Mirrors c_api/src/taichi_core_impl.cpp (entry points, error codes, argument checks) and
// c_api/src/taichi_llvm_impl.cpp (LLVM-arch behaviour: ti_flush no-op, ti_wait = stream sync,
// CUDA interop).  AOT modules are tables of prebuilt kernels instead of JIT-loaded PTX.
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>

#include "b2_host.h"
#include "b2_mpm_simple.h"

struct TiAotModule_t {
  TiRuntime_t *rt;
  int model;
  B2SimpleParams params;
  std::map<std::string, TiKernel_t *> kernels;
  ~TiAotModule_t();
};
struct TiKernel_t {
  TiAotModule_t *module;
  int kernel;
  std::string name;
};
TiAotModule_t::~TiAotModule_t() {
  for (auto &kv : kernels) delete kv.second;
}

static void require(bool ok, TiError code, const char *msg) {
  if (!ok) throw B2Error(code, msg);
}

extern "C" {

uint32_t ti_get_version(void) { return TI_C_API_VERSION; }

void ti_get_available_archs(uint32_t *arch_count, TiArch *archs) {
  if (!arch_count) return;
  int n = 0;
  bool have = cudaGetDeviceCount(&n) == cudaSuccess && n > 0;
  if (!have) cudaGetLastError();
  uint32_t avail = have ? 1 : 0;
  if (archs && *arch_count >= 1 && have) archs[0] = TI_ARCH_CUDA;
  *arch_count = avail;
}

TiRuntime ti_create_runtime(TiArch arch, uint32_t device_index) {
  B2_API_BEGIN
  // the reference returns NOT_SUPPORTED for archs that were not compiled in (taichi_core_impl.cpp:270-330)
  require(arch == TI_ARCH_CUDA, TI_ERROR_NOT_SUPPORTED, "libb200mpm only implements TI_ARCH_CUDA");
  return new TiRuntime_t((int)device_index);
  B2_API_END(TI_NULL_HANDLE)
}
void ti_destroy_runtime(TiRuntime runtime) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  delete runtime;
  B2_API_END_VOID
}

TiMemory ti_allocate_memory(TiRuntime runtime, const TiMemoryAllocateInfo *info) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(info != nullptr, TI_ERROR_ARGUMENT_NULL, "allocate_info");
  require(info->size > 0, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "allocate_info->size");
  B2_CUDA(cudaSetDevice(runtime->device));
  TiMemory_t *m = new TiMemory_t();
  m->size = info->size;
  m->host_access = info->host_read || info->host_write;
  cudaError_t e;
  if (m->host_access) {
    e = cudaMallocHost(&m->dev, info->size);  // pinned, device-visible through UVA
    if (e == cudaSuccess) memset(m->dev, 0, info->size);
  } else {
    e = cudaMalloc(&m->dev, info->size);
    if (e == cudaSuccess) e = cudaMemsetAsync(m->dev, 0, info->size, runtime->stream());  // zero-filled: cuda_device.cpp:40,64-65
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    delete m;
    throw B2Error(TI_ERROR_OUT_OF_MEMORY, "device allocation failed");
  }
  runtime->memories.push_back(m);
  return m;
  B2_API_END(TI_NULL_HANDLE)
}
void ti_free_memory(TiRuntime runtime, TiMemory memory) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(memory != nullptr, TI_ERROR_ARGUMENT_NULL, "memory");
  for (auto &m : runtime->memories)
    if (m == memory) {
      m = nullptr;
      runtime->sync();
      if (!memory->imported && memory->dev) {
        if (memory->host_access)
          cudaFreeHost(memory->dev);
        else
          cudaFree(memory->dev);
      }
      if (memory->staging) cudaFreeHost(memory->staging);
      delete memory;
      return;
    }
  throw B2Error(TI_ERROR_INVALID_ARGUMENT, "memory does not belong to this runtime");
  B2_API_END_VOID
}
void *ti_map_memory(TiRuntime runtime, TiMemory memory) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(memory != nullptr, TI_ERROR_ARGUMENT_NULL, "memory");
  runtime->sync();
  if (memory->host_access) return memory->dev;
  require(memory->staging == nullptr, TI_ERROR_INVALID_STATE, "memory is already mapped");
  B2_CUDA(cudaMallocHost(&memory->staging, memory->size));
  B2_CUDA(cudaMemcpy(memory->staging, memory->dev, memory->size, cudaMemcpyDeviceToHost));
  return memory->staging;
  B2_API_END(nullptr)
}
void ti_unmap_memory(TiRuntime runtime, TiMemory memory) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(memory != nullptr, TI_ERROR_ARGUMENT_NULL, "memory");
  if (memory->host_access) return;
  require(memory->staging != nullptr, TI_ERROR_INVALID_STATE, "memory is not mapped");
  B2_CUDA(cudaMemcpy(memory->dev, memory->staging, memory->size, cudaMemcpyHostToDevice));
  cudaFreeHost(memory->staging);
  memory->staging = nullptr;
  B2_API_END_VOID
}
void ti_copy_memory_device_to_device(TiRuntime runtime, const TiMemorySlice *dst, const TiMemorySlice *src) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(dst != nullptr && dst->memory != nullptr, TI_ERROR_ARGUMENT_NULL, "dst_memory");
  require(src != nullptr && src->memory != nullptr, TI_ERROR_ARGUMENT_NULL, "src_memory");
  require(dst->size == src->size, TI_ERROR_INVALID_ARGUMENT, "dst_memory->size != src_memory->size");
  require(dst->offset + dst->size <= dst->memory->size && src->offset + src->size <= src->memory->size, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "slice out of range");
  B2_CUDA(cudaMemcpyAsync((char *)dst->memory->dev + dst->offset, (char *)src->memory->dev + src->offset, src->size, cudaMemcpyDefault, runtime->stream()));
  B2_API_END_VOID
}
// Capabilities are a SPIR-V notion: the reference's LLVM runtimes report none (its CUDA runtime inherits the empty
// default of taichi_core_impl.cpp); the ext setter is accepted and has nothing to override.
void ti_set_runtime_capabilities_ext(TiRuntime runtime, uint32_t, const TiCapabilityLevelInfo *) {
  B2_API_BEGIN
  if (!runtime) throw B2Error(TI_ERROR_ARGUMENT_NULL, "runtime");
  B2_API_END_VOID
}
void ti_get_runtime_capabilities(TiRuntime runtime, uint32_t *capability_count, TiCapabilityLevelInfo *) {
  B2_API_BEGIN
  if (!runtime) throw B2Error(TI_ERROR_ARGUMENT_NULL, "runtime");
  if (!capability_count) throw B2Error(TI_ERROR_ARGUMENT_NULL, "capability_count");
  *capability_count = 0;
  B2_API_END_VOID
}
// TCM archives hold SPIR-V / LLVM modules of the reference's compiler: not loadable here (kernels are prebuilt tables or
// NVRTC bodies).
TiAotModule ti_create_aot_module(TiRuntime runtime, const void *tcm, uint64_t) {
  B2_API_BEGIN
  if (!runtime) throw B2Error(TI_ERROR_ARGUMENT_NULL, "runtime");
  if (!tcm) throw B2Error(TI_ERROR_ARGUMENT_NULL, "tcm");
  throw B2Error(TI_ERROR_NOT_SUPPORTED, "ti_create_aot_module: TCM modules are not supported (use ti_load_aot_module(\"builtin:...\"))");
  B2_API_END(nullptr)
}
// graphics: unavailable on this box; same answer as the reference gives on LLVM archs (taichi_core_impl.cpp:515-519)
static void not_supported(const char *what) { ti_set_last_error(TI_ERROR_NOT_SUPPORTED, what); }
TiImage ti_allocate_image(TiRuntime, const TiImageAllocateInfo *) {
  not_supported("ti_allocate_image");
  return nullptr;
}
void ti_free_image(TiRuntime, TiImage) { not_supported("ti_free_image"); }
TiSampler ti_create_sampler(TiRuntime, const TiSamplerCreateInfo *) {
  not_supported("ti_create_sampler");
  return nullptr;
}
void ti_destroy_sampler(TiRuntime, TiSampler) { not_supported("ti_destroy_sampler"); }
void ti_copy_image_device_to_device(TiRuntime, const TiImageSlice *, const TiImageSlice *) { not_supported("ti_copy_image_device_to_device"); }
void ti_track_image_ext(TiRuntime, TiImage, uint32_t) { not_supported("ti_track_image_ext"); }
void ti_transition_image(TiRuntime, TiImage, uint32_t) { not_supported("ti_transition_image"); }

void ti_flush(TiRuntime runtime) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");  // no-op on LLVM archs (taichi_llvm_impl.cpp:129-132)
  B2_API_END_VOID
}
void ti_wait(TiRuntime runtime) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  runtime->sync();
  std::string msg;
  if (runtime->check_error(&msg)) throw B2Error(TI_ERROR_INVALID_STATE, msg);  // check_runtime_error
  B2_API_END_VOID
}

void ti_export_cuda_memory(TiRuntime runtime, TiMemory memory, TiCudaMemoryInteropInfo *info) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(memory != nullptr, TI_ERROR_ARGUMENT_NULL, "memory");
  require(info != nullptr, TI_ERROR_ARGUMENT_NULL, "interop_info");
  info->ptr = memory->dev;
  info->size = memory->size;
  B2_API_END_VOID
}
TiMemory ti_import_cuda_memory(TiRuntime runtime, void *ptr, size_t memory_size) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(ptr != nullptr, TI_ERROR_ARGUMENT_NULL, "ptr");
  TiMemory_t *m = new TiMemory_t();
  m->dev = ptr;
  m->size = memory_size;
  m->imported = true;
  runtime->memories.push_back(m);
  return m;
  B2_API_END(TI_NULL_HANDLE)
}

// ---------------------------------------------------------------------------- AOT modules
static float query_param(const std::string &q, const char *key, float dflt) {
  size_t pos = 0;
  std::string k = std::string(key) + "=";
  while (pos < q.size()) {
    size_t end = q.find('&', pos);
    if (end == std::string::npos) end = q.size();
    if (q.compare(pos, k.size(), k) == 0) return (float)atof(q.substr(pos + k.size(), end - pos - k.size()).c_str());
    pos = end + 1;
  }
  return dflt;
}
TiAotModule ti_load_aot_module(TiRuntime runtime, const char *module_path) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(module_path != nullptr, TI_ERROR_ARGUMENT_NULL, "module_path");
  std::string path(module_path), query;
  size_t qm = path.find('?');
  if (qm != std::string::npos) {
    query = path.substr(qm + 1);
    path = path.substr(0, qm);
  }
  while (!path.empty() && path.back() == '/') path.pop_back();
  size_t sl = path.find_last_of("/:");
  std::string base = sl == std::string::npos ? path : path.substr(sl + 1);
  int model = -1;
  if (base.find("mpm128") != std::string::npos)
    model = B2_MODEL_MPM128;
  else if (base.find("mpm88") != std::string::npos)
    model = B2_MODEL_MPM88;
  else if (base.find("mpm3d") != std::string::npos)
    model = B2_MODEL_MPM3D;
  // the reference reports a module it cannot open as TI_ERROR_CORRUPTED_DATA (taichi_llvm_impl.cpp:78-121)
  require(model >= 0, TI_ERROR_CORRUPTED_DATA, "no prebuilt module of that name (builtin:mpm88|mpm3d|mpm128)");
  TiAotModule_t *m = new TiAotModule_t();
  m->rt = runtime;
  m->model = model;
  B2SimpleParams &P = m->params;
  memset(&P, 0, sizeof(P));
  P.n_grid = 0;  // taken from the grid ndarray at launch
  P.p_rho = 1.f;
  P.gravity = 9.8f;
  P.bound = 3;
  if (model == B2_MODEL_MPM88) {
    P.dt = 2e-4f, P.E = 400.f;  // mpm88_graph_aot.py:18-25
  } else if (model == B2_MODEL_MPM3D) {
    P.dt = 4e-4f, P.E = 400.f;  // mpm3d.py:11-23 (n_grid 32 default; 8e-5 at 128)
  } else {
    P.dt = 1e-4f, P.E = 5e3f, P.nu = 0.2f, P.gx = 0.f, P.gy = -1.f;  // mpm128.py:5-11,119
  }
  P.dt = query_param(query, "dt", P.dt);
  P.E = query_param(query, "E", P.E);
  P.nu = query_param(query, "nu", P.nu);
  P.p_rho = query_param(query, "p_rho", P.p_rho);
  P.gravity = query_param(query, "gravity", P.gravity);
  P.bound = (int)query_param(query, "bound", (float)P.bound);
  P.gx = query_param(query, "gx", P.gx);
  P.gy = query_param(query, "gy", P.gy);
  P.ax = query_param(query, "ax", 0.f);
  P.ay = query_param(query, "ay", 0.f);
  P.astrength = query_param(query, "attractor", 0.f);
  static const struct {
    const char *name;
    int k;
  } names[] = {{"init_particles", B2_K_INIT}, {"substep_reset_grid", B2_K_RESET}, {"substep_p2g", B2_K_P2G},
               {"substep_update_grid_v", B2_K_GRIDOP}, {"substep_g2p", B2_K_G2P}};
  for (auto &nk : names) m->kernels[nk.name] = new TiKernel_t{m, nk.k, nk.name};
  return m;
  B2_API_END(TI_NULL_HANDLE)
}
void ti_destroy_aot_module(TiAotModule aot_module) {
  B2_API_BEGIN
  require(aot_module != nullptr, TI_ERROR_ARGUMENT_NULL, "aot_module");
  delete aot_module;
  B2_API_END_VOID
}
TiKernel ti_get_aot_module_kernel(TiAotModule aot_module, const char *name) {
  B2_API_BEGIN
  require(aot_module != nullptr, TI_ERROR_ARGUMENT_NULL, "aot_module");
  require(name != nullptr, TI_ERROR_ARGUMENT_NULL, "name");
  auto it = aot_module->kernels.find(name);
  require(it != aot_module->kernels.end(), TI_ERROR_NAME_NOT_FOUND, name);
  return it->second;
  B2_API_END(TI_NULL_HANDLE)
}
TiComputeGraph ti_get_aot_module_compute_graph(TiAotModule aot_module, const char *name) {
  B2_API_BEGIN
  require(aot_module != nullptr, TI_ERROR_ARGUMENT_NULL, "aot_module");
  require(name != nullptr, TI_ERROR_ARGUMENT_NULL, "name");
  throw B2Error(TI_ERROR_NAME_NOT_FOUND, name);
  B2_API_END(TI_NULL_HANDLE)
}
void ti_launch_compute_graph(TiRuntime runtime, TiComputeGraph graph, uint32_t, const TiNamedArgument *) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(graph != nullptr, TI_ERROR_ARGUMENT_NULL, "compute_graph");
  throw B2Error(TI_ERROR_NOT_SUPPORTED, "compute graphs");
  B2_API_END_VOID
}

static float *nd_f32(const TiArgument &a, const char *what) {
  if (a.type != TI_ARGUMENT_TYPE_NDARRAY) throw B2Error(TI_ERROR_INVALID_ARGUMENT, std::string(what) + ": expected an ndarray argument");
  const TiNdArray &nd = a.value.ndarray;
  if (nd.memory == nullptr) throw B2Error(TI_ERROR_ARGUMENT_NULL, std::string(what) + ".memory");  // taichi_core_impl.cpp:770-787
  if (nd.shape.dim_count > 16 || nd.elem_shape.dim_count > 16) throw B2Error(TI_ERROR_ARGUMENT_OUT_OF_RANGE, std::string(what) + ".shape");
  return (float *)nd.memory->dev;
}
// ndarray argument with `elems` 4-byte elements of type `want` at least (every kernel argument goes through this: an undersized or
// mistyped TiNdArray is TI_ERROR_ARGUMENT_OUT_OF_RANGE / INVALID_ARGUMENT, never an out-of-bounds device access)
static void check_bytes(const TiArgument &a, uint64_t elems, const char *what);
static float *nd_arg(const TiArgument &a, uint64_t elems, const char *what, TiDataType want = TI_DATA_TYPE_F32) {
  float *p = nd_f32(a, what);
  if (a.value.ndarray.elem_type != want) throw B2Error(TI_ERROR_INVALID_ARGUMENT, std::string(what) + (want == TI_DATA_TYPE_I32 ? ": expected an i32 ndarray" : ": expected an f32 ndarray"));
  check_bytes(a, elems, what);
  return p;
}
static void check_bytes(const TiArgument &a, uint64_t elems, const char *what) {
  const TiNdArray &nd = a.value.ndarray;
  uint64_t have = 1;
  for (uint32_t i = 0; i < nd.shape.dim_count; i++) have *= nd.shape.dims[i];
  for (uint32_t i = 0; i < nd.elem_shape.dim_count; i++) have *= nd.elem_shape.dims[i];
  if (have < elems || nd.memory->size < elems * 4) throw B2Error(TI_ERROR_ARGUMENT_OUT_OF_RANGE, std::string(what) + ": ndarray too small");
}

void ti_launch_kernel(TiRuntime runtime, TiKernel kernel, uint32_t arg_count, const TiArgument *args) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(kernel != nullptr, TI_ERROR_ARGUMENT_NULL, "kernel");
  require(args != nullptr || arg_count == 0, TI_ERROR_ARGUMENT_NULL, "args");
  TiAotModule_t *mod = kernel->module;
  B2_CUDA(cudaSetDevice(runtime->device));
  const int model = mod->model, D = model == B2_MODEL_MPM3D ? 3 : 2;
  B2SimpleParams P = mod->params;
  B2SimpleArgs a;
  auto need = [&](uint32_t n) { require(arg_count >= n, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "too few kernel arguments"); };
  auto grid = [&](uint32_t iv, uint32_t im) {
    a.grid_v = nd_f32(args[iv], "grid_v");
    const TiNdArray &g = args[iv].value.ndarray;
    require(g.shape.dim_count == (uint32_t)D, TI_ERROR_INVALID_ARGUMENT, "grid_v: wrong number of dimensions");
    P.n_grid = (int)g.shape.dims[0];
    uint64_t cells = 1;
    for (int d = 0; d < D; d++) {
      require(g.shape.dims[d] == g.shape.dims[0], TI_ERROR_INVALID_ARGUMENT, "grid must be cubic");
      cells *= g.shape.dims[d];
    }
    check_bytes(args[iv], cells * D, "grid_v");
    if (im != ~0u) {
      a.grid_m = nd_f32(args[im], "grid_m");
      check_bytes(args[im], cells, "grid_m");
    }
  };
  auto particles = [&](uint32_t ix) {
    a.x = nd_f32(args[ix], "x");
    const TiNdArray &x = args[ix].value.ndarray;
    require(x.shape.dim_count == 1, TI_ERROR_INVALID_ARGUMENT, "x must be 1-D");
    a.n_particles = x.shape.dims[0];
    check_bytes(args[ix], (uint64_t)a.n_particles * D, "x");
  };
  switch (kernel->kernel) {
    case B2_K_INIT:
      need(3);
      particles(0);
      a.v = nd_arg(args[1], (uint64_t)a.n_particles * D, "v");
      a.J = nd_arg(args[2], (uint64_t)a.n_particles, "J");
      P.n_grid = 1;
      break;
    case B2_K_RESET:
    case B2_K_GRIDOP:
      need(2);
      grid(0, 1);
      if (model == B2_MODEL_MPM128 && kernel->kernel == B2_K_GRIDOP && arg_count >= 7) {
        // optional scalars: gravity.x, gravity.y, attractor.x, attractor.y, attractor_strength
        float *dst[5] = {&P.gx, &P.gy, &P.ax, &P.ay, &P.astrength};
        for (int i = 0; i < 5; i++) {
          require(args[2 + i].type == TI_ARGUMENT_TYPE_F32, TI_ERROR_INVALID_ARGUMENT, "expected f32 scalar");
          *dst[i] = args[2 + i].value.f32;
        }
      }
      break;
    case B2_K_P2G:
      if (model == B2_MODEL_MPM128) {
        need(8);
        particles(0);
        const uint64_t np_ = a.n_particles;
        a.v = nd_arg(args[1], np_ * D, "v"), a.C = nd_arg(args[2], np_ * D * D, "C"), a.F = nd_arg(args[3], np_ * D * D, "F");
        a.material = (int *)nd_arg(args[4], np_, "material", TI_DATA_TYPE_I32), a.Jp = nd_arg(args[5], np_, "Jp");
        grid(6, 7);
      } else {
        need(6);
        particles(0);
        const uint64_t np_ = a.n_particles;
        a.v = nd_arg(args[1], np_ * D, "v"), a.C = nd_arg(args[2], np_ * D * D, "C"), a.J = nd_arg(args[3], np_, "J");
        grid(4, 5);
      }
      break;
    case B2_K_G2P:
      if (model == B2_MODEL_MPM128) {
        need(4);
        particles(0);
        a.v = nd_arg(args[1], (uint64_t)a.n_particles * D, "v"), a.C = nd_arg(args[2], (uint64_t)a.n_particles * D * D, "C");
        grid(3, ~0u);
      } else {
        need(5);
        particles(0);
        const uint64_t np_ = a.n_particles;
        a.v = nd_arg(args[1], np_ * D, "v"), a.C = nd_arg(args[2], np_ * D * D, "C"), a.J = nd_arg(args[3], np_, "J");
        grid(4, ~0u);
        if (arg_count >= 6) a.pos3 = nd_arg(args[5], np_ * 3, "pos");
      }
      break;
    default: throw B2Error(TI_ERROR_INVALID_ARGUMENT, "bad kernel handle");
  }
  b2_simple_launch(runtime, model, kernel->kernel, P, a);
  B2_API_END_VOID
}

}  // extern "C"
