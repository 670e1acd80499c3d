/*
 * include/b200mpm.h -- C ABI of libb200mpm.so: the B200-native replacement for Taichi's
 * sparse-SNode struct-for / MLS-MPM hot path (arch = cuda, sm_100a only).
 *
 * Part 1 (ti_*) keeps the names, argument meaning, struct layouts and error convention of
 * the reference C-API subset this path uses, so a client compiled against
 *   /root/reference/c_api/include/taichi/taichi_core.h   and   taichi_cuda.h
 * links against this library unchanged (reference line numbers given per entry).
 * Part 2 (b2_*) exports what the reference only offers through pybind
 * (taichi/python/export_lang.cpp:498-529 SNode builders; Program::compile_kernel /
 * launch_kernel): SNode-tree building, activation, listgen, struct-for built-ins, GC, and
 * the prebuilt MLS-MPM kernels.
 *
 * Conventions (same as the reference, taichi_core_impl.h:80-92, taichi_core_impl.cpp:63-70):
 * no exception or error code crosses the ABI; failing calls record (TiError, message) in a
 * thread-local slot read with ti_get_last_error and return TI_NULL_HANDLE / void / -1.
 * All launches are asynchronous and in-order on ONE CUDA stream (ti_set_cuda_stream
 * replaces it); ti_wait / b2_sync synchronise.  Pointers named *_any may be host or device
 * pointers (detected like runtime/cuda/kernel_launcher.cpp:7-13 does); host data is staged.
 * There is no CPU fallback: without a CUDA device ti_create_runtime fails.
 */
#ifndef B200MPM_H
#define B200MPM_H

#include <stddef.h>
#include <stdint.h>

#if defined(_WIN32)
#define TI_DLL_EXPORT __declspec(dllexport)
#define TI_API_CALL __stdcall
#else
#define TI_DLL_EXPORT __attribute__((visibility("default")))
#define TI_API_CALL
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ===================================================================================== */
/* Part 1: Taichi C-API subset (taichi_core.h / taichi_cuda.h)                            */
/* Skipped when the reference's own <taichi/taichi.h> (built with -DTI_WITH_CUDA) was     */
/* included first: Part 2 then compiles against the reference's declarations, which is    */
/* how tests/test_cabi_reference.py proves that the two agree.                            */
/* ===================================================================================== */
#ifndef TAICHI_H
#ifndef TI_C_API_VERSION
#define TI_C_API_VERSION 1007000 /* taichi_core.h:229 */
#endif

typedef uint32_t TiBool;  /* taichi_core.h:246 */
typedef uint32_t TiFlags; /* :266 */
#define TI_FALSE 0
#define TI_TRUE 1
#define TI_NULL_HANDLE 0

typedef struct TiRuntime_t *TiRuntime;           /* :280 */
typedef struct TiAotModule_t *TiAotModule;       /* :286 */
typedef struct TiMemory_t *TiMemory;             /* :291 */
typedef struct TiImage_t *TiImage;               /* :296 (graphics: not supported) */
typedef struct TiSampler_t *TiSampler;           /* :303 (graphics: not supported) */
typedef struct TiKernel_t *TiKernel;             /* :308 */
typedef struct TiComputeGraph_t *TiComputeGraph; /* :314 */

typedef enum TiError { /* :319-353 */
  TI_ERROR_SUCCESS = 0,
  TI_ERROR_NOT_SUPPORTED = -1,
  TI_ERROR_CORRUPTED_DATA = -2,
  TI_ERROR_NAME_NOT_FOUND = -3,
  TI_ERROR_INVALID_ARGUMENT = -4,
  TI_ERROR_ARGUMENT_NULL = -5,
  TI_ERROR_ARGUMENT_OUT_OF_RANGE = -6,
  TI_ERROR_ARGUMENT_NOT_FOUND = -7,
  TI_ERROR_INVALID_INTEROP = -8,
  TI_ERROR_INVALID_STATE = -9,
  TI_ERROR_INCOMPATIBLE_MODULE = -10,
  TI_ERROR_OUT_OF_MEMORY = -11,
  TI_ERROR_MAX_ENUM = 0xffffffff,
} TiError;

typedef enum TiArch { /* :358-375 */
  TI_ARCH_RESERVED = 0,
  TI_ARCH_VULKAN = 1,
  TI_ARCH_METAL = 2,
  TI_ARCH_CUDA = 3,
  TI_ARCH_X64 = 4,
  TI_ARCH_ARM64 = 5,
  TI_ARCH_OPENGL = 6,
  TI_ARCH_GLES = 7,
  TI_ARCH_MAX_ENUM = 0xffffffff,
} TiArch;

typedef enum TiDataType { /* :423-450 */
  TI_DATA_TYPE_F16 = 0,
  TI_DATA_TYPE_F32 = 1,
  TI_DATA_TYPE_F64 = 2,
  TI_DATA_TYPE_I8 = 3,
  TI_DATA_TYPE_I16 = 4,
  TI_DATA_TYPE_I32 = 5,
  TI_DATA_TYPE_I64 = 6,
  TI_DATA_TYPE_U1 = 7,
  TI_DATA_TYPE_U8 = 8,
  TI_DATA_TYPE_U16 = 9,
  TI_DATA_TYPE_U32 = 10,
  TI_DATA_TYPE_U64 = 11,
  TI_DATA_TYPE_GEN = 12,
  TI_DATA_TYPE_UNKNOWN = 13,
  TI_DATA_TYPE_MAX_ENUM = 0xffffffff,
} TiDataType;

typedef enum TiArgumentType { /* :455-469 */
  TI_ARGUMENT_TYPE_I32 = 0,
  TI_ARGUMENT_TYPE_F32 = 1,
  TI_ARGUMENT_TYPE_NDARRAY = 2,
  TI_ARGUMENT_TYPE_TEXTURE = 3,
  TI_ARGUMENT_TYPE_SCALAR = 4,
  TI_ARGUMENT_TYPE_TENSOR = 5,
  TI_ARGUMENT_TYPE_MAX_ENUM = 0xffffffff,
} TiArgumentType;

typedef enum TiMemoryUsageFlagBits { /* :475-484 */
  TI_MEMORY_USAGE_STORAGE_BIT = 1 << 0,
  TI_MEMORY_USAGE_UNIFORM_BIT = 1 << 1,
  TI_MEMORY_USAGE_VERTEX_BIT = 1 << 2,
  TI_MEMORY_USAGE_INDEX_BIT = 1 << 3,
} TiMemoryUsageFlagBits;
typedef TiFlags TiMemoryUsageFlags;

typedef struct TiMemoryAllocateInfo { /* :490-503 */
  uint64_t size;
  TiBool host_write;
  TiBool host_read;
  TiBool export_sharing;
  TiMemoryUsageFlags usage;
} TiMemoryAllocateInfo;

typedef struct TiMemorySlice { /* :509-516 */
  TiMemory memory;
  uint64_t offset;
  uint64_t size;
} TiMemorySlice;

typedef struct TiNdShape { /* :522-527 */
  uint32_t dim_count;
  uint32_t dims[16];
} TiNdShape;

typedef struct TiNdArray { /* :532-542 */
  TiMemory memory;
  TiNdShape shape;
  TiNdShape elem_shape;
  TiDataType elem_type;
} TiNdArray;

/* Texture plumbing exists only so that sizeof(TiArgument) matches the reference. */
typedef struct TiImageExtent {
  uint32_t width, height, depth, array_layer_count;
} TiImageExtent;
typedef struct TiTexture { /* :765-777 */
  TiImage image;
  TiSampler sampler;
  uint32_t dimension; /* TiImageDimension */
  TiImageExtent extent;
  uint32_t format; /* TiFormat */
} TiTexture;

typedef union TiScalarValue { /* :788-797 */
  uint8_t x8;
  uint16_t x16;
  uint32_t x32;
  uint64_t x64;
} TiScalarValue;
typedef struct TiScalar {
  TiDataType type;
  TiScalarValue value;
} TiScalar;
typedef union TiTensorValue { /* :810-819 */
  uint8_t x8[128];
  uint16_t x16[64];
  uint32_t x32[32];
  uint64_t x64[16];
} TiTensorValue;
typedef struct TiTensorValueWithLength {
  uint32_t length;
  TiTensorValue data;
} TiTensorValueWithLength;
typedef struct TiTensor {
  TiDataType type;
  TiTensorValueWithLength contents;
} TiTensor;

typedef union TiArgumentValue { /* :840-855 */
  int32_t i32;
  float f32;
  TiNdArray ndarray;
  TiTexture texture;
  TiScalar scalar;
  TiTensor tensor;
} TiArgumentValue;
typedef struct TiArgument { /* :860-865 */
  TiArgumentType type;
  TiArgumentValue value;
} TiArgument;
typedef struct TiNamedArgument { /* :870-875 */
  const char *name;
  TiArgument argument;
} TiNamedArgument;

TI_DLL_EXPORT uint32_t TI_API_CALL ti_get_version(void);                                     /* :881 */
TI_DLL_EXPORT void TI_API_CALL ti_get_available_archs(uint32_t *arch_count, TiArch *archs);  /* :898 */
TI_DLL_EXPORT TiError TI_API_CALL ti_get_last_error(uint64_t *message_size, char *message);  /* :905 */
TI_DLL_EXPORT void TI_API_CALL ti_set_last_error(TiError error, const char *message);        /* :917 */
/* arch must be TI_ARCH_CUDA; device_index selects the CUDA device of THIS process (the
   reference rejects != 0, taichi_core_impl.cpp:270-297; the multi-GPU layer runs one process
   per GPU and passes LOCAL_RANK). */
TI_DLL_EXPORT TiRuntime TI_API_CALL ti_create_runtime(TiArch arch, uint32_t device_index);   /* :927 */
TI_DLL_EXPORT void TI_API_CALL ti_destroy_runtime(TiRuntime runtime);                        /* :937 */
TI_DLL_EXPORT TiMemory TI_API_CALL ti_allocate_memory(TiRuntime runtime,
                                                      const TiMemoryAllocateInfo *allocate_info); /* :960 */
TI_DLL_EXPORT void TI_API_CALL ti_free_memory(TiRuntime runtime, TiMemory memory);           /* :967 */
/* Returns a host pointer: the allocation itself for host_read/host_write memory (pinned,
   device-visible), else a staging copy written back by ti_unmap_memory. */
TI_DLL_EXPORT void *TI_API_CALL ti_map_memory(TiRuntime runtime, TiMemory memory);           /* :974 */
TI_DLL_EXPORT void TI_API_CALL ti_unmap_memory(TiRuntime runtime, TiMemory memory);          /* :982 */
TI_DLL_EXPORT void TI_API_CALL ti_copy_memory_device_to_device(TiRuntime runtime,
                                                               const TiMemorySlice *dst_memory,
                                                               const TiMemorySlice *src_memory); /* :1008 */
TI_DLL_EXPORT void TI_API_CALL ti_launch_kernel(TiRuntime runtime, TiKernel kernel,
                                                uint32_t arg_count, const TiArgument *args);  /* :1044 */
TI_DLL_EXPORT void TI_API_CALL ti_launch_compute_graph(TiRuntime runtime, TiComputeGraph graph,
                                                       uint32_t arg_count,
                                                       const TiNamedArgument *args);         /* :1053 */
TI_DLL_EXPORT void TI_API_CALL ti_flush(TiRuntime runtime);                                  /* :1063 */
TI_DLL_EXPORT void TI_API_CALL ti_wait(TiRuntime runtime);                                   /* :1069 */
/* module_path: "builtin:mpm88" | "builtin:mpm3d" | "builtin:mpm128" (prebuilt kernel tables
   named as tests/cpp/aot/python_scripts/mpm88_graph_aot.py:29-108), or a directory whose
   name ends in one of those (so TAICHI_AOT_FOLDER_PATH-style callers keep working). */
TI_DLL_EXPORT TiAotModule TI_API_CALL ti_load_aot_module(TiRuntime runtime, const char *module_path); /* :1076 */
TI_DLL_EXPORT void TI_API_CALL ti_destroy_aot_module(TiAotModule aot_module);                /* :1091 */
TI_DLL_EXPORT TiKernel TI_API_CALL ti_get_aot_module_kernel(TiAotModule aot_module, const char *name); /* :1098 */
TI_DLL_EXPORT TiComputeGraph TI_API_CALL ti_get_aot_module_compute_graph(TiAotModule aot_module,
                                                                         const char *name);  /* :1106 */

/* taichi_cuda.h:12-32 */
typedef struct TiCudaMemoryInteropInfo {
  void *ptr;
  uint64_t size;
} TiCudaMemoryInteropInfo;
TI_DLL_EXPORT void TI_API_CALL ti_export_cuda_memory(TiRuntime runtime, TiMemory memory,
                                                     TiCudaMemoryInteropInfo *interop_info);
/* Borrowed: never freed by the runtime (rhi/cuda/cuda_device.cpp:106,189-201). */
TI_DLL_EXPORT TiMemory TI_API_CALL ti_import_cuda_memory(TiRuntime runtime, void *ptr, size_t memory_size);
TI_DLL_EXPORT void TI_API_CALL ti_set_cuda_stream(void *stream);
TI_DLL_EXPORT void TI_API_CALL ti_get_cuda_stream(void **stream);

/* Entry points the reference's header-only C++ wrapper (c_api/include/taichi/cpp/taichi.hpp) refers to.
   Capabilities: LLVM archs report none (the reference fills them for SPIR-V devices only); the ext setter is accepted and
   ignored.  ti_create_aot_module takes TCM bytes (:1084): there is no TCM loader here -> TI_ERROR_NOT_SUPPORTED.
   Image / sampler entry points are graphics: TI_ERROR_NOT_SUPPORTED, as the reference answers on LLVM archs
   (taichi_core_impl.cpp:515-519); their argument structs stay opaque. */
typedef enum TiCapability { TI_CAPABILITY_RESERVED = 0, TI_CAPABILITY_MAX_ENUM = 0xffffffff } TiCapability; /* :380-407 */
typedef struct TiCapabilityLevelInfo { /* :413-416 */
  TiCapability capability;
  uint32_t level;
} TiCapabilityLevelInfo;
TI_DLL_EXPORT void TI_API_CALL ti_set_runtime_capabilities_ext(TiRuntime runtime, uint32_t capability_count,
                                                               const TiCapabilityLevelInfo *capabilities); /* :943 */
TI_DLL_EXPORT void TI_API_CALL ti_get_runtime_capabilities(TiRuntime runtime, uint32_t *capability_count,
                                                           TiCapabilityLevelInfo *capabilities);          /* :951 */
TI_DLL_EXPORT TiAotModule TI_API_CALL ti_create_aot_module(TiRuntime runtime, const void *tcm, uint64_t size); /* :1084 */
struct TiImageAllocateInfo;
struct TiSamplerCreateInfo;
struct TiImageSlice;
TI_DLL_EXPORT TiImage TI_API_CALL ti_allocate_image(TiRuntime runtime, const struct TiImageAllocateInfo *allocate_info); /* :989 */
TI_DLL_EXPORT void TI_API_CALL ti_free_image(TiRuntime runtime, TiImage image);                                         /* :995 */
TI_DLL_EXPORT TiSampler TI_API_CALL ti_create_sampler(TiRuntime runtime, const struct TiSamplerCreateInfo *create_info); /* :998 */
TI_DLL_EXPORT void TI_API_CALL ti_destroy_sampler(TiRuntime runtime, TiSampler sampler);
TI_DLL_EXPORT void TI_API_CALL ti_copy_image_device_to_device(TiRuntime runtime, const struct TiImageSlice *dst_image,
                                                              const struct TiImageSlice *src_image);
TI_DLL_EXPORT void TI_API_CALL ti_track_image_ext(TiRuntime runtime, TiImage image, uint32_t layout);
TI_DLL_EXPORT void TI_API_CALL ti_transition_image(TiRuntime runtime, TiImage image, uint32_t layout);
#endif /* TAICHI_H */

/* ===================================================================================== */
/* Part 2: SNode trees, struct-for built-ins, MLS-MPM                                      */
/* ===================================================================================== */
#define B2_MAX_AXES 12 /* taichi/inc/constants.h:5 */

enum { B2_ROOT = 0, B2_DENSE = 1, B2_POINTER = 2, B2_BITMASKED = 3, B2_DYNAMIC = 4, B2_PLACE = 5 };

/* One SNode of a tree, in creation order (index 0 = root).  Mirrors the arguments of
   SNode::dense/pointer/bitmasked/dynamic/place (export_lang.cpp:498-529) and the geometry
   rules of SNode::create_node (taichi/ir/snode.cpp:37-125). */
typedef struct B2SNodeDesc {
  int32_t type;
  int32_t parent; /* index into the array; -1 for the root */
  int32_t num_axes;
  int32_t axes[B2_MAX_AXES];  /* physical axis ids, 0 = i, 1 = j ... */
  int32_t sizes[B2_MAX_AXES]; /* extent along each listed axis */
  int32_t dtype_bytes;        /* place: 4 (f32/i32) or 8 */
  int32_t chunk_size;         /* dynamic */
} B2SNodeDesc;

typedef struct B2Tree_t *B2Tree;
typedef struct B2Mpm_t *B2Mpm;
typedef struct B2JitKernel_t *B2JitKernel;
typedef struct B2Comm_t *B2Comm; /* NCCL communicator of the slab layer */
typedef struct B2Slab_t *B2Slab; /* one rank's slab driver */

/* Program::add_snode_tree + materialize_snode_tree (program.cpp:238-255,
   llvm_runtime_executor.cpp:391-469): layout, zeroed root buffer, one element list per SNode,
   one NodeManager per pointer/dynamic SNode. */
TI_DLL_EXPORT B2Tree b2_create_tree(TiRuntime runtime, const B2SNodeDesc *desc, int32_t n);
TI_DLL_EXPORT void b2_destroy_tree(B2Tree tree);
TI_DLL_EXPORT int64_t b2_snode_num_cells(B2Tree tree, int32_t snode); /* num_cells_per_container */
TI_DLL_EXPORT int64_t b2_snode_cell_size(B2Tree tree, int32_t snode); /* SNode._cell_size_bytes  */
TI_DLL_EXPORT int64_t b2_snode_node_size(B2Tree tree, int32_t snode); /* bytes of one container  */

/* coords: n x 12 int32 (global indices by physical axis), host or device. */
TI_DLL_EXPORT int32_t b2_activate(B2Tree tree, int32_t snode, const int32_t *coords_any, int64_t n);   /* ti.activate, Pointer/Bitmasked/Dynamic_activate */
TI_DLL_EXPORT int32_t b2_deactivate(B2Tree tree, int32_t snode, const int32_t *coords_any, int64_t n); /* ti.deactivate (+gc task, offload.cpp:671-695) */
TI_DLL_EXPORT int32_t b2_is_active(B2Tree tree, int32_t snode, const int32_t *coords_any, int64_t n, uint8_t *out_any); /* ti.is_active */
/* clear_list + element_listgen_root/nonroot for every level root..snode (offload.cpp:204-222);
   returns the number of Elements of `snode`'s list (synchronises). */
TI_DLL_EXPORT int64_t b2_listgen(B2Tree tree, int32_t snode);
/* Same, asynchronous (no read-back). */
TI_DLL_EXPORT int32_t b2_listgen_async(B2Tree tree, int32_t snode);
/* Copies up to cap Elements as 14 x int32 {pcoord[12], loop_lo, loop_hi}; returns the count. */
TI_DLL_EXPORT int64_t b2_get_list(B2Tree tree, int32_t snode, int32_t *out_any, int64_t cap);
/* struct-for over `snode` (listgen + active-cell filter): writes the visited coordinates
   (12 x int32 each) and returns how many cells were visited.  out may be NULL (count only). */
TI_DLL_EXPORT int64_t b2_struct_for_coords(B2Tree tree, int32_t snode, int32_t *out_any, int64_t cap);
TI_DLL_EXPORT int32_t b2_deactivate_all(B2Tree tree, int32_t snode); /* SNode.deactivate_all, snode.py:286-304 */
TI_DLL_EXPORT int32_t b2_gc(B2Tree tree, int32_t snode);             /* gc_parallel_0/1/2 fused */
/* out[4] = {data_list size, free_list size, recycled_list size, free_list_used} */
TI_DLL_EXPORT int32_t b2_alloc_stats(B2Tree tree, int32_t snode, int32_t *out_host);
TI_DLL_EXPORT int64_t b2_num_dynamically_allocated(B2Tree tree, int32_t snode); /* snode.py _num_dynamically_allocated */
/* Field element access through the full access path (lower_access.cpp:227-281): writes
   activate, reads of inactive cells return the ambient zero. */
TI_DLL_EXPORT int32_t b2_write(B2Tree tree, int32_t place, const int32_t *coords_any, int64_t n, const void *vals_any, int32_t atomic_add_f32);
TI_DLL_EXPORT int32_t b2_read(B2Tree tree, int32_t place, const int32_t *coords_any, int64_t n, void *vals_any);
TI_DLL_EXPORT int32_t b2_dynamic_append_i32(B2Tree tree, int32_t dyn, const int32_t *coords_any, int64_t n, const int32_t *vals_any); /* ti.append */
TI_DLL_EXPORT int32_t b2_dynamic_length(B2Tree tree, int32_t dyn, const int32_t *coords_any, int64_t n, int32_t *out_any);          /* ti.length */
/* Whole-field operations = struct-fors with built-in bodies. */
TI_DLL_EXPORT int32_t b2_field_fill(B2Tree tree, int32_t place, uint64_t value_bits);      /* field.fill over ACTIVE cells; the value's bit
                                                                                              pattern (1 / 2 / 4 / 8-byte places), zero-extended */
/* dense row-major image of the field over `shape` (ndim = number of active axes of the place's
   path, ascending axis order); inactive cells read 0 (field.to_numpy, _kernels.py:40-60) */
TI_DLL_EXPORT int32_t b2_field_to_dense(B2Tree tree, int32_t place, void *out_any);
TI_DLL_EXPORT int32_t b2_field_from_dense(B2Tree tree, int32_t place, const void *in_any); /* activates everything (from_numpy) */
/* op: 0 add, 1 min, 2 max; dtype 0 f32, 1 i32.  reduce_* of runtime.cpp:1235-1266 behind a
   thread-local accumulator (make_thread_local.cpp:100-213). Result written to out_host (4 B). */
TI_DLL_EXPORT int32_t b2_field_reduce(B2Tree tree, int32_t place, int32_t op, int32_t dtype, void *out_host);
/* Dense-array reductions (range-for with TLS). */
TI_DLL_EXPORT int32_t b2_reduce(TiRuntime runtime, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_host);
/* same, enqueue only: the 4-byte result lands in device memory (no host round trip) */
TI_DLL_EXPORT int32_t b2_reduce_device(TiRuntime runtime, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_dev);
/* b2_reduce_device for callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
   No counterpart in the reference.  Same arguments and result (4 bytes in device
   memory, enqueued on the runtime's stream); the per-CTA partials and the arrival ticket live in a static __device__ array of
   8448 bytes that every GPU loading the library holds until exit, so the call allocates nothing (no cudaMalloc / cudaFree /
   stream-ordered allocation).  n <= B2_REDUCE_NOALLOC_MAX_N, else TI_ERROR_ARGUMENT_OUT_OF_RANGE.  The partials are folded in
   index order: the result is deterministic for a given n, and differs from b2_reduce_device's only by fp32 summation order.
   All calls share the array: calls that overlap on the GPU (two streams, two host threads on one device) race. */
#define B2_REDUCE_NOALLOC_MAX_N (1ll << 28)
TI_DLL_EXPORT int32_t b2_reduce_noalloc(TiRuntime runtime, const void *data_dev, int64_t n, int32_t op, int32_t dtype, void *out_dev);

/* runtime-level */
TI_DLL_EXPORT void b2_sync(TiRuntime runtime);
TI_DLL_EXPORT int32_t b2_device_index(TiRuntime runtime);
TI_DLL_EXPORT int64_t b2_pool_bytes_used(TiRuntime runtime); /* memory-profiler query */
/* print_memory_profiler_info (llvm_runtime_executor.cpp:314-384) as text: per SNode the element list, the data list and
 * the allocator counters, then the pool usage. Returns the size needed (incl. NUL); synchronises. */
TI_DLL_EXPORT int64_t b2_memory_profiler_info(B2Tree tree, char *out, int64_t cap);
TI_DLL_EXPORT void b2_set_pool_size(uint64_t bytes);         /* ti.init(device_memory_GB=), before ti_create_runtime */
TI_DLL_EXPORT void b2_set_random_seed(int32_t seed);      /* ti.init(random_seed=): seeds the per-thread xorshift128 states of run-time bodies
                                                            (c.rand_u32 / rand_f32 / rand_f64, runtime.cpp:527-541,1790-1827), before the first launch */
/* Kernel launch counter (all kernels this library launched since creation). */
TI_DLL_EXPORT uint64_t b2_launch_count(TiRuntime runtime);
/* device-side error channel (taichi_assert_format / check_runtime_error): returns 0 if none,
   else copies the message. */
TI_DLL_EXPORT int32_t b2_check_runtime_error(TiRuntime runtime, char *msg, uint64_t cap);

/* ---------------------------- MLS-MPM (3-D, mpm3d.py:36-81 model) ---------------------- */
typedef struct B2MpmParams {
  int32_t n_grid;     /* cells per axis (multiple of 4)                                   */
  int32_t layout;     /* 0: dense(n/4).dense(4)   1: pointer(n/32).bitmasked(8).dense(4)
                         2: pointer(n/32).pointer(8).dense(4)                              */
  float dt, E, p_rho, gravity; /* mpm3d.py:13-23                                           */
  int32_t bound;      /* mpm3d.py:23                                                       */
  int32_t block_local;/* 1: shared-memory pad P2G/G2P (ti.block_local), 0: plain atomics   */
  int32_t bls_activate_halo; /* 1: reproduce the CUDA-BLS epilogue's wider active set      */
  int32_t slab_lo, slab_hi;  /* owned range of 4^3-block indices along i ([0,n/4) = all)    */
  int32_t deactivate_every;  /* K > 0 (sparse layouts): every K-th substep starts with deactivate_all on the grid's
                                pointer level + GC (snode.py:286-304, offload.cpp:671-695 -- the reference's sparse-MPM idiom,
                                test_gc.py:40-52); P2G re-activates exactly the touched blocks.  0: the active set only grows */
  int64_t max_particles;
  int64_t max_bins;          /* occupied 4^3 blocks the particle bins can hold at once; 0 = max(4096, max_particles / 64).
                                Vacated bins are recycled by a compaction pass every 64 substeps. */
  int32_t model;             /* 0: mpm3d.py (weakly compressible, J).  1: deformation gradient F + plastic Jp + material per
                                particle with ti.svd -- the mpm128.py:36-63 constitutive model (liquid 0 / jelly 1 / snow 2)
                                carried to 3-D as in ggui_examples/mpm3d_ggui.py:52-88; E, nu give mu_0, lambda_0 */
  float nu;                  /* model 1: Poisson's ratio (mpm128.py:16)                                                     */
  float p_vol;               /* > 0: particle volume (default (dx / 2)^2 as mpm3d.py:17)                                    */
  int32_t pad_;
} B2MpmParams;

TI_DLL_EXPORT B2Mpm b2_mpm_create(TiRuntime runtime, const B2MpmParams *params);
TI_DLL_EXPORT void b2_mpm_destroy(B2Mpm mpm);
TI_DLL_EXPORT B2Tree b2_mpm_tree(B2Mpm mpm);          /* the grid's SNode tree (leaf: dense 4^3 of {vx,vy,vz,m}) */
TI_DLL_EXPORT int32_t b2_mpm_leaf_snode(B2Mpm mpm);   /* id of the dense(4^3) SNode   */
/* x[N][3], v[N][3], C[N][3][3], J[N] in the reference's field layout; host or device.
 * Host arrays are uploaded on a copy stream into double-buffered device staging (the role of the launcher's
 * H2D staging of host ndarray args, kernel_launcher.cpp:78-107) and binned on the compute stream behind an
 * event: with page-locked arrays the call returns at once and the upload overlaps earlier work still running;
 * the arrays must stay unchanged until ti_wait() / b2_mpm_wait_io(). Pageable arrays are read before return. */
TI_DLL_EXPORT int32_t b2_mpm_set_particles(B2Mpm mpm, int64_t n, const float *x_any, const float *v_any, const float *C_any, const float *J_any);
TI_DLL_EXPORT int32_t b2_mpm_add_particles(B2Mpm mpm, int64_t n, const float *x_any, const float *v_any, const float *C_any, const float *J_any, const int32_t *ids_any);
TI_DLL_EXPORT int64_t b2_mpm_num_particles(B2Mpm mpm);
/* model 1: the same with F[N][3][3], Jp[N], material[N] (int32: 0 liquid, 1 jelly, 2 snow); NULL = identity / 1 / 0.
 * Host or device arrays; the get returns F and Jp as the reference holds them after the same number of substeps. */
TI_DLL_EXPORT int32_t b2_mpm_set_particles_model(B2Mpm mpm, int64_t n, const float *x_any, const float *v_any, const float *C_any,
                                                 const float *F_any, const float *Jp_any, const int32_t *material_any);
TI_DLL_EXPORT int32_t b2_mpm_get_particles_model(B2Mpm mpm, float *x_any, float *v_any, float *C_any, float *F_any, float *Jp_any,
                                                 int32_t *material_any);
/* Any pointer may be NULL. Output is in original particle-id order. */
TI_DLL_EXPORT int32_t b2_mpm_get_particles(B2Mpm mpm, float *x_any, float *v_any, float *C_any, float *J_any);
/* Same, but host arrays are filled by a copy stream after return (the D2H half of kernel_launcher.cpp:175-186
 * without its synchronisation): valid after b2_mpm_wait_io(). Two downloads may be in flight. */
TI_DLL_EXPORT int32_t b2_mpm_get_particles_async(B2Mpm mpm, float *x_any, float *v_any, float *C_any, float *J_any);
/* Blocks until every upload and download issued through this solver has finished. */
TI_DLL_EXPORT int32_t b2_mpm_wait_io(B2Mpm mpm);
/* 0-D reductions over the particles (make_thread_local.cpp:100-213 + runtime.cpp:1204-1266 in the reference):
 * out[0] live particles, [1..3] momentum, [4] kinetic energy, [5] sum J, [6] max |v|, [7] max |J - 1|.
 * _device: enqueue only, 8 doubles of device memory. */
TI_DLL_EXPORT int32_t b2_mpm_diagnostics(B2Mpm mpm, double *out8_host);
TI_DLL_EXPORT int32_t b2_mpm_diagnostics_device(B2Mpm mpm, double *out8_dev);
TI_DLL_EXPORT int32_t b2_mpm_substep(B2Mpm mpm, int32_t nsteps);
/* individual phases (mpm88_graph_aot.py kernel split) */
TI_DLL_EXPORT int32_t b2_mpm_clear_grid(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_p2g(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_grid_op(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_g2p(B2Mpm mpm);
/* dense image of the grid, [n][n][n][4] floats {vx,vy,vz,m}; inactive cells 0 */
TI_DLL_EXPORT int32_t b2_mpm_grid_to_dense(B2Mpm mpm, float *out_any);
/* compact image: keys[e] = (bx*nb+by)*nb+bz of the e-th active 4^3 block (leaf list order), cells[e][64][4] its
 * {vx,vy,vz,m} cells.  Returns the number of active blocks (also when keys/cells are NULL); at most cap are written. */
TI_DLL_EXPORT int64_t b2_mpm_grid_blocks(B2Mpm mpm, int32_t *keys_any, float *cells_any, int64_t cap);
/* out_host[8]: {n_particles, n_active_blocks, n_bins, n_work_items, sum_m, sum_px, sum_py, sum_pz} (doubles) */
TI_DLL_EXPORT int32_t b2_mpm_stats(B2Mpm mpm, double *out_host);
/* per-phase device time of the last b2_mpm_substep call with profiling on, ms:
   {clear, p2g, listgen, grid_op, g2p, deactivate_gc} (deactivate_gc averaged over ALL substeps) */
/* The solver keeps the leaf element list of its grid current incrementally (a block that P2G activates appends
 * itself; the reference regenerates every list before every struct-for, transforms/offload.cpp:174-264).  Call this
 * after changing the solver's tree through the b2_* tree API (activate / deactivate / gc): the next phase then
 * regenerates the lists from the root. */
TI_DLL_EXPORT int32_t b2_mpm_invalidate_lists(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_set_compact_every(B2Mpm mpm, int32_t k);    /* substeps between bin-slot compactions (default 64, 0 = never) */
TI_DLL_EXPORT int32_t b2_mpm_set_deactivate_every(B2Mpm mpm, int32_t k); /* B2MpmParams.deactivate_every, changeable between substeps */
TI_DLL_EXPORT int32_t b2_mpm_set_profiling(B2Mpm mpm, int32_t on);
TI_DLL_EXPORT int32_t b2_mpm_phase_ms(B2Mpm mpm, float *out_host6);

/* slab decomposition (new: no reference counterpart). Messages are device buffers
   {u32 count; u32 pad[3]; records[cap]}: block records are 1040 B {i32 key; pad; float4 cells[64]},
   particle records 80 B {float4 x4; u32 id; pad}. All calls are asynchronous; counts stay on the device. */
TI_DLL_EXPORT int32_t b2_mpm_pack_blocks(B2Mpm mpm, int32_t block_layer_i, void *buf_dev, int64_t cap_records);
TI_DLL_EXPORT int32_t b2_mpm_unpack_blocks(B2Mpm mpm, const void *buf_dev, int64_t cap_records, int32_t mode); /* 0 activate+add, 1 overwrite-if-active, 2 add-if-active */
TI_DLL_EXPORT int32_t b2_mpm_extract_migrants(B2Mpm mpm, int32_t side, void *buf_dev, int64_t cap_records);   /* 0: below slab_lo, 1: at/above slab_hi */
TI_DLL_EXPORT int32_t b2_mpm_insert_records(B2Mpm mpm, const void *buf_dev, int64_t cap_records);
/* both directions in one launch (either buffer may be NULL) */
TI_DLL_EXPORT int32_t b2_mpm_extract_migrants2(B2Mpm mpm, void *buf_lower_dev, int64_t cap_lower, void *buf_upper_dev, int64_t cap_upper);
TI_DLL_EXPORT int32_t b2_mpm_insert_records2(B2Mpm mpm, const void *buf_a_dev, int64_t cap_a, const void *buf_b_dev, int64_t cap_b);
/* Drivers that enqueue the phases of a substep themselves (b2_slab.cu) bracket them with begin / end: the cadence work (bin
 * compaction, sparse life cycle) runs in begin, the clear kernel resets the work counters (no memset nodes), message headers are
 * zeroed by ONE call instead of one memset per message. */
TI_DLL_EXPORT int32_t b2_mpm_begin_substep(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_end_substep(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_mpm_zero_headers(B2Mpm mpm, void *a_dev, void *b_dev, void *c_dev, void *d_dev);

/* compact copy of the live particles + their ids; returns how many (synchronises) */
TI_DLL_EXPORT int64_t b2_mpm_export(B2Mpm mpm, float *x_any, float *v_any, float *C_any, float *J_any, int32_t *ids_any, int64_t cap);
/* Same records into HOST arrays of `cap` entries each, filled by the solver's copy stream after return; *count
 * (host) receives the number of records (entries past it are unspecified). Valid after b2_mpm_wait_io(). */
TI_DLL_EXPORT int32_t b2_mpm_export_async(B2Mpm mpm, float *x, float *v, float *C, float *J, int32_t *ids, int64_t cap, int64_t *count);

/* Native slab driver: the whole decomposed substep (kernels + ncclSend/ncclRecv over NVLink) is
   enqueued from C++ on the runtime's stream; no host synchronisation.  NCCL is bound at run time
   (dlopen libnccl.so.2).  Rank 0 creates the id and shares its 128 bytes with the other ranks
   (e.g. torch.distributed.broadcast), then every rank calls b2_comm_init. */
TI_DLL_EXPORT TiRuntime b2_mpm_runtime(B2Mpm mpm);
TI_DLL_EXPORT int32_t b2_comm_unique_id(uint8_t *out128);
TI_DLL_EXPORT B2Comm b2_comm_init(TiRuntime runtime, const uint8_t *id128, int32_t rank, int32_t world);
/* Loopback communicator: all `world` slabs live in this process on one device; a message is a device-to-device copy
 * into the peer's receive buffer (same call order and layout as the NCCL path).  Slabs created on it advance together
 * through b2_slab_substep_lockstep (slabs[r] = rank r).  For single-GPU testing of the native driver. */
TI_DLL_EXPORT B2Comm b2_comm_loopback(TiRuntime runtime, int32_t world);
TI_DLL_EXPORT void b2_comm_destroy(B2Comm comm);
TI_DLL_EXPORT B2Slab b2_slab_create(B2Mpm mpm, B2Comm comm, int32_t rank, int32_t world, int32_t block_lo, int32_t block_hi,
                                    int64_t cap_blocks, int64_t cap_particles);
TI_DLL_EXPORT void b2_slab_destroy(B2Slab slab);
TI_DLL_EXPORT int32_t b2_slab_substep(B2Slab slab, int32_t nsteps);
TI_DLL_EXPORT int32_t b2_slab_substep_lockstep(B2Slab *slabs, int32_t world, int32_t nsteps);
TI_DLL_EXPORT int32_t b2_slab_tune(B2Slab slab); /* message sizes := 2x last counts (synchronises) */
/* Whole-domain particle diagnostics: out[0] live particles, [1..3] momentum, [4] kinetic energy, [5] sum J,
 * [6] max |v| (CFL), [7] max |J - 1|; local reduction + ncclAllReduce (SUM / MAX) over the slabs. Synchronises. */
TI_DLL_EXPORT int32_t b2_slab_diagnostics(B2Slab slab, double *out8_host);

/* ---- struct-for with a run-time compiled body -------------------------------------------
 * Stands in for the reference's per-kernel codegen + CUDA JIT of a struct-for task (codegen_llvm.cpp:2089-2317,
 * runtime/cuda/jit_cuda.cpp:8-240, kernel_launcher.cpp:15-213): `source` is CUDA C++ holding one
 *     B2_BODY(c) { ... }                       // runs once per ACTIVE cell of the loop SNode
 * with c.I[12] = loop index, c.read<T>(f, J) / c.write<T>(f, J, v) / c.ref<T>(f, J) (writes activate, reads of
 * inactive cells see the ambient zero, as in lower_access.cpp:254-279), c.is_active(f, J), c.offset(J, axis, d),
 * c.arg<T>(k); contract in taichi_b200/csrc/b2_jit.cuh.  Compiled once with NVRTC to an sm_100a cubin.
 * A source that does not compile returns NULL with TI_ERROR_INVALID_ARGUMENT and the compiler log as message;
 * a missing libnvrtc.so.12 returns NULL with TI_ERROR_NOT_SUPPORTED. */
TI_DLL_EXPORT B2JitKernel b2_jit_compile(TiRuntime runtime, const char *source, const char *name);
/* compile only, no device needed: cubin size, or -1 with the compiler log as last error */
TI_DLL_EXPORT int64_t b2_jit_check(const char *source, const char *name);
TI_DLL_EXPORT void b2_jit_destroy(B2JitKernel kernel);
TI_DLL_EXPORT int64_t b2_jit_log(B2JitKernel kernel, char *out, int64_t cap); /* diagnostics; returns size needed */
/* listgen of loop_snode (offload.cpp:174-264), then the body over its element list. field_places: <= 8 place
 * SNode ids (field f of the body); args: <= 16 64-bit slots (scalars bit-cast, DEVICE pointers). Asynchronous. */
TI_DLL_EXPORT int32_t b2_jit_launch(B2JitKernel kernel, B2Tree tree, int32_t loop_snode, const int32_t *field_places,
                                    int32_t n_fields, const uint64_t *args, int32_t n_args);
/* ti.block_local for run-time bodies (transforms/make_block_local.cpp:97-365, ir/scratch_pad.h:30-180).  One spec = one pad of
 * field `field` (index into field_places): per pad axis d the loop's block covers `extent[d]` cells of the field's index space
 * (0: the loop SNode's own shape, i.e. dense leaf loops; pid-list loops over a dynamic SNode pass the block size) starting at
 * parent_coordinate * extent, and the body reaches offsets lo[d]..hi[d] (inclusive) around it, so the pad spans
 * [lo, extent + hi) = ScratchPad::bounds.  mode 1: read-only (prologue loads, inactive / out-of-field cells read 0; body:
 * c.bls_read<T>(slot, J)); mode 2 / 3 / 4: accumulate f32 / i32 / f64 (prologue zero-fills, body: c.bls_add<T>(slot, J, v), epilogue
 * adds the pad back through an activating access -- every pad cell if halo != 0 (the reference's CUDA epilogue), else only
 * non-zero sums (its x64 result)).  Also available in every struct-for body: c.reduce_add / reduce_min / reduce_max<T>(k, v), a
 * thread-local accumulator reduced per CTA into the device pointer passed as argument k (make_thread_local.cpp:100-213). */
typedef struct B2BlsSpec {
  int32_t field, mode, ndim, halo;
  int32_t axis[3], extent[3], lo[3], hi[3];
} B2BlsSpec;
TI_DLL_EXPORT int32_t b2_jit_launch_bls(B2JitKernel kernel, B2Tree tree, int32_t loop_snode, const int32_t *field_places, int32_t n_fields,
                                        const uint64_t *args, int32_t n_args, const B2BlsSpec *bls, int32_t n_bls);
/* ScratchPad geometry: block of block_size[d] loop cells, n_access accesses at offsets[a][d]  ->  low[d] = min offset,
 * high[d] = block_size[d] + max offset, pad_size[d] = high - low (any output may be NULL); returns the number of pad cells
 * (known answers: tests/cpp/analysis/bls_analyzer_test.cpp:51-77, tests/cpp/transforms/make_block_local_test.cpp:91-120). */
TI_DLL_EXPORT int64_t b2_bls_pad_geometry(int32_t ndim, const int32_t *block_size, int32_t n_access, const int32_t *offsets, int32_t *low,
                                          int32_t *high, int32_t *pad_size);
/* Range-for (`for i in range(begin, end)`, gpu_parallel_range_for runtime.cpp:1541-1565) over a source holding
 *     B2_RANGE_BODY(c) { ... }                 // c.i = loop index; same field access and arguments
 * e.g. a particle loop scattering into a sparse grid with c.ref<float>(f, J) + atomicAdd (activates). */
TI_DLL_EXPORT int32_t b2_jit_launch_range(B2JitKernel kernel, B2Tree tree, int64_t begin, int64_t end, const int32_t *field_places,
                                          int32_t n_fields, const uint64_t *args, int32_t n_args);

#ifdef __cplusplus
}
#endif
#endif /* B200MPM_H */
