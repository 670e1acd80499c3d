// b2_host.h -- host-side classes of libb200mpm (runtime, memory, SNode tree).
// Replaces, for this path only: LlvmRuntimeExecutor (runtime/llvm/llvm_runtime_executor.cpp:36-756),
// CudaDevice / DeviceMemoryPool (rhi/cuda/cuda_device.cpp, rhi/llvm/device_memory_pool.cpp),
// StructCompilerLLVM (codegen/llvm/struct_llvm.cpp:36-241), SNode::create_node (ir/snode.cpp:37-125).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200mpm.h"
#include "b2_internal.cuh"

struct B2Error : std::runtime_error {
  TiError code;
  B2Error(TiError c, const std::string &m) : std::runtime_error(m), code(c) {}
};

void b2_record_error(TiError code, const std::string &msg);
#define B2_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      throw B2Error(_e == cudaErrorMemoryAllocation ? TI_ERROR_OUT_OF_MEMORY : TI_ERROR_INVALID_STATE, \
                    std::string(#expr) + ": " + cudaGetErrorString(_e));                            \
  } while (0)

#define B2_API_BEGIN try {
#define B2_API_END(ret)                                  \
  }                                                      \
  catch (const B2Error &e) {                             \
    b2_record_error(e.code, e.what());                   \
    return ret;                                          \
  }                                                      \
  catch (const std::exception &e) {                      \
    b2_record_error(TI_ERROR_INVALID_STATE, e.what());   \
    return ret;                                          \
  }
#define B2_API_END_VOID                                  \
  }                                                      \
  catch (const B2Error &e) {                             \
    b2_record_error(e.code, e.what());                   \
  }                                                      \
  catch (const std::exception &e) {                      \
    b2_record_error(TI_ERROR_INVALID_STATE, e.what());   \
  }

cudaStream_t b2_current_stream();

struct TiMemory_t {
  void *dev = nullptr;    // device pointer (or pinned host pointer when host_access)
  uint64_t size = 0;
  bool imported = false;  // borrowed: never freed
  bool host_access = false;
  void *staging = nullptr;  // host staging copy while mapped
};

struct TiRuntime_t {
  int device = 0;
  int num_sms = 0;
  B2Pool *pool_dev = nullptr;  // device copy of the pool header
  unsigned *rand_states = nullptr;  // xorshift128 state {x, y, z, w} per linear thread id of a run-time body (runtime.cpp:527-541)
  int num_rand_states = 0;
  unsigned *ensure_rand_states();   // allocated and seeded at the first use (runtime_initialize_rand_states_cuda, runtime.cpp:973-977)
  void *pool_mem = nullptr;
  uint64_t pool_size = 0;
  uint64_t launches = 0;
  std::vector<TiMemory_t *> memories;

  explicit TiRuntime_t(int device_index);
  ~TiRuntime_t();
  cudaStream_t stream() const { return b2_current_stream(); }
  void sync();
  int grid_for(int64_t work_items, int block, int per_sm = 8) const;
  // device-side error channel
  int check_error(std::string *msg);
  void count_launch(int n = 1) { launches += n; }
};

// Stages a maybe-host buffer on the device for the duration of a call.
struct B2Staged {
  void *dev = nullptr;
  void *host = nullptr;
  size_t bytes = 0;
  bool owned = false;
  bool writeback = false;
  cudaStream_t st = nullptr;
  B2Staged() {}
  B2Staged(const B2Staged &) = delete;
  B2Staged &operator=(const B2Staged &) = delete;
  // in: copy host->device now; out: copy device->host at finish()
  void init(const void *any, size_t nbytes, bool in, bool out, cudaStream_t s);
  void finish();  // writes back (synchronises if needed) and frees
  ~B2Staged();
  template <typename T>
  T *as() const { return (T *)dev; }
};
bool b2_is_device_pointer(const void *p);

struct B2SNodeHost {
  int type = 0, parent = -1, depth = 0;
  std::vector<int> ch;
  int shape[B2_MAX_AXES], acc[B2_MAX_AXES], active[B2_MAX_AXES], from_root[B2_MAX_AXES];
  int64_t num_cells = 1, cell_size = 0, node_size = 0, offset_in_parent_cell = 0, aux_offset = 0, alloc_size = 0;
  int cell_align = 1, node_align = 1, chunk_size = 0, dtype_bytes = 0;
};

struct B2Tree_t {
  TiRuntime_t *rt;
  std::vector<B2SNodeHost> s;
  B2TreeDev host_copy;   // mirror of the device descriptor (pointers are device pointers)
  B2TreeDev *dev = nullptr;
  void *root_mem = nullptr;
  void *lists_mem = nullptr;
  void *mgr_mem = nullptr;
  void *scratch = nullptr;  // small device scratch (counters / results)
  std::vector<void *> ambient;

  B2Tree_t(TiRuntime_t *rt, const B2SNodeDesc *desc, int n);
  ~B2Tree_t();
  void layout(int id);
  B2Path path_to(int target) const;
  std::vector<int> path_ids(int target) const;
  void check_snode(int id, const char *what) const;

  void listgen_async(int target);
  int64_t list_count(int target);  // synchronises
  void gc_async(int sid);
  void deactivate_all(int sid);
  // one level only (listgen of that level + deactivate body + GC if it is a pointer SNode); reset_list: child list := empty
  void deactivate_level_async(int sid);
  // deactivate_all + GC of a pointer SNode with one container under the root, one launch (+ two lists and a bitmap emptied)
  void deactivate_all_gc_async(int sid, unsigned char *container, int list_a, int list_b, void *zero, size_t zero_bytes);
  void reset_list_async(int sid);
};
