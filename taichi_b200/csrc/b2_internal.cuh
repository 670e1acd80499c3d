// b2_internal.cuh -- device-side data structures and access functions of the B200 sparse
// SNode runtime (shared by b2_runtime.cu, b2_mpm.cu, b2_capi.cu).
//
// What it replaces in the reference (semantics, not code):
//   ListManager            taichi/runtime/llvm/runtime_module/runtime.cpp:426-513
//   NodeManager            runtime.cpp:630-707
//   Element                runtime.cpp:517-525
//   Pointer_* / Bitmasked_* / Dynamic_* / Dense_* / Root_*   runtime_module/node_*.h
//   allocate_from_reserved_memory (bump pool)                 runtime.cpp:838-874
// Differences by design (B200-first): no spin locks anywhere -- pointer cells, list chunks and
// allocator chunks are claimed with one CAS (0 -> BUSY) and published with one store; the
// pointer container is {u64 ptr[n]; u32 node_index_plus1[n]} (SoA: listgen scans 8 B/slot and
// recycle needs no ptr2index scan, cf. runtime.cpp:502-512); node sizes are rounded to 128 B
// so that zero-fill and block loads are 128-bit and aligned.
#pragma once
#ifndef __CUDACC_RTC__  // run-time compiled struct-for bodies (b2_jit.cuh) see this header through NVRTC: no host headers there
#include <cuda_runtime.h>
#include <stdint.h>
#endif

#define B2_MAX_AXES 12
#define B2_MAX_SNODES 1024 /* taichi/inc/constants.h:7 taichi_max_num_snodes */
#define B2_MAX_LEVELS 12
#define B2_LISTGEN_MAX_ELEMENT 1024 /* taichi/inc/constants.h:28 */
#define B2_LIST_MAX_CHUNKS 4096
#define B2_ELEM_LOG2_EPC 14 /* 16 Ki Elements (1 MiB) per element-list chunk */
#define B2_IDX_LOG2_EPC 16  /* 64 Ki u32 per index-list chunk                */
#define B2_MGR_MAX_CHUNKS 8192
#define B2_BUSY 1ull

enum { B2T_ROOT = 0, B2T_DENSE = 1, B2T_POINTER = 2, B2T_BITMASKED = 3, B2T_DYNAMIC = 4, B2T_PLACE = 5 };
enum { B2E_NONE = 0, B2E_POOL_OOM = 1, B2E_LIST_OVERFLOW = 2, B2E_MGR_OVERFLOW = 3, B2E_BIN_OVERFLOW = 4, B2E_ASSERT = 5 };

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

// Work item of a struct-for: same 64-byte record as the reference's Element.
struct __align__(16) B2Element {
  u8 *node;  // container the cells [lo,hi) live in
  int lo, hi;
  int pcoord[B2_MAX_AXES];  // coordinates of the parent cell that holds `node`
};
static_assert(sizeof(B2Element) == 64, "Element must be 64 bytes");

struct B2Pool {
  u64 base, size;
  u64 head;  // bump offset (atomic)
  int error; // first device-side error (B2E_*)
  int error_arg;
};

// Chunked append-only list of fixed-size records; chunks come lazily from the pool.
struct B2CList {
  u32 count;
  u32 elem_size;
  u32 log2_epc;
  u32 pad;
  u64 chunks[B2_LIST_MAX_CHUNKS];
};

struct B2NodeMgr {
  u64 elem_size;  // bytes per node (multiple of 128)
  u32 log2_epc;   // nodes per data chunk
  u32 data_count; // nodes ever created      (data_list.size())
  u32 free_count; // size of the free list    (free_list.size())
  u32 free_used;  // consumed since last GC   (free_list_used)
  u32 cur;        // which of free_list[2] is live
  u32 gc_ticket;  // last-CTA-done counter of the fused GC kernel
  u32 gc_bar;     // arrival counter of the grid barrier of k_deactivate_all_gc
  u32 pad_;
  u64 ambient;    // zero node returned for inactive reads
  u64 data_chunks[B2_MGR_MAX_CHUNKS];
  B2CList free_list[2];
  B2CList recycled;
};

struct B2SNodeDev {
  int type, parent, depth, nch;
  int shape[B2_MAX_AXES];
  int acc[B2_MAX_AXES];
  long long num_cells;
  long long cell_size;
  long long node_size;
  long long offset_in_parent_cell;
  long long aux_offset;  // bitmasked: mask words; pointer: u32 index array
  long long alloc_size;  // pointer/dynamic: NodeManager element size
  int chunk_size, dtype_bytes;
  // compact list of axes with shape > 1, ascending
  int nax, ax[B2_MAX_AXES];
  // every shape of this SNode is a power of two: digit_a(l) = (l >> acc_shift[a]) & (shape[a] - 1)
  int pot, acc_shift[B2_MAX_AXES];
};

struct B2TreeDev {
  B2Pool *pool;
  u8 *root;
  int n, pad;
  B2SNodeDev s[B2_MAX_SNODES];
  B2CList *lists[B2_MAX_SNODES];
  B2NodeMgr *mgr[B2_MAX_SNODES];
};

// Access path root..target with the divisors that extract each level's digits.
struct B2Path {
  int n;
  int pot;  // every shape and divisor on the path is a power of two: digits are shift + mask
  int id[B2_MAX_LEVELS];
  int div[B2_MAX_LEVELS][B2_MAX_AXES];
  int shift[B2_MAX_LEVELS][B2_MAX_AXES];
};

struct B2DynNode {  // container of a dynamic SNode (16 B)
  u32 n;
  u32 pad;
  u64 first;  // first chunk
};
struct B2DynChunk {  // header of a dynamic chunk; data follows at +16
  u64 next;
  u32 idx1;
  u32 pad;
};

#ifdef __CUDACC__
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 b2_ld_volatile(const u64 *p) { return *(const volatile u64 *)p; }

__device__ __forceinline__ void b2_set_error(B2Pool *pool, int code, int arg) {
  if (atomicCAS(&pool->error, 0, code) == 0) pool->error_arg = arg;
}

// Bump allocation from the zero-initialised pool. Returns 0 on exhaustion (error recorded).
__device__ __forceinline__ u64 b2_pool_alloc(B2Pool *pool, u64 bytes) {
  bytes = (bytes + 255ull) & ~255ull;
  u64 off = atomicAdd(&pool->head, bytes);
  if (off + bytes > pool->size) {
    b2_set_error(pool, B2E_POOL_OOM, 0);
    return 0;
  }
  return pool->base + off;
}

// Claim-or-wait on a lazily created pointer-sized slot. `make` runs in exactly one thread.
template <typename F>
__device__ __forceinline__ u64 b2_ensure(u64 *slot, F make) {
  // a published pointer never changes again, so a (possibly stale) L1-cached read may be trusted when it
  // shows one; 0 / BUSY go on to the coherent path
  u64 v = *slot;
  if (v > B2_BUSY) return v;
  v = b2_ld_volatile(slot);
  if (v > B2_BUSY) return v;
  if (v == 0) {
    u64 old = atomicCAS(slot, 0ull, B2_BUSY);
    if (old == 0) {
      u64 p = make();
      __threadfence();
      atomicExch(slot, p ? p : 0ull);
      return p;
    }
    if (old > B2_BUSY) return old;
  }
  unsigned ns = 8;
  while ((v = b2_ld_volatile(slot)) == B2_BUSY) {
    __nanosleep(ns);
    if (ns < 256) ns *= 2;
  }
  __threadfence();
  return v;
}

__device__ __forceinline__ u8 *b2_list_slot(B2Pool *pool, B2CList *l, u32 i) {
  u32 c = i >> l->log2_epc;
  if (c >= B2_LIST_MAX_CHUNKS) {
    b2_set_error(pool, B2E_LIST_OVERFLOW, (int)i);
    return nullptr;
  }
  u64 bytes = (u64)l->elem_size << l->log2_epc;
  u64 base = b2_ensure(&l->chunks[c], [&]() { return b2_pool_alloc(pool, bytes); });
  if (!base) return nullptr;
  return (u8 *)base + (u64)(i & ((1u << l->log2_epc) - 1)) * l->elem_size;
}
// Read-only access to an element known to exist.
__device__ __forceinline__ u8 *b2_list_get(const B2CList *l, u32 i) {
  u64 base = l->chunks[i >> l->log2_epc];
  return (u8 *)base + (u64)(i & ((1u << l->log2_epc) - 1)) * l->elem_size;
}
__device__ __forceinline__ void b2_list_append_u32(B2Pool *pool, B2CList *l, u32 v) {
  u32 i = atomicAdd(&l->count, 1u);
  u32 *p = (u32 *)b2_list_slot(pool, l, i);
  if (p) *p = v;
}

// ---- NodeManager (lock-free) --------------------------------------------------------------
__device__ __forceinline__ u8 *b2_mgr_node(B2Pool *pool, B2NodeMgr *m, u32 idx) {
  u32 c = idx >> m->log2_epc;
  if (c >= B2_MGR_MAX_CHUNKS) {
    b2_set_error(pool, B2E_MGR_OVERFLOW, (int)idx);
    return nullptr;
  }
  u64 bytes = m->elem_size << m->log2_epc;
  u64 base = b2_ensure(&m->data_chunks[c], [&]() { return b2_pool_alloc(pool, bytes); });
  if (!base) return nullptr;
  return (u8 *)base + (u64)(idx & ((1u << m->log2_epc) - 1)) * m->elem_size;
}
// NodeManager::allocate (runtime.cpp:666-677): reuse a GC'd node if the free list still has
// one, else create a fresh (zero) node.
__device__ __forceinline__ u8 *b2_mgr_alloc(B2Pool *pool, B2NodeMgr *m, u32 *idx_out) {
  u32 cur = atomicAdd(&m->free_used, 1u);
  u32 idx;
  if (cur < *(volatile u32 *)&m->free_count)
    idx = *(u32 *)b2_list_get(&m->free_list[m->cur], cur);
  else
    idx = atomicAdd(&m->data_count, 1u);
  *idx_out = idx;
  return b2_mgr_node(pool, m, idx);
}
__device__ __forceinline__ void b2_mgr_recycle(B2Pool *pool, B2NodeMgr *m, u32 idx) {
  b2_list_append_u32(pool, &m->recycled, idx);
}

// ---- per-type cell operations ---------------------------------------------------------------
__device__ __forceinline__ u32 *b2_mask_words(const B2SNodeDev &s, u8 *node) { return (u32 *)(node + s.aux_offset); }
__device__ __forceinline__ u32 *b2_ptr_idx(const B2SNodeDev &s, u8 *node) { return (u32 *)(node + s.aux_offset); }

__device__ __forceinline__ bool b2_is_active(const B2SNodeDev &s, u8 *node, int i) {
  switch (s.type) {
    case B2T_POINTER: return b2_ld_volatile((u64 *)node + i) > B2_BUSY;
    case B2T_BITMASKED: return (b2_mask_words(s, node)[i >> 5] >> (i & 31)) & 1u;
    case B2T_DYNAMIC: return (u32)i < *(volatile u32 *)&((B2DynNode *)node)->n;
    default: return true;
  }
}

__device__ __forceinline__ u8 *b2_pointer_activate(B2TreeDev *T, int sid, u8 *node, int i) {
  const B2SNodeDev &s = T->s[sid];
  u64 p = b2_ensure((u64 *)node + i, [&]() {
    u32 idx;
    u8 *q = b2_mgr_alloc(T->pool, T->mgr[sid], &idx);
    b2_ptr_idx(s, node)[i] = idx + 1;
    return (u64)q;
  });
  return (u8 *)p;
}

// Makes chunk number k of a dynamic node exist and returns it.
__device__ __forceinline__ B2DynChunk *b2_dynamic_chunk(B2TreeDev *T, int sid, B2DynNode *d, int k) {
  u64 *pc = &d->first;
  B2DynChunk *c = nullptr;
  for (int q = 0; q <= k; q++) {
    u64 p = b2_ensure(pc, [&]() {
      u32 idx;
      B2DynChunk *nc = (B2DynChunk *)b2_mgr_alloc(T->pool, T->mgr[sid], &idx);
      if (nc) {
        nc->next = 0;  // recycled nodes are zero-filled, fresh ones too; explicit for clarity
        nc->idx1 = idx + 1;
      }
      return (u64)nc;
    });
    c = (B2DynChunk *)p;
    if (!c) return nullptr;
    pc = &c->next;
  }
  return c;
}

// <Type>_activate
__device__ __forceinline__ void b2_activate_cell(B2TreeDev *T, int sid, u8 *node, int i) {
  const B2SNodeDev &s = T->s[sid];
  switch (s.type) {
    case B2T_POINTER: b2_pointer_activate(T, sid, node, i); break;
    case B2T_BITMASKED: {
      u32 *w = b2_mask_words(s, node) + (i >> 5);
      u32 bit = 1u << (i & 31);
      if (!(*(volatile u32 *)w & bit)) atomicOr(w, bit);
      break;
    }
    case B2T_DYNAMIC: {  // Dynamic_activate (node_dynamic.h:16-42)
      B2DynNode *d = (B2DynNode *)node;
      atomicMax(&d->n, (u32)(i + 1));
      b2_dynamic_chunk(T, sid, d, i / s.chunk_size);
      break;
    }
    default: break;
  }
}

// <Type>_deactivate
__device__ __forceinline__ void b2_deactivate_cell(B2TreeDev *T, int sid, u8 *node, int i) {
  const B2SNodeDev &s = T->s[sid];
  switch (s.type) {
    case B2T_POINTER: {
      u64 old = b2_ld_volatile((u64 *)node + i);
      if (old > B2_BUSY && atomicCAS((u64 *)node + i, old, 0ull) == old) {
        u32 idx1 = b2_ptr_idx(s, node)[i];
        b2_mgr_recycle(T->pool, T->mgr[sid], idx1 - 1);
      }
      break;
    }
    case B2T_BITMASKED: {
      u32 *w = b2_mask_words(s, node) + (i >> 5);
      u32 bit = 1u << (i & 31);
      if (*(volatile u32 *)w & bit) atomicAnd(w, ~bit);
      break;
    }
    case B2T_DYNAMIC: {  // Dynamic_deactivate (node_dynamic.h:44-59): whole list, i ignored
      B2DynNode *d = (B2DynNode *)node;
      u32 n = atomicExch(&d->n, 0u);
      if (n > 0) {
        u64 c = atomicExch(&d->first, 0ull);
        while (c > B2_BUSY) {
          B2DynChunk *ch = (B2DynChunk *)c;
          u64 nx = ch->next;
          b2_mgr_recycle(T->pool, T->mgr[sid], ch->idx1 - 1);
          c = nx;
        }
      }
      break;
    }
    default: break;
  }
}

// <Type>_lookup_element (non-activating): returns the cell's storage (for pointer cells: the
// child node, or the ambient zero node if inactive).
__device__ __forceinline__ u8 *b2_lookup(B2TreeDev *T, int sid, u8 *node, int i) {
  const B2SNodeDev &s = T->s[sid];
  switch (s.type) {
    case B2T_ROOT: return node;
    case B2T_POINTER: {
      u64 p = b2_ld_volatile((u64 *)node + i);
      return p > B2_BUSY ? (u8 *)p : (u8 *)T->mgr[sid]->ambient;
    }
    case B2T_DYNAMIC: {
      B2DynNode *d = (B2DynNode *)node;
      if ((u32)i >= *(volatile u32 *)&d->n) return (u8 *)T->mgr[sid]->ambient + 16;
      u64 c = b2_ld_volatile(&d->first);
      for (int k = i / s.chunk_size; k > 0 && c > B2_BUSY; k--) c = b2_ld_volatile(&((B2DynChunk *)c)->next);
      if (c <= B2_BUSY) return (u8 *)T->mgr[sid]->ambient + 16;
      return (u8 *)c + 16 + (long long)(i % s.chunk_size) * s.cell_size;
    }
    default: return node + s.cell_size * i;
  }
}

// <Type>_get_num_elements
__device__ __forceinline__ int b2_num_elements(const B2SNodeDev &s, u8 *node) {
  if (s.type == B2T_ROOT) return 1;
  if (s.type == B2T_DYNAMIC) return (int)*(volatile u32 *)&((B2DynNode *)node)->n;
  return (int)s.num_cells;
}

// refine_coordinates (struct_llvm.cpp:171-188): out[a] = in[a]*shape_a + digit_a(l)
__device__ __forceinline__ void b2_refine(const B2SNodeDev &s, const int *in, int *out, int l) {
#pragma unroll
  for (int a = 0; a < B2_MAX_AXES; a++) out[a] = in[a] * s.shape[a];
  if (s.pot) {
    for (int k = 0; k < s.nax; k++) {
      int a = s.ax[k];
      out[a] += (int)(((u32)l >> s.acc_shift[a]) & (u32)(s.shape[a] - 1));
    }
    return;
  }
  for (int k = 0; k < s.nax; k++) {
    int a = s.ax[k];
    out[a] += (int)(((u32)l / (u32)s.acc[a]) % (u32)s.shape[a]);
  }
}

// linearised cell index of global coordinate I at level l of the path
// (scalar_pointer_lowerer.cpp:32-86 + LinearizeStmt)
__device__ __forceinline__ int b2_level_index(const B2TreeDev *T, const B2Path &P, int l, const int *I) {
  const B2SNodeDev &s = T->s[P.id[l]];
  int v = 0;
  if (P.pot) {
    for (int k = 0; k < s.nax; k++) {
      int a = s.ax[k];
      v = v * s.shape[a] + (int)(((u32)I[a] >> P.shift[l][a]) & (u32)(s.shape[a] - 1));
    }
    return v;
  }
  for (int k = 0; k < s.nax; k++) {
    int a = s.ax[k];
    int e = (int)(((u32)I[a] / (u32)P.div[l][a]) % (u32)s.shape[a]);
    v = v * s.shape[a] + e;
  }
  return v;
}

// Walks root..target. ACT: activate every level above the target (lower_access.cpp:254-279).
// Where stores go after an allocation failure (never read back).
static __device__ unsigned long long b2_sink[8];
// Returns the container of the target SNode (or an ambient node) and the cell index in it.
// *inactive is set when a non-activating walk crossed an inactive cell.
template <bool ACT>
__device__ __forceinline__ u8 *b2_walk(B2TreeDev *T, const B2Path &P, const int *I, int *cell, bool *inactive) {
  u8 *node = T->root;
  bool dead = false;
  for (int l = 0; l + 1 < P.n; l++) {
    int sid = P.id[l];
    int i = b2_level_index(T, P, l, I);
    if (ACT && !dead) {
      if (T->s[sid].type == B2T_POINTER) {
        node = b2_pointer_activate(T, sid, node, i);
        if (!node) {  // pool exhausted (error already recorded): the rest of the walk only LOOKS UP inside the ambient node --
          node = (u8 *)T->mgr[sid]->ambient;  // nothing is activated or stored through it; callers test *inactive and
          dead = true;                        // send their store to b2_sink, so later reads of inactive cells stay zero
        }
      } else {
        b2_activate_cell(T, sid, node, i);
        node = b2_lookup(T, sid, node, i);
      }
    } else {
      if (!b2_is_active(T->s[sid], node, i)) dead = true;
      node = b2_lookup(T, sid, node, i);
    }
    node += T->s[P.id[l + 1]].offset_in_parent_cell;
  }
  *cell = b2_level_index(T, P, P.n - 1, I);
  if (inactive) *inactive = dead;
  return node;
}
#endif  // __CUDACC__
