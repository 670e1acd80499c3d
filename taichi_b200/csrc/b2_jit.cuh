// b2_jit.cuh -- device side of run-time compiled struct-for bodies.
//
// Stands in for what the reference generates per kernel: the struct-for task body
// (codegen_llvm.cpp:2089-2317), its loop-index / global-pointer lowering (lower_access.cpp:42-281,
// scalar_pointer_lowerer.cpp:32-86) and the JIT session (runtime/cuda/jit_cuda.cpp:8-240).  The user
// writes the body as CUDA C++:
//
//     B2_BODY(c) {                                   // for I in ti.grouped(loop_snode):
//       float v = c.read<float>(0, c.I);             //   v = x[I]
//       int J[B2_MAX_AXES]; c.offset(J, 0, +1);      //   J = I + (1, 0, ..)
//       c.write<float>(1, J, 2.0f * v);              //   y[J] = 2 v        (activates, like the reference)
//       atomicAdd(c.arg<float *>(0), v);             //   s[None] += v
//     }
//
// and / or a particle loop (`for i in range(n)`, the reference's range-for task):
//
//     B2_RANGE_BODY(c) {                             // for p in range(n):
//       const float3 xp = c.arg<const float3 *>(0)[c.i];
//       int J[B2_MAX_AXES] = {int(xp.x * 64), int(xp.y * 64), int(xp.z * 64)};
//       atomicAdd(c.ref<float>(0, J), 1.0f);         //   m[J] += 1   (activates the cell)
//     }
//
// The library wraps it with this header (and b2_internal.cuh), compiles it for sm_100a with NVRTC
// and launches b2_jit_struct_for over the element list of the loop SNode / b2_jit_range_for grid-strided.  Field f = f-th entry of the
// place list given at launch; arg k = k-th 64-bit launch argument (scalars bit-cast, device pointers).
#pragma once
#define B2_JIT_MAX_FIELDS 8
#define B2_JIT_MAX_ARGS 16

struct B2JitField {
  B2Path path;     // root .. the SNode that holds the place
  long long off;   // byte offset of the place inside that SNode's cell
  int bytes, holder;
};
struct B2JitTable {
  int n_fields, n_args;
  unsigned long long args[B2_JIT_MAX_ARGS];
  B2JitField f[B2_JIT_MAX_FIELDS];
};

#ifdef __CUDACC_RTC__  // everything below exists only in the run-time compiled translation unit
// Field access shared by both loop kinds (global pointer lowering, lower_access.cpp:42-281).
struct B2Access {
  B2TreeDev *T;
  const B2JitTable *tab;

  template <typename V>
  __device__ __forceinline__ V arg(int k) const {
    static_assert(sizeof(V) <= 8, "launch arguments are 64-bit slots");
    union {
      unsigned long long u;
      V v;
    } x;
    x.u = tab->args[k];
    return x.v;
  }
  // Address of field f at index J.  ACT: activate every level on the way like a write access in the
  // reference (lower_access.cpp:254-279); otherwise inactive cells resolve to the ambient zero element.
  template <typename V, bool ACT>
  __device__ __forceinline__ V *walk(int f, const int *J) const {
    const B2JitField &F = tab->f[f];
    int cell;
    u8 *node = b2_walk<ACT>(T, F.path, J, &cell, nullptr);
    if (ACT) b2_activate_cell(T, F.holder, node, cell);
    return (V *)(b2_lookup(T, F.holder, node, cell) + F.off);
  }
  // ti.is_active(ancestor, ti.rescale_index(field, ancestor, J)): `up` levels above the SNode that holds field f
  // (up = 0: the holder itself, 1: its parent SNode, ...), J at the field's own resolution
  __device__ __forceinline__ bool is_active_up(int f, const int *J, int up) const {
    const B2Path &P = tab->f[f].path;
    const int q = P.n - 1 - up;  // level of the queried SNode on the path (level 0 = root)
    if (q < 1) return true;
    u8 *node = T->root;
    for (int l = 0; l < q; l++) {
      const int sid = P.id[l], i = b2_level_index(T, P, l, J);
      if (!b2_is_active(T->s[sid], node, i)) return false;
      node = b2_lookup(T, sid, node, i) + T->s[P.id[l + 1]].offset_in_parent_cell;
    }
    return b2_is_active(T->s[P.id[q]], node, b2_level_index(T, P, q, J));
  }
  // ti.deactivate(ancestor, ti.rescale_index(field, ancestor, J)): non-recursive, like the reference (sparse.md:345-351).
  // Pointer cells hand their node to the recycle list; the caller runs the SNode's GC after the kernel, which is
  // what the reference's offloader inserts behind every task that deactivates (offload.cpp:671-695).
  __device__ __forceinline__ void deactivate_up(int f, const int *J, int up) const {
    const B2Path &P = tab->f[f].path;
    const int q = P.n - 1 - up;
    if (q < 1) return;
    u8 *node = T->root;
    for (int l = 0; l < q; l++) {
      const int sid = P.id[l], i = b2_level_index(T, P, l, J);
      if (!b2_is_active(T->s[sid], node, i)) return;  // nothing below an inactive cell
      node = b2_lookup(T, sid, node, i) + T->s[P.id[l + 1]].offset_in_parent_cell;
    }
    b2_deactivate_cell(T, P.id[q], node, b2_level_index(T, P, q, J));
  }
  // ti.append(holder of field f, J, v) for a field under a dynamic SNode (node_dynamic.h:61-88): J addresses the parent
  // cell (the dynamic axis is ignored), ancestors are activated, returns the index the value landed at.
  template <typename V>
  __device__ __forceinline__ int append(int f, const int *J, V v) const {
    const B2JitField &F = tab->f[f];
    const B2SNodeDev &s = T->s[F.holder];
    int cell;
    B2DynNode *d = (B2DynNode *)b2_walk<true>(T, F.path, J, &cell, nullptr);
    const u32 i = atomicAdd(&d->n, 1u);
    B2DynChunk *ch = b2_dynamic_chunk(T, F.holder, d, (int)(i / (u32)s.chunk_size));
    if (ch) *(V *)((u8 *)ch + 16 + (long long)(i % (u32)s.chunk_size) * s.cell_size + F.off) = v;
    return (int)i;
  }
  // ti.length(holder of field f, J)
  __device__ __forceinline__ int length(int f, const int *J) const {
    int cell;
    const B2DynNode *d = (const B2DynNode *)b2_walk<false>(T, tab->f[f].path, J, &cell, nullptr);
    return (int)*(const volatile u32 *)&d->n;
  }
  // ti.is_active(holder of field f, J)
  __device__ __forceinline__ bool is_active(int f, const int *J) const {
    const B2JitField &F = tab->f[f];
    int cell;
    bool dead;
    u8 *node = b2_walk<false>(T, F.path, J, &cell, &dead);
    return !dead && b2_is_active(T->s[F.holder], node, cell);
  }
};

// struct-for: `for I in ti.grouped(loop_snode)`
struct B2Ctx : B2Access {
  const B2SNodeDev *ts;  // the loop SNode
  const B2Element *E;    // the work item this cell belongs to
  int sid, j;            // loop SNode id, linear cell index inside E->node
  int I[B2_MAX_AXES];    // loop index (refine_coordinates, struct_llvm.cpp:171-188)

  template <typename V, bool ACT>
  __device__ __forceinline__ V *ptr(int f, const int *J) const {
    const B2JitField &F = tab->f[f];
    if (!ACT && F.holder == sid) {  // the loop's own cell at the loop index: no walk (a struct-for never re-activates its own path)
      bool same = true;
#pragma unroll
      for (int a = 0; a < B2_MAX_AXES; a++) same = same && J[a] == I[a];
      if (same) return (V *)(b2_lookup(T, sid, E->node, j) + F.off);
    }
    return walk<V, ACT>(f, J);
  }
  template <typename V>
  __device__ __forceinline__ V read(int f, const int *J) const { return *ptr<V, false>(f, J); }
  template <typename V>
  __device__ __forceinline__ void write(int f, const int *J, V v) const { *ptr<V, true>(f, J) = v; }
  template <typename V>
  __device__ __forceinline__ V *ref(int f, const int *J) const { return ptr<V, true>(f, J); }  // for atomics / +=
  // J = I with axis a shifted by d
  __device__ __forceinline__ void offset(int *J, int a, int d) const {
#pragma unroll
    for (int q = 0; q < B2_MAX_AXES; q++) J[q] = I[q];
    J[a] += d;
  }
};

// range-for: `for i in range(begin, end)` (particle loops; gpu_parallel_range_for, runtime.cpp:1541-1565)
struct B2RangeCtx : B2Access {
  long long i;
  template <typename V>
  __device__ __forceinline__ V read(int f, const int *J) const { return *walk<V, false>(f, J); }
  template <typename V>
  __device__ __forceinline__ void write(int f, const int *J, V v) const { *walk<V, true>(f, J) = v; }
  template <typename V>
  __device__ __forceinline__ V *ref(int f, const int *J) const { return walk<V, true>(f, J); }
};

#define B2_BODY(c) __device__ __forceinline__ void b2_body(B2Ctx &c)
#define B2_RANGE_BODY(c) __device__ __forceinline__ void b2_range_body(B2RangeCtx &c)

#ifdef B2_JIT_HAS_BODY
__device__ __forceinline__ void b2_body(B2Ctx &c);
// parallel_struct_for (runtime.cpp:1422-1464): CTA <-> Element, thread <-> cell
extern "C" __global__ void __launch_bounds__(256) b2_jit_struct_for(B2TreeDev *T, int sid, const B2JitTable *tab) {
  const B2SNodeDev &ts = T->s[sid];
  B2CList *L = T->lists[sid];
  const u32 ne = L->count;
  for (u32 e = blockIdx.x; e < ne; e += gridDim.x) {
    const B2Element E = *(const B2Element *)b2_list_get(L, e);
    for (int j0 = E.lo; j0 < E.hi; j0 += blockDim.x) {
      const int j = j0 + threadIdx.x;
      if (j < E.hi && b2_is_active(ts, E.node, j)) {
        B2Ctx c;
        c.T = T, c.tab = tab, c.ts = &ts, c.E = &E, c.sid = sid, c.j = j;
        b2_refine(ts, E.pcoord, c.I, j);
        b2_body(c);
      }
    }
  }
}
#endif
#ifdef B2_JIT_HAS_RANGE
__device__ __forceinline__ void b2_range_body(B2RangeCtx &c);
extern "C" __global__ void __launch_bounds__(256) b2_jit_range_for(B2TreeDev *T, long long begin, long long end, const B2JitTable *tab) {
  for (long long i = begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (long long)gridDim.x * blockDim.x) {
    B2RangeCtx c;
    c.T = T, c.tab = tab, c.i = i;
    b2_range_body(c);
  }
}
#endif
#endif  // __CUDACC_RTC__
