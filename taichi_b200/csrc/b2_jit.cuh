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
// One block-local pad = ti.block_local(field) on a struct-for (transforms/make_block_local.cpp:97-365, ir/scratch_pad.h:30-180).
// The loop's block is the Element's parent cell; per axis d of the pad (physical axis `axis[d]`) it covers
// `extent[d]` cells of the FIELD's index space starting at corner = parent_coordinate * extent (BlockCornerIndexStmt,
// make_block_local.cpp:196-203), and the body reaches offsets lo..hi around its cells, so the pad holds
// bounds [lo, extent + hi) -- ScratchPad::bounds -- i.e. size = extent + hi - lo cells.
// mode 1: read-only (prologue loads the pad; inactive / out-of-field cells read 0)
// mode 2..4: accumulate f32 / i32 / f64 (prologue zero-fills, the body atomically adds into the pad, the epilogue adds the pad
//            to the field through an ACTIVATING access: every pad cell when `halo`, else only cells that received a non-zero
//            sum -- the reference's x64 result)
#define B2_JIT_MAX_BLS 4
struct B2JitBls {
  int field, mode, ndim, halo;
  int axis[3], extent[3], lo[3], hi[3];
  int size[3], field_extent[3];
  int count, bytes, smem_off, pad;
};
struct B2JitTable {
  int n_fields, n_args;
  unsigned long long args[B2_JIT_MAX_ARGS];
  B2JitField f[B2_JIT_MAX_FIELDS];
  int n_bls, pad;
  B2JitBls bls[B2_JIT_MAX_BLS];
  unsigned long long rand_states;  // uint4 {x, y, z, w} per linear thread id
  int num_rand, pad2;
};

#ifdef __CUDACC_RTC__  // everything below exists only in the run-time compiled translation unit
__device__ __forceinline__ void b2_atomic_minmax(int *p, int v, bool mx) { mx ? atomicMax(p, v) : atomicMin(p, v); }
__device__ __forceinline__ void b2_atomic_minmax(float *p, float v, bool mx) {  // CAS loop as atomic.h:75 (f32 min / max)
  int old = *(volatile int *)p;
  for (;;) {
    const float f = __int_as_float(old);
    if (mx ? !(v > f) : !(v < f)) return;
    const int seen = atomicCAS((int *)p, old, __float_as_int(v));
    if (seen == old) return;
    old = seen;
  }
}
__device__ __forceinline__ void b2_atomic_minmax(double *p, double v, bool mx) {
  unsigned long long old = *(volatile unsigned long long *)p;
  for (;;) {
    const double f = __longlong_as_double((long long)old);
    if (mx ? !(v > f) : !(v < f)) return;
    const unsigned long long seen = atomicCAS((unsigned long long *)p, old, (unsigned long long)__double_as_longlong(v));
    if (seen == old) return;
    old = seen;
  }
}
// Field access shared by both loop kinds (global pointer lowering, lower_access.cpp:42-281).
struct B2Access {
  B2TreeDev *T;
  const B2JitTable *tab;

  // ti.random: xorshift128 with one persistent state per linear thread id of the launch, output multiplied by a prime
  // (rand_u32 / rand_f32 / rand_f64 / rand_i32, runtime.cpp:1790-1827).  The stream a loop index sees depends on the
  // launch geometry, as in the reference.
  __device__ __forceinline__ unsigned rand_u32() const {
    uint4 *st = (uint4 *)tab->rand_states + (blockIdx.x * blockDim.x + threadIdx.x) % (unsigned)tab->num_rand;
    uint4 s = *st;
    const unsigned t = s.x ^ (s.x << 11);
    s.x = s.y, s.y = s.z, s.z = s.w;
    s.w = (s.w ^ (s.w >> 19)) ^ (t ^ (t >> 8));
    *st = s;
    return s.w * 1000000007u;
  }
  __device__ __forceinline__ int rand_i32() const { return (int)rand_u32(); }
  __device__ __forceinline__ float rand_f32() const { return (float)(rand_u32() >> 8) * (1.0f / 16777216.0f); }
  __device__ __forceinline__ double rand_f64() const {
    const unsigned long long hi = rand_u32(), lo = rand_u32();
    return (double)(((hi << 32) + lo) >> 11) * (1.0 / 9007199254740992.0);
  }

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
    bool dead;
    u8 *node = b2_walk<ACT>(T, F.path, J, &cell, &dead);
    if (ACT) {
      if (dead) return (V *)b2_sink;  // out of memory (flag raised): the store lands in the sink, never in the ambient node
      b2_activate_cell(T, F.holder, node, cell);
    }
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
    bool dead;
    B2DynNode *d = (B2DynNode *)b2_walk<true>(T, F.path, J, &cell, &dead);
    if (dead) return -1;
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

// Thread-local reduction slots (make_thread_local.cpp:100-213: a 0-D destination becomes a thread-local accumulator; the
// epilogue reduces the CTA -- warp shuffles, one shared-memory round -- and issues ONE atomic per CTA, codegen_cuda.cpp:358-392).
// Slot k reduces into the device pointer passed as launch argument k.
#define B2_TLS_SLOTS 4
struct B2Tls {
  unsigned long long bits[B2_TLS_SLOTS];  // the accumulator's bit pattern (f32 / i32 / f64)
  int tag[B2_TLS_SLOTS];                  // 0 unused, else type (1 f32, 2 i32, 3 f64) | op << 4 (0 add, 1 min, 2 max)
};
template <typename V>
struct B2TlsType;
template <>
struct B2TlsType<float> { static constexpr int id = 1; };
template <>
struct B2TlsType<int> { static constexpr int id = 2; };
template <>
struct B2TlsType<double> { static constexpr int id = 3; };
template <typename V>
__device__ __forceinline__ V b2_tls_combine(V a, V b, int op) {
  return op == 0 ? a + b : op == 1 ? (b < a ? b : a) : (b > a ? b : a);
}

// struct-for: `for I in ti.grouped(loop_snode)`
struct B2Ctx : B2Access {
  const B2SNodeDev *ts;  // the loop SNode
  const B2Element *E;    // the work item this cell belongs to
  int sid, j;            // loop SNode id, linear cell index inside E->node
  int I[B2_MAX_AXES];    // loop index (refine_coordinates, struct_llvm.cpp:171-188)
  B2Tls *tls;            // this thread's reduction slots
  unsigned char *smem;   // the Element's block-local pads

  // ---- ti.block_local: pad `slot` at field index J (J must lie inside the declared bounds) ----
  template <typename V>
  __device__ __forceinline__ V *bls_ptr(int slot, const int *J) const {
    const B2JitBls &B = tab->bls[slot];
    int idx = 0;
#pragma unroll
    for (int d = 0; d < 3; d++)
      if (d < B.ndim) idx = idx * B.size[d] + (J[B.axis[d]] - E->pcoord[B.axis[d]] * B.extent[d] - B.lo[d]);
    return (V *)(smem + B.smem_off) + idx;
  }
  template <typename V>
  __device__ __forceinline__ V bls_read(int slot, const int *J) const { return *bls_ptr<V>(slot, J); }
  template <typename V>
  __device__ __forceinline__ void bls_add(int slot, const int *J, V v) const { atomicAdd(bls_ptr<V>(slot, J), v); }

  // ---- thread-local reductions: dst = launch argument k (a device pointer to V) ----
  template <typename V>
  __device__ __forceinline__ void reduce(int k, V v, int op) const {
    union {
      unsigned long long u;
      V x;
    } a;
    a.u = 0;
    if (tls->tag[k] == 0) {
      a.x = v;
      tls->tag[k] = B2TlsType<V>::id | op << 4;
    } else {
      a.u = tls->bits[k];
      a.x = b2_tls_combine<V>(a.x, v, op);
    }
    tls->bits[k] = a.u;
  }
  template <typename V>
  __device__ __forceinline__ void reduce_add(int k, V v) const { reduce<V>(k, v, 0); }
  template <typename V>
  __device__ __forceinline__ void reduce_min(int k, V v) const { reduce<V>(k, v, 1); }
  template <typename V>
  __device__ __forceinline__ void reduce_max(int k, V v) const { reduce<V>(k, v, 2); }

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
// pad cell p of pad B of the Element with parent coordinates pc -> field index J; false if it lies outside the field
__device__ __forceinline__ bool b2_bls_coord(const B2JitBls &B, const int *pc, int p, int *J) {
#pragma unroll
  for (int a = 0; a < B2_MAX_AXES; a++) J[a] = 0;
  bool inside = true;
  for (int d = B.ndim - 1; d >= 0; d--) {
    const int q = p % B.size[d];
    p /= B.size[d];
    const int x = pc[B.axis[d]] * B.extent[d] + B.lo[d] + q;
    J[B.axis[d]] = x;
    inside = inside && x >= 0 && x < B.field_extent[d];
  }
  return inside;
}
template <typename V>
__device__ __forceinline__ void b2_tls_finish(const B2Access &A, B2Tls &tls, int k, int tag, unsigned long long *part) {
  const int op = tag >> 4, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  union {
    unsigned long long u;
    V x;
  } a, b;
  a.u = tls.bits[k];
  bool has = tls.tag[k] != 0;
  // warp reduction over the lanes that hold a value, then one shared-memory round over the warps
  for (int d = 16; d > 0; d >>= 1) {
    b.u = __shfl_down_sync(0xffffffffu, a.u, d);
    const bool bh = __shfl_down_sync(0xffffffffu, (int)has, d) != 0;
    if (bh) a.x = has ? b2_tls_combine<V>(a.x, b.x, op) : b.x, has = true;
  }
  __syncthreads();
  if (lane == 0) part[warp] = a.u, part[32 + warp] = has;
  __syncthreads();
  if (warp == 0) {
    a.u = lane < nw ? part[lane] : 0ull;
    has = lane < nw && part[32 + lane] != 0;
    for (int d = 16; d > 0; d >>= 1) {
      b.u = __shfl_down_sync(0xffffffffu, a.u, d);
      const bool bh = __shfl_down_sync(0xffffffffu, (int)has, d) != 0;
      if (bh) a.x = has ? b2_tls_combine<V>(a.x, b.x, op) : b.x, has = true;
    }
    if (lane == 0 && has) {
      V *dst = A.arg<V *>(k);
      if (op == 0)
        atomicAdd(dst, a.x);
      else
        b2_atomic_minmax(dst, a.x, op == 2);
    }
  }
}
// parallel_struct_for (runtime.cpp:1422-1464): CTA <-> Element, thread <-> cell; with block-local pads the CTA stages them
// before its Element's cells run and flushes the accumulating ones afterwards (make_block_local.cpp:97-365)
extern "C" __global__ void __launch_bounds__(256) b2_jit_struct_for(B2TreeDev *T, int sid, const B2JitTable *tab) {
  extern __shared__ __align__(16) unsigned char b2_bls_smem[];
  __shared__ unsigned long long b2_tls_part[64];
  const B2SNodeDev &ts = T->s[sid];
  B2CList *L = T->lists[sid];
  const u32 ne = L->count;
  const int nbls = tab->n_bls;
  B2Tls tls;
#pragma unroll
  for (int k = 0; k < B2_TLS_SLOTS; k++) tls.bits[k] = 0ull, tls.tag[k] = 0;
  B2Access A;
  A.T = T, A.tab = tab;
  for (u32 e = blockIdx.x; e < ne; e += gridDim.x) {
    const B2Element E = *(const B2Element *)b2_list_get(L, e);
    if (nbls) {  // prologue
      __syncthreads();
      for (int b = 0; b < nbls; b++) {
        const B2JitBls &B = tab->bls[b];
        for (int p = threadIdx.x; p < B.count; p += blockDim.x) {
          int J[B2_MAX_AXES];
          unsigned long long v = 0ull;
          if (B.mode == 1 && b2_bls_coord(B, E.pcoord, p, J)) {
            if (B.bytes == 8)
              v = *A.walk<unsigned long long, false>(B.field, J);
            else
              v = *A.walk<unsigned int, false>(B.field, J);
          }
          if (B.bytes == 8)
            ((unsigned long long *)(b2_bls_smem + B.smem_off))[p] = v;
          else
            ((unsigned int *)(b2_bls_smem + B.smem_off))[p] = (unsigned int)v;
        }
      }
      __syncthreads();
    }
    for (int j0 = E.lo; j0 < E.hi; j0 += blockDim.x) {
      const int j = j0 + threadIdx.x;
      if (j < E.hi && b2_is_active(ts, E.node, j)) {
        B2Ctx c;
        c.T = T, c.tab = tab, c.ts = &ts, c.E = &E, c.sid = sid, c.j = j, c.tls = &tls, c.smem = b2_bls_smem;
        b2_refine(ts, E.pcoord, c.I, j);
        b2_body(c);
      }
    }
    if (nbls) {  // epilogue
      __syncthreads();
      for (int b = 0; b < nbls; b++) {
        const B2JitBls &B = tab->bls[b];
        if (B.mode < 2) continue;
        for (int p = threadIdx.x; p < B.count; p += blockDim.x) {
          int J[B2_MAX_AXES];
          if (!b2_bls_coord(B, E.pcoord, p, J)) continue;
          if (B.mode == 2) {
            const float v = ((const float *)(b2_bls_smem + B.smem_off))[p];
            if (v != 0.0f || B.halo) atomicAdd(A.walk<float, true>(B.field, J), v);
          } else if (B.mode == 3) {
            const int v = ((const int *)(b2_bls_smem + B.smem_off))[p];
            if (v != 0 || B.halo) atomicAdd(A.walk<int, true>(B.field, J), v);
          } else {
            const double v = ((const double *)(b2_bls_smem + B.smem_off))[p];
            if (v != 0.0 || B.halo) atomicAdd(A.walk<double, true>(B.field, J), v);
          }
        }
      }
    }
  }
  // thread-local reductions: one atomic per CTA and slot (the tag is the same in every thread that used the slot)
  __shared__ int b2_tls_tag[B2_TLS_SLOTS];
  if (threadIdx.x < B2_TLS_SLOTS) b2_tls_tag[threadIdx.x] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < B2_TLS_SLOTS; k++)
    if (tls.tag[k]) b2_tls_tag[k] = tls.tag[k];  // (every user of a slot writes the same value)
  __syncthreads();
#pragma unroll
  for (int k = 0; k < B2_TLS_SLOTS; k++) {
    const int tag = b2_tls_tag[k];
    if (tag == 0) continue;
    const int ty = tag & 15;
    if (ty == 1)
      b2_tls_finish<float>(A, tls, k, tag, b2_tls_part);
    else if (ty == 2)
      b2_tls_finish<int>(A, tls, k, tag, b2_tls_part);
    else
      b2_tls_finish<double>(A, tls, k, tag, b2_tls_part);
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
