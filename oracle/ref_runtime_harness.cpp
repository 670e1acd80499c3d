/*
 * oracle/ref_runtime_harness.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * Drives the REFERENCE's own device runtime, compiled unchanged for x64 from
 * where it lies (/root/reference/taichi/runtime/llvm/runtime_module/runtime.cpp;
 * no reference source is copied into this repo).  Build: oracle/Makefile ->
 * oracle/_ref/libref_runtime.so (git-ignored, travels with gpurun).
 *
 * What is the reference here: ListManager, NodeManager, Element,
 * element_listgen_root/nonroot, clear_list, Pointer_/Bitmasked_/Dynamic_/Dense_/Root_*,
 * node_gc, runtime_initialize*.  What this file restates by hand (the LLVM-generated
 * glue that does not exist as C++ in the reference): StructMeta filling
 * (codegen_llvm.cpp:259-311), refine_coordinates / get_ch (struct_llvm.cpp:146-233),
 * per-level index extraction (scalar_pointer_lowerer.cpp:32-86), activation along an
 * access path (lower_access.cpp:254-279), the clear+listgen task plan
 * (offload.cpp:174-222) and the deactivate_all recursion (python/taichi/lang/snode.py:286-304).
 */
#define ARCH_x64 1
#include "taichi/runtime/llvm/runtime_module/runtime.cpp"

#include <cstdio>
#include <cstdlib>
#include <vector>
#include <utility>

extern "C" {
#include "snode_layout.h"
}

namespace {

OrcTree g_tree;
LLVMRuntime *g_rt = nullptr;
RuntimeContext g_ctx;
uint64 g_result[taichi_result_buffer_entries];
Ptr g_root = nullptr;
std::vector<void *> g_allocs;

union AnyMeta {
  StructMeta base;
  DenseMeta dense;
  PointerMeta pointer;
  BitmaskedMeta bitmasked;
  DynamicMeta dynamic;
  RootMeta root;
  AnyMeta() {
  }
};
AnyMeta g_meta[ORC_MAX_SNODES];

void *host_alloc(void *, std::size_t size, std::size_t align) {
  void *p = nullptr;
  if (align < sizeof(void *))
    align = sizeof(void *);
  if (posix_memalign(&p, align, size) != 0)
    return nullptr;
  std::memset(p, 0, size);
  g_allocs.push_back(p);
  return p;
}

int quiet_printf(const char *, ...) {
  return 0;
}
template <int ID>
void refine_tpl(PhysicalCoordinates *in, PhysicalCoordinates *out, int l) {
  orc_refine(&g_tree.s[ID], in->val, out->val, l);
}
template <int ID>
Ptr getch_tpl(Ptr parent_cell) {
  return parent_cell + g_tree.s[ID].offset_in_parent_cell;
}

using refine_fn = void (*)(PhysicalCoordinates *, PhysicalCoordinates *, int);
using getch_fn = Ptr (*)(Ptr);
refine_fn g_refine[ORC_MAX_SNODES];
getch_fn g_getch[ORC_MAX_SNODES];

template <int... I>
void fill_tables(std::integer_sequence<int, I...>) {
  ((g_refine[I] = &refine_tpl<I>, g_getch[I] = &getch_tpl<I>), ...);
}

void fill_meta(int id) {
  OrcSNode &s = g_tree.s[id];
  StructMeta &m = g_meta[id].base;
  std::memset(&g_meta[id], 0, sizeof(AnyMeta));
  m.snode_id = id;
  m.element_size = (std::size_t)s.element_size;
  m.max_num_elements = s.num_cells;
  m.context = &g_ctx;
  m.refine_coordinates = g_refine[id];
  m.from_parent_element = g_getch[id];
  switch (s.type) {
    case ORC_ROOT:
      m.lookup_element = Root_lookup_element;
      m.is_active = Root_is_active;
      m.get_num_elements = Root_get_num_elements;
      break;
    case ORC_DENSE:
      m.lookup_element = Dense_lookup_element;
      m.is_active = Dense_is_active;
      m.get_num_elements = Dense_get_num_elements;
      break;
    case ORC_POINTER:
      m.lookup_element = Pointer_lookup_element;
      m.is_active = Pointer_is_active;
      m.get_num_elements = Pointer_get_num_elements;
      break;
    case ORC_BITMASKED:
      m.lookup_element = Bitmasked_lookup_element;
      m.is_active = Bitmasked_is_active;
      m.get_num_elements = Bitmasked_get_num_elements;
      break;
    case ORC_DYNAMIC:
      m.lookup_element = Dynamic_lookup_element;
      m.is_active = Dynamic_is_active;
      m.get_num_elements = Dynamic_get_num_elements;
      g_meta[id].dynamic.chunk_size = s.chunk_size;
      break;
    default:
      break;
  }
}

void do_activate(int id, Ptr node, int i) {
  Ptr meta = (Ptr)&g_meta[id];
  switch (g_tree.s[id].type) {
    case ORC_POINTER:
      Pointer_activate(meta, node, i);
      break;
    case ORC_BITMASKED:
      Bitmasked_activate(meta, node, i);
      break;
    case ORC_DYNAMIC:
      Dynamic_activate(meta, node, i);
      break;
    default:
      break;
  }
}

/* Walk root -> container of `target`, as a lowered global pointer would
   (SNodeLookupStmt[+activate] + GetChStmt per level).  Returns the container of
   `target` and the linear cell index inside it.  With activate=false inactive
   pointer cells resolve to the ambient element, like the reference. */
Ptr walk(int target, const int32_t *I, bool activate, int *cell_out, bool *through_inactive) {
  int path[ORC_MAX_SNODES], lin[ORC_MAX_SNODES];
  int plen = orc_path(&g_tree, target, path);
  orc_level_indices(&g_tree, path, plen, I, lin);
  Ptr node = g_root;
  if (through_inactive)
    *through_inactive = false;
  for (int l = 0; l + 1 < plen; l++) {
    int id = path[l];
    StructMeta &m = g_meta[id].base;
    if (activate)
      do_activate(id, node, lin[l]);
    else if (through_inactive && !m.is_active((Ptr)&m, node, lin[l]))
      *through_inactive = true;
    Ptr cell = m.lookup_element((Ptr)&m, node, lin[l]);
    node = g_meta[path[l + 1]].base.from_parent_element(cell);
  }
  *cell_out = lin[plen - 1];
  return node;
}

int listgen_to(int target) {
  int path[ORC_MAX_SNODES];
  int plen = orc_path(&g_tree, target, path);
  for (int i = 1; i < plen; i++) {
    StructMeta *parent = &g_meta[path[i - 1]].base;
    StructMeta *child = &g_meta[path[i]].base;
    clear_list(g_rt, parent, child);
    if (i == 1)
      element_listgen_root(g_rt, parent, child);
    else
      element_listgen_nonroot(g_rt, parent, child);
  }
  return g_rt->element_lists[target]->size();
}

}  // namespace

extern "C" {

int refrt_create(const OrcSNodeDesc *desc, int n) {
  for (void *p : g_allocs)
    free(p);
  g_allocs.clear();
  if (orc_build_tree(&g_tree, desc, n) != 0)
    return -1;
  fill_tables(std::make_integer_sequence<int, ORC_MAX_SNODES>{});
  std::memset(g_result, 0, sizeof(g_result));
  runtime_initialize((Ptr)g_result, nullptr, 0, nullptr, 1, (void *)host_alloc, (void *)quiet_printf, (void *)vsnprintf);
  g_rt = (LLVMRuntime *)g_result[taichi_result_buffer_ret_value_id];
  g_ctx = RuntimeContext{};
  g_ctx.runtime = g_rt;
  g_ctx.result_buffer = g_result;
  std::size_t root_size = (std::size_t)g_tree.s[0].node_size;
  std::size_t rounded = taichi::iroundup(root_size ? root_size : 1, taichi_page_size);
  g_root = (Ptr)host_alloc(nullptr, rounded, 4096);
  runtime_initialize_snodes(g_rt, root_size, 0, n, 0, rounded, g_root, false);
  for (int i = 0; i < n; i++) {
    fill_meta(i);
    int ty = g_tree.s[i].type;
    if (ty == ORC_POINTER || ty == ORC_DYNAMIC) {
      runtime_NodeAllocator_initialize(g_rt, i, (std::size_t)g_tree.s[i].alloc_node_size);
      runtime_allocate_ambient(g_rt, i, (std::size_t)g_tree.s[i].alloc_node_size);
    }
  }
  return 0;
}

/* layout queries (so tests can pin the layout restatement itself) */
int64_t refrt_node_size(int id) { return g_tree.s[id].node_size; }
int64_t refrt_cell_size(int id) { return g_tree.s[id].cell_size; }
int64_t refrt_num_cells(int id) { return g_tree.s[id].num_cells; }

/* coords: n x 12 int32, indexed by physical axis.  Activates every level above
   and including `target` the way a write to a place below `target` would. */
int refrt_activate(int target, const int32_t *coords, int n) {
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, true, &cell, nullptr);
    do_activate(target, node, cell);
  }
  return 0;
}

int refrt_is_active(int target, const int32_t *coords, int n, uint8_t *out) {
  for (int k = 0; k < n; k++) {
    int cell;
    bool inactive_above;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, false, &cell, &inactive_above);
    StructMeta &m = g_meta[target].base;
    out[k] = (!inactive_above && m.is_active((Ptr)&m, node, cell)) ? 1 : 0;
  }
  return 0;
}

/* ti.deactivate(snode, I): non-recursive (docs/lang/articles/basic/sparse.md:345-351) */
int refrt_deactivate(int target, const int32_t *coords, int n) {
  for (int k = 0; k < n; k++) {
    int cell;
    bool inactive_above;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, false, &cell, &inactive_above);
    if (inactive_above)
      continue;
    Ptr meta = (Ptr)&g_meta[target];
    switch (g_tree.s[target].type) {
      case ORC_POINTER:
        Pointer_deactivate(meta, node, cell);
        break;
      case ORC_BITMASKED:
        Bitmasked_deactivate(meta, node, cell);
        break;
      case ORC_DYNAMIC:
        Dynamic_deactivate(meta, node);
        break;
      default:
        break;
    }
  }
  return 0;
}

int refrt_listgen(int target) { return listgen_to(target); }

/* out: cnt x 14 int32 = pcoord[12], loop_bounds[2] */
int refrt_get_list(int target, int32_t *out, int cap) {
  ListManager *l = g_rt->element_lists[target];
  int cnt = l->size();
  for (int k = 0; k < cnt && k < cap; k++) {
    Element &e = l->get<Element>(k);
    for (int a = 0; a < 12; a++)
      out[k * 14 + a] = e.pcoord.val[a];
    out[k * 14 + 12] = e.loop_bounds[0];
    out[k * 14 + 13] = e.loop_bounds[1];
  }
  return cnt;
}

/* struct-for over `target` (parallel_struct_for + the generated body scaffold,
   codegen_llvm.cpp:2089-2317): clear+listgen down to target, then visit every cell
   j in [lb0,lb1) of every Element, refined coordinates, skipping inactive cells of
   pointer/bitmasked/dynamic leaves.  out: cnt x 12 int32 (may be NULL to count). */
int refrt_struct_for_coords(int target, int32_t *out, int cap) {
  int cnt = listgen_to(target);
  ListManager *l = g_rt->element_lists[target];
  StructMeta &m = g_meta[target].base;
  int k = 0;
  for (int e_ = 0; e_ < cnt; e_++) {
    Element e = l->get<Element>(e_);
    for (int j = e.loop_bounds[0]; j < e.loop_bounds[1]; j++) {
      if (!m.is_active((Ptr)&m, e.element, j))
        continue;
      if (out && k < cap) {
        PhysicalCoordinates rc;
        m.refine_coordinates(&e.pcoord, &rc, j);
        for (int a = 0; a < 12; a++)
          out[k * 12 + a] = rc.val[a];
      }
      k++;
    }
  }
  return k;
}

int refrt_gc(int id) {
  if (!g_rt->node_allocators[id])
    return -1;
  node_gc(g_rt, id);
  return 0;
}

/* snode_deactivate kernel (python/taichi/_kernels.py:264-273) + gc task after it
   (offload.cpp:671-695), recursing children first (snode.py:286-304). */
static void deactivate_all_rec(int id) {
  OrcSNode &s = g_tree.s[id];
  for (int c = 0; c < s.nch; c++)
    deactivate_all_rec(s.ch[c]);
  if (s.type == ORC_POINTER || s.type == ORC_BITMASKED) {
    int cnt = listgen_to(id);
    ListManager *l = g_rt->element_lists[id];
    StructMeta &m = g_meta[id].base;
    for (int k = 0; k < cnt; k++) {
      Element e = l->get<Element>(k);
      for (int j = e.loop_bounds[0]; j < e.loop_bounds[1]; j++) {
        if (!m.is_active((Ptr)&m, e.element, j))
          continue;
        if (s.type == ORC_POINTER)
          Pointer_deactivate((Ptr)&m, e.element, j);
        else
          Bitmasked_deactivate((Ptr)&m, e.element, j);
      }
    }
    if (s.type == ORC_POINTER)
      node_gc(g_rt, id);
  } else if (s.type == ORC_DYNAMIC) {
    /* struct-for over the parent's cells, deactivate(b, I) on each */
    int par = s.parent;
    int cnt = listgen_to(par);
    ListManager *l = g_rt->element_lists[par];
    StructMeta &pm = g_meta[par].base;
    StructMeta &m = g_meta[id].base;
    for (int k = 0; k < cnt; k++) {
      Element e = l->get<Element>(k);
      for (int j = e.loop_bounds[0]; j < e.loop_bounds[1]; j++) {
        if (!pm.is_active((Ptr)&pm, e.element, j))
          continue;
        Ptr cell = pm.lookup_element((Ptr)&pm, e.element, j);
        Dynamic_deactivate((Ptr)&m, m.from_parent_element(cell));
      }
    }
    node_gc(g_rt, id);
  }
}
int refrt_deactivate_all(int id) {
  deactivate_all_rec(id);
  return 0;
}

/* out = {data_list size (= _num_dynamically_allocated), free_list size,
          recycled_list size, free_list_used} */
int refrt_alloc_stats(int id, int32_t *out) {
  NodeManager *nm = g_rt->node_allocators[id];
  if (!nm)
    return -1;
  out[0] = nm->data_list->size();
  out[1] = nm->free_list->size();
  out[2] = nm->recycled_list->size();
  out[3] = nm->free_list_used;
  return 0;
}

/* place accessors; `place` is a place snode id.  write activates, read does not. */
int refrt_write(int place, const int32_t *coords, int n, const void *vals, int accumulate_f32) {
  int par = g_tree.s[place].parent;
  int sz = g_tree.s[place].dtype_bytes;
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(par, coords + k * ORC_MAX_AXES, true, &cell, nullptr);
    do_activate(par, node, cell);
    StructMeta &m = g_meta[par].base;
    Ptr c = m.lookup_element((Ptr)&m, node, cell) + g_tree.s[place].offset_in_parent_cell;
    if (accumulate_f32)
      *(float *)c += ((const float *)vals)[k];
    else
      std::memcpy(c, (const char *)vals + (size_t)k * sz, sz);
  }
  return 0;
}
int refrt_read(int place, const int32_t *coords, int n, void *vals) {
  int par = g_tree.s[place].parent;
  int sz = g_tree.s[place].dtype_bytes;
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(par, coords + k * ORC_MAX_AXES, false, &cell, nullptr);
    StructMeta &m = g_meta[par].base;
    Ptr c = m.lookup_element((Ptr)&m, node, cell) + g_tree.s[place].offset_in_parent_cell;
    std::memcpy((char *)vals + (size_t)k * sz, c, sz);
  }
  return 0;
}

/* Dynamic: ti.append(parent cell coords, value) / ti.length */
int refrt_dynamic_append_i32(int dyn, const int32_t *coords, int n, const int32_t *vals) {
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(dyn, coords + k * ORC_MAX_AXES, true, &cell, nullptr);
    i32 len;
    Ptr p = Dynamic_allocate((Ptr)&g_meta[dyn], node, &len);
    *(int32_t *)p = vals[k];
  }
  return 0;
}
int refrt_dynamic_length(int dyn, const int32_t *coords, int n, int32_t *out) {
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(dyn, coords + k * ORC_MAX_AXES, false, &cell, nullptr);
    out[k] = Dynamic_get_num_elements((Ptr)&g_meta[dyn], node);
  }
  return 0;
}

/* The reference's own NodeManager known-answer tests
   (taichi/runtime/llvm/runtime_module/internal_functions.h:65-155), run as shipped. */
int refrt_selftest_list_manager() { return test_list_manager(&g_ctx); }
int refrt_selftest_node_allocator() { return test_node_allocator(&g_ctx); }
int refrt_selftest_node_allocator_gc() { return test_node_allocator_gc_cpu(&g_ctx); }

}  // extern "C"
