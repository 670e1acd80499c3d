/*
 * oracle/snode_oracle.c -- TEST INFRASTRUCTURE ONLY (checker; never linked into or
 * called by the product path; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it).
 *
 * Plain-C restatement of the reference's sparse-SNode runtime as its x64 backend
 * executes it (serial listgen, runtime.cpp:1299-1302,1350-1355):
 *   ListManager append/clear          runtime.cpp:426-513,1664-1689
 *   NodeManager allocate/recycle/gc   runtime.cpp:630-707
 *   clear_list / element_listgen_*    runtime.cpp:1270-1383
 *   Pointer_*    node_pointer.h:41-99      Bitmasked_*  node_bitmasked.h:10-43
 *   Dynamic_*    node_dynamic.h:16-119     Dense_/Root_ node_dense.h, node_root.h
 * Geometry / layout / index lowering: snode_layout.h.
 *
 * Parity pinned: tests/test_oracle_vs_reference.py runs this file and
 * oracle/_ref/libref_runtime.so (the reference's own runtime.cpp) on identical inputs
 * and requires identical lists / active sets / allocator counts, and checks the
 * reference's golden counts (tests/python/test_bitmasked.py, test_sparse_deactivate.py,
 * internal_functions.h:80-155).
 */
#include "snode_layout.h"
#include <stdio.h>
#include <stdlib.h>

typedef unsigned char *Ptr;

typedef struct {
  Ptr element;
  int32_t lb[2];
  int32_t pcoord[ORC_MAX_AXES];
} OElement; /* 64 bytes, runtime.cpp:517-521 */

typedef struct {
  OElement *e;
  int n, cap;
} OList;

typedef struct {
  int64_t element_size;
  Ptr *nodes; /* data_list: one malloc per node is enough for the oracle */
  int n_nodes, cap_nodes;
  int32_t *free_list;
  int n_free, cap_free;
  int32_t *recycled;
  int n_rec, cap_rec;
  int free_used;
  Ptr ambient;
} OAlloc;

static OrcTree T;
static OList L[ORC_MAX_SNODES];
static OAlloc A[ORC_MAX_SNODES];
static Ptr ROOT;

static void list_push(OList *l, const OElement *e) {
  if (l->n == l->cap) {
    l->cap = l->cap ? l->cap * 2 : 256;
    l->e = (OElement *)realloc(l->e, sizeof(OElement) * (size_t)l->cap);
  }
  l->e[l->n++] = *e;
}
static void ipush(int32_t **v, int *n, int *cap, int32_t x) {
  if (*n == *cap) {
    *cap = *cap ? *cap * 2 : 64;
    *v = (int32_t *)realloc(*v, sizeof(int32_t) * (size_t)*cap);
  }
  (*v)[(*n)++] = x;
}

/* NodeManager::allocate (runtime.cpp:666-677) */
static Ptr alloc_node(OAlloc *a) {
  int cur = a->free_used++;
  int idx;
  if (cur >= a->n_free) {
    if (a->n_nodes == a->cap_nodes) {
      a->cap_nodes = a->cap_nodes ? a->cap_nodes * 2 : 64;
      a->nodes = (Ptr *)realloc(a->nodes, sizeof(Ptr) * (size_t)a->cap_nodes);
    }
    idx = a->n_nodes++;
    a->nodes[idx] = (Ptr)calloc(1, (size_t)a->element_size);
  } else {
    idx = a->free_list[cur];
  }
  return a->nodes[idx];
}
/* NodeManager::recycle = locate + append (runtime.cpp:679-686) */
static void recycle_node(OAlloc *a, Ptr p) {
  for (int i = 0; i < a->n_nodes; i++)
    if (a->nodes[i] == p) {
      ipush(&a->recycled, &a->n_rec, &a->cap_rec, i);
      return;
    }
}
/* NodeManager::gc_serial (runtime.cpp:688-706) */
static void gc_serial(OAlloc *a) {
  for (int i = a->free_used; i < a->n_free; i++) a->free_list[i - a->free_used] = a->free_list[i];
  int unused = a->n_free - a->free_used;
  if (unused < 0) unused = 0;
  a->free_used = 0;
  a->n_free = unused;
  for (int i = 0; i < a->n_rec; i++) {
    int idx = a->recycled[i];
    memset(a->nodes[idx], 0, (size_t)a->element_size);
    ipush(&a->free_list, &a->n_free, &a->cap_free, idx);
  }
  a->n_rec = 0;
}

/* ---- per-type cell operations ---- */
static Ptr *ptr_slot(const OrcSNode *s, Ptr node, int i) { return (Ptr *)(node + 8 * (s->num_cells + i)); }
static uint32_t *mask_base(const OrcSNode *s, Ptr node) { return (uint32_t *)(node + s->element_size * s->num_cells); }
typedef struct {
  int32_t lock, n;
  Ptr ptr;
} ODyn;

static int is_active(const OrcSNode *s, Ptr node, int i) {
  switch (s->type) {
    case ORC_POINTER: return *ptr_slot(s, node, i) != NULL;
    case ORC_BITMASKED: return (mask_base(s, node)[i / 32] >> (i % 32)) & 1u;
    case ORC_DYNAMIC: return i < ((ODyn *)node)->n;
    default: return 1;
  }
}
static void activate(const OrcSNode *s, Ptr node, int i) {
  switch (s->type) {
    case ORC_POINTER: {
      Ptr *slot = ptr_slot(s, node, i);
      if (*slot == NULL) *slot = alloc_node(&A[s->id]);
      break;
    }
    case ORC_BITMASKED: mask_base(s, node)[i / 32] |= 1u << (i % 32); break;
    case ORC_DYNAMIC: { /* Dynamic_activate, node_dynamic.h:16-42 */
      ODyn *d = (ODyn *)node;
      if (d->n < i + 1) d->n = i + 1;
      Ptr *pc = &d->ptr;
      int start = 0;
      for (;;) {
        if (*pc == NULL) *pc = alloc_node(&A[s->id]);
        if (i < start + s->chunk_size) break;
        pc = (Ptr *)*pc;
        start += s->chunk_size;
      }
      break;
    }
    default: break;
  }
}
static void deactivate(const OrcSNode *s, Ptr node, int i) {
  switch (s->type) {
    case ORC_POINTER: {
      Ptr *slot = ptr_slot(s, node, i);
      if (*slot) {
        recycle_node(&A[s->id], *slot);
        *slot = NULL;
      }
      break;
    }
    case ORC_BITMASKED: mask_base(s, node)[i / 32] &= ~(1u << (i % 32)); break;
    case ORC_DYNAMIC: { /* Dynamic_deactivate, node_dynamic.h:44-59 */
      ODyn *d = (ODyn *)node;
      if (d->n > 0) {
        d->n = 0;
        Ptr *pc = &d->ptr;
        while (*pc) {
          recycle_node(&A[s->id], *pc);
          pc = (Ptr *)*pc;
        }
        d->ptr = NULL;
      }
      break;
    }
    default: break;
  }
}
static Ptr lookup(const OrcSNode *s, Ptr node, int i) {
  switch (s->type) {
    case ORC_ROOT: return node;
    case ORC_POINTER: {
      Ptr p = *ptr_slot(s, node, i);
      return p ? p : A[s->id].ambient;
    }
    case ORC_DYNAMIC: { /* Dynamic_lookup_element, node_dynamic.h:94-114 */
      ODyn *d = (ODyn *)node;
      if (i >= d->n) return A[s->id].ambient;
      Ptr c = d->ptr;
      int start = 0;
      while (i >= start + s->chunk_size) {
        c = *(Ptr *)c;
        start += s->chunk_size;
      }
      return c + 8 + (int64_t)(i - start) * s->element_size;
    }
    default: return node + s->element_size * i; /* dense, bitmasked */
  }
}
static int num_elements(const OrcSNode *s, Ptr node) {
  if (s->type == ORC_ROOT) return 1;
  if (s->type == ORC_DYNAMIC) return ((ODyn *)node)->n;
  return (int)s->num_cells;
}

/* ---- access path walk (lower_access.cpp:227-281; scalar_pointer_lowerer.cpp) ---- */
static Ptr walk(int target, const int32_t *I, int act, int *cell_out, int *inactive_above) {
  int path[ORC_MAX_SNODES], lin[ORC_MAX_SNODES];
  int plen = orc_path(&T, target, path);
  orc_level_indices(&T, path, plen, I, lin);
  Ptr node = ROOT;
  if (inactive_above) *inactive_above = 0;
  for (int l = 0; l + 1 < plen; l++) {
    const OrcSNode *s = &T.s[path[l]];
    if (act)
      activate(s, node, lin[l]);
    else if (inactive_above && !is_active(s, node, lin[l]))
      *inactive_above = 1;
    node = lookup(s, node, lin[l]) + T.s[path[l + 1]].offset_in_parent_cell;
  }
  *cell_out = lin[plen - 1];
  return node;
}

/* element_listgen_root / _nonroot in their x64 (serial) form */
static void emit_children(OList *out, const OrcSNode *child, Ptr ch_node, const int32_t *pcoord) {
  int n = num_elements(child, ch_node);
  int step = n < ORC_LISTGEN_MAX_ELEMENT ? n : ORC_LISTGEN_MAX_ELEMENT;
  for (int lo = 0; lo < n; lo += step) {
    OElement e;
    e.element = ch_node;
    e.lb[0] = lo;
    e.lb[1] = lo + step < n ? lo + step : n;
    memcpy(e.pcoord, pcoord, sizeof(e.pcoord));
    list_push(out, &e);
  }
}
static int listgen_to(int target) {
  int path[ORC_MAX_SNODES];
  int plen = orc_path(&T, target, path);
  for (int i = 1; i < plen; i++) {
    const OrcSNode *par = &T.s[path[i - 1]], *ch = &T.s[path[i]];
    OList *pl = &L[par->id], *cl = &L[ch->id];
    cl->n = 0; /* clear_list */
    for (int k = 0; k < pl->n; k++) {
      OElement e = pl->e[k];
      for (int j = e.lb[0]; j < e.lb[1]; j++) {
        int32_t rc[ORC_MAX_AXES];
        orc_refine(par, e.pcoord, rc, j);
        if (!is_active(par, e.element, j)) continue;
        Ptr chn = lookup(par, e.element, j) + ch->offset_in_parent_cell;
        emit_children(cl, ch, chn, rc);
      }
    }
  }
  return L[target].n;
}

/* ================= exported API (mirrors refrt_* of ref_runtime_harness.cpp) ================= */
int orc_create(const OrcSNodeDesc *desc, int n) {
  /* oracle trees are small and short-lived: old storage is leaked on purpose */
  if (orc_build_tree(&T, desc, n) != 0) return -1;
  memset(L, 0, sizeof(L));
  memset(A, 0, sizeof(A));
  int64_t sz = T.s[0].node_size;
  ROOT = (Ptr)calloc(1, (size_t)orc_roundup(sz ? sz : 1, 4096));
  for (int i = 0; i < n; i++)
    if (T.s[i].type == ORC_POINTER || T.s[i].type == ORC_DYNAMIC) {
      A[i].element_size = T.s[i].alloc_node_size;
      A[i].ambient = (Ptr)calloc(1, (size_t)A[i].element_size);
    }
  OElement e;
  memset(&e, 0, sizeof(e));
  e.element = ROOT;
  e.lb[1] = 1;
  list_push(&L[0], &e); /* runtime_initialize_snodes seeds the root list (runtime.cpp:1006-1016) */
  return 0;
}
int64_t orc_node_size(int id) { return T.s[id].node_size; }
int64_t orc_cell_size(int id) { return T.s[id].cell_size; }
int64_t orc_num_cells(int id) { return T.s[id].num_cells; }

int orc_activate(int target, const int32_t *coords, int n) {
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, 1, &cell, NULL);
    activate(&T.s[target], node, cell);
  }
  return 0;
}
int orc_is_active(int target, const int32_t *coords, int n, uint8_t *out) {
  for (int k = 0; k < n; k++) {
    int cell, ia;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, 0, &cell, &ia);
    out[k] = (!ia && is_active(&T.s[target], node, cell)) ? 1 : 0;
  }
  return 0;
}
int orc_deactivate(int target, const int32_t *coords, int n) {
  for (int k = 0; k < n; k++) {
    int cell, ia;
    Ptr node = walk(target, coords + k * ORC_MAX_AXES, 0, &cell, &ia);
    if (!ia) deactivate(&T.s[target], node, cell);
  }
  return 0;
}
int orc_listgen(int target) { return listgen_to(target); }
int orc_get_list(int target, int32_t *out, int cap) {
  OList *l = &L[target];
  for (int k = 0; k < l->n && k < cap; k++) {
    memcpy(out + k * 14, l->e[k].pcoord, 48);
    out[k * 14 + 12] = l->e[k].lb[0];
    out[k * 14 + 13] = l->e[k].lb[1];
  }
  return l->n;
}
/* struct-for over `target`: see refrt_struct_for_coords */
int orc_struct_for_coords(int target, int32_t *out, int cap) {
  int cnt = listgen_to(target);
  const OrcSNode *s = &T.s[target];
  int k = 0;
  for (int q = 0; q < cnt; q++) {
    OElement e = L[target].e[q];
    for (int j = e.lb[0]; j < e.lb[1]; j++) {
      if (!is_active(s, e.element, j)) continue;
      if (out && k < cap) orc_refine(s, e.pcoord, out + k * 12, j);
      k++;
    }
  }
  return k;
}
int orc_gc(int id) {
  if (T.s[id].type != ORC_POINTER && T.s[id].type != ORC_DYNAMIC) return -1;
  gc_serial(&A[id]);
  return 0;
}
static void deactivate_all_rec(int id) {
  const OrcSNode *s = &T.s[id];
  for (int c = 0; c < s->nch; c++) deactivate_all_rec(s->ch[c]);
  if (s->type == ORC_POINTER || s->type == ORC_BITMASKED) {
    int cnt = listgen_to(id);
    for (int k = 0; k < cnt; k++) {
      OElement e = L[id].e[k];
      for (int j = e.lb[0]; j < e.lb[1]; j++)
        if (is_active(s, e.element, j)) deactivate(s, e.element, j);
    }
    if (s->type == ORC_POINTER) gc_serial(&A[id]);
  } else if (s->type == ORC_DYNAMIC) {
    const OrcSNode *p = &T.s[s->parent];
    int cnt = listgen_to(p->id);
    for (int k = 0; k < cnt; k++) {
      OElement e = L[p->id].e[k];
      for (int j = e.lb[0]; j < e.lb[1]; j++)
        if (is_active(p, e.element, j)) deactivate(s, lookup(p, e.element, j) + s->offset_in_parent_cell, 0);
    }
    gc_serial(&A[id]);
  }
}
int orc_deactivate_all(int id) {
  deactivate_all_rec(id);
  return 0;
}
int orc_alloc_stats(int id, int32_t *out) {
  if (T.s[id].type != ORC_POINTER && T.s[id].type != ORC_DYNAMIC) return -1;
  out[0] = A[id].n_nodes;
  out[1] = A[id].n_free;
  out[2] = A[id].n_rec;
  out[3] = A[id].free_used;
  return 0;
}
int orc_write(int place, const int32_t *coords, int n, const void *vals, int accumulate_f32) {
  int par = T.s[place].parent, sz = T.s[place].dtype_bytes;
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(par, coords + k * ORC_MAX_AXES, 1, &cell, NULL);
    activate(&T.s[par], node, cell);
    Ptr c = lookup(&T.s[par], node, cell) + T.s[place].offset_in_parent_cell;
    if (accumulate_f32)
      *(float *)c += ((const float *)vals)[k];
    else
      memcpy(c, (const char *)vals + (size_t)k * sz, (size_t)sz);
  }
  return 0;
}
int orc_read(int place, const int32_t *coords, int n, void *vals) {
  int par = T.s[place].parent, sz = T.s[place].dtype_bytes;
  for (int k = 0; k < n; k++) {
    int cell;
    Ptr node = walk(par, coords + k * ORC_MAX_AXES, 0, &cell, NULL);
    Ptr c = lookup(&T.s[par], node, cell) + T.s[place].offset_in_parent_cell;
    memcpy((char *)vals + (size_t)k * sz, c, (size_t)sz);
  }
  return 0;
}
/* Dynamic_allocate (node_dynamic.h:61-88): ti.append */
int orc_dynamic_append_i32(int dyn, const int32_t *coords, int n, const int32_t *vals) {
  const OrcSNode *s = &T.s[dyn];
  for (int k = 0; k < n; k++) {
    int cell;
    ODyn *d = (ODyn *)walk(dyn, coords + k * ORC_MAX_AXES, 1, &cell, NULL);
    int i = d->n++;
    Ptr *pc = &d->ptr;
    int start = 0;
    for (;;) {
      if (*pc == NULL) *pc = alloc_node(&A[dyn]);
      if (i < start + s->chunk_size) break;
      pc = (Ptr *)*pc;
      start += s->chunk_size;
    }
    *(int32_t *)(*pc + 8 + (int64_t)(i - start) * s->element_size) = vals[k];
  }
  return 0;
}
int orc_dynamic_length(int dyn, const int32_t *coords, int n, int32_t *out) {
  for (int k = 0; k < n; k++) {
    int cell;
    ODyn *d = (ODyn *)walk(dyn, coords + k * ORC_MAX_AXES, 0, &cell, NULL);
    out[k] = d->n;
  }
  return 0;
}
