/*
 * oracle/snode_layout.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the reference's SNode-tree geometry and byte layout:
 *   - axis extractors / num_cells_per_container: taichi/ir/snode.cpp:37-125
 *   - node byte layout (natural LLVM struct layout, x86-64/nvptx64 data layout):
 *       taichi/codegen/llvm/struct_llvm.cpp:36-144
 *   - refine_coordinates: taichi/codegen/llvm/struct_llvm.cpp:146-190
 *   - get_ch (child offset inside the parent cell): struct_llvm.cpp:192-233
 *   - StructMeta.element_size rules: taichi/codegen/llvm/codegen_llvm.cpp:259-311
 *   - per-level index extraction: taichi/transforms/scalar_pointer_lowerer.cpp:32-86
 *   - NodeManager node sizes: taichi/runtime/llvm/llvm_runtime_executor.cpp:446-467
 *
 * Plain C99, header-only. Shared by ref_runtime_harness.cpp (which drives the
 * reference's own runtime.cpp) and snode_oracle.c (the hand restatement).
 */
#ifndef ORACLE_SNODE_LAYOUT_H
#define ORACLE_SNODE_LAYOUT_H

#include <stdint.h>
#include <string.h>

#define ORC_MAX_AXES 12   /* taichi/inc/constants.h:5 */
#define ORC_MAX_SNODES 64 /* oracle trees are small */
#define ORC_LISTGEN_MAX_ELEMENT 1024 /* taichi/inc/constants.h:28 */

enum { ORC_ROOT = 0, ORC_DENSE = 1, ORC_POINTER = 2, ORC_BITMASKED = 3, ORC_DYNAMIC = 4, ORC_PLACE = 5 };

/* Same byte layout as B2SNodeDesc in include/b200mpm.h (tests pass one buffer to both). */
typedef struct OrcSNodeDesc {
  int32_t type;
  int32_t parent; /* index into the array, -1 for the root (index 0) */
  int32_t num_axes;
  int32_t axes[ORC_MAX_AXES];
  int32_t sizes[ORC_MAX_AXES];
  int32_t dtype_bytes; /* place: size of the scalar */
  int32_t chunk_size;  /* dynamic */
} OrcSNodeDesc;

typedef struct OrcSNode {
  int type, parent, id;
  int nch, ch[16];
  int shape[ORC_MAX_AXES];     /* extractors[a].shape */
  int acc_shape[ORC_MAX_AXES]; /* extractors[a].acc_shape */
  int active[ORC_MAX_AXES];
  int from_root[ORC_MAX_AXES]; /* num_elements_from_root */
  int64_t num_cells;           /* num_cells_per_container == max_num_elements() */
  int chunk_size;
  int dtype_bytes;
  /* byte layout */
  int64_t cell_size;      /* size of ch_type (struct of child containers)        */
  int cell_align;
  int64_t node_size;      /* size of one container of this snode                 */
  int node_align;
  int64_t offset_in_parent_cell;
  int64_t element_size;   /* StructMeta.element_size (codegen_llvm.cpp:259-311)  */
  int64_t alloc_node_size;/* NodeManager element (pointer/dynamic)               */
} OrcSNode;

typedef struct OrcTree {
  int n;
  OrcSNode s[ORC_MAX_SNODES];
} OrcTree;

static int64_t orc_roundup(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

static void orc_layout_node(OrcTree *t, int id) {
  OrcSNode *s = &t->s[id];
  /* children first */
  int64_t off = 0;
  int align = 1;
  for (int c = 0; c < s->nch; c++) {
    orc_layout_node(t, s->ch[c]);
    OrcSNode *ch = &t->s[s->ch[c]];
    off = orc_roundup(off, ch->node_align);
    ch->offset_in_parent_cell = off;
    off += ch->node_size;
    if (ch->node_align > align) align = ch->node_align;
  }
  s->cell_size = orc_roundup(off, align);
  s->cell_align = align;
  switch (s->type) {
    case ORC_PLACE:
      s->node_size = s->dtype_bytes;
      s->node_align = s->dtype_bytes;
      s->element_size = s->dtype_bytes;
      break;
    case ORC_ROOT:
      s->node_size = s->cell_size;
      s->node_align = s->cell_align;
      s->element_size = s->cell_size;
      break;
    case ORC_DENSE:
      s->node_size = s->cell_size * s->num_cells;
      s->node_align = s->cell_align;
      s->element_size = s->cell_size;
      break;
    case ORC_BITMASKED: {
      /* LLVM type is {i32 mask[(n+31)/32], cell body[n]}; only its total size
         is used, the runtime puts the mask AFTER the data (node_bitmasked.h:14-39) */
      int64_t words = (s->num_cells + 31) / 32;
      int a = align > 4 ? align : 4;
      int64_t sz = orc_roundup(words * 4, align) + s->cell_size * s->num_cells;
      s->node_size = orc_roundup(sz, a);
      s->node_align = a;
      s->element_size = s->cell_size;
      break;
    }
    case ORC_POINTER:
      s->node_size = 16 * s->num_cells; /* i64 lock[n], i8* ptr[n] */
      s->node_align = 8;
      /* element_size = node type of ch[0] (codegen_llvm.cpp:267-270) */
      s->element_size = s->nch ? t->s[s->ch[0]].node_size : 0;
      s->alloc_node_size = s->cell_size;
      break;
    case ORC_DYNAMIC:
      s->node_size = 16; /* {i32 lock, i32 n}, i8* */
      s->node_align = 8;
      s->element_size = s->cell_size;
      s->alloc_node_size = 8 + s->cell_size * s->chunk_size;
      break;
  }
}

/* Build geometry + layout from a descriptor array. Returns 0 on success. */
static int orc_build_tree(OrcTree *t, const OrcSNodeDesc *d, int n) {
  if (n <= 0 || n > ORC_MAX_SNODES || d[0].type != ORC_ROOT) return -1;
  memset(t, 0, sizeof(*t));
  t->n = n;
  for (int i = 0; i < n; i++) {
    OrcSNode *s = &t->s[i];
    s->id = i;
    s->type = d[i].type;
    s->parent = d[i].parent;
    s->chunk_size = d[i].chunk_size;
    s->dtype_bytes = d[i].dtype_bytes;
    for (int a = 0; a < ORC_MAX_AXES; a++) {
      s->shape[a] = 1;
      s->acc_shape[a] = 1;
      s->from_root[a] = (i == 0) ? 1 : t->s[s->parent].from_root[a];
    }
    if (i > 0) {
      if (s->parent < 0 || s->parent >= i) return -2;
      OrcSNode *p = &t->s[s->parent];
      if (p->nch >= 16) return -3;
      p->ch[p->nch++] = i;
    }
    for (int k = 0; k < d[i].num_axes; k++) {
      int a = d[i].axes[k];
      if (a < 0 || a >= ORC_MAX_AXES || d[i].sizes[k] <= 0) return -4;
      s->active[a] = 1;
      s->shape[a] = d[i].sizes[k];
      s->from_root[a] *= d[i].sizes[k];
    }
    int64_t acc = 1;
    for (int a = ORC_MAX_AXES - 1; a >= 0; a--) {
      s->acc_shape[a] = (int)acc;
      acc *= s->shape[a];
    }
    s->num_cells = acc;
  }
  orc_layout_node(t, 0);
  return 0;
}

/* refined[a] = in[a]*shape_a + (l % (acc_a*shape_a)) / acc_a  (struct_llvm.cpp:171-188) */
static void orc_refine(const OrcSNode *s, const int32_t *in, int32_t *out, int l) {
  for (int a = 0; a < ORC_MAX_AXES; a++) {
    int add = 0;
    if (s->shape[a] > 1)
      add = (int)(((uint32_t)l % (uint32_t)(s->acc_shape[a] * s->shape[a])) / (uint32_t)s->acc_shape[a]);
    out[a] = in[a] * s->shape[a] + add;
  }
}

/* Path root..target (inclusive). Returns length. */
static int orc_path(const OrcTree *t, int target, int *path) {
  int tmp[ORC_MAX_SNODES], n = 0;
  for (int p = target; p >= 0; p = t->s[p].parent) tmp[n++] = p;
  for (int i = 0; i < n; i++) path[i] = tmp[n - 1 - i];
  return n;
}

/* Linearised cell index at every level of `path` for global coordinate `I`
   (indexed by physical axis). scalar_pointer_lowerer.cpp:32-86 + LinearizeStmt
   (codegen_llvm.cpp:1736-1744). */
static void orc_level_indices(const OrcTree *t, const int *path, int plen, const int32_t *I, int *lin) {
  int total[ORC_MAX_AXES], first[ORC_MAX_AXES];
  for (int a = 0; a < ORC_MAX_AXES; a++) {
    total[a] = 1;
    first[a] = 1;
    for (int i = 0; i < plen; i++) total[a] *= t->s[path[i]].shape[a];
  }
  for (int i = 0; i < plen; i++) {
    const OrcSNode *s = &t->s[path[i]];
    int v = 0;
    for (int a = 0; a < ORC_MAX_AXES; a++) {
      if (!s->active[a]) continue;
      int prev = total[a];
      total[a] /= s->shape[a];
      int next = total[a];
      int e = first[a] ? I[a] : (I[a] % prev);
      e = e / next;
      first[a] = 0;
      v = v * s->shape[a] + e;
    }
    lin[i] = v;
  }
}

#endif
