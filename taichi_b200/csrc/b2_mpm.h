// b2_mpm.h -- device-visible state of the block-binned 3-D MLS-MPM solver (b2_mpm.cu)
#pragma once
#include "b2_internal.cuh"

#define MPM_CHUNK 256            /* particles per bin chunk                      */
#define MPM_PLANE_BYTES 4096     /* 256 x float4                                 */
#define MPM_CHUNK_BYTES 18176    /* 4 float4 planes + 256 x u32 particle ids + 256 x u16 cell-sorted order + 256 x u8 cell */
#define MPM_XCHUNK_BYTES 12288    /* model 1: bytes per chunk of the side arena (3 float4 planes) */
#define MPM_PERM_OFF 17408       /* byte offset of the u16 order plane (written by P2G, read by G2P) */
#define MPM_CELL_OFF 17920       /* byte offset of the u8 cell plane: cell of the entry inside its 4^3 block, 0xff = tombstone (written with the
                                    position by insert / G2P, read by P2G's count phase instead of the 16-byte position) */
#define MPM_PASS 1024            /* bin entries per P2G pass (4 chunks) = per G2P pass when it follows P2G */
#define MPM_MAXPASS 8            /* passes per bin: maxc * 256 / MPM_PASS                                  */
#define MPM_SB 2048              /* particles cell-sorted per pass (shared-memory keys) */
#define MPM_TOMB (-1.0e30f)      /* p0.x of a bin entry whose particle moved to another block */

// Everything kernels need, passed by value.
struct MpmDev {
  // physics (mpm3d.py:13-23)
  int n_grid, nb;  // cells / 4^3-blocks per axis
  int pot;         // n_grid is a power of two (x*inv_dx is then exact = x/dx)
  int nb_shift;    // log2(nb) when pot
  float dx, inv_dx, dt, p_vol, p_mass, E, gravity, stress_k;  // stress_k = -dt*4*E*p_vol/dx^2
  int bound;
  int model;         // 0: mpm3d.py (J);  1: F / Jp / material with SVD (mpm128.py:36-63 in 3-D)
  float mu_0, lambda_0;
  int activate_halo;  // reproduce the CUDA-BLS epilogue's wider active set
  int slab_lo, slab_hi;
  // grid
  B2TreeDev *T;
  B2Path leaf;  // path to the dense(4^3) SNode holding {vx,vy,vz,m}
  int leaf_sid;
  // fast non-activating block lookup (layout known at create time)
  int layout, np;       // np = pointer cells per axis (layouts 1, 2)
  u8 *top;              // the container directly under root
  long long mask_off;   // layout 1: byte offset of the bitmasked container's mask words
  // block slots (a block gets a slot the first time a particle lands in it)
  int *slot_of;    // [nb^3] -> slot, -1 none, -2 being created
  int *slot_key;   // [max_slots]
  u32 *nslots;
  int max_slots, maxc;
  // two particle generations (read g, write g^1)
  u32 *count[2];     // [max_slots]
  u32 *chunk_of[2];  // [max_slots * maxc] chunk id + 1
  u64 *nbptr;        // [max_slots * 8] the bin's neighbour blocks {0,1}^3 as P2G resolved them (valid until the G2P of the substep)
  u32 *live[2];      // [max_slots * MPM_MAXPASS] live (non-tombstone) entries per pass, written by P2G
  u8 *arena[2];
  u8 *xarena[2];     // model 1: three more float4 planes per chunk {F00..F03 | F10.. | F22, Jp, material, -}, same chunk ids
  u32 *arena_next[2];
  u32 arena_cap;
  u32 *work_counter;  // [0..7] dynamic slot fetch for the persistent kernels; [8..23] lengths of the 16 size classes of glist
  u32 *glist;         // [16 * max_slots] the bins P2G has processed, by size class (0 = largest): the order G2P takes them in
  u32 *listed;        // [nb^3 / 32] bit per block: the block is in the leaf element list (kept current incrementally)
  int dbg;              // B2_DBG experiment switches (0 in production)
  double *stats;      // [8]
};
