// b2_mpm.cu -- block-binned 3-D MLS-MPM (mpm3d.py:36-81 physics) over a sparse SNode grid, sm_100a.
//
// What this replaces in the reference: the JIT-generated struct-for bodies of a BLS MPM program
// (bls_test_template.py:185-238 idiom): make_block_local's shared-memory pad prologue/epilogue
// (transforms/make_block_local.cpp:97-365), the per-block dynamic particle-id lists
// (node_dynamic.h), the per-access SNode walk with activation (lower_access.cpp:227-281) and the
// 27x4 float atomics per particle (codegen_llvm.cpp:1495-1569).
//
// B200-first design:
//  * Particles live in per-4^3-block bins: chains of 256-particle AoSoA chunks (four float4 planes
//    {x,J} {v,C00} {C01..C11} {C12..C22} + a u32 id plane). G2P writes every particle straight
//    into the bin of the block it now belongs to (next generation), so the particles stay
//    block-sorted with no sort pass and no extra HBM traffic.
//  * P2G: one CTA per occupied block.  The CTA counting-sorts each chunk by cell in shared memory,
//    then thread (cell, i-plane) accumulates its 9 stencil nodes x {mv, m} in REGISTERS over the
//    cell's particles -- no shared or global float atomics in the inner loop -- a gather pass sums
//    the per-cell partials into the (4+2)^3 pad, and the pad is flushed with one 128-bit
//    red.global.add.v4.f32 per node into the <=8 destination blocks, which are activated
//    (lock-free CAS / atomicOr) only if a particle really touches them (x64 semantics;
//    activate_halo=1 reproduces the CUDA-BLS epilogue).
//  * G2P: one CTA per occupied block: pad of post-grid-op velocities staged once, particles
//    cell-sorted so that same-cell lanes broadcast their shared-memory reads.
//  * grid clear / grid op: struct-for over the leaf Element list, one float4 per cell, 1 KiB per
//    block fully coalesced.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "b2_host.h"
#include "b2_mpm.h"
#include "b2_svd.cuh"

// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void red_add_v4(float4 *addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ int ld_volatile_i32(const int *p) { return *(const volatile int *)p; }
#ifndef MPM_SPIN_LIMIT
#define MPM_SPIN_LIMIT (1u << 24) /* polls (a few seconds) before a wait on another CTA's publication gives up and flags an error */
#endif
__device__ __forceinline__ u32 ld_volatile_u32(const u32 *p) { return *(const volatile u32 *)p; }

// Container of the 4^3 block (bx,by,bz); ACT activates every level above it.  nullptr if the
// block is outside the grid / the owned slab (+ghost layer), or (non-activating) inactive.
template <bool ACT>
__device__ __forceinline__ float4 *mpm_block_ptr(const MpmDev &M, int bx, int by, int bz) {
  if ((unsigned)bx >= (unsigned)M.nb || (unsigned)by >= (unsigned)M.nb || (unsigned)bz >= (unsigned)M.nb) return nullptr;
  // fast path: the three layouts' address arithmetic straight from kernel parameters (no
  // descriptor loads, at most two dependent reads); anything inactive falls through
  if (M.layout == 0) return (float4 *)(M.top + ((u64)((bx * M.nb + by) * M.nb + bz) << 10));
  {
    const int pc = ((bx >> 3) * M.np + (by >> 3)) * M.np + (bz >> 3);
    const u64 c = b2_ld_volatile((const u64 *)M.top + pc);
    const int li = ((bx & 7) << 6) | ((by & 7) << 3) | (bz & 7);
    if (c > B2_BUSY) {
      if (M.layout == 1) {
        const u32 w = *(const volatile u32 *)((const u8 *)c + M.mask_off + (li >> 5) * 4);
        if ((w >> (li & 31)) & 1u) return (float4 *)((u8 *)c + ((u64)li << 10));
      } else {
        const u64 c2 = b2_ld_volatile((const u64 *)c + li);
        if (c2 > B2_BUSY) return (float4 *)c2;
      }
    }
  }
  if (!ACT) return nullptr;
  int I[B2_MAX_AXES];
#pragma unroll
  for (int a = 0; a < B2_MAX_AXES; a++) I[a] = 0;
  I[0] = bx * 4, I[1] = by * 4, I[2] = bz * 4;
  int cell;
  bool dead;
  u8 *node = b2_walk<true>(M.T, M.leaf, I, &cell, &dead);
  if (dead) return nullptr;
  // Keep the leaf element list current (the reference regenerates every list before every struct-for,
  // offload.cpp:174-264; here a block is appended once, by the thread that wins its bit).
  {
    const int key = (bx * M.nb + by) * M.nb + bz;
    const u32 bit = 1u << (key & 31);
    if (!(atomicOr(&M.listed[key >> 5], bit) & bit)) {
      B2CList *L = M.T->lists[M.leaf_sid];
      uint4 *d = (uint4 *)b2_list_slot(M.T->pool, L, atomicAdd(&L->count, 1u));
      if (d) {
        const u64 p = (u64)node;
        d[0] = make_uint4((u32)p, (u32)(p >> 32), 0u, 64u);
        d[1] = make_uint4((u32)bx, (u32)by, (u32)bz, 0u);
        d[2] = make_uint4(0u, 0u, 0u, 0u);
        d[3] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  }
  return (float4 *)node;
}

struct MpmStencil {
  int base[3];
  float fx[3];
  float w[3][3];  // w[offset][axis]
};
__device__ __forceinline__ void mpm_stencil(const MpmDev &M, float x, float y, float z, MpmStencil &s) {
  const float xp[3] = {x, y, z};
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float X = M.pot ? xp[d] * M.inv_dx : xp[d] / M.dx;  // Xp = x / dx (mpm3d.py:41)
    s.base[d] = (int)(X - 0.5f);
    s.fx[d] = X - (float)s.base[d];
    s.w[0][d] = 0.5f * (1.5f - s.fx[d]) * (1.5f - s.fx[d]);
    s.w[1][d] = 0.75f - (s.fx[d] - 1.0f) * (s.fx[d] - 1.0f);
    s.w[2][d] = 0.5f * (s.fx[d] - 0.5f) * (s.fx[d] - 0.5f);
  }
}
__device__ __forceinline__ int mpm_key(const MpmDev &M, int bx, int by, int bz) { return (bx * M.nb + by) * M.nb + bz; }
__device__ __forceinline__ void mpm_unkey(const MpmDev &M, int key, int &bx, int &by, int &bz) {
  if (M.pot) {
    bz = key & (M.nb - 1), by = (key >> M.nb_shift) & (M.nb - 1), bx = key >> (2 * M.nb_shift);
  } else {
    bz = key % M.nb, by = (key / M.nb) % M.nb, bx = key / (M.nb * M.nb);
  }
}

__device__ __forceinline__ u8 *mpm_chunk_ptr(const MpmDev &M, int gen, u32 id1) { return M.arena[gen] + (u64)(id1 - 1) * MPM_CHUNK_BYTES; }
__device__ __forceinline__ u8 *mpm_xchunk_ptr(const MpmDev &M, int gen, u32 id1) { return M.xarena[gen] + (u64)(id1 - 1) * MPM_XCHUNK_BYTES; }
// model 1: the particle's F (row-major), Jp and material as three float4
struct MpmX {
  float4 a, b, c;  // {F00,F01,F02,F10} {F11,F12,F20,F21} {F22,Jp,material bits,0}
};
__device__ __forceinline__ MpmX mpm_x_load(const u8 *xch, u32 i) {
  MpmX r;
  r.a = ((const float4 *)xch)[i], r.b = ((const float4 *)(xch + MPM_PLANE_BYTES))[i], r.c = ((const float4 *)(xch + 2 * MPM_PLANE_BYTES))[i];
  return r;
}
__device__ __forceinline__ void mpm_x_store(u8 *xch, u32 i, const MpmX &r) {
  ((float4 *)xch)[i] = r.a, ((float4 *)(xch + MPM_PLANE_BYTES))[i] = r.b, ((float4 *)(xch + 2 * MPM_PLANE_BYTES))[i] = r.c;
}

// Appends one particle to the bin of block `key` in generation `gen`.  Must be called by all 32
// lanes of the warp (valid = false for idle lanes): lanes that go to the same bin share one
// atomicAdd (match_any), the lane that opens a new chunk allocates it, the others wait for it.
__device__ __forceinline__ void mpm_insert(const MpmDev &M, int gen, int key, int cell, bool valid, float4 p0, float4 p1, float4 p2, float4 p3, u32 id,
                                           const MpmX *px = nullptr) {
  const int lane = threadIdx.x & 31;
  int slot = -1;
  if (valid) {
    slot = ld_volatile_i32(&M.slot_of[key]);
    if (slot == -1) {
      if (atomicCAS(&M.slot_of[key], -1, -2) == -1) {  // this lane creates the slot
        const u32 s = atomicAdd(M.nslots, 1u);
        if (s >= (u32)M.max_slots) {
          // no bin left: the particle is DROPPED and the error flag raised (never aliased onto another block's bin);
          // the block stays without a slot, so it can get one after the next compaction
          atomicSub(M.nslots, 1u);
          b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, key);
          __threadfence();
          *(volatile int *)&M.slot_of[key] = -1;
          slot = -3;
        } else {
          M.slot_key[s] = key;
          __threadfence();
          *(volatile int *)&M.slot_of[key] = (int)s;
          slot = (int)s;
        }
      }
    }
  }
  __syncwarp();
  if (valid && (slot == -1 || slot == -2)) {  // somebody else is creating the slot: wait for the outcome
    unsigned ns = 8, polls = 0;
    while ((slot = ld_volatile_i32(&M.slot_of[key])) == -2) {
      __nanosleep(ns);
      if (ns < 256) ns *= 2;
      if (++polls > MPM_SPIN_LIMIT) {  // (the creator is gone: flag it, drop the particle)
        b2_set_error(M.T->pool, B2E_ASSERT, -1002);
        slot = -3;
        break;
      }
    }
    __threadfence();
  }
  if (valid && slot < 0) {  // creation failed (here or in the creating lane): dropped, flag raised
    b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, key);
    valid = false;
  }
  const unsigned peers = __match_any_sync(0xffffffffu, valid ? slot : -1 - lane);
  const int leader = __ffs(peers) - 1;
  u32 basepos = 0;
  if (valid && lane == leader) basepos = atomicAdd(&M.count[gen][slot], (u32)__popc(peers));
  basepos = __shfl_sync(0xffffffffu, basepos, leader);
  u32 *cslot = nullptr;
  u32 i = 0;
  if (valid) {
    const u32 pos = basepos + __popc(peers & ((1u << lane) - 1u));
    const u32 c = pos / MPM_CHUNK;
    i = pos % MPM_CHUNK;
    if (c >= (u32)M.maxc) {
      b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, key);
      valid = false;
    } else {
      cslot = &M.chunk_of[gen][(u64)slot * M.maxc + c];
      if (i == 0) {
        u32 cid = atomicAdd(M.arena_next[gen], 1u);
        if (cid >= M.arena_cap) {  // out of chunks: the data goes to the sink chunk (owned by no bin), flag raised
          b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, -1);
          cid = M.arena_cap;
        }
        __threadfence();
        *(volatile u32 *)cslot = cid + 1;
      }
    }
  }
  __syncwarp();
  if (valid) {
    u32 id1;
    unsigned ns = 8, polls = 0;
    while ((id1 = ld_volatile_u32(cslot)) == 0) {
      __nanosleep(ns);
      if (ns < 256) ns *= 2;
      if (++polls > MPM_SPIN_LIMIT) {  // (the chunk's holder is gone: flag it, the data goes to the sink chunk)
        b2_set_error(M.T->pool, B2E_ASSERT, -1001);
        id1 = M.arena_cap + 1;
        break;
      }
    }
    u8 *ch = mpm_chunk_ptr(M, gen, id1);
    ((float4 *)ch)[i] = p0;
    ((float4 *)(ch + MPM_PLANE_BYTES))[i] = p1;
    ((float4 *)(ch + 2 * MPM_PLANE_BYTES))[i] = p2;
    ((float4 *)(ch + 3 * MPM_PLANE_BYTES))[i] = p3;
    ((u32 *)(ch + 4 * MPM_PLANE_BYTES))[i] = id;
    (ch + MPM_CELL_OFF)[i] = (u8)cell;
    if (px) mpm_x_store(mpm_xchunk_ptr(M, gen, id1), i, *px);
  }
}

// block key of a position; *cell = its cell inside the block ((base & 3) per axis, x-major)
__device__ __forceinline__ int mpm_key_of_position(const MpmDev &M, float x, float y, float z, bool *ok, int *cell) {
  const float xp[3] = {x, y, z};
  int b[3], l[3];
  // one unsigned compare per axis covers base < 0 and base + 2 >= n_grid; a NaN converts to base 0 and is caught once,
  // on the sum, below (this runs for every particle of G2P: 40 instructions before, 25 now)
  const unsigned lim = (unsigned)(M.n_grid - 2);
  bool good = (x + y + z) == (x + y + z);
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float X = M.pot ? xp[d] * M.inv_dx : xp[d] / M.dx;
    int base = (int)(X - 0.5f);
    if ((unsigned)base >= lim) {  // outside the grid
      good = false;
      base = base < 0 ? 0 : M.n_grid - 3;
    }
    b[d] = base >> 2;
    l[d] = base & 3;
  }
  *ok = good;
  *cell = (l[0] << 4) | (l[1] << 2) | l[2];
  return mpm_key(M, b[0], b[1], b[2]);
}

// ------------------------------------------------------------------------------------------
// particle I/O
// ------------------------------------------------------------------------------------------
// x[N][3] v[N][3] C[N][3][3] J[N] (reference field layouts) -> bins of generation gen
__global__ void __launch_bounds__(256) k_mpm_insert(MpmDev M, int gen, const float *__restrict__ x, const float *__restrict__ v,
                                                     const float *__restrict__ C, const float *__restrict__ J, const int *__restrict__ ids,
                                                     long long n, long long id0, const float *__restrict__ Fm = nullptr,
                                                     const float *__restrict__ Jp = nullptr, const int *__restrict__ mat = nullptr) {
  const long long nround = (n + 31) / 32 * 32;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < nround; p += (long long)gridDim.x * blockDim.x) {
    const bool valid = p < n;
    float4 p0 = make_float4(0, 0, 0, 1), p1 = make_float4(0, 0, 0, 0), p2 = p1, p3 = p1;
    int key = 0, cell = 0;
    u32 id = 0;
    if (valid) {
      p0 = make_float4(x[p * 3], x[p * 3 + 1], x[p * 3 + 2], J ? J[p] : 1.f);
      if (v) p1.x = v[p * 3], p1.y = v[p * 3 + 1], p1.z = v[p * 3 + 2];
      if (C) {
        const float *c = C + p * 9;
        p1.w = c[0];
        p2 = make_float4(c[1], c[2], c[3], c[4]);
        p3 = make_float4(c[5], c[6], c[7], c[8]);
      }
      id = ids ? (u32)ids[p] : (u32)(id0 + p);
      bool ok;
      key = mpm_key_of_position(M, p0.x, p0.y, p0.z, &ok, &cell);
      if (!ok) b2_set_error(M.T->pool, B2E_ASSERT, (int)p);
    }
    if (M.model) {  // F = identity, Jp = 1, material 0 unless given
      MpmX X;
      X.a = make_float4(1.f, 0.f, 0.f, 0.f), X.b = make_float4(1.f, 0.f, 0.f, 0.f), X.c = make_float4(1.f, 1.f, 0.f, 0.f);
      if (valid) {
        if (Fm) {
          const float *f = Fm + p * 9;
          X.a = make_float4(f[0], f[1], f[2], f[3]), X.b = make_float4(f[4], f[5], f[6], f[7]), X.c.x = f[8];
        }
        if (Jp) X.c.y = Jp[p];
        if (mat) X.c.z = __int_as_float(mat[p]);
      }
      mpm_insert(M, gen, key, cell, valid, p0, p1, p2, p3, id, &X);
    } else {
      mpm_insert(M, gen, key, cell, valid, p0, p1, p2, p3, id);
    }
  }
}

__global__ void __launch_bounds__(256) k_mpm_extract(MpmDev M, int gen, float *x, float *v, float *C, float *J, long long nmax, float *Fm = nullptr,
                                                      float *Jp = nullptr, int *mat = nullptr) {
  const u32 ns = *M.nslots;
  for (u32 slot = blockIdx.x; slot < ns; slot += gridDim.x) {
    const u32 cnt = M.count[gen][slot];
    for (u32 q = threadIdx.x; q < cnt; q += blockDim.x) {
      const u32 id1 = M.chunk_of[gen][(u64)slot * M.maxc + q / MPM_CHUNK];
      const u8 *ch = mpm_chunk_ptr(M, gen, id1);
      const u32 i = q % MPM_CHUNK;
      const float4 p0 = ((const float4 *)ch)[i];
      if (p0.x == MPM_TOMB) continue;
      const u32 id = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[i];
      if ((long long)id >= nmax) continue;
      const float4 p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[i];
      const float4 p2 = ((const float4 *)(ch + 2 * MPM_PLANE_BYTES))[i];
      const float4 p3 = ((const float4 *)(ch + 3 * MPM_PLANE_BYTES))[i];
      if (x) x[(u64)id * 3] = p0.x, x[(u64)id * 3 + 1] = p0.y, x[(u64)id * 3 + 2] = p0.z;
      if (J) J[id] = p0.w;
      if (v) v[(u64)id * 3] = p1.x, v[(u64)id * 3 + 1] = p1.y, v[(u64)id * 3 + 2] = p1.z;
      if (C) {
        float *c = C + (u64)id * 9;
        c[0] = p1.w, c[1] = p2.x, c[2] = p2.y, c[3] = p2.z, c[4] = p2.w, c[5] = p3.x, c[6] = p3.y, c[7] = p3.z, c[8] = p3.w;
      }
      if (M.model && (Fm || Jp || mat)) {
        const MpmX X = mpm_x_load(mpm_xchunk_ptr(M, gen, id1), i);
        if (Fm) {
          float *f = Fm + (u64)id * 9;
          f[0] = X.a.x, f[1] = X.a.y, f[2] = X.a.z, f[3] = X.a.w, f[4] = X.b.x, f[5] = X.b.y, f[6] = X.b.z, f[7] = X.b.w, f[8] = X.c.x;
        }
        if (Jp) Jp[id] = X.c.y;
        if (mat) mat[id] = __float_as_int(X.c.z);
      }
    }
  }
}

// model 1, before P2G (mpm128.py:36-63 carried to 3-D, ggui_examples/mpm3d_ggui.py:52-88): F <- (I + dt C) F, hardening from Jp,
// ti.svd, snow plasticity (singular values clamped, Jp updated), liquid reset, fixed-corotated stress.  The kernel leaves
// affine / p_mass = C + stress / p_mass in the particle's C planes and J = 1, so that the P2G kernel, which scatters
// p_mass v + (stress_k (J - 1) I + p_mass C) dpos, needs no variant; G2P rewrites C from the grid afterwards.
#ifndef MAT_T
#define MAT_T 128
#endif
#ifndef MAT_MINB
#define MAT_MINB 8 /* 64 registers: 0.74 ms on the C3 cloud against 0.84 ms at 6 x 80 */
#endif
__global__ void __launch_bounds__(MAT_T, MAT_MINB) k_mpm_material(MpmDev M, int gen) {
  const u32 ns = *M.nslots;
  for (u32 slot = blockIdx.x; slot < ns; slot += gridDim.x) {
    const u32 cnt = M.count[gen][slot];
    for (u32 q = threadIdx.x; q < cnt; q += blockDim.x) {
      const u32 id1 = M.chunk_of[gen][(u64)slot * M.maxc + q / MPM_CHUNK];
      u8 *ch = mpm_chunk_ptr(M, gen, id1);
      const u32 i = q % MPM_CHUNK;
      if ((ch + MPM_CELL_OFF)[i] == 0xffu) continue;  // tombstone
      u8 *xch = mpm_xchunk_ptr(M, gen, id1);
      float4 p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[i];
      const float4 p2 = ((const float4 *)(ch + 2 * MPM_PLANE_BYTES))[i], p3 = ((const float4 *)(ch + 3 * MPM_PLANE_BYTES))[i];
      MpmX X = mpm_x_load(xch, i);
      const float C[9] = {p1.w, p2.x, p2.y, p2.z, p2.w, p3.x, p3.y, p3.z, p3.w};
      const float Fo[9] = {X.a.x, X.a.y, X.a.z, X.a.w, X.b.x, X.b.y, X.b.z, X.b.w, X.c.x};
      const int material = __float_as_int(X.c.z);
      float F[9];
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
          float a = 0.f;
          for (int m = 0; m < 3; m++) a += ((r == m ? 1.f : 0.f) + M.dt * C[r * 3 + m]) * Fo[m * 3 + c];
          F[r * 3 + c] = a;
        }
      float jp = X.c.y;
      float h = fmaxf(0.1f, fminf(5.f, expf(10.f * (1.f - jp))));
      if (material == 1) h = 0.3f;
      float mu = M.mu_0 * h;
      const float la = M.lambda_0 * h;
      if (material == 0) mu = 0.f;
      float U[9], sig[3], V[9];
      b2_svd3(F, U, sig, V);
      float J = 1.f;
      for (int d = 0; d < 3; d++) {
        float nsig = sig[d];
        if (material == 2) nsig = fminf(fmaxf(sig[d], 1.f - 2.5e-2f), 1.f + 4.5e-3f);
        jp *= sig[d] / nsig, sig[d] = nsig, J *= nsig;
      }
      float R[9];
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
          R[r * 3 + c] = U[r * 3] * V[c * 3] + U[r * 3 + 1] * V[c * 3 + 1] + U[r * 3 + 2] * V[c * 3 + 2];
          if (material == 0)
            F[r * 3 + c] = r == c ? cbrtf(J) : 0.f;
          else if (material == 2)
            F[r * 3 + c] = U[r * 3] * sig[0] * V[c * 3] + U[r * 3 + 1] * sig[1] * V[c * 3 + 1] + U[r * 3 + 2] * sig[2] * V[c * 3 + 2];
        }
      const float k = -M.dt * M.p_vol * 4.f * M.inv_dx * M.inv_dx / M.p_mass;
      float A[9];
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
          float a = 0.f;
          for (int m = 0; m < 3; m++) a += (F[r * 3 + m] - R[r * 3 + m]) * F[c * 3 + m];
          A[r * 3 + c] = k * (2.f * mu * a + (r == c ? la * J * (J - 1.f) : 0.f)) + C[r * 3 + c];
        }
      X.a = make_float4(F[0], F[1], F[2], F[3]), X.b = make_float4(F[4], F[5], F[6], F[7]), X.c.x = F[8], X.c.y = jp;
      mpm_x_store(xch, i, X);
      p1.w = A[0];
      ((float4 *)(ch + MPM_PLANE_BYTES))[i] = p1;
      ((float4 *)(ch + 2 * MPM_PLANE_BYTES))[i] = make_float4(A[1], A[2], A[3], A[4]);
      ((float4 *)(ch + 3 * MPM_PLANE_BYTES))[i] = make_float4(A[5], A[6], A[7], A[8]);
      ((float *)ch)[i * 4 + 3] = 1.0f;  // J: the scalar stress of the mpm3d.py model is switched off
    }
  }
}

// Bin-slot compaction: a block gets a slot the first time a particle lands in it and material that moves on
// leaves empty slots behind; every kernel walks all slots, and the table is finite.  Between two substeps only generation
// `gen` holds particles: occupied slots are renumbered densely into the spare tables, vacated blocks lose their slot.
__global__ void __launch_bounds__(256) k_mpm_compact(MpmDev M, int gen, u32 *count2, u32 *chunk2, int *key2, u32 *nslots2) {
  const u32 ns = *M.nslots;
  const int lane = threadIdx.x & 31;
  for (u32 s0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; s0 < ns; s0 += gridDim.x * blockDim.x) {
    const u32 s = s0 + lane;
    const u32 cnt = s < ns ? M.count[gen][s] : 0u;
    const unsigned m = __ballot_sync(0xffffffffu, cnt > 0);
    u32 base = 0;
    if (lane == 0 && m) base = atomicAdd(nslots2, (u32)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (s < ns) {
      const int key = M.slot_key[s];
      if (cnt > 0) {
        const u32 n = base + __popc(m & ((1u << lane) - 1u));
        count2[n] = cnt;
        key2[n] = key;
        M.slot_of[key] = (int)n;
        const u32 nch = (cnt + MPM_CHUNK - 1) / MPM_CHUNK;
        for (u32 c = 0; c < nch && c < (u32)M.maxc; c++) chunk2[(u64)n * M.maxc + c] = M.chunk_of[gen][(u64)s * M.maxc + c];
      } else {
        M.slot_of[key] = -1;
      }
    }
  }
}

// stats[0] live particles, [1] slots, [2] occupied slots, [3] chunks in use, [5] bin entries incl. tombstones
__global__ void __launch_bounds__(256) k_mpm_stats(MpmDev M, int gen) {
  const u32 ns = *M.nslots;
  unsigned long long np = 0, occ = 0, ent = 0;
  for (u32 s = blockIdx.x; s < ns; s += gridDim.x) {
    const u32 c = M.count[gen][s];
    if (threadIdx.x == 0) occ += c > 0, ent += c;
    for (u32 q = threadIdx.x; q < c; q += blockDim.x) {
      const u8 *ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)s * M.maxc + q / MPM_CHUNK]);
      np += ((const float4 *)ch)[q % MPM_CHUNK].x != MPM_TOMB;
    }
  }
  for (int d = 16; d > 0; d >>= 1) {
    np += __shfl_down_sync(0xffffffffu, np, d);
    occ += __shfl_down_sync(0xffffffffu, occ, d);
    ent += __shfl_down_sync(0xffffffffu, ent, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&M.stats[0], (double)np);
    atomicAdd(&M.stats[2], (double)occ);
    atomicAdd(&M.stats[5], (double)ent);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    M.stats[1] = (double)ns;
    M.stats[3] = (double)*M.arena_next[gen];
  }
}

// Particle diagnostics (the 0-D reductions an MPM driver asks for every few steps: the reference lowers
// `s[None] += ...` / ti.atomic_max to a thread-local accumulator + warp reduce + one atomic per warp,
// make_thread_local.cpp:100-213, runtime.cpp:1204-1266).  out[0] live particles, [1..3] momentum p_mass sum v,
// [4] kinetic energy, [5] sum J, [6] max |v| (CFL), [7] max |J - 1|.  Sums in fp64; the two maxima are
// non-negative, so their IEEE bit patterns order like integers (atomicMax on the bits).
__global__ void __launch_bounds__(256) k_mpm_diag(MpmDev M, int gen, double *out) {
  const u32 ns = *M.nslots;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  float vmax = 0.f, jmax = 0.f;
  for (u32 s = blockIdx.x; s < ns; s += gridDim.x) {
    const u32 c = M.count[gen][s];
    for (u32 q = threadIdx.x; q < c; q += blockDim.x) {
      const u8 *ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)s * M.maxc + q / MPM_CHUNK]);
      const float4 p0 = ((const float4 *)ch)[q % MPM_CHUNK];
      if (p0.x == MPM_TOMB) continue;
      const float4 p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[q % MPM_CHUNK];
      const double v2 = (double)p1.x * p1.x + (double)p1.y * p1.y + (double)p1.z * p1.z;
      acc[0] += 1.0, acc[1] += p1.x, acc[2] += p1.y, acc[3] += p1.z, acc[4] += v2, acc[5] += p0.w;
      vmax = fmaxf(vmax, sqrtf((float)v2));
      jmax = fmaxf(jmax, fabsf(p0.w - 1.0f));
    }
  }
#pragma unroll
  for (int k = 0; k < 6; k++)
    for (int d = 16; d > 0; d >>= 1) acc[k] += __shfl_down_sync(0xffffffffu, acc[k], d);
  for (int d = 16; d > 0; d >>= 1) {
    vmax = fmaxf(vmax, __shfl_down_sync(0xffffffffu, vmax, d));
    jmax = fmaxf(jmax, __shfl_down_sync(0xffffffffu, jmax, d));
  }
  if ((threadIdx.x & 31) == 0 && acc[0] > 0.0) {
    atomicAdd(&out[0], acc[0]);
    atomicAdd(&out[1], acc[1] * (double)M.p_mass), atomicAdd(&out[2], acc[2] * (double)M.p_mass), atomicAdd(&out[3], acc[3] * (double)M.p_mass);
    atomicAdd(&out[4], 0.5 * (double)M.p_mass * acc[4]);
    atomicAdd(&out[5], acc[5]);
    atomicMax((unsigned long long *)&out[6], (unsigned long long)__double_as_longlong((double)vmax));
    atomicMax((unsigned long long *)&out[7], (unsigned long long)__double_as_longlong((double)jmax));
  }
}

// ------------------------------------------------------------------------------------------
// grid kernels: struct-for over the leaf Element list (one Element = one 4^3 block = 1 KiB)
// ------------------------------------------------------------------------------------------
// zero_a / zero_b (optional): the substep's work counters and the chunk-arena cursor of the generation G2P will
// write -- reset here, by the first kernel of the substep, instead of by three memset nodes between the kernels
__global__ void __launch_bounds__(256) k_mpm_clear(B2TreeDev *T, int leaf, u32 *zero_a, u32 *zero_b) {
  if (blockIdx.x == 0 && threadIdx.x < 24) {
    if (zero_a) zero_a[threadIdx.x] = 0;  // (work counters and the size-class lengths of the G2P order)
    if (zero_b && threadIdx.x < 4) zero_b[threadIdx.x] = 0;
  }
  B2CList *L = T->lists[leaf];
  const u32 ne = L->count;
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  for (u32 e = blockIdx.x * 4 + sub; e < ne; e += gridDim.x * 4) {
    const B2Element *E = (const B2Element *)b2_list_get(L, e);
    float4 *blk = (float4 *)E->node;
    blk[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// mpm3d.py:56-61: v /= m (m > 0); v.y -= dt*g; walls
__global__ void __launch_bounds__(256) k_mpm_gridop(MpmDev M) {
  B2CList *L = M.T->lists[M.leaf_sid];
  const u32 ne = L->count;
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  for (u32 e = blockIdx.x * 4 + sub; e < ne; e += gridDim.x * 4) {
    const B2Element *E = (const B2Element *)b2_list_get(L, e);
    float4 *blk = (float4 *)E->node;
    const int I[3] = {E->pcoord[0] * 4 + (t >> 4), E->pcoord[1] * 4 + ((t >> 2) & 3), E->pcoord[2] * 4 + (t & 3)};
    float4 g = blk[t];
    float vv[3] = {g.x, g.y, g.z};
    if (g.w > 0.f) {
      vv[0] /= g.w;
      vv[1] /= g.w;
      vv[2] /= g.w;
    }
    vv[1] -= M.dt * M.gravity;
#pragma unroll
    for (int d = 0; d < 3; d++) {
      const bool cond = (I[d] < M.bound && vv[d] < 0.f) || (I[d] > M.n_grid - M.bound && vv[d] > 0.f);
      if (cond) vv[d] = 0.f;
    }
    blk[t] = make_float4(vv[0], vv[1], vv[2], g.w);
  }
}

// grid -> dense [n][n][n][4] image
__global__ void __launch_bounds__(256) k_mpm_grid_to_dense(MpmDev M, float4 *out) {
  B2CList *L = M.T->lists[M.leaf_sid];
  const u32 ne = L->count;
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  for (u32 e = blockIdx.x * 4 + sub; e < ne; e += gridDim.x * 4) {
    const B2Element *E = (const B2Element *)b2_list_get(L, e);
    const float4 *blk = (const float4 *)E->node;
    const long long i = E->pcoord[0] * 4 + (t >> 4), j = E->pcoord[1] * 4 + ((t >> 2) & 3), k = E->pcoord[2] * 4 + (t & 3);
    out[(i * M.n_grid + j) * M.n_grid + k] = blk[t];
  }
}

// grid -> compact list of active blocks: key (bx*nb+by)*nb+bz and the 64 cells of each (list order)
__global__ void __launch_bounds__(256) k_mpm_grid_blocks(MpmDev M, int *keys, float4 *cells, long long cap) {
  B2CList *L = M.T->lists[M.leaf_sid];
  const u32 ne = (u32)min((long long)L->count, cap);
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  for (u32 e = blockIdx.x * 4 + sub; e < ne; e += gridDim.x * 4) {
    const B2Element *E = (const B2Element *)b2_list_get(L, e);
    if (t == 0) keys[e] = (E->pcoord[0] * M.nb + E->pcoord[1]) * M.nb + E->pcoord[2];
    cells[(u64)e * 64 + t] = ((const float4 *)E->node)[t];
  }
}

// after a full listgen: every listed block gets its bit
__global__ void __launch_bounds__(256) k_mpm_mark_listed(MpmDev M) {
  B2CList *L = M.T->lists[M.leaf_sid];
  const u32 ne = L->count;
  for (u32 e = blockIdx.x * blockDim.x + threadIdx.x; e < ne; e += gridDim.x * blockDim.x) {
    const B2Element *E = (const B2Element *)b2_list_get(L, e);
    const int key = (E->pcoord[0] * M.nb + E->pcoord[1]) * M.nb + E->pcoord[2];
    atomicOr(&M.listed[key >> 5], 1u << (key & 31));
  }
}

#include "b2_f32x2.cuh"
#include "b2_mpm_bls.cuh"
#include "b2_mpm_p2gw.cuh"
#include "b2_mpm_g2po.cuh"

// Plain path (ti.block_local off): every particle walks the SNode tree with activation and issues
// one vector red per node -- the reference's non-BLS access pattern (K15) on the binned particles.
__global__ void __launch_bounds__(256) k_mpm_p2g_plain(MpmDev M, int gen) {
  const u32 nslots = *M.nslots;
  for (u32 slot = blockIdx.x; slot < nslots; slot += gridDim.x) {
    const u32 total = M.count[gen][slot];
    for (u32 q = threadIdx.x; q < total; q += blockDim.x) {
      const u8 *ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)slot * M.maxc + q / MPM_CHUNK]);
      const u32 t = q % MPM_CHUNK;
      const float4 p0 = ((const float4 *)ch)[t];
      if (p0.x == MPM_TOMB) continue;
      const float4 p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[t];
      const float4 p2 = ((const float4 *)(ch + 2 * MPM_PLANE_BYTES))[t];
      const float4 p3 = ((const float4 *)(ch + 3 * MPM_PLANE_BYTES))[t];
      MpmStencil s;
      mpm_stencil(M, p0.x, p0.y, p0.z, s);
      const float stress = M.stress_k * (p0.w - 1.0f);
      const float m = M.p_mass;
      const float A[3][3] = {{stress + m * p1.w, m * p2.x, m * p2.y}, {m * p2.z, stress + m * p2.w, m * p3.x}, {m * p3.y, m * p3.z, stress + m * p3.w}};
      const float vp[3] = {p1.x, p1.y, p1.z};
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          for (int k = 0; k < 3; k++) {
            const float dpos[3] = {((float)i - s.fx[0]) * M.dx, ((float)j - s.fx[1]) * M.dx, ((float)k - s.fx[2]) * M.dx};
            const float w = s.w[i][0] * s.w[j][1] * s.w[k][2];
            const int gx = s.base[0] + i, gy = s.base[1] + j, gz = s.base[2] + k;
            float4 *blk = mpm_block_ptr<true>(M, gx >> 2, gy >> 2, gz >> 2);
            if (!blk) continue;
            float4 val;
            val.x = w * (m * vp[0] + (A[0][0] * dpos[0] + A[0][1] * dpos[1] + A[0][2] * dpos[2]));
            val.y = w * (m * vp[1] + (A[1][0] * dpos[0] + A[1][1] * dpos[1] + A[1][2] * dpos[2]));
            val.z = w * (m * vp[2] + (A[2][0] * dpos[0] + A[2][1] * dpos[1] + A[2][2] * dpos[2]));
            val.w = w * m;
            red_add_v4(blk + (((gx & 3) << 4) | ((gy & 3) << 2) | (gz & 3)), val);
          }
    }
  }
}

// ------------------------------------------------------------------------------------------
// G2P + re-binning
// ------------------------------------------------------------------------------------------
#define G2P_PLAIN_THREADS 256
template <bool BLS>
__global__ void __launch_bounds__(G2P_PLAIN_THREADS, 3) k_mpm_g2p(MpmDev M, int gen) {
  __shared__ float4 sPad[216];
  __shared__ float4 sP0[MPM_CHUNK];
  __shared__ u32 sId[MPM_CHUNK];
  __shared__ int sHist[64];
  __shared__ int sStart[65];
  __shared__ int sSlot;
  __shared__ float4 *sNb[8];
  const int tid = threadIdx.x;
  const u32 nslots = *M.nslots;

  for (;;) {
    __syncthreads();
    if (tid == 0) sSlot = (int)atomicAdd(&M.work_counter[1], 1u);
    __syncthreads();
    const u32 slot = (u32)sSlot;
    if (slot >= nslots) break;
    const u32 total = M.count[gen][slot];
    if (total == 0) continue;
    const int key = M.slot_key[slot];
    const int bz = key % M.nb, by = (key / M.nb) % M.nb, bx = key / (M.nb * M.nb);

    if (BLS) {
      if (tid < 8) sNb[tid] = mpm_block_ptr<false>(M, bx + (tid >> 2), by + ((tid >> 1) & 1), bz + (tid & 1));
      __syncthreads();
      if (tid < 216) {
        const int px = tid / 36, py = (tid / 6) % 6, pz = tid % 6;
        const float4 *src = sNb[((px >> 2) << 2) | ((py >> 2) << 1) | (pz >> 2)];
        sPad[tid] = src ? src[((px & 3) << 4) | ((py & 3) << 2) | (pz & 3)] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    const u32 nchunks = (total + MPM_CHUNK - 1) / MPM_CHUNK;
    for (u32 c = 0; c < nchunks; c++) {
      const u8 *ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)slot * M.maxc + c]);
      const int n = (int)min((u32)MPM_CHUNK, total - c * MPM_CHUNK);
      if (tid < 64) sHist[tid] = 0;
      __syncthreads();
      float4 p0 = make_float4(0.f, 0.f, 0.f, 1.f);
      u32 id = 0;
      int cidx = 0, rank = 0;
      if (tid < n) {
        p0 = ((const float4 *)ch)[tid];
        id = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[tid];
        const float xp[3] = {p0.x, p0.y, p0.z};
        int l[3];
#pragma unroll
        for (int d = 0; d < 3; d++) {
          const float X = M.pot ? xp[d] * M.inv_dx : xp[d] / M.dx;
          l[d] = ((int)(X - 0.5f)) & 3;
        }
        cidx = (l[0] << 4) | (l[1] << 2) | l[2];
        rank = atomicAdd(&sHist[cidx], 1);
      }
      __syncthreads();
      if (tid < 64) {
        const int lane = tid & 31;
        int inc = sHist[tid];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int tv = __shfl_up_sync(0xffffffffu, inc, d);
          if (lane >= d) inc += tv;
        }
        sStart[tid + 1] = inc;
      }
      __syncthreads();
      if (tid >= 32 && tid < 64) sStart[tid + 1] += sStart[32];
      if (tid == 0) sStart[0] = 0;
      __syncthreads();
      if (tid < n) {
        const int q = sStart[cidx] + rank;
        sP0[q] = p0;
        sId[q] = id;
      }
      __syncthreads();
      const bool valid = tid < n;
      float4 o0 = make_float4(0.f, 0.f, 0.f, 1.f), o1 = make_float4(0.f, 0.f, 0.f, 0.f), o2 = o1, o3 = o1;
      int nkey = 0, ncell = 0;
      if (valid) {
        p0 = sP0[tid];
        id = sId[tid];
        MpmStencil s;
        mpm_stencil(M, p0.x, p0.y, p0.z, s);
        float v0 = 0.f, v1 = 0.f, v2 = 0.f;
        float Sx0 = 0.f, Sx1 = 0.f, Sx2 = 0.f, Sy0 = 0.f, Sy1 = 0.f, Sy2 = 0.f, Sz0 = 0.f, Sz1 = 0.f, Sz2 = 0.f;
        const int l0 = s.base[0] & 3, l1 = s.base[1] & 3, l2 = s.base[2] & 3;
        const float wz0 = s.w[0][2], wz1 = s.w[1][2], wz2 = s.w[2][2];
#pragma unroll
        for (int i = 0; i < 3; i++) {
#pragma unroll
          for (int j = 0; j < 3; j++) {
            const float wxy = s.w[i][0] * s.w[j][1];
            float4 g0, g1, g2;
            if (BLS) {
              const int pidx = ((l0 + i) * 6 + (l1 + j)) * 6 + l2;
              g0 = sPad[pidx], g1 = sPad[pidx + 1], g2 = sPad[pidx + 2];
            } else {
              float4 g[3];
#pragma unroll
              for (int k = 0; k < 3; k++) {
                const int gx = s.base[0] + i, gy = s.base[1] + j, gz = s.base[2] + k;
                const float4 *blk = mpm_block_ptr<false>(M, gx >> 2, gy >> 2, gz >> 2);
                g[k] = blk ? blk[((gx & 3) << 4) | ((gy & 3) << 2) | (gz & 3)] : make_float4(0.f, 0.f, 0.f, 0.f);
              }
              g0 = g[0], g1 = g[1], g2 = g[2];
            }
            // row = sum_k wz_k g_k ; rz = sum_k k wz_k g_k
            const float t1x = wz1 * g1.x, t1y = wz1 * g1.y, t1z = wz1 * g1.z;
            const float t2x = wz2 * g2.x, t2y = wz2 * g2.y, t2z = wz2 * g2.z;
            const float rx = fmaf(wz0, g0.x, t1x + t2x), ry = fmaf(wz0, g0.y, t1y + t2y), rzv = fmaf(wz0, g0.z, t1z + t2z);
            const float zx = fmaf(2.f, t2x, t1x), zy = fmaf(2.f, t2y, t1y), zz = fmaf(2.f, t2z, t1z);
            v0 = fmaf(wxy, rx, v0), v1 = fmaf(wxy, ry, v1), v2 = fmaf(wxy, rzv, v2);
            Sz0 = fmaf(wxy, zx, Sz0), Sz1 = fmaf(wxy, zy, Sz1), Sz2 = fmaf(wxy, zz, Sz2);
            if (j > 0) {
              const float wj = (float)j * wxy;
              Sy0 = fmaf(wj, rx, Sy0), Sy1 = fmaf(wj, ry, Sy1), Sy2 = fmaf(wj, rzv, Sy2);
            }
            if (i > 0) {
              const float wi = (float)i * wxy;
              Sx0 = fmaf(wi, rx, Sx0), Sx1 = fmaf(wi, ry, Sx1), Sx2 = fmaf(wi, rzv, Sx2);
            }
          }
        }
        // C[a][b] = 4/dx * (sum w g_a o_b - fx_b v_a)   (mpm3d.py:75 with dpos = (o - fx) dx)
        const float k4 = 4.0f * M.inv_dx;
        const float C00 = k4 * (Sx0 - s.fx[0] * v0), C01 = k4 * (Sy0 - s.fx[1] * v0), C02 = k4 * (Sz0 - s.fx[2] * v0);
        const float C10 = k4 * (Sx1 - s.fx[0] * v1), C11 = k4 * (Sy1 - s.fx[1] * v1), C12 = k4 * (Sz1 - s.fx[2] * v1);
        const float C20 = k4 * (Sx2 - s.fx[0] * v2), C21 = k4 * (Sy2 - s.fx[1] * v2), C22 = k4 * (Sz2 - s.fx[2] * v2);
        const float nx = p0.x + M.dt * v0, ny = p0.y + M.dt * v1, nz = p0.z + M.dt * v2;
        const float nJ = p0.w * (1.0f + M.dt * (C00 + C11 + C22));
        o0 = make_float4(nx, ny, nz, nJ);
        o1 = make_float4(v0, v1, v2, C00);
        o2 = make_float4(C01, C02, C10, C11);
        o3 = make_float4(C12, C20, C21, C22);
        bool ok;
        nkey = mpm_key_of_position(M, nx, ny, nz, &ok, &ncell);
        if (!ok) b2_set_error(M.T->pool, B2E_ASSERT, (int)id);
      }
      mpm_insert(M, gen ^ 1, nkey, ncell, valid, o0, o1, o2, o3, id);
      __syncthreads();
    }
    // this generation's bin is consumed: leave it empty for the generation after next
    if (tid < (int)nchunks) M.chunk_of[gen][(u64)slot * M.maxc + tid] = 0;
    for (int t = tid + G2P_PLAIN_THREADS; t < (int)nchunks; t += G2P_PLAIN_THREADS) M.chunk_of[gen][(u64)slot * M.maxc + t] = 0;
    if (tid == 0) M.count[gen][slot] = 0;
  }
}

#include "b2_mpm_slab.cuh"

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct B2Mpm_t {
  TiRuntime_t *rt;
  B2MpmParams prm;
  B2Tree_t *tree = nullptr;
  MpmDev M;
  int gen = 0;
  int64_t n_particles = 0;
  bool lists_valid = false;
  bool order_valid = false;  // P2G has left the cell order of the current generation in the chunks (G2P skips its sort)
  bool sparse = false;
  bool profiling = false;
  long long substep_index = 0;  // substeps since the particles were set (deactivate_every counts on it)
  u32 *count_sp = nullptr, *chunk_sp = nullptr, *nslots_sp = nullptr;  // spare slot tables for compact_bins()
  int *key_sp = nullptr;
  int compact_every = 64;       // substeps between two bin-slot compactions
  size_t listed_bytes = 0;  // size of M.listed (multiple of 16)
  bool in_substep = false;  // inside substep(): k_mpm_clear resets the work counters / arena cursor for the phases that follow
  std::vector<cudaEvent_t> prof_events;
  double phase_sum[6] = {0, 0, 0, 0, 0, 0};
  long long phase_count = 0;
  std::vector<void *> allocs;

  // Host <-> device particle I/O runs on two copy streams with double-buffered device staging, so that the
  // upload of the next state and the download of the last result overlap the substeps in between.
  struct Io {
    cudaStream_t h2d = nullptr, d2h = nullptr;
    void *stage[2] = {nullptr, nullptr}, *out[2] = {nullptr, nullptr};
    size_t stage_cap[2] = {0, 0}, out_cap[2] = {0, 0};
    cudaEvent_t up_done[2], stage_free[2], out_ready[2], out_free[2];
    bool ready = false;
    unsigned up_i = 0, out_i = 0;
    void ensure();
    void *reserve(void **buf, size_t *cap, size_t need);
    void wait();
    void destroy();
  } io;

  template <typename T>
  T *dalloc(size_t count, int fill = 0) {
    void *p = nullptr;
    B2_CUDA(cudaMalloc(&p, std::max<size_t>(count * sizeof(T), 16)));
    B2_CUDA(cudaMemsetAsync(p, fill, std::max<size_t>(count * sizeof(T), 16), rt->stream()));
    allocs.push_back(p);
    return (T *)p;
  }
  B2Mpm_t(TiRuntime_t *rt_, const B2MpmParams &p);
  ~B2Mpm_t();
  void reset_bins();
  void compact_bins();
  void ensure_lists();
  void clear_grid();
  void p2g();
  void grid_op();
  void g2p();
  void begin_substep();
  void end_substep();
  void substep(int n);
};

static void add_axes3(B2SNodeDesc &d, int size) {
  d.num_axes = 3;
  for (int a = 0; a < 3; a++) d.axes[a] = a, d.sizes[a] = size;
}

B2Mpm_t::B2Mpm_t(TiRuntime_t *rt_, const B2MpmParams &p) : rt(rt_), prm(p) {
  if (p.n_grid < 8 || p.n_grid % 4) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "n_grid must be a multiple of 4, >= 8");
  if (p.layout < 0 || p.layout > 2) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "layout");
  if (p.layout > 0 && p.n_grid % 32) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "sparse layouts need n_grid % 32 == 0");
  if (p.max_particles <= 0) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "max_particles");
  if (p.model < 0 || p.model > 1) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "model");
  if (p.model == 1 && !(p.nu > -1.f && p.nu < 0.5f)) throw B2Error(TI_ERROR_INVALID_ARGUMENT, "model 1 needs Poisson's ratio nu in (-1, 0.5)");
  if (p.model == 1 && !p.block_local) throw B2Error(TI_ERROR_NOT_SUPPORTED, "model 1 runs on the block-local kernels only");
  if (p.model == 1 && (p.slab_lo != 0 || (p.slab_hi != 0 && p.slab_hi != p.n_grid / 4)))
    throw B2Error(TI_ERROR_NOT_SUPPORTED, "model 1 is not slab-decomposed (migrant records carry no F / Jp / material)");
  const int nb = p.n_grid / 4;
  // grid SNode tree; leaf cell = {vx, vy, vz, m}
  std::vector<B2SNodeDesc> d;
  auto node = [&](int type, int parent) {
    B2SNodeDesc x;
    memset(&x, 0, sizeof(x));
    x.type = type;
    x.parent = parent;
    d.push_back(x);
    return (int)d.size() - 1;
  };
  node(B2_ROOT, -1);
  int leaf;
  if (p.layout == 0) {
    int a = node(B2_DENSE, 0);
    add_axes3(d[a], nb);
    leaf = node(B2_DENSE, a);
  } else {
    int a = node(B2_POINTER, 0);
    add_axes3(d[a], nb / 8);
    int b = node(p.layout == 1 ? B2_BITMASKED : B2_POINTER, a);
    add_axes3(d[b], 8);
    leaf = node(B2_DENSE, b);
  }
  add_axes3(d[leaf], 4);
  for (int k = 0; k < 4; k++) {
    int pl = node(B2_PLACE, leaf);
    d[pl].dtype_bytes = 4;
  }
  sparse = p.layout != 0;
  tree = new B2Tree_t(rt, d.data(), (int)d.size());
  if (tree->s[leaf].cell_size != 16 || tree->s[leaf].node_size != 1024)
    throw B2Error(TI_ERROR_INVALID_STATE, "leaf block must be 64 x float4");

  memset(&M, 0, sizeof(M));
  M.n_grid = p.n_grid;
  M.nb = nb;
  M.pot = (p.n_grid & (p.n_grid - 1)) == 0;
  M.nb_shift = 0;
  while ((1 << M.nb_shift) < nb) M.nb_shift++;
  M.dx = 1.0f / (float)p.n_grid;
  M.inv_dx = (float)p.n_grid;
  M.dt = p.dt;
  M.p_vol = p.p_vol > 0.f ? p.p_vol : (M.dx * 0.5f) * (M.dx * 0.5f);
  M.model = p.model;
  M.mu_0 = p.E / (2.f * (1.f + p.nu)), M.lambda_0 = p.E * p.nu / ((1.f + p.nu) * (1.f - 2.f * p.nu));
  M.p_mass = M.p_vol * p.p_rho;
  M.E = p.E;
  M.gravity = p.gravity;
  M.stress_k = -p.dt * 4.0f * p.E * M.p_vol / (M.dx * M.dx);
  M.bound = p.bound;
  M.activate_halo = p.bls_activate_halo;
  M.slab_lo = p.slab_lo;
  M.slab_hi = p.slab_hi > 0 ? p.slab_hi : nb;
  M.T = tree->dev;
  M.leaf = tree->path_to(leaf);
  M.leaf_sid = leaf;
  M.layout = p.layout;
  M.np = nb / 8;
  M.top = tree->host_copy.root + tree->s[1].offset_in_parent_cell;
  M.mask_off = tree->s[2].aux_offset;
  if (tree->s[2].offset_in_parent_cell != 0 || tree->s[leaf].offset_in_parent_cell != 0 || (p.layout == 1 && tree->s[2].cell_size != 1024) ||
      (p.layout == 0 && tree->s[1].cell_size != 1024))
    throw B2Error(TI_ERROR_INVALID_STATE, "unexpected grid layout");
  const int64_t nblocks = (int64_t)nb * nb * nb;
  M.max_slots = (int)std::min<int64_t>(nblocks, p.max_bins > 0 ? p.max_bins : std::max<int64_t>(4096, p.max_particles / 64));
  M.maxc = 32;
  M.arena_cap = (u32)((p.max_particles + MPM_CHUNK - 1) / MPM_CHUNK + M.max_slots + 64);
  M.slot_of = dalloc<int>(nblocks, 0xFF);
  M.slot_key = dalloc<int>(M.max_slots);
  M.nslots = dalloc<u32>(4);
  M.nbptr = dalloc<u64>((size_t)M.max_slots * 8);
  for (int g = 0; g < 2; g++) {
    M.count[g] = dalloc<u32>(M.max_slots);
    M.chunk_of[g] = dalloc<u32>((size_t)M.max_slots * M.maxc);
    M.live[g] = dalloc<u32>((size_t)M.max_slots * MPM_MAXPASS);
    void *a = nullptr;
    B2_CUDA(cudaMalloc(&a, ((size_t)M.arena_cap + 1) * MPM_CHUNK_BYTES));  // (+ the sink chunk of an exhausted arena)
    allocs.push_back(a);
    M.arena[g] = (u8 *)a;
    M.xarena[g] = nullptr;
    if (p.model == 1) {  // F, Jp, material: a side arena with the same chunk ids (the default model's chunks stay as they are)
      void *xa = nullptr;
      B2_CUDA(cudaMalloc(&xa, ((size_t)M.arena_cap + 1) * MPM_XCHUNK_BYTES));
      allocs.push_back(xa);
      M.xarena[g] = (u8 *)xa;
    }
    M.arena_next[g] = dalloc<u32>(4);
  }
  count_sp = dalloc<u32>(M.max_slots), chunk_sp = dalloc<u32>((size_t)M.max_slots * M.maxc), key_sp = dalloc<int>(M.max_slots), nslots_sp = dalloc<u32>(4);
  if (const char *e = getenv("B2_COMPACT_EVERY")) compact_every = atoi(e);
  M.dbg = getenv("B2_DBG") ? atoi(getenv("B2_DBG")) : 0;
  M.work_counter = dalloc<u32>(24);
  M.glist = dalloc<u32>((size_t)16 * M.max_slots);
  listed_bytes = ((size_t)(nblocks + 31) / 32 * 4 + 15) / 16 * 16;
  M.listed = dalloc<u32>(listed_bytes / 4);
  M.stats = dalloc<double>(8);
  B2_CUDA(cudaFuncSetAttribute(k_mpm_p2g<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(P2W_WARPS * sizeof(P2WWarp))));
  B2_CUDA(cudaFuncSetAttribute(k_mpm_p2g<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(P2W_WARPS * sizeof(P2WWarp))));
  B2_CUDA(cudaFuncSetAttribute(k_mpm_p2g<false, P2W_CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(P2W_WARPS * sizeof(P2WWarp))));
  rt->sync();
}
void B2Mpm_t::Io::ensure() {
  if (ready) return;
  B2_CUDA(cudaStreamCreateWithFlags(&h2d, cudaStreamNonBlocking));
  B2_CUDA(cudaStreamCreateWithFlags(&d2h, cudaStreamNonBlocking));
  for (int b = 0; b < 2; b++)
    for (cudaEvent_t *e : {&up_done[b], &stage_free[b], &out_ready[b], &out_free[b]}) B2_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  ready = true;
}
void *B2Mpm_t::Io::reserve(void **buf, size_t *cap, size_t need) {
  if (*cap < need) {  // grows on first use only; everything in flight on the old buffer is drained first
    B2_CUDA(cudaDeviceSynchronize());
    if (*buf) B2_CUDA(cudaFree(*buf));
    *buf = nullptr, *cap = 0;
    B2_CUDA(cudaMalloc(buf, need));
    *cap = need;
  }
  return *buf;
}
void B2Mpm_t::Io::wait() {
  if (!ready) return;
  B2_CUDA(cudaStreamSynchronize(h2d));
  B2_CUDA(cudaStreamSynchronize(d2h));
}
void B2Mpm_t::Io::destroy() {
  if (!ready) return;
  cudaStreamSynchronize(h2d), cudaStreamSynchronize(d2h);
  for (int b = 0; b < 2; b++) {
    cudaFree(stage[b]), cudaFree(out[b]);
    cudaEventDestroy(up_done[b]), cudaEventDestroy(stage_free[b]), cudaEventDestroy(out_ready[b]), cudaEventDestroy(out_free[b]);
  }
  cudaStreamDestroy(h2d), cudaStreamDestroy(d2h);
  ready = false;
}
B2Mpm_t::~B2Mpm_t() {
  cudaStreamSynchronize(rt->stream());
  io.destroy();
  for (void *p : allocs) cudaFree(p);
  for (auto &e : prof_events) cudaEventDestroy(e);
  delete tree;
}
void B2Mpm_t::reset_bins() {
  cudaStream_t st = rt->stream();
  const int64_t nblocks = (int64_t)M.nb * M.nb * M.nb;
  B2_CUDA(cudaMemsetAsync(M.slot_of, 0xFF, nblocks * 4, st));
  B2_CUDA(cudaMemsetAsync(M.nslots, 0, 16, st));
  for (int g = 0; g < 2; g++) {
    B2_CUDA(cudaMemsetAsync(M.count[g], 0, (size_t)M.max_slots * 4, st));
    B2_CUDA(cudaMemsetAsync(M.chunk_of[g], 0, (size_t)M.max_slots * M.maxc * 4, st));
    B2_CUDA(cudaMemsetAsync(M.arena_next[g], 0, 16, st));
  }
  gen = 0;
  n_particles = 0;
  order_valid = false;
  substep_index = 0;
}
void B2Mpm_t::compact_bins() {
  cudaStream_t st = rt->stream();
  B2_CUDA(cudaMemsetAsync(nslots_sp, 0, 16, st));
  k_mpm_compact<<<rt->num_sms * 4, 256, 0, st>>>(M, gen, count_sp, chunk_sp, key_sp, nslots_sp);
  rt->count_launch();
  B2_CUDA(cudaGetLastError());
  // the old tables become the (zeroed) spares
  B2_CUDA(cudaMemsetAsync(M.count[gen], 0, (size_t)M.max_slots * 4, st));
  B2_CUDA(cudaMemsetAsync(M.chunk_of[gen], 0, (size_t)M.max_slots * M.maxc * 4, st));
  std::swap(M.count[gen], count_sp), std::swap(M.chunk_of[gen], chunk_sp), std::swap(M.slot_key, key_sp), std::swap(M.nslots, nslots_sp);
}
void B2Mpm_t::ensure_lists() {
  if (!lists_valid) {  // full regeneration (first use, or after the tree was changed behind the solver's back)
    const int64_t nblocks = (int64_t)M.nb * M.nb * M.nb;
    B2_CUDA(cudaMemsetAsync(M.listed, 0, (size_t)(nblocks + 31) / 32 * 4, rt->stream()));
    tree->listgen_async(M.leaf_sid);
    k_mpm_mark_listed<<<rt->num_sms * 4, 256, 0, rt->stream()>>>(M);
    rt->count_launch();
    lists_valid = true;
  }
}
void B2Mpm_t::clear_grid() {
  ensure_lists();
  k_mpm_clear<<<rt->num_sms * 8, 256, 0, rt->stream()>>>(tree->dev, M.leaf_sid, in_substep ? M.work_counter : nullptr,
                                                        in_substep ? M.arena_next[gen ^ 1] : nullptr);
  rt->count_launch();
}
void B2Mpm_t::p2g() {
  cudaStream_t st = rt->stream();
  if (M.model == 1) {  // the reference does this at the top of its P2G loop (mpm128.py:36-63)
    k_mpm_material<<<rt->num_sms * MAT_MINB * 2, MAT_T, 0, st>>>(M, gen);
    rt->count_launch();
  }
  if (!in_substep) B2_CUDA(cudaMemsetAsync(M.work_counter, 0, 96, st));  // (P2G takes slots from counter 0, G2P from counter 1; [8..23]: G2P order)
  if (prm.block_local)
  {
    static const int ctas = getenv("B2_P2G_CTAS") ? atoi(getenv("B2_P2G_CTAS")) : P2W_CTAS;  // experiment switch
    const size_t smem = P2W_WARPS * sizeof(P2WWarp);
    if (!M.pot)
      k_mpm_p2g<false, P2W_CTAS><<<rt->num_sms * P2W_CTAS, P2W_WARPS * 32, smem, st>>>(M, gen);
    else if (ctas == 2)
      k_mpm_p2g<true, 2><<<rt->num_sms * 2, P2W_WARPS * 32, smem, st>>>(M, gen);
    else
      k_mpm_p2g<true, 3><<<rt->num_sms * 3, P2W_WARPS * 32, smem, st>>>(M, gen);
    order_valid = true;
  }
  else
    k_mpm_p2g_plain<<<rt->num_sms * 8, 256, 0, st>>>(M, gen);
  rt->count_launch();
  // blocks that P2G activates append themselves to the leaf element list (mpm_block_ptr<true>)
}
void B2Mpm_t::grid_op() {
  ensure_lists();
  k_mpm_gridop<<<rt->num_sms * 8, 256, 0, rt->stream()>>>(M);
  rt->count_launch();
}
void B2Mpm_t::g2p() {
  cudaStream_t st = rt->stream();
  // the generation being written starts with an empty chunk arena
  if (!in_substep) {
    B2_CUDA(cudaMemsetAsync(M.arena_next[gen ^ 1], 0, 16, st));
    B2_CUDA(cudaMemsetAsync(M.work_counter, 0, 32, st));
  }
  if (prm.block_local) {
    static const bool no_order = getenv("B2_G2P_SORT") != nullptr;  // experiment switch: always sort
    static const bool r1_form = getenv("B2_G2P_R1") != nullptr;  // experiment switch: the r1 bin prologue
    if (order_valid && !no_order && r1_form)
      k_mpm_g2p_bls<G2PK_T, G2PK_MINB, true><<<rt->num_sms * G2PK_MINB, G2PK_T, 0, st>>>(M, gen);
    else if (M.model == 1) {
      if (!order_valid) throw B2Error(TI_ERROR_INVALID_STATE, "model 1: g2p needs the p2g of the same substep before it");
      k_mpm_g2p_ord<64, G2PM_MINB, true><<<rt->num_sms * G2PM_MINB, 64, 0, st>>>(M, gen);
    } else if (order_valid && !no_order)
      k_mpm_g2p_ord<G2PK_T, G2PK_MINB, false><<<rt->num_sms * G2PK_MINB, G2PK_T, 0, st>>>(M, gen);
    else
      k_mpm_g2p_bls<64, 12, false><<<rt->num_sms * 12, 64, 0, st>>>(M, gen);
  } else {
    k_mpm_g2p<false><<<rt->num_sms * 3, G2P_PLAIN_THREADS, 0, st>>>(M, gen);
  }
  rt->count_launch();
  gen ^= 1;
  order_valid = false;
}
// one substep's fixed points for drivers that enqueue the phases themselves (b2_slab.cu): cadence work (bin compaction,
// sparse life cycle), counters reset by the clear kernel instead of memset nodes
void B2Mpm_t::begin_substep() {
  if (compact_every > 0 && substep_index > 0 && substep_index % compact_every == 0) compact_bins();
  if (sparse && prm.deactivate_every > 0 && substep_index % prm.deactivate_every == 0) {
    tree->deactivate_all(1);
    lists_valid = false;
  }
  in_substep = true;
}
void B2Mpm_t::end_substep() {
  in_substep = false;
  substep_index++;
}
void B2Mpm_t::substep(int n) {
  cudaStream_t st = rt->stream();
  constexpr int NE = 7;  // events per substep: [deactivate + GC] [clear] [p2g] [listgen] [grid-op] [g2p]
  if (profiling) {
    while ((int)prof_events.size() < NE * n) {
      cudaEvent_t e;
      B2_CUDA(cudaEventCreate(&e));
      prof_events.push_back(e);
    }
  }
  in_substep = true;  // clear resets the counters of the whole substep: no memset nodes between the kernels
  struct Leave {
    bool &f;
    ~Leave() { f = false; }
  } leave{in_substep};
  for (int s = 0; s < n; s++, substep_index++) {
    cudaEvent_t *e = profiling ? &prof_events[NE * s] : nullptr;
    if (e) B2_CUDA(cudaEventRecord(e[0], st));
    if (compact_every > 0 && substep_index > 0 && substep_index % compact_every == 0) compact_bins();
    if (sparse && prm.deactivate_every > 0 && substep_index % prm.deactivate_every == 0) {
      // the reference's sparse life cycle, all enqueued: children first, then the pointer level, recycled containers
      // zero-filled by the fused GC; the leaf list is regenerated (empty) and P2G activates what it touches
      if (prm.layout == 1) {
        // pointer -> bitmasked: the bitmasked container (cells + mask) IS the pointer level's NodeManager element, and the
        // fused GC zero-fills recycled elements whole -- mask included -- so the children need no pass of their own.
        // ONE launch: deactivate every cell of the pointer container, grid barrier, GC; the lists of the two levels below
        // and the leaf bitmap are emptied by the same kernel (r2 v3 used 3 launches + 3 memsets: 0.042 ms per substep)
        static const bool unfused = getenv("B2_LIFECYCLE_UNFUSED") != nullptr;  // experiment switch
        if (unfused) {
          tree->deactivate_level_async(1);
          tree->reset_list_async(2), tree->reset_list_async(M.leaf_sid);
          B2_CUDA(cudaMemsetAsync(M.listed, 0, listed_bytes, st));
        } else {
          tree->deactivate_all_gc_async(1, M.top, 2, M.leaf_sid, M.listed, listed_bytes);
        }
        lists_valid = true;
      } else {
        tree->deactivate_all(1);
        lists_valid = false;
        ensure_lists();  // (full regeneration: counted with the life cycle, not with the clear)
      }
    }
    if (e) B2_CUDA(cudaEventRecord(e[1], st));
    clear_grid();
    if (e) B2_CUDA(cudaEventRecord(e[2], st));
    p2g();
    if (e) B2_CUDA(cudaEventRecord(e[3], st));
    ensure_lists();
    if (e) B2_CUDA(cudaEventRecord(e[4], st));
    grid_op();
    if (e) B2_CUDA(cudaEventRecord(e[5], st));
    g2p();
    if (e) B2_CUDA(cudaEventRecord(e[6], st));
  }
  if (profiling && n > 0) {  // one synchronisation per call, then accumulate every substep's phases
    B2_CUDA(cudaEventSynchronize(prof_events[NE * (n - 1) + NE - 1]));
    for (int s = 0; s < n; s++)
      for (int k = 0; k < NE - 1; k++) {
        float ms;
        B2_CUDA(cudaEventElapsedTime(&ms, prof_events[NE * s + k], prof_events[NE * s + k + 1]));
        phase_sum[k == 0 ? 5 : k - 1] += ms;  // slots: clear, p2g, listgen, grid_op, g2p, deactivate_gc
      }
    phase_count += n;
  }
  B2_CUDA(cudaGetLastError());
}

static void require(bool ok, TiError code, const char *msg) {
  if (!ok) throw B2Error(code, msg);
}

extern "C" {
B2Mpm b2_mpm_create(TiRuntime runtime, const B2MpmParams *params) {
  B2_API_BEGIN
  require(runtime != nullptr, TI_ERROR_ARGUMENT_NULL, "runtime");
  require(params != nullptr, TI_ERROR_ARGUMENT_NULL, "params");
  B2_CUDA(cudaSetDevice(runtime->device));
  return new B2Mpm_t(runtime, *params);
  B2_API_END(nullptr)
}
void b2_mpm_destroy(B2Mpm mpm) { delete mpm; }
B2Tree b2_mpm_tree(B2Mpm mpm) { return mpm ? mpm->tree : nullptr; }
TiRuntime b2_mpm_runtime(B2Mpm mpm) { return mpm ? mpm->rt : nullptr; }
int32_t b2_mpm_leaf_snode(B2Mpm mpm) { return mpm ? mpm->M.leaf_sid : -1; }

int32_t b2_mpm_add_particles(B2Mpm m, int64_t n, const float *x, const float *v, const float *C, const float *J, const int32_t *ids) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  if (n <= 0) return 0;
  require(x != nullptr, TI_ERROR_ARGUMENT_NULL, "x");
  require(m->n_particles + n <= m->prm.max_particles, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "more particles than max_particles");
  cudaStream_t st = m->rt->stream();
  auto on_host = [](const void *p) { return p == nullptr || !b2_is_device_pointer(p); };
  if (!b2_is_device_pointer(x) && on_host(v) && on_host(C) && on_host(J) && on_host(ids)) {
    // host arrays: upload on the copy stream into the staging buffer whose last insert has finished
    B2Mpm_t::Io &io = m->io;
    io.ensure();
    const int b = (int)(io.up_i++ & 1u);
    char *base = (char *)io.reserve(&io.stage[b], &io.stage_cap[b], (size_t)n * 68);
    B2_CUDA(cudaStreamWaitEvent(io.h2d, io.stage_free[b], 0));
    float *dx = (float *)base, *dv = (float *)(base + (size_t)n * 12), *dC = (float *)(base + (size_t)n * 24), *dJ = (float *)(base + (size_t)n * 60);
    int *di = (int *)(base + (size_t)n * 64);
    B2_CUDA(cudaMemcpyAsync(dx, x, (size_t)n * 12, cudaMemcpyHostToDevice, io.h2d));
    if (v) B2_CUDA(cudaMemcpyAsync(dv, v, (size_t)n * 12, cudaMemcpyHostToDevice, io.h2d));
    if (C) B2_CUDA(cudaMemcpyAsync(dC, C, (size_t)n * 36, cudaMemcpyHostToDevice, io.h2d));
    if (J) B2_CUDA(cudaMemcpyAsync(dJ, J, (size_t)n * 4, cudaMemcpyHostToDevice, io.h2d));
    if (ids) B2_CUDA(cudaMemcpyAsync(di, ids, (size_t)n * 4, cudaMemcpyHostToDevice, io.h2d));
    B2_CUDA(cudaEventRecord(io.up_done[b], io.h2d));
    B2_CUDA(cudaStreamWaitEvent(st, io.up_done[b], 0));
    k_mpm_insert<<<m->rt->grid_for(n, 256, 8), 256, 0, st>>>(m->M, m->gen, dx, v ? dv : nullptr, C ? dC : nullptr, J ? dJ : nullptr, ids ? di : nullptr, n, m->n_particles);
    m->rt->count_launch();
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(io.stage_free[b], st));
    m->n_particles += n;
    m->order_valid = false;
    return 0;
  }
  B2Staged sx, sv, sc, sj, si;
  sx.init(x, (size_t)n * 12, true, false, st);
  if (v) sv.init(v, (size_t)n * 12, true, false, st);
  if (C) sc.init(C, (size_t)n * 36, true, false, st);
  if (J) sj.init(J, (size_t)n * 4, true, false, st);
  if (ids) si.init(ids, (size_t)n * 4, true, false, st);
  k_mpm_insert<<<m->rt->grid_for(n, 256, 8), 256, 0, st>>>(m->M, m->gen, sx.as<float>(), sv.as<float>(), sc.as<float>(), sj.as<float>(), si.as<int>(), n, m->n_particles);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  m->n_particles += n;
  m->order_valid = false;
  sx.finish(), sv.finish(), sc.finish(), sj.finish(), si.finish();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_set_particles(B2Mpm m, int64_t n, const float *x, const float *v, const float *C, const float *J) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->reset_bins();
  B2_API_END(-1)
  return b2_mpm_add_particles(m, n, x, v, C, J, nullptr);
}
int64_t b2_mpm_num_particles(B2Mpm m) { return m ? m->n_particles : -1; }
int32_t b2_mpm_set_particles_model(B2Mpm m, int64_t n, const float *x, const float *v, const float *C, const float *Fm, const float *Jp, const int32_t *mat) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  require(m->M.model == 1, TI_ERROR_INVALID_STATE, "the solver was created with model 0 (B2MpmParams.model)");
  m->reset_bins();
  if (n <= 0) return 0;
  require(x != nullptr, TI_ERROR_ARGUMENT_NULL, "x");
  require(n <= m->prm.max_particles, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "more particles than max_particles");
  cudaStream_t st = m->rt->stream();
  B2Staged sx, sv, sc, sf, sj, sm;
  sx.init(x, (size_t)n * 12, true, false, st);
  if (v) sv.init(v, (size_t)n * 12, true, false, st);
  if (C) sc.init(C, (size_t)n * 36, true, false, st);
  if (Fm) sf.init(Fm, (size_t)n * 36, true, false, st);
  if (Jp) sj.init(Jp, (size_t)n * 4, true, false, st);
  if (mat) sm.init(mat, (size_t)n * 4, true, false, st);
  k_mpm_insert<<<m->rt->grid_for(n, 256, 8), 256, 0, st>>>(m->M, m->gen, sx.as<float>(), sv.as<float>(), sc.as<float>(), nullptr, nullptr, n, 0,
                                                            sf.as<float>(), sj.as<float>(), sm.as<int>());
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  m->n_particles = n;
  m->order_valid = false;
  sx.finish(), sv.finish(), sc.finish(), sf.finish(), sj.finish(), sm.finish();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_get_particles_model(B2Mpm m, float *x, float *v, float *C, float *Fm, float *Jp, int32_t *mat) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  require(m->M.model == 1, TI_ERROR_INVALID_STATE, "the solver was created with model 0 (B2MpmParams.model)");
  const int64_t n = m->n_particles;
  if (n == 0) return 0;
  cudaStream_t st = m->rt->stream();
  B2Staged sx, sv, sc, sf, sj, sm;
  if (x) sx.init(x, (size_t)n * 12, false, true, st);
  if (v) sv.init(v, (size_t)n * 12, false, true, st);
  if (C) sc.init(C, (size_t)n * 36, false, true, st);
  if (Fm) sf.init(Fm, (size_t)n * 36, false, true, st);
  if (Jp) sj.init(Jp, (size_t)n * 4, false, true, st);
  if (mat) sm.init(mat, (size_t)n * 4, false, true, st);
  k_mpm_extract<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, m->gen, sx.as<float>(), sv.as<float>(), sc.as<float>(), nullptr, n, sf.as<float>(),
                                                    sj.as<float>(), sm.as<int>());
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  sx.finish(), sv.finish(), sc.finish(), sf.finish(), sj.finish(), sm.finish();
  m->rt->sync();
  return 0;
  B2_API_END(-1)
}
static int32_t mpm_get_particles(B2Mpm m, float *x, float *v, float *C, float *J, bool wait) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  const int64_t n = m->n_particles;
  if (n == 0) return 0;
  cudaStream_t st = m->rt->stream();
  auto on_host = [](const void *p) { return p == nullptr || !b2_is_device_pointer(p); };
  if (on_host(x) && on_host(v) && on_host(C) && on_host(J)) {
    // host arrays: extract into the staging buffer whose last download has finished, download on the copy stream
    B2Mpm_t::Io &io = m->io;
    io.ensure();
    const int b = (int)(io.out_i++ & 1u);
    const size_t ox = 0, ov = ox + (x ? (size_t)n * 12 : 0), oc = ov + (v ? (size_t)n * 12 : 0), oj = oc + (C ? (size_t)n * 36 : 0);
    const size_t need = oj + (J ? (size_t)n * 4 : 0);
    if (need == 0) return 0;
    char *base = (char *)io.reserve(&io.out[b], &io.out_cap[b], need);
    B2_CUDA(cudaStreamWaitEvent(st, io.out_free[b], 0));
    B2_CUDA(cudaMemsetAsync(base, 0, need, st));
    k_mpm_extract<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, m->gen, x ? (float *)(base + ox) : nullptr, v ? (float *)(base + ov) : nullptr,
                                                      C ? (float *)(base + oc) : nullptr, J ? (float *)(base + oj) : nullptr, n);
    m->rt->count_launch();
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(io.out_ready[b], st));
    B2_CUDA(cudaStreamWaitEvent(io.d2h, io.out_ready[b], 0));
    if (x) B2_CUDA(cudaMemcpyAsync(x, base + ox, (size_t)n * 12, cudaMemcpyDeviceToHost, io.d2h));
    if (v) B2_CUDA(cudaMemcpyAsync(v, base + ov, (size_t)n * 12, cudaMemcpyDeviceToHost, io.d2h));
    if (C) B2_CUDA(cudaMemcpyAsync(C, base + oc, (size_t)n * 36, cudaMemcpyDeviceToHost, io.d2h));
    if (J) B2_CUDA(cudaMemcpyAsync(J, base + oj, (size_t)n * 4, cudaMemcpyDeviceToHost, io.d2h));
    B2_CUDA(cudaEventRecord(io.out_free[b], io.d2h));
    if (wait) B2_CUDA(cudaStreamSynchronize(io.d2h));
    return 0;
  }
  B2Staged sx, sv, sc, sj;
  if (x) sx.init(x, (size_t)n * 12, false, true, st);
  if (v) sv.init(v, (size_t)n * 12, false, true, st);
  if (C) sc.init(C, (size_t)n * 36, false, true, st);
  if (J) sj.init(J, (size_t)n * 4, false, true, st);
  k_mpm_extract<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, m->gen, sx.as<float>(), sv.as<float>(), sc.as<float>(), sj.as<float>(), n);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  sx.finish(), sv.finish(), sc.finish(), sj.finish();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_get_particles(B2Mpm m, float *x, float *v, float *C, float *J) { return mpm_get_particles(m, x, v, C, J, true); }
int32_t b2_mpm_get_particles_async(B2Mpm m, float *x, float *v, float *C, float *J) { return mpm_get_particles(m, x, v, C, J, false); }
int32_t b2_mpm_wait_io(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->io.wait();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_substep(B2Mpm m, int32_t nsteps) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->substep(nsteps);
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_clear_grid(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->clear_grid();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_p2g(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->p2g();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_grid_op(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->grid_op();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_g2p(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->g2p();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_begin_substep(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->begin_substep();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_end_substep(B2Mpm m) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  m->end_substep();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_grid_to_dense(B2Mpm m, float *out) {
  B2_API_BEGIN
  require(m != nullptr && out != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm/out");
  cudaStream_t st = m->rt->stream();
  const size_t bytes = (size_t)m->M.n_grid * m->M.n_grid * m->M.n_grid * 16;
  B2Staged o;
  o.init(out, bytes, false, true, st);
  if (!o.owned) B2_CUDA(cudaMemsetAsync(o.dev, 0, bytes, st));
  m->ensure_lists();
  k_mpm_grid_to_dense<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, o.as<float4>());
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  o.finish();
  return 0;
  B2_API_END(-1)
}
int64_t b2_mpm_grid_blocks(B2Mpm m, int32_t *keys, float *cells, int64_t cap) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  cudaStream_t st = m->rt->stream();
  m->ensure_lists();
  m->rt->sync();
  const int64_t n = (int64_t)m->tree->list_count(m->M.leaf_sid);
  if (keys == nullptr || cells == nullptr || cap <= 0) return n;
  const int64_t k = std::min(n, cap);
  B2Staged sk, sc;
  sk.init(keys, (size_t)k * 4, false, true, st);
  sc.init(cells, (size_t)k * 1024, false, true, st);
  k_mpm_grid_blocks<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, sk.as<int>(), sc.as<float4>(), k);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  sk.finish(), sc.finish();
  return n;
  B2_API_END(-1)
}
int32_t b2_mpm_stats(B2Mpm m, double *out) {
  B2_API_BEGIN
  require(m != nullptr && out != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm/out");
  cudaStream_t st = m->rt->stream();
  B2_CUDA(cudaMemsetAsync(m->M.stats, 0, 64, st));
  k_mpm_stats<<<m->rt->num_sms * 4, 256, 0, st>>>(m->M, m->gen);
  m->rt->count_launch();
  B2_CUDA(cudaMemcpyAsync(out, m->M.stats, 64, cudaMemcpyDeviceToHost, st));
  m->rt->sync();
  m->ensure_lists();
  out[4] = (double)m->tree->list_count(m->M.leaf_sid);
  return 0;
  B2_API_END(-1)
}
// Enqueues the reduction into 8 doubles of device memory (zeroed first); the slab driver all-reduces them.
int32_t b2_mpm_diagnostics_device(B2Mpm m, double *out8_dev) {
  B2_API_BEGIN
  require(m != nullptr && out8_dev != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm/out");
  cudaStream_t st = m->rt->stream();
  B2_CUDA(cudaMemsetAsync(out8_dev, 0, 64, st));
  k_mpm_diag<<<m->rt->num_sms * 4, 256, 0, st>>>(m->M, m->gen, out8_dev);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_diagnostics(B2Mpm m, double *out8) {
  B2_API_BEGIN
  require(m != nullptr && out8 != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm/out");
  if (b2_mpm_diagnostics_device(m, m->M.stats) != 0) return -1;
  B2_CUDA(cudaMemcpyAsync(out8, m->M.stats, 64, cudaMemcpyDeviceToHost, m->rt->stream()));
  m->rt->sync();
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_invalidate_lists(B2Mpm m) {
  if (!m) return -1;
  m->lists_valid = false;
  return 0;
}
int32_t b2_mpm_set_compact_every(B2Mpm m, int32_t k) {
  if (!m || k < 0) return -1;
  m->compact_every = k;
  return 0;
}
int32_t b2_mpm_set_deactivate_every(B2Mpm m, int32_t k) {
  if (!m || k < 0) return -1;
  m->prm.deactivate_every = k;
  return 0;
}
int32_t b2_mpm_set_profiling(B2Mpm m, int32_t on) {
  if (!m) return -1;
  m->profiling = on != 0;
  for (auto &v : m->phase_sum) v = 0;
  m->phase_count = 0;
  return 0;
}
int32_t b2_mpm_phase_ms(B2Mpm m, float *out) {
  if (!m || !out) return -1;
  for (int k = 0; k < 6; k++) out[k] = m->phase_count ? (float)(m->phase_sum[k] / (double)m->phase_count) : 0.f;
  return 0;
}
// ---- slab decomposition (see b2_mpm_slab.cuh) ----------------------------------------------
static void check_msg(const void *buf, int64_t cap) {
  require(buf != nullptr, TI_ERROR_ARGUMENT_NULL, "buf");
  require(cap >= 0, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "cap");
  require(b2_is_device_pointer(buf), TI_ERROR_INVALID_ARGUMENT, "message buffers must be device memory");
}
int32_t b2_mpm_pack_blocks(B2Mpm m, int32_t layer, void *buf, int64_t cap) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  check_msg(buf, cap);
  cudaStream_t st = m->rt->stream();
  if (!m->in_substep) B2_CUDA(cudaMemsetAsync(buf, 0, MPM_MSG_HEADER, st));  // (a driver inside begin/end_substep zeroes its headers itself)
  m->ensure_lists();
  k_mpm_pack_blocks<<<m->rt->num_sms * 4, 256, 0, st>>>(m->M, layer, (u8 *)buf, cap);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_unpack_blocks(B2Mpm m, const void *buf, int64_t cap, int32_t mode) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  check_msg(buf, cap);
  require(mode >= 0 && mode <= 2, TI_ERROR_INVALID_ARGUMENT, "mode");
  k_mpm_unpack_blocks<<<m->rt->num_sms * 4, 256, 0, m->rt->stream()>>>(m->M, (const u8 *)buf, cap, mode);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_extract_migrants2(B2Mpm m, void *bufL, int64_t capL, void *bufR, int64_t capR) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  if (bufL) check_msg(bufL, capL);
  if (bufR) check_msg(bufR, capR);
  if (!bufL && !bufR) return 0;
  cudaStream_t st = m->rt->stream();
  if (!m->in_substep) {
    if (bufL) B2_CUDA(cudaMemsetAsync(bufL, 0, MPM_MSG_HEADER, st));
    if (bufR) B2_CUDA(cudaMemsetAsync(bufR, 0, MPM_MSG_HEADER, st));
  }
  k_mpm_extract_migrants<<<m->rt->num_sms * 4, 256, 0, st>>>(m->M, m->gen, (u8 *)bufL, capL, (u8 *)bufR, capR);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_extract_migrants(B2Mpm m, int32_t side, void *buf, int64_t cap) {
  if (side != 0 && side != 1) {
    ti_set_last_error(TI_ERROR_INVALID_ARGUMENT, "side");
    return -1;
  }
  return side == 0 ? b2_mpm_extract_migrants2(m, buf, cap, nullptr, 0) : b2_mpm_extract_migrants2(m, nullptr, 0, buf, cap);
}
int32_t b2_mpm_insert_records2(B2Mpm m, const void *bufA, int64_t capA, const void *bufB, int64_t capB) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  if (bufA) check_msg(bufA, capA);
  if (bufB) check_msg(bufB, capB);
  if (!bufA && !bufB) return 0;
  m->order_valid = false;
  const int64_t work = std::max<int64_t>((bufA ? capA : 0) + (bufB ? capB : 0) + 32, 1);
  k_mpm_insert_records<<<m->rt->grid_for(work, 256, 4), 256, 0, m->rt->stream()>>>(m->M, m->gen, (const u8 *)bufA, capA, (const u8 *)bufB, capB);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_insert_records(B2Mpm m, const void *buf, int64_t cap) { return b2_mpm_insert_records2(m, buf, cap, nullptr, 0); }
int32_t b2_mpm_zero_headers(B2Mpm m, void *a, void *b, void *c, void *d) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  k_mpm_zero_headers<<<1, 4, 0, m->rt->stream()>>>((u32 *)a, (u32 *)b, (u32 *)c, (u32 *)d);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  return 0;
  B2_API_END(-1)
}
int32_t b2_mpm_export_async(B2Mpm m, float *x, float *v, float *C, float *J, int32_t *ids, int64_t cap, int64_t *count) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  require(count != nullptr, TI_ERROR_ARGUMENT_NULL, "count");
  require(cap > 0, TI_ERROR_ARGUMENT_OUT_OF_RANGE, "cap");
  auto on_host = [](const void *p) { return p == nullptr || !b2_is_device_pointer(p); };
  require(on_host(x) && on_host(v) && on_host(C) && on_host(J) && on_host(ids) && !b2_is_device_pointer(count), TI_ERROR_INVALID_ARGUMENT,
          "export_async fills host arrays");
  cudaStream_t st = m->rt->stream();
  B2Mpm_t::Io &io = m->io;
  io.ensure();
  const int b = (int)(io.out_i++ & 1u);
  const size_t ox = 16, ov = ox + (x ? (size_t)cap * 12 : 0), oc = ov + (v ? (size_t)cap * 12 : 0), oj = oc + (C ? (size_t)cap * 36 : 0);
  const size_t oi = oj + (J ? (size_t)cap * 4 : 0), need = oi + (ids ? (size_t)cap * 4 : 0);
  char *base = (char *)io.reserve(&io.out[b], &io.out_cap[b], need);
  B2_CUDA(cudaStreamWaitEvent(st, io.out_free[b], 0));
  B2_CUDA(cudaMemsetAsync(base, 0, 16, st));  // the record counter lives in front of the arrays
  k_mpm_export<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, m->gen, x ? (float *)(base + ox) : nullptr, v ? (float *)(base + ov) : nullptr,
                                                   C ? (float *)(base + oc) : nullptr, J ? (float *)(base + oj) : nullptr,
                                                   ids ? (int *)(base + oi) : nullptr, cap, (unsigned long long *)base);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaEventRecord(io.out_ready[b], st));
  B2_CUDA(cudaStreamWaitEvent(io.d2h, io.out_ready[b], 0));
  B2_CUDA(cudaMemcpyAsync(count, base, 8, cudaMemcpyDeviceToHost, io.d2h));
  if (x) B2_CUDA(cudaMemcpyAsync(x, base + ox, (size_t)cap * 12, cudaMemcpyDeviceToHost, io.d2h));
  if (v) B2_CUDA(cudaMemcpyAsync(v, base + ov, (size_t)cap * 12, cudaMemcpyDeviceToHost, io.d2h));
  if (C) B2_CUDA(cudaMemcpyAsync(C, base + oc, (size_t)cap * 36, cudaMemcpyDeviceToHost, io.d2h));
  if (J) B2_CUDA(cudaMemcpyAsync(J, base + oj, (size_t)cap * 4, cudaMemcpyDeviceToHost, io.d2h));
  if (ids) B2_CUDA(cudaMemcpyAsync(ids, base + oi, (size_t)cap * 4, cudaMemcpyDeviceToHost, io.d2h));
  B2_CUDA(cudaEventRecord(io.out_free[b], io.d2h));
  return 0;
  B2_API_END(-1)
}
int64_t b2_mpm_export(B2Mpm m, float *x, float *v, float *C, float *J, int32_t *ids, int64_t cap) {
  B2_API_BEGIN
  require(m != nullptr, TI_ERROR_ARGUMENT_NULL, "mpm");
  cudaStream_t st = m->rt->stream();
  B2Staged sx, sv, sc, sj, si;
  if (x) sx.init(x, (size_t)cap * 12, false, true, st);
  if (v) sv.init(v, (size_t)cap * 12, false, true, st);
  if (C) sc.init(C, (size_t)cap * 36, false, true, st);
  if (J) sj.init(J, (size_t)cap * 4, false, true, st);
  if (ids) si.init(ids, (size_t)cap * 4, false, true, st);
  unsigned long long *counter = (unsigned long long *)(m->M.stats + 7);
  B2_CUDA(cudaMemsetAsync(counter, 0, 8, st));
  k_mpm_export<<<m->rt->num_sms * 8, 256, 0, st>>>(m->M, m->gen, sx.as<float>(), sv.as<float>(), sc.as<float>(), sj.as<float>(), si.as<int>(), cap, counter);
  m->rt->count_launch();
  B2_CUDA(cudaGetLastError());
  unsigned long long n = 0;
  B2_CUDA(cudaMemcpyAsync(&n, counter, 8, cudaMemcpyDeviceToHost, st));
  m->rt->sync();
  sx.finish(), sv.finish(), sc.finish(), sj.finish(), si.finish();
  return (int64_t)n;
  B2_API_END(-1)
}
}  // extern "C"
