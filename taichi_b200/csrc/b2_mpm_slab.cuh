// b2_mpm_slab.cuh -- device side of the slab decomposition (new: the reference is single-GPU,
// rhi/cuda/cuda_context.cpp:22).  Included at the end of b2_mpm.cu.
//
// A rank owns the 4^3-blocks with slab_lo <= bx < slab_hi and the particles whose base cell lies
// in them.  A particle only touches nodes base..base+2, so P2G spills into exactly one foreign
// block layer, bx == slab_hi (the ghost layer).  Per substep:
//   A  after P2G      ghost layer (bx == hi)      -> rank+1, which ADDS it into its owned blocks
//   B  after grid-op  owned layer  (bx == lo)     -> rank-1, which OVERWRITES its ghost layer
//   C  after G2P      particles binned outside [lo,hi) -> rank-1 / rank+1, which insert them
// Messages are fixed-capacity device buffers {u32 count; pad[3]; records...}: the receiver reads
// the count on the device, so no host synchronisation is needed anywhere in the exchange.
#pragma once

struct MpmBlockRec {  // 1040 B
  int key;
  int pad[3];
  float4 cells[64];
};
struct MpmParticleRec {  // 80 B
  float4 p0, p1, p2, p3;
  u32 id;
  u32 pad[3];
};
#define MPM_MSG_HEADER 16

// active leaf blocks with pcoord[0] == layer  ->  records (struct-for over the leaf list)
__global__ void __launch_bounds__(256) k_mpm_pack_blocks(MpmDev M, int layer, u8 *buf, long long cap) {
  __shared__ long long sRec[4];
  B2CList *L = M.T->lists[M.leaf_sid];
  const u32 ne = L->count;
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  u32 *count = (u32 *)buf;
  MpmBlockRec *recs = (MpmBlockRec *)(buf + MPM_MSG_HEADER);
  for (u32 e0 = blockIdx.x * 4; e0 < ne; e0 += gridDim.x * 4) {
    const u32 e = e0 + sub;
    const B2Element *E = e < ne ? (const B2Element *)b2_list_get(L, e) : nullptr;
    const bool take = E && E->pcoord[0] == layer;
    if (t == 0) {
      long long r = -1;
      if (take) {
        r = (long long)atomicAdd(count, 1u);
        if (r >= cap) {
          b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, -2);
          r = -1;
        }
      }
      sRec[sub] = r;
    }
    __syncthreads();
    const long long r = sRec[sub];
    if (take && r >= 0) {
      if (t == 0) recs[r].key = mpm_key(M, E->pcoord[0], E->pcoord[1], E->pcoord[2]);
      recs[r].cells[t] = ((const float4 *)E->node)[t];
    }
    __syncthreads();
  }
}
// Clamp the count to the capacity (a sender that overflowed has already raised the error flag).
__device__ __forceinline__ u32 mpm_msg_count(const u8 *buf, long long cap) {
  const u32 c = *(const u32 *)buf;
  return (long long)c > cap ? (u32)cap : c;
}
// mode 0: activate + add (ghost contributions arriving at their owner); mode 1: overwrite blocks that are active here;
// mode 2: add into blocks that are active here (the owner's contributions arriving at a ghost layer)
__global__ void __launch_bounds__(256) k_mpm_unpack_blocks(MpmDev M, const u8 *buf, long long cap, int mode) {
  __shared__ float4 *sDst[4];
  const u32 n = mpm_msg_count(buf, cap);
  const MpmBlockRec *recs = (const MpmBlockRec *)(buf + MPM_MSG_HEADER);
  const int sub = threadIdx.x >> 6, t = threadIdx.x & 63;
  for (u32 r0 = blockIdx.x * 4; r0 < n; r0 += gridDim.x * 4) {
    const u32 r = r0 + sub;
    if (t == 0) {
      float4 *dst = nullptr;
      if (r < n) {
        int bx, by, bz;
        mpm_unkey(M, recs[r].key, bx, by, bz);
        dst = mode == 0 ? mpm_block_ptr<true>(M, bx, by, bz) : mpm_block_ptr<false>(M, bx, by, bz);  // modes 1, 2: existing blocks only
      }
      sDst[sub] = dst;
    }
    __syncthreads();
    float4 *dst = sDst[sub];
    if (dst) {
      const float4 v = recs[r].cells[t];
      if (mode != 1) {
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) red_add_v4(dst + t, v);
      } else {
        dst[t] = v;
      }
    }
    __syncthreads();
  }
}

// side 0: bins with bx < slab_lo, side 1: bins with bx >= slab_hi  ->  particle records; the bins
// are emptied.  One CTA per slot.
// (bufL, capL): message for the lower neighbour (bins with bx < slab_lo), (bufR, capR): for the upper one; either may be null
__global__ void __launch_bounds__(256) k_mpm_extract_migrants(MpmDev M, int gen, u8 *bufL, long long capL, u8 *bufR, long long capR) {
  __shared__ long long sBase;
  const u32 ns = *M.nslots;
  for (u32 slot = blockIdx.x; slot < ns; slot += gridDim.x) {
    const u32 total = M.count[gen][slot];
    if (total == 0) continue;
    int bx, by, bz;
    mpm_unkey(M, M.slot_key[slot], bx, by, bz);
    u8 *buf = bx < M.slab_lo ? bufL : bx >= M.slab_hi ? bufR : nullptr;
    if (!buf) continue;
    const long long cap = bx < M.slab_lo ? capL : capR;
    u32 *count = (u32 *)buf;
    MpmParticleRec *recs = (MpmParticleRec *)(buf + MPM_MSG_HEADER);
    __syncthreads();
    if (threadIdx.x == 0) {
      // reserve [b, b + total) with one atomicAdd (a CAS loop on this single counter serialises the hundreds of CTAs that hold a
      // migrant bin: 94 us measured).  A bin that does not fit stays where it is -- its particles travel with a later substep,
      // nothing is lost -- and the part of its reservation that lies inside the message is filled with tombstones, which the
      // receiver skips; the receiver clamps the count to the capacity.
      long long b = (long long)atomicAdd(count, total);
      if (b + (long long)total > cap) {
        b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, -3);
        b = -1 - b;  // (< 0: overflow; the reserved start is kept for the tombstones)
      }
      sBase = b;
    }
    __syncthreads();
    const long long base = sBase;
    if (base < 0) {  // (uniform: sBase is shared)
      const long long b0 = -1 - base;
      for (long long q = b0 + threadIdx.x; q < cap && q < b0 + (long long)total; q += blockDim.x) recs[q].p0 = make_float4(MPM_TOMB, 0.f, 0.f, 1.f);
      continue;
    }
    for (u32 q = threadIdx.x; q < total; q += blockDim.x) {
      const u8 *ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)slot * M.maxc + q / MPM_CHUNK]);
      const u32 i = q % MPM_CHUNK;
      {
        MpmParticleRec rec;
        rec.p0 = ((const float4 *)ch)[i];
        rec.p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[i];
        rec.p2 = ((const float4 *)(ch + 2 * MPM_PLANE_BYTES))[i];
        rec.p3 = ((const float4 *)(ch + 3 * MPM_PLANE_BYTES))[i];
        rec.id = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[i];
        rec.pad[0] = rec.pad[1] = rec.pad[2] = 0;
        recs[base + q] = rec;  // a tombstone travels as a tombstone and is skipped by the receiver
      }
    }
    __syncthreads();
    for (u32 c = threadIdx.x; c < (total + MPM_CHUNK - 1) / MPM_CHUNK; c += blockDim.x) M.chunk_of[gen][(u64)slot * M.maxc + c] = 0;
    if (threadIdx.x == 0) M.count[gen][slot] = 0;
  }
}
// records of up to two messages (either may be null) in one launch
__global__ void __launch_bounds__(256) k_mpm_insert_records(MpmDev M, int gen, const u8 *bufA, long long capA, const u8 *bufB, long long capB) {
  const u32 nA = bufA ? mpm_msg_count(bufA, capA) : 0u, nB = bufB ? mpm_msg_count(bufB, capB) : 0u;
  const u32 nAr = (nA + 31) / 32 * 32, n = nAr + nB;  // (message B starts on a warp boundary)
  const u32 nround = (n + 31) / 32 * 32;
  for (u32 p = blockIdx.x * blockDim.x + threadIdx.x; p < nround; p += gridDim.x * blockDim.x) {
    bool valid = p < nA || (p >= nAr && p < n);
    MpmParticleRec rec;
    rec.p0 = make_float4(0, 0, 0, 1), rec.p1 = rec.p2 = rec.p3 = make_float4(0, 0, 0, 0), rec.id = 0;
    int key = 0, cell = 0;
    if (valid) {
      rec = p < nA ? ((const MpmParticleRec *)(bufA + MPM_MSG_HEADER))[p] : ((const MpmParticleRec *)(bufB + MPM_MSG_HEADER))[p - nAr];
      if (rec.p0.x == MPM_TOMB) {
        valid = false;
      } else {
        bool ok;
        key = mpm_key_of_position(M, rec.p0.x, rec.p0.y, rec.p0.z, &ok, &cell);
        if (!ok) b2_set_error(M.T->pool, B2E_ASSERT, (int)rec.id);
      }
    }
    mpm_insert(M, gen, key, cell, valid, rec.p0, rec.p1, rec.p2, rec.p3, rec.id);
  }
}
__global__ void k_mpm_zero_headers(u32 *a, u32 *b, u32 *c, u32 *d) {
  u32 *p = threadIdx.x == 0 ? a : threadIdx.x == 1 ? b : threadIdx.x == 2 ? c : d;
  if (p) *p = 0u;
}
// compact export of the live particles with their ids (original order is the caller's business)
__global__ void __launch_bounds__(256) k_mpm_export(MpmDev M, int gen, float *x, float *v, float *C, float *J, int *ids, long long cap, unsigned long long *counter) {
  const u32 ns = *M.nslots;
  const int lane = threadIdx.x & 31;
  for (u32 slot = blockIdx.x; slot < ns; slot += gridDim.x) {
    const u32 cnt = M.count[gen][slot];
    const u32 cround = (cnt + blockDim.x - 1) / blockDim.x * blockDim.x;
    for (u32 q = threadIdx.x; q < cround; q += blockDim.x) {
      bool live = false;
      const u8 *ch = nullptr;
      u32 i = 0;
      float4 p0 = make_float4(0, 0, 0, 0);
      if (q < cnt) {
        ch = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)slot * M.maxc + q / MPM_CHUNK]);
        i = q % MPM_CHUNK;
        p0 = ((const float4 *)ch)[i];
        live = p0.x != MPM_TOMB;
      }
      const unsigned m = __ballot_sync(0xffffffffu, live);
      if (!m) continue;
      unsigned long long base = 0;
      const int leader = __ffs(m) - 1;
      if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (live) {
        const long long o = (long long)(base + __popc(m & ((1u << lane) - 1u)));
        if (o < cap) {
          const float4 p1 = ((const float4 *)(ch + MPM_PLANE_BYTES))[i];
          const float4 p2 = ((const float4 *)(ch + 2 * MPM_PLANE_BYTES))[i];
          const float4 p3 = ((const float4 *)(ch + 3 * MPM_PLANE_BYTES))[i];
          if (x) x[o * 3] = p0.x, x[o * 3 + 1] = p0.y, x[o * 3 + 2] = p0.z;
          if (J) J[o] = p0.w;
          if (v) v[o * 3] = p1.x, v[o * 3 + 1] = p1.y, v[o * 3 + 2] = p1.z;
          if (C) {
            float *c = C + o * 9;
            c[0] = p1.w, c[1] = p2.x, c[2] = p2.y, c[3] = p2.z, c[4] = p2.w, c[5] = p3.x, c[6] = p3.y, c[7] = p3.z, c[8] = p3.w;
          }
          if (ids) ids[o] = (int)((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[i];
        }
      }
    }
  }
}
