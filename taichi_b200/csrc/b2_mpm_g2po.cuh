// b2_mpm_g2po.cuh -- G2P + re-binning right after an ordered P2G (r2), included by b2_mpm.cu.
//
// Same processing as k_mpm_g2p_bls<PERM = true> (b2_mpm_bls.cuh: cell-major order from P2G's order plane, stayers compacted
// into one reserved range of the next generation's bin, leavers through mpm_insert, tombstoned tail), with the start of a bin
// taken off the critical path.  The r1 form began every bin with eleven DEPENDENT global round trips (work counter -> entries
// -> chunk ids -> two tree levels per neighbour block -> pad -> live count -> order plane -> reservation -> chunk allocation
// -> first particle); the ncu source view charges a fifth of the kernel's warp samples to the instructions that wait for them
// (profiles/r2_ncu_g2p_f2.txt).  Here:
//   * the slot of the NEXT bin is looked up while the current one is processed;
//   * entries, key, first live count, chunk ids and the eight neighbour-block pointers (left in M.nbptr by P2G: no tree walk)
//     are independent loads of one round trip;
//   * pad and order plane travel global -> shared with cp.async while thread 0 reserves the range in the next generation;
//   * the first particles are loaded before the chunks of the reservation are allocated, and the barrier that publishes
//     those chunks to the CTA sits after the first particle's arithmetic, right before its store.
// Three or four exposed round trips per bin instead of eleven.
#pragma once
#ifndef G2PM_MINB
#define G2PM_MINB 12 /* CTAs per SM of the materials-model variant (80 registers) */
#endif

__device__ __forceinline__ void g2o_cp16(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((u32)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}

// MODEL: the particle's F / Jp / material (side arena, model 1) travel with it.
template <int NT, int MINB, bool MODEL>
__global__ void __launch_bounds__(NT, MINB) k_mpm_g2p_ord(MpmDev M, int gen) {
  __shared__ float4 sPad[216];
  __shared__ __align__(16) unsigned short sOrder[MPM_PASS];
  __shared__ const u8 *sIn[32];  // the bin's chunks (maxc <= 32)
  __shared__ u8 *sOut[8];
  __shared__ u8 *sOutX[MODEL ? 8 : 1];
  __shared__ const u8 *sInX[MODEL ? 32 : 1];
  __shared__ float4 *sNb[8];
  __shared__ u32 sBase, sStay, sSlot;
  const int tid = threadIdx.x, lane = tid & 31;
  const unsigned FULL = 0xffffffffu;
  const u32 nslots = *M.nslots;
  const int g1 = gen ^ 1;
  u32 next = 0;  // (thread 0) the slot after the current one
  // Work order: the bins P2G processed, largest size class first, dealt round-robin to the CTAs (item w = blockIdx.x + k *
  // gridDim.x of the concatenated classes): every CTA gets the same mix and they finish together.
  __shared__ u32 sCls[17];  // prefix of the class lengths
  if (tid == 0) {
    u32 acc = 0;
    for (int c = 0; c < 16; c++) sCls[c] = acc, acc += M.work_counter[8 + c];
    sCls[16] = acc;
  }
  __syncthreads();
  auto item = [&](u32 w) -> u32 {  // (thread 0) slot of item w, nslots past the end
    if (w >= sCls[16]) return nslots;
    int c = 0;
    while (w >= sCls[c + 1]) c++;
    return M.glist[(u64)c * M.max_slots + (w - sCls[c])];
  };
  u32 kitem = blockIdx.x;
  if (tid == 0) sSlot = item(kitem);

  for (;;) {
    __syncthreads();  // everybody has left the previous bin, and thread 0 has left the next slot in sSlot (every thread reads
                      // it before the bin's second barrier, thread 0 writes it again after the bin's last one)
    const u32 slot = sSlot;
    if (slot >= nslots) break;
    // ---- header: independent loads ------------------------------------------------------------------------------------
    const u32 total = M.count[gen][slot];
    const int key = M.slot_key[slot];
    const u32 live0 = M.live[gen][(u64)slot * MPM_MAXPASS];
    if (tid < 32) {
      const u32 cid = max(M.chunk_of[gen][(u64)slot * M.maxc + min((u32)tid, (u32)M.maxc - 1u)], 1u);
      sIn[tid] = mpm_chunk_ptr(M, gen, cid);
      if (MODEL) sInX[tid] = mpm_xchunk_ptr(M, gen, cid);
    }
    if (tid >= NT - 8) sNb[tid - (NT - 8)] = (float4 *)M.nbptr[(u64)slot * 8 + (u32)(tid - (NT - 8))];
    if (total == 0) {  // (uniform; rare: slots are compacted)
      __syncthreads();  // (everybody has read sSlot)
      kitem += gridDim.x;
      if (tid == 0) sSlot = item(kitem);
      continue;
    }
    __syncthreads();
    for (u32 q0 = 0; q0 < total; q0 += MPM_PASS) {
      const int n = (int)min((u32)MPM_PASS, total - q0);
      if (q0 > 0) __syncthreads();  // the previous pass is done with sOrder / sOut / sBase
      // ---- order plane (+ pad) -> shared memory, asynchronously; the reservation meanwhile -------------------------------
      for (int w = tid; w * 8 < n; w += NT) {
        const u32 qs = q0 + 8u * (u32)w;
        g2o_cp16(&sOrder[w * 8], sIn[qs / MPM_CHUNK] + MPM_PERM_OFF + (qs % MPM_CHUNK) * 2);
      }
      if (q0 == 0) {
        for (int nidx = tid; nidx < 216; nidx += NT) {
          const int px = nidx / 36, py = (nidx / 6) % 6, pz = nidx % 6;
          const float4 *src = sNb[((px >> 2) << 2) | ((py >> 2) << 1) | (pz >> 2)];
          if (src)
            g2o_cp16(&sPad[nidx], src + (((px & 3) << 4) | ((py & 3) << 2) | (pz & 3)));
          else
            sPad[nidx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      const int nlive = (int)(q0 == 0 ? live0 : M.live[gen][(u64)slot * MPM_MAXPASS + q0 / MPM_PASS]);
      if (tid == 0) {
        sStay = 0;
        if (nlive > 0) sBase = atomicAdd(&M.count[g1][slot], (u32)nlive);
      }
      asm volatile("cp.async.wait_all;" ::: "memory");
      __syncthreads();  // order, pad, base
      // the slot after this one (a dependent load off the size-ordered list), consumed one bin later
      if (q0 == 0) {
        kitem += gridDim.x;
        if (tid == 0) next = item(kitem);
      }
      if (nlive == 0) continue;  // (uniform)
      const u32 base = sBase;
      const u32 c0 = base / MPM_CHUNK;
      u32 c1 = (base + (u32)nlive - 1) / MPM_CHUNK;
      if (c1 >= (u32)M.maxc) {
        if (tid == 0) b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, key);
        c1 = (u32)M.maxc - 1;
      }
      // ---- the first particles are on their way while the chunks of the reservation are allocated ------------------------
      float4 np0 = make_float4(0.f, 0.f, 0.f, 1.f);
      u32 nid = 0;
      MpmX nX;
      auto load = [&](int t) {
        if (t < nlive) {
          const u32 q = q0 + sOrder[t];
          const u8 *ch = sIn[q / MPM_CHUNK];
          np0 = ((const float4 *)ch)[q % MPM_CHUNK];
          nid = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[q % MPM_CHUNK];
          if (MODEL) nX = mpm_x_load(sInX[q / MPM_CHUNK], q % MPM_CHUNK);
        }
      };
      load(tid);
      for (u32 c = c0 + tid; c <= c1; c += NT) {  // the holder of position c * 256 allocates chunk c
        u32 *cs = &M.chunk_of[g1][(u64)slot * M.maxc + c];
        u32 id1;
        if (c * MPM_CHUNK >= base) {
          u32 cid = atomicAdd(M.arena_next[g1], 1u);
          if (cid >= M.arena_cap) {
            b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, -1);
            cid = M.arena_cap;  // the sink chunk
          }
          id1 = cid + 1;
          __threadfence();
          *(volatile u32 *)cs = id1;
        } else {
          unsigned ns = 8, polls = 0;
          while ((id1 = ld_volatile_u32(cs)) == 0) {
            __nanosleep(ns);
            if (ns < 256) ns *= 2;
            if (++polls > MPM_SPIN_LIMIT) {  // (the holder is gone: flag it, the data goes to the sink chunk)
              b2_set_error(M.T->pool, B2E_ASSERT, -1003);
              id1 = M.arena_cap + 1;
              break;
            }
          }
        }
        sOut[c - c0] = mpm_chunk_ptr(M, g1, id1);
        if (MODEL) sOutX[c - c0] = mpm_xchunk_ptr(M, g1, id1);
      }
      for (int t0 = 0; t0 < nlive; t0 += NT) {
        const int t = t0 + tid;
        const bool valid = t < nlive;
        float4 o0 = make_float4(0.f, 0.f, 0.f, 1.f), o1 = make_float4(0.f, 0.f, 0.f, 0.f), o2 = o1, o3 = o1;
        int nkey = -1, ncell = 0;
        const float4 p0 = np0;
        const u32 id = nid;
        MpmX X;
        if (MODEL) X = nX;
        load(t + NT);  // software pipeline: the next particle is loaded while this one is computed
        if (valid) mpm_g2p_particle(M, sPad, p0, id, o0, o1, o2, o3, nkey, ncell);
        if (t0 == 0) __syncthreads();  // sOut (the arithmetic above did not need it)
        const bool stay = valid && nkey == key;
        const unsigned smask = __ballot_sync(FULL, stay);
        if (smask) {
          u32 off = 0;
          const int leader = __ffs(smask) - 1;
          if (lane == leader) off = atomicAdd(&sStay, (u32)__popc(smask));
          off = __shfl_sync(FULL, off, leader);
          if (stay) {
            const u32 pos = base + off + __popc(smask & ((1u << lane) - 1u));
            const u32 c = pos / MPM_CHUNK, i = pos % MPM_CHUNK;
            if (c <= c1) {
              u8 *ch = sOut[c - c0];
              ((float4 *)ch)[i] = o0;
              ((float4 *)(ch + MPM_PLANE_BYTES))[i] = o1;
              ((float4 *)(ch + 2 * MPM_PLANE_BYTES))[i] = o2;
              ((float4 *)(ch + 3 * MPM_PLANE_BYTES))[i] = o3;
              ((u32 *)(ch + 4 * MPM_PLANE_BYTES))[i] = id;
              (ch + MPM_CELL_OFF)[i] = (u8)ncell;
              if (MODEL) mpm_x_store(sOutX[c - c0], i, X);
            }
          }
        }
        if (__any_sync(FULL, valid && !stay)) mpm_insert(M, g1, nkey, ncell, valid && !stay, o0, o1, o2, o3, id, MODEL ? &X : nullptr);
      }
      __syncthreads();
      // unused tail of the reservation (one entry per particle that left the block)
      for (u32 pos = base + sStay + tid; pos < base + (u32)nlive; pos += NT) {
        const u32 c = pos / MPM_CHUNK;
        if (c <= c1) {
          ((float4 *)sOut[c - c0])[pos % MPM_CHUNK] = make_float4(MPM_TOMB, 0.f, 0.f, 1.f);
          (sOut[c - c0] + MPM_CELL_OFF)[pos % MPM_CHUNK] = 0xffu;
        }
      }
    }
    // this generation's bin is consumed: leave it empty for the generation after next
    const u32 nchunks = (total + MPM_CHUNK - 1) / MPM_CHUNK;
    for (int t = tid; t < (int)nchunks; t += NT) M.chunk_of[gen][(u64)slot * M.maxc + t] = 0;
    if (tid == 0) {
      M.count[gen][slot] = 0;
      sSlot = next;
    }
  }
}
