// b2_mpm_bls.cuh -- the block-local (ti.block_local) G2P kernel, included by b2_mpm.cu (P2G: b2_mpm_p2g.cuh).
//
// The kernel processes one occupied 4^3 block per CTA (persistent CTAs fetch block slots from an
// atomic work counter) and start with the same step: a counting sort of the block's particles by
// cell in shared memory (64 integer bins; ranks come from the native integer ATOMS.ADD), which
// gives   sStart[cell], sCnt[cell]   and   sPerm[start + rank] = particle index.
// Tombstones (p0.x == MPM_TOMB, left behind by particles that moved to another block) sort into
// bin 64 and are skipped.
#pragma once
#ifndef G2PK_T
#define G2PK_T 64    /* threads per CTA of the ordered (PERM) G2P */
#endif
#ifndef G2PK_MINB
#define G2PK_MINB 20 /* CTAs per SM */
#endif
#ifndef G2P_F2
#define G2P_F2 1 /* packed fp32x2 gather of the x / y components */
#endif

template <int SB>
struct MpmSortSmemT {
  unsigned short perm[SB]; // cell-sorted -> index inside the pass
  int cnt[65];
  int start[66];
  int maxcnt;
  u32 touched;
  const u8 *chunk[32];         // chunk base pointers of the bin (maxc <= 32)
};

__device__ __forceinline__ int mpm_cell_of(const MpmDev &M, const float4 p0) {
  const float xp[3] = {p0.x, p0.y, p0.z};
  int l[3];
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float X = M.pot ? xp[d] * M.inv_dx : xp[d] / M.dx;
    l[d] = ((int)(X - 0.5f)) & 3;
  }
  return (l[0] << 4) | (l[1] << 2) | l[2];
}

// Sorts particles [q0, q0+n) of the bin. All threads of the CTA must call it. NT = blockDim.x.
// On return (after its final barrier): S.cnt / S.start / S.perm / S.maxcnt / S.touched are valid;
// S.start[64] = number of live particles in the pass.
template <int NT, int U, int SB>
__device__ __forceinline__ void mpm_sort_pass(const MpmDev &M, MpmSortSmemT<SB> &S, u32 *key, u32 q0, int n, int cbase = 0, int ncell = 64) {  // key[SB]: cell << 16 | rank
  const int tid = threadIdx.x;
  if (tid < 65) S.cnt[tid] = 0;
  __syncthreads();
  u32 tmask = 0;
  for (int tb = tid; tb < n; tb += U * NT) {  // U independent 16-B loads in flight per thread
    float4 pv[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int t = tb + u * NT;
      if (t < n) {
        const u32 q = q0 + t;
        pv[u] = ((const float4 *)S.chunk[q / MPM_CHUNK])[q % MPM_CHUNK];
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int t = tb + u * NT;
      if (t >= n) break;
      const float4 p0 = pv[u];
      int c = 64;
      if (p0.x != MPM_TOMB) {
        c = mpm_cell_of(M, p0);
        const int h = (((c >> 4) & 3) >> 1) << 2 | (((c >> 2) & 3) >> 1) << 1 | ((c & 3) >> 1);
        // neighbour blocks {0,1}^3 whose nodes this particle's stencil reaches
        c -= cbase;  // cells outside this CTA's window sort into the dead bin
        if ((unsigned)c >= (unsigned)ncell)
          c = 64;
        else
          tmask |= (h == 0) ? 0x01u : (h == 1) ? 0x03u : (h == 2) ? 0x05u : (h == 3) ? 0x0fu : (h == 4) ? 0x11u : (h == 5) ? 0x33u : (h == 6) ? 0x55u : 0xffu;
      }
      const int r = atomicAdd(&S.cnt[c], 1);
      key[t] = ((u32)c << 16) | (u32)r;
    }
  }
  tmask = __reduce_or_sync(0xffffffffu, tmask);
  if ((tid & 31) == 0 && tmask) atomicOr(&S.touched, tmask);
  __syncthreads();
  if (tid < 32) {  // exclusive scan of 65 counts: lanes own entries 2l, 2l+1; entry 64 appended
    const int a = S.cnt[2 * tid], b = S.cnt[2 * tid + 1];
    int inc = a + b;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int tv = __shfl_up_sync(0xffffffffu, inc, d);
      if (tid >= d) inc += tv;
    }
    S.start[2 * tid] = inc - a - b;
    S.start[2 * tid + 1] = inc - b;
    int mx = a > b ? a : b;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    if (tid == 31) {
      S.start[64] = inc;
      S.start[65] = inc + S.cnt[64];
      S.maxcnt = mx;
    }
  }
  __syncthreads();
  for (int t = tid; t < n; t += NT) {
    const u32 k = key[t];
    const int c = (int)(k >> 16);
    if (c < 64) S.perm[S.start[c] + (int)(k & 0xffffu)] = (unsigned short)t;
  }
  __syncthreads();
}

// One particle of G2P (mpm3d.py:62-79): gather v and the affine matrix C from the 27 nodes of the block's (4+2)^3 pad, advect,
// and find the particle's new block / cell.  id only labels the error of a particle that leaves the grid.
__device__ __forceinline__ void mpm_g2p_particle(const MpmDev &M, const float4 *pad, const float4 p0, u32 id, float4 &o0, float4 &o1,
                                                 float4 &o2, float4 &o3, int &nkey, int &ncell) {
  MpmStencil s;
  mpm_stencil(M, p0.x, p0.y, p0.z, s);
  float v0, v1, v2, Sx0, Sx1, Sx2, Sy0, Sy1, Sy2, Sz0, Sz1, Sz2;
  const int l0 = s.base[0] & 3, l1 = s.base[1] & 3, l2 = s.base[2] & 3;
  const float wz0 = s.w[0][2], wz1 = s.w[1][2], wz2 = s.w[2][2];
#if G2P_F2
  {  // x and y components travel as one packed fp32x2 (FFMA2 / FMUL2 / FADD2: the same roundings as the scalar form,
     // half the issue slots); the pad's float4 {vx, vy, vz, m} hands (vx, vy) over as an aligned register pair
    u64 vxy = 0ull, Sxxy = 0ull, Syxy = 0ull, Szxy = 0ull;
    v2 = 0.f, Sx2 = 0.f, Sy2 = 0.f, Sz2 = 0.f;
    const u64 wz0p = f2_pack(wz0, wz0), wz1p = f2_pack(wz1, wz1), wz2p = f2_pack(wz2, wz2), twop = f2_pack(2.f, 2.f);
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const float wxy = s.w[i][0] * s.w[j][1];
        const u64 wp = f2_pack(wxy, wxy);
        const int pidx = ((l0 + i) * 6 + (l1 + j)) * 6 + l2;
        const float4 g0 = pad[pidx], g1v = pad[pidx + 1], g2 = pad[pidx + 2];
        const u64 t1 = f2_mul(wz1p, f2_pack(g1v.x, g1v.y)), t2 = f2_mul(wz2p, f2_pack(g2.x, g2.y));
        const u64 r = f2_fma(wz0p, f2_pack(g0.x, g0.y), f2_add(t1, t2)), z = f2_fma(twop, t2, t1);
        const float t1z = wz1 * g1v.z, t2z = wz2 * g2.z;
        const float rzv = fmaf(wz0, g0.z, t1z + t2z), zz = fmaf(2.f, t2z, t1z);
        vxy = f2_fma(wp, r, vxy), v2 = fmaf(wxy, rzv, v2);
        Szxy = f2_fma(wp, z, Szxy), Sz2 = fmaf(wxy, zz, Sz2);
        if (j > 0) {
          const float wj = (float)j * wxy;
          Syxy = f2_fma(f2_pack(wj, wj), r, Syxy), Sy2 = fmaf(wj, rzv, Sy2);
        }
        if (i > 0) {
          const float wi = (float)i * wxy;
          Sxxy = f2_fma(f2_pack(wi, wi), r, Sxxy), Sx2 = fmaf(wi, rzv, Sx2);
        }
      }
    }
    f2_unpack(vxy, v0, v1), f2_unpack(Sxxy, Sx0, Sx1), f2_unpack(Syxy, Sy0, Sy1), f2_unpack(Szxy, Sz0, Sz1);
  }
#else
  v0 = 0.f, v1 = 0.f, v2 = 0.f;
  Sx0 = 0.f, Sx1 = 0.f, Sx2 = 0.f, Sy0 = 0.f, Sy1 = 0.f, Sy2 = 0.f, Sz0 = 0.f, Sz1 = 0.f, Sz2 = 0.f;
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const float wxy = s.w[i][0] * s.w[j][1];
      const int pidx = ((l0 + i) * 6 + (l1 + j)) * 6 + l2;
      const float4 g0 = pad[pidx], g1v = pad[pidx + 1], g2 = pad[pidx + 2];
      // row = sum_k wz_k g_k ; rz = sum_k k wz_k g_k
      const float t1x = wz1 * g1v.x, t1y = wz1 * g1v.y, t1z = wz1 * g1v.z;
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
#endif
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

// ------------------------------------------------------------------------------------------
// G2P + re-binning.  Particles are processed in cell-sorted order (same-cell lanes read the same
// pad addresses).  A particle that stays in its block is written to the next generation's bin of
// the SAME slot at a position inside a range the CTA reserved with ONE atomicAdd per pass; only
// particles that change block go through the generic warp-aggregated append (mpm_insert).  The
// unused tail of the reserved range (= number of leavers) is filled with tombstones.
// ------------------------------------------------------------------------------------------
// PERM = true: P2G has just run on this generation and left the cell-major order of every pass of
// MPM_PASS entries in the chunks' order plane (+ the live count per pass): no sort here.
template <int G2P_THREADS, int G2P_MINB, bool PERM>
__global__ void __launch_bounds__(G2P_THREADS, G2P_MINB) k_mpm_g2p_bls(MpmDev M, int gen) {
  constexpr int SB = PERM ? MPM_PASS : MPM_SB;  // entries per pass
  __shared__ float4 sPad[216];
  __shared__ MpmSortSmemT<PERM ? 2 : MPM_SB> S;
  __shared__ u32 sKey[PERM ? 2 : MPM_SB];
  __shared__ __align__(4) unsigned short sOrder[PERM ? MPM_PASS : 2];  // the pass's order plane (one coalesced read)
  __shared__ u8 *sOut[65];
  __shared__ u32 sBase, sStay;
  __shared__ int sSlot;
  __shared__ float4 *sNb[8];
  const int tid = threadIdx.x, lane = tid & 31;
  const u32 nslots = *M.nslots;
  const int g1 = gen ^ 1;

  for (;;) {
    __syncthreads();
    if (tid == 0) sSlot = (int)atomicAdd(&M.work_counter[1], 1u);
    __syncthreads();
    const u32 slot = (u32)sSlot;
    if (slot >= nslots) break;
    const u32 total = M.count[gen][slot];
    if (total == 0) continue;
    const int key = M.slot_key[slot];
    int bx, by, bz;
    mpm_unkey(M, key, bx, by, bz);
    const u32 nchunks = (total + MPM_CHUNK - 1) / MPM_CHUNK;
    for (int t = tid; t < (int)nchunks; t += G2P_THREADS) S.chunk[t] = mpm_chunk_ptr(M, gen, M.chunk_of[gen][(u64)slot * M.maxc + t]);
    constexpr int NB0 = G2P_THREADS >= 64 ? 32 : 0;  // second warp if there is one
    if (tid >= NB0 && tid < NB0 + 8) {
      const int nbh = tid - NB0;
      sNb[nbh] = mpm_block_ptr<false>(M, bx + (nbh >> 2), by + ((nbh >> 1) & 1), bz + (nbh & 1));
    }
    __syncthreads();
    for (int nidx = tid; nidx < 216; nidx += G2P_THREADS) {
      const int px = nidx / 36, py = (nidx / 6) % 6, pz = nidx % 6;
      const float4 *src = sNb[((px >> 2) << 2) | ((py >> 2) << 1) | (pz >> 2)];
      sPad[nidx] = src ? src[((px & 3) << 4) | ((py & 3) << 2) | (pz & 3)] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (PERM) __syncthreads();  // pad complete (the sort pass's barriers do this otherwise)
    auto sorted_entry = [&](u32 q0, int j) -> u32 {  // j-th live entry of the pass in cell order
      if constexpr (PERM) {
        return q0 + sOrder[j];
      } else {
        return q0 + S.perm[j];
      }
    };
    for (u32 q0 = 0; q0 < total; q0 += SB) {
      int nlive;
      if constexpr (PERM) {
        nlive = (int)M.live[gen][(u64)slot * MPM_MAXPASS + q0 / MPM_PASS];
        __syncthreads();  // the previous pass is done with sOrder
        for (int w = tid; 2 * w < nlive; w += G2P_THREADS) {
          const u32 qs = q0 + 2u * (u32)w;
          ((u32 *)sOrder)[w] = ((const u32 *)(S.chunk[qs / MPM_CHUNK] + MPM_PERM_OFF))[(qs % MPM_CHUNK) / 2];
        }
      } else {
        const int n = (int)min((u32)SB, total - q0);
        mpm_sort_pass<G2P_THREADS, 4>(M, S, sKey, q0, n);
        nlive = S.start[64];
      }
      if (nlive == 0) continue;  // uniform
      if (tid == 0) {
        sBase = atomicAdd(&M.count[g1][slot], (u32)nlive);
        sStay = 0;
      }
      __syncthreads();
      const u32 base = sBase;
      const u32 c0 = base / MPM_CHUNK;
      u32 c1 = (base + (u32)nlive - 1) / MPM_CHUNK;
      if (c1 >= (u32)M.maxc) {
        if (tid == 0) b2_set_error(M.T->pool, B2E_BIN_OVERFLOW, key);
        c1 = (u32)M.maxc - 1;
      }
      // chunks of the reserved range: the holder of position c*256 allocates chunk c
      for (u32 c = c0 + tid; c <= c1; c += G2P_THREADS) {
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
          unsigned ns = 8;
          while ((id1 = ld_volatile_u32(cs)) == 0) {
            __nanosleep(ns);
            if (ns < 256) ns *= 2;
          }
        }
        sOut[c - c0] = mpm_chunk_ptr(M, g1, id1);
      }
      __syncthreads();
      float4 np0 = make_float4(0.f, 0.f, 0.f, 1.f);
      u32 nid = 0;
      if (tid < nlive) {  // software pipeline: particle t0+NT is loaded while particle t0 is computed
        const u32 q = sorted_entry(q0, tid);
        const u8 *ch = S.chunk[q / MPM_CHUNK];
        np0 = ((const float4 *)ch)[q % MPM_CHUNK];
        nid = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[q % MPM_CHUNK];
      }
      for (int t0 = 0; t0 < nlive; t0 += G2P_THREADS) {
        const int t = t0 + tid;
        const bool valid = t < nlive;
        float4 o0 = make_float4(0.f, 0.f, 0.f, 1.f), o1 = make_float4(0.f, 0.f, 0.f, 0.f), o2 = o1, o3 = o1;
        int nkey = -1, ncell = 0;
        const float4 p0 = np0;
        u32 id = nid;
        if (t + G2P_THREADS < nlive) {
          const u32 q = sorted_entry(q0, t + G2P_THREADS);
          const u8 *ch = S.chunk[q / MPM_CHUNK];
          np0 = ((const float4 *)ch)[q % MPM_CHUNK];
          nid = ((const u32 *)(ch + 4 * MPM_PLANE_BYTES))[q % MPM_CHUNK];
        }
        if (valid) mpm_g2p_particle(M, sPad, p0, id, o0, o1, o2, o3, nkey, ncell);
        const bool stay = valid && nkey == key;
        const unsigned smask = __ballot_sync(0xffffffffu, stay);
        if (smask) {
          u32 off = 0;
          const int leader = __ffs(smask) - 1;
          if (lane == leader) off = atomicAdd(&sStay, (u32)__popc(smask));
          off = __shfl_sync(0xffffffffu, off, leader);
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
            }
          }
        }
        if (__any_sync(0xffffffffu, valid && !stay)) mpm_insert(M, g1, nkey, ncell, valid && !stay, o0, o1, o2, o3, id);
      }
      __syncthreads();
      // unused tail of the reservation (one entry per particle that left the block)
      for (u32 pos = base + sStay + tid; pos < base + (u32)nlive; pos += G2P_THREADS) {
        const u32 c = pos / MPM_CHUNK;
        if (c <= c1) {
          ((float4 *)sOut[c - c0])[pos % MPM_CHUNK] = make_float4(MPM_TOMB, 0.f, 0.f, 1.f);
          (sOut[c - c0] + MPM_CELL_OFF)[pos % MPM_CHUNK] = 0xffu;
        }
      }
    }
    // this generation's bin is consumed: leave it empty for the generation after next
    for (int t = tid; t < (int)nchunks; t += G2P_THREADS) M.chunk_of[gen][(u64)slot * M.maxc + t] = 0;
    if (tid == 0) M.count[gen][slot] = 0;
  }
}
