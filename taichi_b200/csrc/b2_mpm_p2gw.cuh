// b2_mpm_p2gw.cuh -- the block-local (ti.block_local) P2G kernel, warp-autonomous form (r2), included by b2_mpm.cu.
//
// Replaces make_block_local's pad prologue/epilogue (transforms/make_block_local.cpp:97-365) and the
// 27 x 4 float atomics per particle of the reference's P2G (codegen_llvm.cpp:1495-1569).
//
// r1's kernel spent a third of its warp samples at CTA barriers between its five phases and moved every
// particle through shared memory twice.  Here ONE WARP owns one occupied 4^3 block's bin from the first
// load to the last red: there is no CTA-wide barrier anywhere in the kernel.
//   count    the warp reads the u8 cell plane of a pass (<= MPM_PASS = 1024 bin entries = 4 whole chunks; the cell of
//            an entry is written with its position by whoever bins it) and takes every entry's rank inside its cell
//            from the native integer ATOMS.ADD;
//   order    a warp scan of the 64 counts gives the cell-major order of the entries (kept in shared
//            memory, and written to the chunks' u16 order plane for the G2P that follows);
//   rank     the cells (cut into units of <= P2W_UNIT particles) are ranked by length with a counting sort, so that
//            the 32 lanes of a round loop over nearly equal trip counts: round r hands unit rank 32 r + lane to lane;
//   gather   lane = cell: the lane walks its cell's particles through the order.  The four 16-byte planes
//            of a particle travel global -> shared with cp.async (SASS LDGSTS) into a per-lane ring
//            P2W_DEPTH particles deep, so no register holds data in flight (an L2 bulk prefetch of the next
//            bin, as r1 had, doubles the DRAM reads here and was measured slower: 0.381 vs 0.367 ms).  All 27 stencil
//            nodes x {mv, m} are accumulated in 54 packed-fp32x2 REGISTERS: no atomics, no shuffles;
//   pad      the round's 32 cells are distinct, so for a fixed stencil offset the 32 lanes add into 32
//            different nodes of the warp's private (4+2)^3 pad: 27 plain LDS/FADD2/STS steps;
//   flush    once per bin: 216 pad nodes -> one red.global.add.v4.f32 each into the <= 8 destination
//            blocks, which are activated only if a particle really touches them (x64 semantics;
//            activate_halo reproduces the CUDA-BLS epilogue).
#pragma once

#ifndef P2W_WARPS
#define P2W_WARPS 4 /* warps per CTA (the CTA is only a container: warps never synchronise with each other) */
#endif
#ifndef P2W_CTAS
#define P2W_CTAS 3  /* CTAs per SM: 12 warps at 168 registers (2 x 4 warps at 248: 0.368 ms on C3, 3 x 4: 0.335 ms) */
#endif
#ifndef P2W_COUNT_U
#define P2W_COUNT_U 4 /* position loads in flight per lane in the count phase */
#endif
#ifndef P2W_UNIT
#define P2W_UNIT 32 /* longest run of one cell's particles a lane accumulates before it goes to the pad (< 64) */
#endif
#ifndef P2W_DEPTH
#define P2W_DEPTH 2 /* particles in flight per lane (power of two); 4 is not faster and costs 4 KB of L1 per warp */
#endif
static_assert(MPM_PASS == 4 * MPM_CHUNK, "a pass is four whole chunks");
static_assert(P2W_DEPTH * 4 * 32 * 16 >= MPM_PASS * 2, "ck fits the ring");

struct P2WWarp {                      // private state of one warp
  union {
    float4 ring[P2W_DEPTH][4][32];    // gather: [slot][plane][lane] particles in flight
    unsigned short ck[MPM_PASS];      // count / order: entry -> cell << 10 | rank inside the cell (live entries)
  };
  float4 pad[216];                    // (4+2)^3 nodes {sum w mv_x, sum w mv_y, sum w mv_z, sum w}
  unsigned short order[MPM_PASS];     // cell-major position -> entry of the pass
  int cnt[64];                        // particles per cell
  int hist[64];                       // units per length
  unsigned short cstart[64];          // first cell-major position of the cell
  unsigned short nk[64];              // units longer than the index
  unsigned short unit[64 + MPM_PASS / P2W_UNIT];  // rank by length -> cell | segment << 6
  u8 *ch[4];                          // the pass's chunks
  float4 *nb[8];                      // destination blocks {0,1}^3
};

// bins of the particle's stencil neighbours {0,1}^3 reached from cell c (bit nbh = dx*4 + dy*2 + dz)
__device__ __forceinline__ u32 p2w_touch_mask(int c) {
  const int h = ((c >> 5) & 1) << 2 | ((c >> 3) & 1) << 1 | ((c >> 1) & 1);
  return (u32)((0xff553311'0f050301ull >> (h * 8)) & 0xffu);
}
__device__ __forceinline__ void p2w_cp16(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((u32)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void p2w_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void p2w_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <bool POT>
__device__ __forceinline__ void p2w_stencil(const MpmDev &M, const float4 p0, float fx[3], float w[3][3]) {
  const float xp[3] = {p0.x, p0.y, p0.z};
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float X = POT ? xp[d] * M.inv_dx : xp[d] / M.dx;  // Xp = x / dx (mpm3d.py:41)
    fx[d] = X - (float)(int)(X - 0.5f);
    w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
    w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
    w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
  }
}
template <bool POT, int CTAS>
__global__ void __launch_bounds__(P2W_WARPS * 32, CTAS) k_mpm_p2g(MpmDev M, int gen) {
  extern __shared__ __align__(16) unsigned char p2w_smem[];
  P2WWarp &S = ((P2WWarp *)p2w_smem)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  const u32 nslots = *M.nslots;
  const int cA = 2 * lane, cB = cA + 1;

  // header of a bin, one word per lane: lane 0 entries, lane 1 block key, lanes 2..5 the first pass's chunk ids.  The
  // header of the NEXT bin is requested before this bin is processed: the dependent chain work counter -> count /
  // key -> chunk ids never sits on the critical path (it was 9 % of the warp samples).
  auto header = [&](u32 sl) -> u32 {
    if (sl >= nslots || lane >= 6) return 0u;
    if (lane == 0) return M.count[gen][sl];
    if (lane == 1) return (u32)M.slot_key[sl];
    return M.chunk_of[gen][(u64)sl * M.maxc + (u32)(lane - 2)];
  };
  u32 next = 0;
  if (lane == 0) next = atomicAdd(&M.work_counter[0], 1u);
  next = __shfl_sync(FULL, next, 0);
  u32 hnext = header(next);
  for (;;) {
    const u32 slot = next, hdr = hnext;
    if (slot >= nslots) break;
    if (lane == 0) next = atomicAdd(&M.work_counter[0], 1u);  // (read after the count phase: its latency is covered)
    const u32 total = __shfl_sync(FULL, hdr, 0);
    if (total == 0) {
      next = __shfl_sync(FULL, next, 0);
      hnext = header(next);
      continue;
    }
    const int key = (int)__shfl_sync(FULL, hdr, 1);
    // G2P takes the bins largest first, dealt round-robin to its CTAs (they then finish together): the bin is appended to its
    // size class; the position is requested here and used at the end of the bin (the warp never waits for the atomic)
    const u32 gcls = 15u - min(15u, (total - 1u) >> 6);
    u32 gpos = 0;
    if (lane == 0) gpos = atomicAdd(&M.work_counter[8 + gcls], 1u);
    int bx, by, bz;
    mpm_unkey(M, key, bx, by, bz);
    for (int i = lane; i < 216; i += 32) S.pad[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    u32 touched = M.activate_halo ? 0xffu : 0u;

    for (u32 pass = 0; pass * MPM_PASS < total; pass++) {
      const u32 q0 = pass * MPM_PASS;
      const int n = (int)min((u32)MPM_PASS, total - q0);
      __syncwarp();
      {
        u32 cid = __shfl_sync(FULL, hdr, (lane & 3) + 2);
        if (pass > 0 && lane < 4) cid = M.chunk_of[gen][(u64)slot * M.maxc + min(pass * 4u + (u32)lane, (u32)M.maxc - 1u)];
        if (lane < 4) S.ch[lane] = mpm_chunk_ptr(M, gen, max(cid, 1u));
      }
      S.cnt[cA] = 0, S.cnt[cB] = 0, S.hist[cA] = 0, S.hist[cB] = 0;
      __syncwarp();
      auto entry = [&](int t) -> const u8 * { return S.ch[t >> 8] + (t & (MPM_CHUNK - 1)) * 16; };
      // ---- count: rank of every entry inside its cell.  The cells come from the chunks' u8 cell plane (written with the
      // positions): one 16-byte load per lane covers 16 entries, the whole pass is two loads deep.
      uint4 cbv[2];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int t = h * 512 + lane * 16;
        cbv[h] = make_uint4(~0u, ~0u, ~0u, ~0u);
        if (t < n) cbv[h] = *(const uint4 *)(S.ch[t >> 8] + MPM_CELL_OFF + (t & (MPM_CHUNK - 1)));
      }
      auto cell_at = [&](int h, int e) -> int {  // cell of entry h * 512 + lane * 16 + e; >= 64: tombstone or past the end
        const u32 wv = e < 4 ? cbv[h].x : e < 8 ? cbv[h].y : e < 12 ? cbv[h].z : cbv[h].w;
        const int c = (int)((wv >> ((e & 3) * 8)) & 0xffu);
        return h * 512 + lane * 16 + e < n ? c : 255;
      };
#pragma unroll
      for (int h = 0; h < 2; h++) {
        if (h * 512 < n) {
#pragma unroll
          for (int e = 0; e < 16; e++) {
            const int c = cell_at(h, e);
            if (c < 64) S.ck[h * 512 + lane * 16 + e] = (unsigned short)((c << 10) | atomicAdd(&S.cnt[c], 1));
          }
        }
      }
      __syncwarp();
      if (pass == 0) {  // the next bin's header travels while this bin is gathered
        next = __shfl_sync(FULL, next, 0);
        hnext = header(next);
      }
      // ---- scan: cell-major start of every cell; rank of the cells by count ------------------------------------
      const int nA = S.cnt[cA], nB = S.cnt[cB];
      touched |= __reduce_or_sync(FULL, (nA > 0 ? p2w_touch_mask(cA) : 0u) | (nB > 0 ? p2w_touch_mask(cB) : 0u));
      int cs = nA + nB;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(FULL, cs, d);
        if (lane >= d) cs += v;
      }
      S.cstart[cA] = (unsigned short)(cs - nA - nB), S.cstart[cB] = (unsigned short)(cs - nB);
      const int nlive = __shfl_sync(FULL, cs, 31);
      if (lane == 0) M.live[gen][(u64)slot * MPM_MAXPASS + pass] = (u32)nlive;
      // units: a cell's particles in runs of <= P2W_UNIT (long runs would cost fp32 accuracy and leave 31 lanes idle);
      // counting sort of the units by length: full units first, then the cells' last units by decreasing length
      const int uA = (nA + P2W_UNIT - 1) / P2W_UNIT, uB = (nB + P2W_UNIT - 1) / P2W_UNIT;
      const int rA = nA - (uA - 1) * P2W_UNIT, rB = nB - (uB - 1) * P2W_UNIT;  // length of the last unit (cells with particles)
      const int fA = uA > 1 ? atomicAdd(&S.hist[P2W_UNIT], uA - 1) : 0, fB = uB > 1 ? atomicAdd(&S.hist[P2W_UNIT], uB - 1) : 0;
      const int wA = nA > 0 ? atomicAdd(&S.hist[rA], 1) : 0, wB = nB > 0 ? atomicAdd(&S.hist[rB], 1) : 0;
      const int nunits = __reduce_add_sync(FULL, uA + uB);
      __syncwarp();
      {
        const int hA = S.hist[cA], hB = S.hist[cB];  // units of length 2 lane / 2 lane + 1
        int sfx = hA + hB;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int v = __shfl_down_sync(FULL, sfx, d);
          if (lane + d < 32) sfx += v;
        }
        S.nk[cB] = (unsigned short)(sfx - hA - hB), S.nk[cA] = (unsigned short)(sfx - hA);
      }
      __syncwarp();
      if (nA > 0) S.unit[S.nk[rA] + wA] = (unsigned short)(cA | (uA - 1) << 6);
      if (nB > 0) S.unit[S.nk[rB] + wB] = (unsigned short)(cB | (uB - 1) << 6);
      for (int sg = 0; sg < uA - 1; sg++) S.unit[fA + sg] = (unsigned short)(cA | sg << 6);  // (no unit is longer than a full one)
      for (int sg = 0; sg < uB - 1; sg++) S.unit[fB + sg] = (unsigned short)(cB | sg << 6);
      // ---- order: cell-major position -> entry; the order plane of the chunks is what G2P will read ---------------
#pragma unroll
      for (int h = 0; h < 2; h++) {
        if (h * 512 < n) {
#pragma unroll
          for (int e = 0; e < 16; e++) {
            const int t = h * 512 + lane * 16 + e;
            if (cell_at(h, e) < 64) {
              const int ck = S.ck[t];
              S.order[S.cstart[ck >> 10] + (ck & 1023)] = (unsigned short)t;
            }
          }
        }
      }
      __syncwarp();
      for (int w = lane; 2 * w < nlive; w += 32) ((u32 *)(S.ch[w >> 7] + MPM_PERM_OFF))[w & 127] = ((const u32 *)S.order)[w];
      if (nlive == 0) continue;
      // ---- rounds: lane = cell of rank r0 + lane -------------------------------------------------------------------
      for (int r0 = 0; r0 < nunits; r0 += 32) {
        const bool has = r0 + lane < nunits;
        const int ut = has ? (int)S.unit[r0 + lane] : 0;
        const int cell = ut & 63, sg = ut >> 6;
        const int np = has ? min(P2W_UNIT, S.cnt[cell] - sg * P2W_UNIT) : 0;
        const int st = S.cstart[cell] + sg * P2W_UNIT;
        const int steps = __reduce_max_sync(FULL, np);
        // units of one cell that share the round take turns at the pad (only cells with more than P2W_UNIT particles have several)
        const unsigned peers = __match_any_sync(FULL, has ? cell : 64 + lane);
        const int turn = __popc(peers & ((1u << lane) - 1u));
        const int turns = __reduce_max_sync(FULL, turn) + 1;
        u64 a01[27], a2m[27];  // per node {sum w mv_x, sum w mv_y}, {sum w mv_z, sum w}
#pragma unroll
        for (int q = 0; q < 27; q++) a01[q] = 0ull, a2m[q] = 0ull;
        auto issue = [&](int kk) {  // particle kk of this lane's cell -> ring slot kk % DEPTH (one commit group, possibly empty)
          if (kk < np) {
            const u8 *ent = entry(S.order[st + kk]);
            float4 *dst = &S.ring[kk & (P2W_DEPTH - 1)][0][lane];
            p2w_cp16(dst, ent), p2w_cp16(dst + 32, ent + MPM_PLANE_BYTES);
            p2w_cp16(dst + 64, ent + 2 * MPM_PLANE_BYTES), p2w_cp16(dst + 96, ent + 3 * MPM_PLANE_BYTES);
          }
          p2w_commit();
        };
#pragma unroll
        for (int d = 0; d < P2W_DEPTH; d++) issue(d);
#pragma unroll 1
        for (int k = 0; k < steps; k++) {
          p2w_wait<P2W_DEPTH - 1>();
          const float4 *src = &S.ring[k & (P2W_DEPTH - 1)][0][lane];
          const float4 p0 = src[0], p1 = src[32], p2 = src[64], p3 = src[96];
          issue(k + P2W_DEPTH);
          if (k < np) {
            float fx[3], w[3][3];
            p2w_stencil<POT>(M, p0, fx, w);
            const float stress = M.stress_k * (p0.w - 1.0f);
            // A' = (stress I + m C) dx ; b = m v - A' fx ; node value = b + A' o  (mpm3d.py:46-55, re-associated)
            const float m = M.p_mass;
            const float A00 = (stress + m * p1.w) * M.dx, A01 = m * p2.x * M.dx, A02 = m * p2.y * M.dx;
            const float A10 = m * p2.z * M.dx, A11 = (stress + m * p2.w) * M.dx, A12 = m * p3.x * M.dx;
            const float A20 = m * p3.y * M.dx, A21 = m * p3.z * M.dx, A22 = (stress + m * p3.w) * M.dx;
            const float b0 = m * p1.x - (A00 * fx[0] + A01 * fx[1] + A02 * fx[2]);
            const float b1 = m * p1.y - (A10 * fx[0] + A11 * fx[1] + A12 * fx[2]);
            const float b2 = m * p1.z - (A20 * fx[0] + A21 * fx[1] + A22 * fx[2]);
            u64 r01 = f2_pack(b0, b1), r2o = f2_pack(b2, 1.0f);
            const u64 i01 = f2_pack(A00, A10), i2 = f2_pack(A20, 0.0f);
            const u64 j01 = f2_pack(A01, A11), j2 = f2_pack(A21, 0.0f);
            const u64 k01 = f2_pack(A02, A12), k2 = f2_pack(A22, 0.0f);
            const u64 wy01 = f2_pack(w[0][1], w[1][1]), wz01 = f2_pack(w[0][2], w[1][2]);
#pragma unroll
            for (int i = 0; i < 3; i++) {
              u64 q01 = r01, q2o = r2o;
              float wxy[3];  // the weight products of two nodes come from one packed multiply
              f2_unpack(f2_mul(f2_pack(w[i][0], w[i][0]), wy01), wxy[0], wxy[1]);
              wxy[2] = w[i][0] * w[2][1];
#pragma unroll
              for (int j = 0; j < 3; j++) {
                u64 c01 = q01, c2o = q2o;
                float wn[3];
                f2_unpack(f2_mul(f2_pack(wxy[j], wxy[j]), wz01), wn[0], wn[1]);
                wn[2] = wxy[j] * w[2][2];
#pragma unroll
                for (int kk = 0; kk < 3; kk++) {
                  const u64 w2 = f2_pack(wn[kk], wn[kk]);
                  a01[(i * 3 + j) * 3 + kk] = f2_fma(w2, c01, a01[(i * 3 + j) * 3 + kk]);
                  a2m[(i * 3 + j) * 3 + kk] = f2_fma(w2, c2o, a2m[(i * 3 + j) * 3 + kk]);
                  if (kk < 2) c01 = f2_add(c01, k01), c2o = f2_add(c2o, k2);
                }
                if (j < 2) q01 = f2_add(q01, j01), q2o = f2_add(q2o, j2);
              }
              if (i < 2) r01 = f2_add(r01, i01), r2o = f2_add(r2o, i2);
            }
          }
        }
        p2w_wait<0>();
        // ---- pad: for a fixed offset the round's cells map to distinct nodes: plain read-modify-write ----------------
        ulonglong2 *const P = (ulonglong2 *)S.pad + (((cell >> 4) * 6 + ((cell >> 2) & 3)) * 6 + (cell & 3));
#pragma unroll 1
        for (int tn = 0; tn < turns; tn++) {
#pragma unroll
          for (int o = 0; o < 27; o++) {
            if (np > 0 && turn == tn) {
              ulonglong2 *d = P + ((o / 9) * 36 + ((o / 3) % 3) * 6 + (o % 3));
              const ulonglong2 v = *d;
              *d = make_ulonglong2(f2_add(v.x, a01[o]), f2_add(v.y, a2m[o]));
            }
            __syncwarp();
          }
        }
      }
    }
    // ---- flush: destination blocks resolved once per bin (activated only where a particle reaches), one red per node ------
    if (lane < 8) {
      float4 *p = nullptr;
      if ((touched >> lane) & 1u) {
        const int nx = bx + (lane >> 2), ny = by + ((lane >> 1) & 1), nz = bz + (lane & 1);
        p = mpm_block_ptr<false>(M, nx, ny, nz);
        if (!p) p = mpm_block_ptr<true>(M, nx, ny, nz);
      }
      S.nb[lane] = p;
      M.nbptr[(u64)slot * 8 + (u32)lane] = (u64)p;  // G2P finds the bin's neighbour blocks without walking the tree again
    }
    __syncwarp();
    for (int nidx = lane; nidx < 216; nidx += 32) {
      float4 v = S.pad[nidx];
      v.w *= M.p_mass;
      const int px = nidx / 36, py = (nidx / 6) % 6, pz = nidx % 6;
      float4 *dst = S.nb[((px >> 2) << 2) | ((py >> 2) << 1) | (pz >> 2)];
      if (dst && (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f)) red_add_v4(dst + (((px & 3) << 4) | ((py & 3) << 2) | (pz & 3)), v);
    }
    if (lane == 0) M.glist[(u64)gcls * M.max_slots + gpos] = slot;
    __syncwarp();
  }
}
