// b2_mpm_p2g.cuh -- the block-local (ti.block_local) P2G kernel, included by b2_mpm.cu.
//
// Replaces make_block_local's pad prologue/epilogue (transforms/make_block_local.cpp:97-365) and the
// 27 x 4 float atomics per particle of the reference's P2G (codegen_llvm.cpp:1495-1569).
//
// A persistent CTA = 7 worker warps + 1 helper warp.  A work unit is <= P2G_PASS entries of one occupied
// 4^3 block's particle bin.  While the workers process unit n, the helper prepares unit n+1, so the
// workers never wait for a dependent chain of global loads or for a serial section:
//   helper   fetches the next bin from an atomic counter (two bins ahead), prefetches its particle planes
//            into L2 (cp.async.bulk.prefetch), finds every entry's cell and takes its rank inside the cell
//            from the native integer ATOMS.ADD, then cuts every cell's particle list into UNITS of <= L
//            particles (L = the smallest length that needs <= 74 units) and ranks the units by length.
//            Thread triple u = tid / 3 of the workers owns unit u, so the lanes of a warp loop over
//            (nearly) equal trip counts and no warp waits for a fat cell.
//   stage    (workers) per particle: stencil weights and the affine momentum b + A'.o
//            (A' = (stress I + m C) dx, b = m v - A' fx) go to shared memory in a TRANSPOSED order: row k
//            holds the k-th particle of every unit that has one, in rank order.  Rows are padded to odd
//            lengths so that the 5 STS.128 of one cell's consecutive particles hit distinct bank groups,
//            and the LDS.128 of a warp in the next phase are contiguous.
//   scatter  thread (unit, i-plane) accumulates its 9 nodes x {mv, m} in 18 packed fp32x2 registers with
//            FMUL2/FFMA2/FADD2 (Blackwell packed fp32: 5 issue slots per node instead of 8); no atomics.
//   flush    per-cell partials -> shared (a cell's last unit stores, its full units add in turn); every
//            node of the (4+2)^3 pad sums its <= 27 partials through a host-built source table ordered by
//            work (warps get equal trip counts) and issues one red.global.add.v4.f32 into the <= 8
//            destination blocks.  Destination blocks are resolved once per unit by walking the SNode
//            tree; only blocks that a particle really touches are activated (x64 semantics;
//            activate_halo reproduces the CUDA-BLS epilogue).
#pragma once

#ifdef P2G_WIDE                    /* experiment (make EXTRA=-DP2G_WIDE): 2 CTAs of 384 threads per SM; 0.618 ms vs 0.489 ms on C3 */
#define P2G_NT 384
#define P2G_NW 352
#define P2G_UNITS 117
#define P2G_CTAS 2
#else
#define P2G_NT 256
#define P2G_NW 224                 /* worker threads (7 warps); the last warp is the helper      */
#define P2G_UNITS 74               /* worker thread triples                                      */
#define P2G_CTAS 3
#endif
#define P2G_STR2(x) #x
#define P2G_STR(x) P2G_STR2(x)
#define P2G_PASS MPM_PASS          /* bin entries per work unit                                  */
#define P2G_ROUNDS ((P2G_PASS + P2G_NW - 1) / P2G_NW) /* entries per worker thread */
#define P2G_XUNITS (P2G_UNITS - 64)/* units beyond one per cell                                  */
#define P2G_CAP (P2G_PASS + 32)    /* staging slots: the first 32 rows are padded to odd lengths */
#ifndef P2G_NPL
#define P2G_NPL 5                  /* staging planes: 5 = all weights staged; 4 (make EXTRA=-DP2G_NPL=4) = the scatter
                                      recomputes the y/z weights, 44 KB per CTA, 132 KB carve-out: 0.506 ms vs 0.483 ms on C3 */
#endif
#define P2G_TASKS 280              /* flush tasks: pad nodes, the heavy ones split in groups of <= 9 sources */
#define P2G_ZERO 1728              /* partial slot that is kept zero (padding of short tasks)    */

struct P2GTab {                    // flush tasks ordered by decreasing source count
  unsigned short task[P2G_TASKS][10];  // [0..8] partial index (cell * 27 + i * 9 + j * 3 + k) or P2G_ZERO; [9] pad node
};

struct P2GUnit {                   // one work unit; header by the helper, ck by the workers, tables by the helper
  u32 slot, total;                 // slot >= nslots: no more work
  int key;
  u32 q0;                          // first bin entry of the unit
  u32 cid[4];                      // chunk ids (+1) covering [q0, q0 + PASS)
  int L, nunits, maxfull;
  float invL;
  u32 touched;
  int rowstart[64];
  u32 cinfo[64];                   // by cell: full units | rank of the first full unit << 8 | rank of the last unit << 16
  u32 unit[P2G_UNITS + 2];         // by rank: cell | length << 8 | turn in the partial sum << 16
  unsigned short cstart[64];       // by cell: first position in cell-major order (the order G2P will use)
  unsigned short ck[P2G_PASS];     // by entry: cell << 10 | rank inside the cell; 0xffff = tombstone
};

struct P2GShared {
  float4 stg[P2G_NPL][P2G_CAP];        // staging planes; the 64 x 27 per-cell partials (+ zero slot) alias them later
  P2GUnit unit[2];
  int cnt[64];                     // particles per cell (counted by the workers, consumed and zeroed by the helper)
  int hist[64];                    // helper: cells per last-unit length
  unsigned short nk[64];           // helper: last units longer than k
  float4 *nb[8];
};

__device__ __forceinline__ void bar_sync_64() { asm volatile("bar.sync 1, 64;" ::: "memory"); }

// bins of the particle's stencil neighbours {0,1}^3 reached from cell c (bit nbh = dx*4 + dy*2 + dz)
__device__ __forceinline__ u32 p2g_touch_mask(int c) {
  const int h = ((c >> 5) & 1) << 2 | ((c >> 3) & 1) << 1 | ((c >> 1) & 1);
  return (u32)((0xff553311'0f050301ull >> (h * 8)) & 0xffu);
}
__device__ __forceinline__ void bar_workers() { asm volatile("bar.sync 1, " P2G_STR(P2G_NW) ";" ::: "memory"); }
__device__ __forceinline__ void bar_count_arrive() { asm volatile("bar.arrive 2, " P2G_STR(P2G_NT) ";" ::: "memory"); }
__device__ __forceinline__ void bar_count_wait() { asm volatile("bar.sync 2, " P2G_STR(P2G_NT) ";" ::: "memory"); }


// ---- helper warp -----------------------------------------------------------------------------------
// Header of a work unit, one word per lane (lanes 0..7 = the first 8 words of P2GUnit).
__device__ __forceinline__ u32 p2g_desc_field(const MpmDev &M, int gen, u32 nslots, u32 slot, u32 q0, int f) {
  if (f == 0) return slot;
  if (slot >= nslots) return 0u;
  if (f == 1) return M.count[gen][slot];
  if (f == 2) return (u32)M.slot_key[slot];
  if (f == 3) return q0;
  const u32 ci = min(q0 / MPM_CHUNK + (u32)(f - 4), (u32)M.maxc - 1u);
  return M.chunk_of[gen][(u64)slot * M.maxc + ci];
}
__device__ __forceinline__ int p2g_unit_entries(u32 total, u32 q0) { return total > q0 ? (int)min((u32)P2G_PASS, total - q0) : 0; }
__device__ __forceinline__ const u8 *p2g_entry_ptr(const MpmDev &M, int gen, u32 off0, u32 cid0, u32 cid1, u32 cid2, int t) {
  const u32 q = off0 + (u32)t, ci = q / MPM_CHUNK;
  return mpm_chunk_ptr(M, gen, ci == 0 ? cid0 : ci == 1 ? cid1 : cid2) + (q % MPM_CHUNK) * 16;
}

// Publishes the header `dsc` of a unit and starts the L2 prefetch of its four particle planes.
__device__ __forceinline__ void p2g_publish(const MpmDev &M, int gen, u32 nslots, P2GUnit &U, u32 dsc) {
  const int lane = threadIdx.x & 31;
  const u32 slot = __shfl_sync(0xffffffffu, dsc, 0), total = __shfl_sync(0xffffffffu, dsc, 1), q0 = __shfl_sync(0xffffffffu, dsc, 3);
  const u32 cid0 = __shfl_sync(0xffffffffu, dsc, 4), cid1 = __shfl_sync(0xffffffffu, dsc, 5), cid2 = __shfl_sync(0xffffffffu, dsc, 6);
  if (lane < 8) ((u32 *)&U)[lane] = dsc;
  if (slot >= nslots) return;
  const int n = p2g_unit_entries(total, q0);
  const u32 off0 = q0 % MPM_CHUNK;  // the unit starts at entry off0 of chunk cid0
  if (lane < 12) {                  // 4 planes x <= 3 chunks, one bulk prefetch each
    const int ch = lane >> 2, pl = lane & 3;
    const int first = ch == 0 ? (int)off0 : 0;
    const int before = ch == 0 ? 0 : (MPM_CHUNK - (int)off0) + (ch - 1) * MPM_CHUNK;
    const int ne = min(MPM_CHUNK - first, n - before);
    if (ne > 0) {
      const u8 *src = mpm_chunk_ptr(M, gen, ch == 0 ? cid0 : ch == 1 ? cid1 : cid2) + pl * MPM_PLANE_BYTES + first * 16;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(ne * 16) : "memory");
    }
  }
}

// Cuts the cells (counts in S.cnt) into units of <= L particles and ranks the units by length.
#ifdef P2G_TIMING
#define OCK(i) do { const long long _n = clock64(); ot[i] += _n - *ol; *ol = _n; } while (0)
#else
#define OCK(i) do { } while (0)
#endif
__device__ __forceinline__ void p2g_order(const MpmDev &M, int gen, P2GShared &S, P2GUnit &U, long long *ot = nullptr, long long *ol = nullptr) {
  const int lane = threadIdx.x & 31;
  const int cA = 2 * lane, cB = cA + 1;
  const int nA = S.cnt[cA], nB = S.cnt[cB];
  u32 tm = (nA > 0 ? p2g_touch_mask(cA) : 0u) | (nB > 0 ? p2g_touch_mask(cB) : 0u);
  tm = __reduce_or_sync(0xffffffffu, tm);
  const int nlive = __reduce_add_sync(0xffffffffu, nA + nB);
  OCK(5);
  int L = max(1, (nlive + P2G_UNITS - 1) / P2G_UNITS), fA, fB, F;
  float invL;
  for (;; L++) {  // smallest L whose full units fit the spare triples; L <= 53 for n <= 576
    invL = 1.0f / (float)L;
    fA = nA > 0 ? (int)(((float)(nA - 1) + 0.5f) * invL) : 0;
    fB = nB > 0 ? (int)(((float)(nB - 1) + 0.5f) * invL) : 0;
    F = __reduce_add_sync(0xffffffffu, fA + fB);
    if (F <= P2G_XUNITS) break;
  }
  OCK(6);
  const int remA = nA - L * fA, remB = nB - L * fB;  // length of the cell's last unit: 1..L (0: empty cell)
  int inc = fA + fB;  // ranks of the full units: cell order
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += v;
  }
  const int baseA = inc - fA - fB, baseB = baseA + fA;
  int cs = nA + nB;  // cell-major order of the live entries (handed to G2P through the chunks' order plane)
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, cs, d);
    if (lane >= d) cs += v;
  }
  U.cstart[cA] = (unsigned short)(cs - nA - nB), U.cstart[cB] = (unsigned short)(cs - nB);
  if (lane == 0) M.live[gen][(u64)U.slot * MPM_MAXPASS + U.q0 / P2G_PASS] = (u32)nlive;
  const int maxfull = __reduce_max_sync(0xffffffffu, max(fA, fB));
  const int wA = atomicAdd(&S.hist[remA], 1), wB = atomicAdd(&S.hist[remB], 1);
  __syncwarp();
  OCK(7);
  const int hA = S.hist[cA], hB = S.hist[cB];  // lengths 2*lane, 2*lane+1
  int sfx = hA + hB;  // inclusive suffix sum over lanes
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_down_sync(0xffffffffu, sfx, d);
    if (lane + d < 32) sfx += v;
  }
  const int nkB = sfx - hA - hB, nkA = nkB + hB;  // last units longer than 2*lane+1 / 2*lane
  S.nk[cA] = (unsigned short)nkA, S.nk[cB] = (unsigned short)nkB;
  // row k: every full unit + the last units longer than k, padded to an odd length
  const int padr = lane < 16 ? 1 : 0;  // rows 0..31 only: P2G_CAP
  const int lenA = cA < L ? ((F + nkA) | padr) : 0, lenB = cB < L ? ((F + nkB) | padr) : 0;
  int ps = lenA + lenB;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, ps, d);
    if (lane >= d) ps += v;
  }
  U.rowstart[cA] = ps - lenA - lenB, U.rowstart[cB] = ps - lenB;
  __syncwarp();
  OCK(8);
  const int rkA = F + S.nk[remA] + wA, rkB = F + S.nk[remB] + wB;
  U.cinfo[cA] = (u32)fA | (u32)baseA << 8 | (u32)rkA << 16;
  U.cinfo[cB] = (u32)fB | (u32)baseB << 8 | (u32)rkB << 16;
  U.unit[rkA] = (u32)cA | (u32)remA << 8;
  U.unit[rkB] = (u32)cB | (u32)remB << 8;
  for (int sgm = 0; sgm < fA; sgm++) U.unit[baseA + sgm] = (u32)cA | (u32)L << 8 | (u32)(sgm + 1) << 16;
  for (int sgm = 0; sgm < fB; sgm++) U.unit[baseB + sgm] = (u32)cB | (u32)L << 8 | (u32)(sgm + 1) << 16;
  if (lane == 0) {
    U.L = L, U.invL = invL, U.nunits = F + 64, U.maxfull = maxfull;
    U.touched = (M.activate_halo ? 0xffu : 0u) | tm;
  }
  __syncwarp();
  S.cnt[cA] = 0, S.cnt[cB] = 0, S.hist[cA] = 0, S.hist[cB] = 0;  // ready for the next unit
}

// Workers: cell and rank inside the cell of every entry of unit U (header only) -> U.ck, S.cnt
__device__ __forceinline__ void p2g_count(const MpmDev &M, int gen, u32 nslots, P2GShared &S, P2GUnit &U) {
  const int tid = threadIdx.x;
  if (U.slot >= nslots) return;
  const int n = p2g_unit_entries(U.total, U.q0);
  const u32 off0 = U.q0 % MPM_CHUNK, cid0 = U.cid[0], cid1 = U.cid[1], cid2 = U.cid[2];
  // entries in REVERSED thread order: the stage phase gives the ragged third round (entries >= 448) to the first
  // warps, so the count that follows it without a barrier gives its third round to the last ones
  const int rt = P2G_NW - 1 - tid;
  float4 p0[P2G_ROUNDS];
#pragma unroll
  for (int e = 0; e < P2G_ROUNDS; e++) {
    const int t = rt + e * P2G_NW;
    if (t < n) p0[e] = *(const float4 *)p2g_entry_ptr(M, gen, off0, cid0, cid1, cid2, t);
  }
#pragma unroll
  for (int e = 0; e < P2G_ROUNDS; e++) {
    const int t = rt + e * P2G_NW;
    if (t < n) {
      unsigned short ck = 0xffffu;
      if (p0[e].x != MPM_TOMB) {
        const int c = mpm_cell_of(M, p0[e]);
        ck = (unsigned short)((c << 10) | atomicAdd(&S.cnt[c], 1));
      }
      U.ck[t] = ck;
    }
  }
}

// one flush task: sum <= 9 partials of one pad node, one vector red into the destination block
__device__ __forceinline__ void p2g_flush_task(const MpmDev &M, const P2GShared &S, const ulonglong2 *PP, int task) {
  const u32 *row = (const u32 *)M.p2g_tab + task * 5;  // 20-B rows, read-only: stays in L1
  const uint4 t0 = make_uint4(__ldg(row), __ldg(row + 1), __ldg(row + 2), __ldg(row + 3));  // sources 0..7
  const u32 t1 = __ldg(row + 4);                                                             // source 8, pad node
  const ulonglong2 r0 = PP[t0.x & 0xffffu], r1 = PP[t0.x >> 16], r2 = PP[t0.y & 0xffffu], r3 = PP[t0.y >> 16], r4 = PP[t0.z & 0xffffu];
  const ulonglong2 r5 = PP[t0.z >> 16], r6 = PP[t0.w & 0xffffu], r7 = PP[t0.w >> 16], r8 = PP[t1 & 0xffffu];
  const u64 sxy = f2_add(f2_add(f2_add(r0.x, r1.x), f2_add(r2.x, r3.x)), f2_add(f2_add(r4.x, r5.x), f2_add(f2_add(r6.x, r7.x), r8.x)));
  const u64 szw = f2_add(f2_add(f2_add(r0.y, r1.y), f2_add(r2.y, r3.y)), f2_add(f2_add(r4.y, r5.y), f2_add(f2_add(r6.y, r7.y), r8.y)));
  float4 s4;
  f2_unpack(sxy, s4.x, s4.y);
  f2_unpack(szw, s4.z, s4.w);
  s4.w *= M.p_mass;
  const int nidx = (int)(t1 >> 16);
  const int px = nidx / 36, py = (nidx / 6) % 6, pz = nidx % 6;
  float4 *dst = S.nb[((px >> 2) << 2) | ((py >> 2) << 1) | (pz >> 2)];
  if (dst && (s4.x != 0.f || s4.y != 0.f || s4.z != 0.f || s4.w != 0.f)) red_add_v4(dst + (((px & 3) << 4) | ((py & 3) << 2) | (pz & 3)), s4);
}

#ifdef P2G_TIMING
#define TCK(i) do { const long long _n = clock64(); tacc[i] += _n - tlast; tlast = _n; } while (0)
#else
#define TCK(i) do { } while (0)
#endif
static_assert(sizeof(P2GShared) <= 55 * 1024, "4 CTAs per SM");
static_assert(P2G_NPL == 5 || sizeof(P2GShared) + 1024 <= 132 * 1024 / 3, "three CTAs inside the 132 KB carve-out");
__global__ void __launch_bounds__(P2G_NT, P2G_CTAS) k_mpm_p2g(MpmDev M, int gen) {
#ifdef P2G_TIMING
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
  int nun = 0;
#endif
  extern __shared__ __align__(16) unsigned char p2g_smem[];
  P2GShared &S = *(P2GShared *)p2g_smem;
  const int tid = threadIdx.x, lane = tid & 31;
  const u32 nslots = *M.nslots;
  if (tid < 64) S.cnt[tid] = 0, S.hist[tid] = 0;

  if (tid >= P2G_NW) {
    // ================================ helper warp ================================================
    // dsc: header of unit k (lanes 0..7); tok: the bin after the ones already turned into headers (lane 0)
    u32 tok = 0, first = 0;
    if (lane == 0) first = atomicAdd(&M.work_counter[0], 1u), tok = atomicAdd(&M.work_counter[0], 1u);
    first = __shfl_sync(0xffffffffu, first, 0);
    u32 dsc = lane < 8 ? p2g_desc_field(M, gen, nslots, first, 0u, lane) : 0u;
    p2g_publish(M, gen, nslots, S.unit[0], dsc);
    __syncthreads();  // header 0 is visible: the workers count unit 0
    for (int k = 0;; k++) {
      const u32 slot = __shfl_sync(0xffffffffu, dsc, 0), total = __shfl_sync(0xffffffffu, dsc, 1), q0 = __shfl_sync(0xffffffffu, dsc, 3);
      // header k + 1 is requested now and consumed after the order phase
      u32 nxt = lane == 0 ? 0xffffffffu : 0u;
      if (slot < nslots) {
        if (q0 + P2G_PASS < total) {
          if (lane < 8) nxt = lane < 3 ? dsc : p2g_desc_field(M, gen, nslots, slot, q0 + P2G_PASS, lane);  // same bin, next pass
        } else {
          const u32 s2 = __shfl_sync(0xffffffffu, tok, 0);
          if (lane == 0) tok = atomicAdd(&M.work_counter[0], 1u);
          if (lane < 8) nxt = p2g_desc_field(M, gen, nslots, s2, 0u, lane);
        }
      }
      TCK(0);
      bar_count_wait();  // the workers have counted unit k
      TCK(1);
#ifdef P2G_TIMING
      if (slot < nslots) p2g_order(M, gen, S, S.unit[k & 1], tacc, &tlast);
#else
      if (slot < nslots) p2g_order(M, gen, S, S.unit[k & 1]);
#endif
      TCK(2);
      p2g_publish(M, gen, nslots, S.unit[(k + 1) & 1], nxt);
      TCK(3);
      __syncthreads();  // round k starts: tables k and header k + 1 are visible
      TCK(4);
      if (slot >= nslots) break;
      dsc = nxt;
    }
#ifdef P2G_TIMING
    if (blockIdx.x == 7 && lane == 0) printf("helper: request %lld countwait %lld order[cnt %lld Lloop %lld scan1 %lld scan23 %lld tables %lld] publish %lld topbar %lld\n", tacc[0], tacc[1], tacc[5], tacc[6], tacc[7], tacc[8], tacc[2], tacc[3], tacc[4]);
#endif
    return;
  }
  // ================================== workers ======================================================
  const int u = tid / 3, plane = tid - 3 * u;
  // quadratic B-spline weight of this thread's x-offset: w_i(f) = wa (f - wb)^2 + wg  (mpm3d.py:43)
  const float fpl = (float)plane;
  const float wa = plane == 1 ? -1.0f : 0.5f, wb = plane == 0 ? 1.5f : plane == 1 ? 1.0f : 0.5f, wg = plane == 1 ? 0.75f : 0.0f;
  float4 *const P = &S.stg[0][0];  // partials [64][27] + zero slot
  __syncthreads();
  p2g_count(M, gen, nslots, S, S.unit[0]);
  bar_count_arrive();

  for (int it = 0;; it++) {
    TCK(9);
    __syncthreads();  // tables of unit `it` and the header of unit it + 1 are published
    TCK(0);
    const P2GUnit &U = S.unit[it & 1];
    if (U.slot >= nslots) break;
    const int n = p2g_unit_entries(U.total, U.q0);
    int bx, by, bz;
    mpm_unkey(M, U.key, bx, by, bz);
    if (n > 0) {
      // ---- loads + stage: the four planes of two entries are in flight at any time (a third entry only
      // exists for threads < n - 448; its loads are issued once the first entry's registers are free) ----
      const u32 off0 = U.q0 % MPM_CHUNK, cid0 = U.cid[0], cid1 = U.cid[1], cid2 = U.cid[2];
      const int L = U.L;
      const float invL = U.invL;
      struct Ent {
        float4 p0, p1, p2, p3;
        int ck;
      };
      auto load = [&](int e, Ent &E) {
        const int t = tid + e * P2G_NW;
        E.ck = t < n ? (int)U.ck[t] : 0xffff;
        if (E.ck != 0xffff) {
          const u8 *ent = p2g_entry_ptr(M, gen, off0, cid0, cid1, cid2, t);
          E.p0 = *(const float4 *)ent;
          E.p1 = *(const float4 *)(ent + MPM_PLANE_BYTES);
          E.p2 = *(const float4 *)(ent + 2 * MPM_PLANE_BYTES);
          E.p3 = *(const float4 *)(ent + 3 * MPM_PLANE_BYTES);
        }
      };
      auto stage = [&](int e, const Ent &E) {
        if (E.ck == 0xffff) return;
        const int c = E.ck >> 10, r = E.ck & 1023;
        const u32 info = U.cinfo[c];
        const int sgm = (int)(((float)r + 0.5f) * invL), k = r - sgm * L;
        const int uu = sgm < (int)(info & 0xffu) ? (int)((info >> 8) & 0xffu) + sgm : (int)(info >> 16);
        const int pos = U.rowstart[k] + uu;
        {  // cell-major order for G2P: entry (cstart[c] + r) of the unit's order plane names this entry
          const u32 qs = off0 + (u32)U.cstart[c] + (u32)r, ci = qs / MPM_CHUNK;
          u8 *chn = mpm_chunk_ptr(M, gen, ci == 0 ? cid0 : ci == 1 ? cid1 : cid2);
          ((unsigned short *)(chn + MPM_PERM_OFF))[qs % MPM_CHUNK] = (unsigned short)(tid + e * P2G_NW);
        }
        MpmStencil s;
        mpm_stencil(M, E.p0.x, E.p0.y, E.p0.z, s);
        const float stress = M.stress_k * (E.p0.w - 1.0f);
        // A' = (stress I + m C) dx ; b = m v - A' fx ; node value = b + A' o
        const float m = M.p_mass;
        const float A00 = (stress + m * E.p1.w) * M.dx, A01 = m * E.p2.x * M.dx, A02 = m * E.p2.y * M.dx;
        const float A10 = m * E.p2.z * M.dx, A11 = (stress + m * E.p2.w) * M.dx, A12 = m * E.p3.x * M.dx;
        const float A20 = m * E.p3.y * M.dx, A21 = m * E.p3.z * M.dx, A22 = (stress + m * E.p3.w) * M.dx;
        const float b0 = m * E.p1.x - (A00 * s.fx[0] + A01 * s.fx[1] + A02 * s.fx[2]);
        const float b1 = m * E.p1.y - (A10 * s.fx[0] + A11 * s.fx[1] + A12 * s.fx[2]);
        const float b2 = m * E.p1.z - (A20 * s.fx[0] + A21 * s.fx[1] + A22 * s.fx[2]);
#if P2G_NPL == 5
        S.stg[0][pos] = make_float4(s.w[0][1], s.w[1][1], s.w[2][1], s.w[0][2]);
        S.stg[1][pos] = make_float4(s.w[1][2], s.w[2][2], s.fx[0], 0.0f);
        S.stg[2][pos] = make_float4(b0, b1, A00, A10);
        S.stg[3][pos] = make_float4(A01, A11, A02, A12);
        S.stg[4][pos] = make_float4(b2, A20, A21, A22);
#else
        S.stg[0][pos] = make_float4(s.fx[0], s.fx[1], s.fx[2], b2);
        S.stg[1][pos] = make_float4(b0, b1, A00, A10);
        S.stg[2][pos] = make_float4(A01, A11, A02, A12);
        S.stg[3][pos] = make_float4(A20, A21, A22, 0.0f);
#endif
      };
      Ent Ea, Eb;
      load(0, Ea);
      load(1, Eb);
      // neighbour blocks that already exist are resolved now (latency hidden behind the stage phase);
      // touched-but-inactive ones are activated before the flush
      if (tid >= P2G_NW - 8) {
        const int nbh = tid - (P2G_NW - 8);
        S.nb[nbh] = mpm_block_ptr<false>(M, bx + (nbh >> 2), by + ((nbh >> 1) & 1), bz + (nbh & 1));
      }
      stage(0, Ea);
      if (P2G_ROUNDS > 2) load(2, Ea);
      stage(1, Eb);
      if (P2G_ROUNDS > 2) stage(2, Ea);
    }
    TCK(1);
    // ---- count the NEXT unit; the helper orders it while this unit is scattered and flushed -----------
    p2g_count(M, gen, nslots, S, S.unit[(it + 1) & 1]);
    bar_count_arrive();
    TCK(2);
    if (n == 0) continue;
    bar_workers();
    TCK(3);
    // ---- scatter: thread (unit u, plane) over the unit's particles ------------------------------------
    const u32 unit = u < U.nunits ? U.unit[u] : 0xffff0000u;
    const int cell = (int)(unit & 0xffu), np = (int)((unit >> 8) & 0xffu), turn = (int)(unit >> 16);
    u64 a01[9], a2m[9];  // {sum w mv_x, sum w mv_y}, {sum w mv_z, sum w}
#pragma unroll
    for (int q = 0; q < 9; q++) a01[q] = 0ull, a2m[q] = 0ull;
#pragma unroll 1
    for (int k = 0; k < np; k++) {
      const int pos = U.rowstart[k] + u;
#if P2G_NPL == 5
      const float4 h0 = S.stg[0][pos], h1 = S.stg[1][pos], h2 = S.stg[2][pos], h3 = S.stg[3][pos], h4 = S.stg[4][pos];
      const float tx = h1.z - wb;
      const float wx = fmaf(wa * tx, tx, wg);
      const u64 wx2 = f2_pack(wx, wx);
      const u64 wxy2[3] = {f2_mul(wx2, f2_pack(h0.x, h0.x)), f2_mul(wx2, f2_pack(h0.y, h0.y)), f2_mul(wx2, f2_pack(h0.z, h0.z))};
      const u64 wz2[3] = {f2_pack(h0.w, h0.w), f2_pack(h1.x, h1.x), f2_pack(h1.y, h1.y)};
#else
      // four planes: the (y, z) weights are recomputed here as one packed pair per offset (same formulas as mpm_stencil)
      const float4 hf = S.stg[0][pos], h2 = S.stg[1][pos], h3 = S.stg[2][pos], hA = S.stg[3][pos];
      const float4 h4 = make_float4(hf.w, hA.x, hA.y, hA.z);
      const float tx = hf.x - wb;
      const float wx = fmaf(wa * tx, tx, wg);
      const u64 wx2 = f2_pack(wx, wx);
      const u64 fyz = f2_pack(hf.y, hf.z), half2 = f2_pack(0.5f, 0.5f);
      const u64 t0 = f2_fma(fyz, f2_pack(-1.0f, -1.0f), f2_pack(1.5f, 1.5f));
      const u64 t1 = f2_add(fyz, f2_pack(-1.0f, -1.0f)), t2 = f2_add(fyz, f2_pack(-0.5f, -0.5f));
      const u64 wyz0 = f2_mul(f2_mul(half2, t0), t0);
      const u64 wyz1 = f2_fma(f2_mul(t1, f2_pack(-1.0f, -1.0f)), t1, f2_pack(0.75f, 0.75f));
      const u64 wyz2 = f2_mul(f2_mul(half2, t2), t2);
      float wy0, wy1, wy2, wz0, wz1, wz2s;
      f2_unpack(wyz0, wy0, wz0), f2_unpack(wyz1, wy1, wz1), f2_unpack(wyz2, wy2, wz2s);
      const u64 wxy2[3] = {f2_mul(wx2, f2_pack(wy0, wy0)), f2_mul(wx2, f2_pack(wy1, wy1)), f2_mul(wx2, f2_pack(wy2, wy2))};
      const u64 wz2[3] = {f2_pack(wz0, wz0), f2_pack(wz1, wz1), f2_pack(wz2s, wz2s)};
#endif
      // value at node (i = plane, 0, 0) = b + i A'[:,0]; walks along j with A'[:,1], along k with A'[:,2]
      u64 r01 = f2_fma(f2_pack(fpl, fpl), f2_pack(h2.z, h2.w), f2_pack(h2.x, h2.y));
      u64 r2o = f2_pack(fmaf(fpl, h4.y, h4.x), 1.0f);
      const u64 j01 = f2_pack(h3.x, h3.y), j2 = f2_pack(h4.z, 0.0f);
      const u64 k01 = f2_pack(h3.z, h3.w), k2 = f2_pack(h4.w, 0.0f);
#pragma unroll
      for (int j = 0; j < 3; j++) {
        u64 c01 = r01, c2o = r2o;
#pragma unroll
        for (int kk = 0; kk < 3; kk++) {
          const u64 w2 = f2_mul(wxy2[j], wz2[kk]);
          a01[j * 3 + kk] = f2_fma(w2, c01, a01[j * 3 + kk]);
          a2m[j * 3 + kk] = f2_fma(w2, c2o, a2m[j * 3 + kk]);
          if (kk < 2) c01 = f2_add(c01, k01), c2o = f2_add(c2o, k2);
        }
        if (j < 2) r01 = f2_add(r01, j01), r2o = f2_add(r2o, j2);
      }
    }
    TCK(4);
    bar_workers();  // staging is dead: the partials take its place
    TCK(5);
    // a cell's last unit stores its partials (every cell has one), the cell's full units add theirs in turn
    ulonglong2 *const P2 = (ulonglong2 *)&P[cell * 27 + plane * 9];
    if (turn == 0) {
#pragma unroll
      for (int q = 0; q < 9; q++) P2[q] = make_ulonglong2(a01[q], a2m[q]);
    }
    if (tid == P2G_NW - 1) P[P2G_ZERO] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int maxfull = U.maxfull;
    for (int tn = 1; tn <= maxfull; tn++) {
      bar_workers();
      if (turn == tn) {
#pragma unroll
        for (int q = 0; q < 9; q++) {
          const ulonglong2 o = P2[q];
          P2[q] = make_ulonglong2(f2_add(a01[q], o.x), f2_add(a2m[q], o.y));
        }
      }
    }
    // touched neighbour blocks that do not exist yet are activated now (rare in steady state)
    if (tid >= P2G_NW - 8) {
      const int nbh = tid - (P2G_NW - 8);
      if (!((U.touched >> nbh) & 1u))
        S.nb[nbh] = nullptr;
      else if (S.nb[nbh] == nullptr)
        S.nb[nbh] = mpm_block_ptr<true>(M, bx + (nbh >> 2), by + ((nbh >> 1) & 1), bz + (nbh & 1));
    }
    TCK(6);
    bar_workers();
    TCK(7);
    // ---- flush: every task sums <= 9 partials of one pad node and issues one vector red -----------------
    if (tid < P2G_TASKS) p2g_flush_task(M, S, (const ulonglong2 *)P, tid);
    if (P2G_TASKS > P2G_NW && tid >= 2 * P2G_NW - P2G_TASKS) p2g_flush_task(M, S, (const ulonglong2 *)P, P2G_NW + (P2G_NW - 1 - tid));
    TCK(8);
#ifdef P2G_TIMING
    nun++;
#endif
  }
#ifdef P2G_TIMING
  if (blockIdx.x == 7 && (tid == 0 || tid == 100 || tid == 200))
    printf("worker %d units %d: topbar %lld loads+stage %lld count %lld bar1 %lld scatter %lld bar2 %lld partials %lld bar3 %lld flush %lld end %lld\n", tid, nun, tacc[0], tacc[1], tacc[2], tacc[3], tacc[4], tacc[5], tacc[6], tacc[7], tacc[8], tacc[9]);
#endif
}
