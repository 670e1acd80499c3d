/*
 * oracle/mpm_oracle.c -- TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline; never
 * linked into or called by the product path).
 *
 * fp32 C restatement of the reference's user-level MLS-MPM kernels (they exist only as
 * Python in the reference, so no compiled reference exists for this half of the path):
 *   mpm3d  : python/taichi/examples/simulation/mpm3d.py:36-81      (3-D, J model)
 *   mpm88  : tests/python/test_mpm88.py:34-80                        (2-D, J model)
 *   mpm128 : python/taichi/examples/simulation/mpm128.py:28-102      (2-D, F/SVD model)
 *            + python/taichi/_funcs.py:53-157 (2x2 polar decomposition, SVD)
 * Statement order and float casts follow the Python sources; `int()`/.cast(int) = C
 * truncation.  Built with -ffp-contract=off.
 *
 * Parity pinned: mpm88 reproduces the reference's regression vector
 * tests/python/test_mpm88.py:94-101 (tests/test_mpm_oracle.py).  mpm128's SVD/snow branch
 * has no numeric regression test in the reference: "parity unpinned" for that sub-path.
 *
 * P2G accumulation modes: 0 = serial, fp64 accumulators (the checker: order-free up to
 * fp64 rounding); 1 = multi-threaded fp32 CAS atomics, as the reference's x64 backend
 * does (taichi/runtime/llvm/runtime_module/atomic.h:75) -- used for the timed baseline.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------- tiny parallel-for over [0,n) with pthreads ---------------- */
typedef void (*range_fn)(void *ctx, int64_t lo, int64_t hi);
typedef struct {
  range_fn fn;
  void *ctx;
  int64_t lo, hi;
} PJob;
static void *pjob_main(void *p) {
  PJob *j = (PJob *)p;
  j->fn(j->ctx, j->lo, j->hi);
  return NULL;
}
static void parallel_for(int64_t n, int nthreads, range_fn fn, void *ctx) {
  if (nthreads <= 1 || n < 1024) {
    fn(ctx, 0, n);
    return;
  }
  pthread_t th[256];
  PJob jobs[256];
  if (nthreads > 256) nthreads = 256;
  int64_t chunk = (n + nthreads - 1) / nthreads;
  int started = 0;
  for (int t = 0; t < nthreads; t++) {
    int64_t lo = t * chunk, hi = lo + chunk < n ? lo + chunk : n;
    if (lo >= hi) break;
    jobs[t] = (PJob){fn, ctx, lo, hi};
    pthread_create(&th[t], NULL, pjob_main, &jobs[t]);
    started++;
  }
  for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
}
static void atomic_add_f32(float *p, float v) { /* CAS loop, atomic.h:75 */
  uint32_t *ip = (uint32_t *)p, old, neu;
  do {
    old = __atomic_load_n(ip, __ATOMIC_RELAXED);
    float f;
    memcpy(&f, &old, 4);
    f += v;
    memcpy(&neu, &f, 4);
  } while (!__atomic_compare_exchange_n(ip, &old, neu, 0, __ATOMIC_SEQ_CST, __ATOMIC_RELAXED));
}

/* ============================ 3-D J model (mpm3d.py) ============================ */
typedef struct {
  int n_grid;
  int64_t N;
  float dt, dx, p_vol, p_mass, E, gravity;
  int bound;
  float *x, *v, *C, *J;     /* [N][3], [N][3], [N][3][3], [N] */
  float *gv, *gm;           /* [n^3][3], [n^3] (fp32 views)   */
  double *acc;              /* [n^3][4] fp64 accumulators (mode 0) or NULL */
  int mode;
} M3;

static void m3_p2g(void *c_, int64_t lo, int64_t hi) {
  M3 *c = (M3 *)c_;
  const int n = c->n_grid;
  const float dx = c->dx, dt = c->dt;
  for (int64_t p = lo; p < hi; p++) {
    float Xp[3], fx[3], w[3][3];
    int base[3];
    for (int d = 0; d < 3; d++) {
      Xp[d] = c->x[p * 3 + d] / dx;
      base[d] = (int)(Xp[d] - 0.5f);
      fx[d] = Xp[d] - (float)base[d];
      w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
      w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
      w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
    }
    float stress = -dt * 4.0f * c->E * c->p_vol * (c->J[p] - 1.0f) / (dx * dx);
    float A[3][3];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) A[i][j] = (i == j ? stress : 0.0f) + c->p_mass * c->C[p * 9 + i * 3 + j];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) {
          float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)k - fx[2]) * dx};
          float weight = 1.0f;
          weight *= w[i][0];
          weight *= w[j][1];
          weight *= w[k][2];
          int64_t g = ((int64_t)(base[0] + i) * n + (base[1] + j)) * n + (base[2] + k);
          float mv[3];
          for (int d = 0; d < 3; d++)
            mv[d] = weight * (c->p_mass * c->v[p * 3 + d] + (A[d][0] * dpos[0] + A[d][1] * dpos[1] + A[d][2] * dpos[2]));
          float m = weight * c->p_mass;
          if (c->mode == 0) {
            for (int d = 0; d < 3; d++) c->acc[g * 4 + d] += (double)mv[d];
            c->acc[g * 4 + 3] += (double)m;
          } else {
            for (int d = 0; d < 3; d++) atomic_add_f32(&c->gv[g * 3 + d], mv[d]);
            atomic_add_f32(&c->gm[g], m);
          }
        }
  }
}
static void m3_gridop(void *c_, int64_t lo, int64_t hi) {
  M3 *c = (M3 *)c_;
  const int n = c->n_grid;
  for (int64_t g = lo; g < hi; g++) {
    int I[3] = {(int)(g / ((int64_t)n * n)), (int)((g / n) % n), (int)(g % n)};
    float *v = &c->gv[g * 3];
    float m = c->gm[g];
    if (m > 0.0f)
      for (int d = 0; d < 3; d++) v[d] /= m;
    v[1] -= c->dt * c->gravity;
    for (int d = 0; d < 3; d++) {
      int cond = ((I[d] < c->bound) && (v[d] < 0.0f)) || ((I[d] > n - c->bound) && (v[d] > 0.0f));
      if (cond) v[d] = 0.0f;
    }
  }
}
static void m3_g2p(void *c_, int64_t lo, int64_t hi) {
  M3 *c = (M3 *)c_;
  const int n = c->n_grid;
  const float dx = c->dx, dt = c->dt;
  for (int64_t p = lo; p < hi; p++) {
    float Xp[3], fx[3], w[3][3];
    int base[3];
    for (int d = 0; d < 3; d++) {
      Xp[d] = c->x[p * 3 + d] / dx;
      base[d] = (int)(Xp[d] - 0.5f);
      fx[d] = Xp[d] - (float)base[d];
      w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
      w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
      w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
    }
    float nv[3] = {0, 0, 0}, nC[9] = {0};
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) {
          float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)k - fx[2]) * dx};
          float weight = 1.0f;
          weight *= w[i][0];
          weight *= w[j][1];
          weight *= w[k][2];
          int64_t g = ((int64_t)(base[0] + i) * n + (base[1] + j)) * n + (base[2] + k);
          const float *gv = &c->gv[g * 3];
          for (int a = 0; a < 3; a++) {
            nv[a] += weight * gv[a];
            for (int b = 0; b < 3; b++) nC[a * 3 + b] += 4.0f * weight * (gv[a] * dpos[b]) / (dx * dx);
          }
        }
    for (int d = 0; d < 3; d++) {
      c->v[p * 3 + d] = nv[d];
      c->x[p * 3 + d] += dt * nv[d];
    }
    c->J[p] *= 1.0f + dt * (nC[0] + nC[4] + nC[8]);
    memcpy(&c->C[p * 9], nC, sizeof(nC));
  }
}
static void m3_finish_acc(void *c_, int64_t lo, int64_t hi) {
  M3 *c = (M3 *)c_;
  for (int64_t g = lo; g < hi; g++) {
    for (int d = 0; d < 3; d++) c->gv[g * 3 + d] = (float)c->acc[g * 4 + d];
    c->gm[g] = (float)c->acc[g * 4 + 3];
  }
}

/* One or more substeps of mpm3d.  grid_v/grid_m hold the post-grid-op state of the
   LAST substep on return; if p2g_v/p2g_m are non-NULL they receive the raw P2G
   mass/momentum of the last substep (before normalisation). */
int orc_mpm3d_substeps(int n_grid, int64_t N, float *x, float *v, float *C, float *J, float dt, float E,
                       float p_rho, float gravity, int bound, int nsteps, int mode, int nthreads,
                       float *grid_v, float *grid_m, float *p2g_v, float *p2g_m) {
  M3 c;
  memset(&c, 0, sizeof(c));
  int64_t cells = (int64_t)n_grid * n_grid * n_grid;
  c.n_grid = n_grid;
  c.N = N;
  c.dt = dt;
  c.dx = 1.0f / (float)n_grid;
  c.p_vol = (c.dx * 0.5f) * (c.dx * 0.5f);
  c.p_mass = c.p_vol * p_rho;
  c.E = E;
  c.gravity = gravity;
  c.bound = bound;
  c.x = x, c.v = v, c.C = C, c.J = J;
  c.gv = grid_v, c.gm = grid_m;
  c.mode = mode;
  if (mode == 0) c.acc = (double *)malloc(sizeof(double) * 4 * (size_t)cells);
  for (int s = 0; s < nsteps; s++) {
    memset(grid_v, 0, sizeof(float) * 3 * (size_t)cells);
    memset(grid_m, 0, sizeof(float) * (size_t)cells);
    if (mode == 0) {
      memset(c.acc, 0, sizeof(double) * 4 * (size_t)cells);
      m3_p2g(&c, 0, N);
      parallel_for(cells, nthreads, m3_finish_acc, &c);
    } else {
      parallel_for(N, nthreads, m3_p2g, &c);
    }
    if (s == nsteps - 1 && p2g_v) memcpy(p2g_v, grid_v, sizeof(float) * 3 * (size_t)cells);
    if (s == nsteps - 1 && p2g_m) memcpy(p2g_m, grid_m, sizeof(float) * (size_t)cells);
    parallel_for(cells, nthreads, m3_gridop, &c);
    parallel_for(N, nthreads, m3_g2p, &c);
  }
  free(c.acc);
  return 0;
}

/* ============== 3-D J model on a SPARSE grid (blocked active set, the reference's ti.cpu cost model) ==============
 * Same arithmetic as m3_p2g / m3_gridop / m3_g2p above (mpm3d.py:36-81), but the grid is what a
 * pointer -> bitmasked -> dense(4^3) SNode tree gives the reference: storage exists for ACTIVE 4^3 blocks only;
 * a write to an inactive block activates it (lower_access.cpp:254-279: claim, zero-filled node, publish), a read
 * of an inactive block sees zeros (the ambient element, node_pointer.h:90-98); clear / grid-op touch the active
 * blocks only (struct-for over the leaf list); `deactivate_every` = K > 0 reproduces the reference's sparse-MPM
 * idiom of deactivating the grid every K-th substep (snode.py:286-304 deactivate_all + GC, test_gc.py:40-52),
 * after which P2G re-activates exactly the touched blocks.
 * mode 0: fp64 accumulators (CAS on the bit pattern, any thread count: order-free to fp64 rounding) = the checker;
 * mode 1: fp32 CAS atomics as the reference's x64 backend (atomic.h:75) = the timed CPU baseline.            */
typedef struct {
  int n_grid, nb;
  int64_t N;
  float dt, dx, p_vol, p_mass, E, gravity;
  int bound;
  float *x, *v, *C, *J;
  int32_t *blk_of;  /* [nb^3] -> block index, -1 inactive, -2 being activated */
  int32_t *blk_key; /* [cap]  */
  int32_t nblk, cap;
  float *val;  /* [cap][64][4] {vx, vy, vz, m} */
  double *acc; /* [cap][64][4] (mode 0)        */
  int mode, overflow;
} M3S;

static void atomic_add_f64(double *p, double v) {
  uint64_t *ip = (uint64_t *)p, old, neu;
  do {
    old = __atomic_load_n(ip, __ATOMIC_RELAXED);
    double f;
    memcpy(&f, &old, 8);
    f += v;
    memcpy(&neu, &f, 8);
  } while (!__atomic_compare_exchange_n(ip, &old, neu, 0, __ATOMIC_SEQ_CST, __ATOMIC_RELAXED));
}
/* block index of 4^3 block (bx,by,bz); activate = 1 claims and zero-fills a new block */
static inline int32_t m3s_block(M3S *c, int bx, int by, int bz, int activate) {
  if (bx >= c->nb || by >= c->nb || bz >= c->nb) return -1; /* (only reached by the non-touching corner lookups of G2P) */
  const int64_t key = ((int64_t)bx * c->nb + by) * c->nb + bz;
  int32_t b = __atomic_load_n(&c->blk_of[key], __ATOMIC_ACQUIRE);
  if (b >= 0 || !activate) return b;
  if (b == -1) {
    int32_t expect = -1;
    if (__atomic_compare_exchange_n(&c->blk_of[key], &expect, -2, 0, __ATOMIC_ACQ_REL, __ATOMIC_ACQUIRE)) {
      int32_t idx = __atomic_fetch_add(&c->nblk, 1, __ATOMIC_RELAXED);
      if (idx >= c->cap) {
        c->overflow = 1;
        idx = c->cap - 1;
      }
      c->blk_key[idx] = (int32_t)key;
      memset(c->val + (size_t)idx * 256, 0, 1024);
      if (c->acc) memset(c->acc + (size_t)idx * 256, 0, 2048);
      __atomic_store_n(&c->blk_of[key], idx, __ATOMIC_RELEASE);
      return idx;
    }
  }
  while ((b = __atomic_load_n(&c->blk_of[key], __ATOMIC_ACQUIRE)) < 0) {
  }
  return b;
}
static void m3s_clear(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  memset(c->val + (size_t)lo * 256, 0, (size_t)(hi - lo) * 1024);
  if (c->acc) memset(c->acc + (size_t)lo * 256, 0, (size_t)(hi - lo) * 2048);
}
static void m3s_deactivate(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  for (int64_t b = lo; b < hi; b++) c->blk_of[c->blk_key[b]] = -1;
}
static void m3s_p2g(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  const float dx = c->dx, dt = c->dt;
  for (int64_t p = lo; p < hi; p++) {
    float Xp[3], fx[3], w[3][3];
    int base[3];
    for (int d = 0; d < 3; d++) {
      Xp[d] = c->x[p * 3 + d] / dx;
      base[d] = (int)(Xp[d] - 0.5f);
      fx[d] = Xp[d] - (float)base[d];
      w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
      w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
      w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
    }
    float stress = -dt * 4.0f * c->E * c->p_vol * (c->J[p] - 1.0f) / (dx * dx);
    float A[3][3];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) A[i][j] = (i == j ? stress : 0.0f) + c->p_mass * c->C[p * 9 + i * 3 + j];
    /* the stencil reaches <= 2 blocks per axis: resolve (and activate) the touched ones once per particle */
    size_t blk[8];
    const int hi[3] = {((base[0] + 2) >> 2) - (base[0] >> 2), ((base[1] + 2) >> 2) - (base[1] >> 2), ((base[2] + 2) >> 2) - (base[2] >> 2)};
    for (int q = 0; q < 8; q++) {
      const int qx = q >> 2, qy = (q >> 1) & 1, qz = q & 1;
      blk[q] = (qx <= hi[0] && qy <= hi[1] && qz <= hi[2])
                   ? (size_t)m3s_block(c, (base[0] >> 2) + qx, (base[1] >> 2) + qy, (base[2] >> 2) + qz, 1) * 64
                   : 0;
    }
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) {
          float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)k - fx[2]) * dx};
          float weight = 1.0f;
          weight *= w[i][0];
          weight *= w[j][1];
          weight *= w[k][2];
          const int gx = base[0] + i, gy = base[1] + j, gz = base[2] + k;
          const int q = ((gx >> 2) - (base[0] >> 2)) * 4 + ((gy >> 2) - (base[1] >> 2)) * 2 + ((gz >> 2) - (base[2] >> 2));
          const size_t g = blk[q] + (size_t)(((gx & 3) << 4) | ((gy & 3) << 2) | (gz & 3));
          float mv[3];
          for (int d = 0; d < 3; d++)
            mv[d] = weight * (c->p_mass * c->v[p * 3 + d] + (A[d][0] * dpos[0] + A[d][1] * dpos[1] + A[d][2] * dpos[2]));
          float m = weight * c->p_mass;
          if (c->mode == 0) {
            for (int d = 0; d < 3; d++) atomic_add_f64(&c->acc[g * 4 + d], (double)mv[d]);
            atomic_add_f64(&c->acc[g * 4 + 3], (double)m);
          } else {
            for (int d = 0; d < 3; d++) atomic_add_f32(&c->val[g * 4 + d], mv[d]);
            atomic_add_f32(&c->val[g * 4 + 3], m);
          }
        }
  }
}
static void m3s_finish_acc(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  for (int64_t g = lo * 256; g < hi * 256; g++) c->val[g] = (float)c->acc[g];
}
static void m3s_gridop(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  const int n = c->n_grid, nb = c->nb;
  for (int64_t b = lo; b < hi; b++) {
    const int32_t key = c->blk_key[b];
    const int bx = key / (nb * nb), by = (key / nb) % nb, bz = key % nb;
    for (int t = 0; t < 64; t++) {
      int I[3] = {bx * 4 + (t >> 4), by * 4 + ((t >> 2) & 3), bz * 4 + (t & 3)};
      float *v = &c->val[((size_t)b * 64 + t) * 4];
      float m = v[3];
      if (m > 0.0f)
        for (int d = 0; d < 3; d++) v[d] /= m;
      v[1] -= c->dt * c->gravity;
      for (int d = 0; d < 3; d++) {
        int cond = ((I[d] < c->bound) && (v[d] < 0.0f)) || ((I[d] > n - c->bound) && (v[d] > 0.0f));
        if (cond) v[d] = 0.0f;
      }
    }
  }
}
static void m3s_g2p(void *c_, int64_t lo, int64_t hi) {
  M3S *c = (M3S *)c_;
  const float dx = c->dx, dt = c->dt;
  static const float zero4[4] = {0, 0, 0, 0};
  for (int64_t p = lo; p < hi; p++) {
    float Xp[3], fx[3], w[3][3];
    int base[3];
    for (int d = 0; d < 3; d++) {
      Xp[d] = c->x[p * 3 + d] / dx;
      base[d] = (int)(Xp[d] - 0.5f);
      fx[d] = Xp[d] - (float)base[d];
      w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
      w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
      w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
    }
    float nv[3] = {0, 0, 0}, nC[9] = {0};
    const float *blk[8]; /* inactive blocks read the ambient zeros */
    for (int q = 0; q < 8; q++) {
      const int32_t b = m3s_block(c, (base[0] >> 2) + (q >> 2), (base[1] >> 2) + ((q >> 1) & 1), (base[2] >> 2) + (q & 1), 0);
      blk[q] = b >= 0 ? &c->val[(size_t)b * 256] : NULL;
    }
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) {
          float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)k - fx[2]) * dx};
          float weight = 1.0f;
          weight *= w[i][0];
          weight *= w[j][1];
          weight *= w[k][2];
          const int gx = base[0] + i, gy = base[1] + j, gz = base[2] + k;
          const int q = ((gx >> 2) - (base[0] >> 2)) * 4 + ((gy >> 2) - (base[1] >> 2)) * 2 + ((gz >> 2) - (base[2] >> 2));
          const float *gv = blk[q] ? blk[q] + (size_t)(((gx & 3) << 4) | ((gy & 3) << 2) | (gz & 3)) * 4 : zero4;
          for (int a = 0; a < 3; a++) {
            nv[a] += weight * gv[a];
            for (int bb = 0; bb < 3; bb++) nC[a * 3 + bb] += 4.0f * weight * (gv[a] * dpos[bb]) / (dx * dx);
          }
        }
    for (int d = 0; d < 3; d++) {
      c->v[p * 3 + d] = nv[d];
      c->x[p * 3 + d] += dt * nv[d];
    }
    c->J[p] *= 1.0f + dt * (nC[0] + nC[4] + nC[8]);
    memcpy(&c->C[p * 9], nC, sizeof(nC));
  }
}
/* nsteps substeps on the sparse grid.  Particles are updated in place.  On return out_keys[0..*out_nblk) are the
   linear keys ((bx*nb+by)*nb+bz) of the active blocks (activation order: compare as sorted sets), out_grid their
   post-grid-op {vx,vy,vz,m} cells [64][4] of the LAST substep and out_p2g the raw P2G momentum/mass of the last
   substep (any of the three may be NULL).  Returns 0, or -1 if more than cap blocks became active. */
int orc_mpm3d_sparse_substeps(int n_grid, int64_t N, float *x, float *v, float *C, float *J, float dt, float E, float p_rho,
                              float gravity, int bound, int nsteps, int mode, int nthreads, int deactivate_every, int32_t cap,
                              int32_t *out_nblk, int32_t *out_keys, float *out_p2g, float *out_grid) {
  M3S c;
  memset(&c, 0, sizeof(c));
  c.n_grid = n_grid, c.nb = n_grid / 4, c.N = N;
  c.dt = dt;
  c.dx = 1.0f / (float)n_grid;
  c.p_vol = (c.dx * 0.5f) * (c.dx * 0.5f);
  c.p_mass = c.p_vol * p_rho;
  c.E = E, c.gravity = gravity, c.bound = bound;
  c.x = x, c.v = v, c.C = C, c.J = J;
  c.mode = mode, c.cap = cap;
  const size_t nblocks = (size_t)c.nb * c.nb * c.nb;
  c.blk_of = (int32_t *)malloc(nblocks * 4);
  c.blk_key = (int32_t *)malloc((size_t)cap * 4);
  c.val = (float *)malloc((size_t)cap * 1024);
  if (mode == 0) c.acc = (double *)malloc((size_t)cap * 2048);
  memset(c.blk_of, 0xFF, nblocks * 4);
  for (int s = 0; s < nsteps; s++) {
    if (deactivate_every > 0 && s % deactivate_every == 0) {
      parallel_for(c.nblk, nthreads, m3s_deactivate, &c); /* deactivate_all + GC: nodes are zero-filled when reused */
      c.nblk = 0;
    } else {
      parallel_for(c.nblk, c.nblk >= 1024 ? nthreads : 1, m3s_clear, &c);
    }
    parallel_for(N, nthreads, m3s_p2g, &c);
    if (c.overflow) break;
    if (mode == 0) parallel_for(c.nblk, c.nblk >= 1024 ? nthreads : 1, m3s_finish_acc, &c);
    if (s == nsteps - 1 && out_p2g) memcpy(out_p2g, c.val, (size_t)c.nblk * 1024);
    parallel_for(c.nblk, c.nblk >= 1024 ? nthreads : 1, m3s_gridop, &c);
    parallel_for(N, nthreads, m3s_g2p, &c);
  }
  if (out_nblk) *out_nblk = c.nblk;
  if (!c.overflow) {
    if (out_keys) memcpy(out_keys, c.blk_key, (size_t)c.nblk * 4);
    if (out_grid) memcpy(out_grid, c.val, (size_t)c.nblk * 1024);
  }
  const int rc = c.overflow ? -1 : 0;
  free(c.blk_of), free(c.blk_key), free(c.val), free(c.acc);
  return rc;
}

/* ============================ 2-D J model (test_mpm88.py) ============================ */
typedef struct {
  int n_grid;
  int64_t N;
  float dt, dx, inv_dx, p_vol, p_mass, E;
  float *x, *v, *C, *J; /* [N][2], [N][2], [N][2][2], [N] */
  float *gv, *gm;
  double *acc; /* [n^2][3] */
} M88;

static void m88_weights(const float *xp, float inv_dx, int *base, float *fx, float w[3][2]) {
  for (int d = 0; d < 2; d++) {
    float s = xp[d] * inv_dx;
    base[d] = (int)(s - 0.5f);
    fx[d] = s - (float)base[d];
    w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
    w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
    w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
  }
}
static void m88_p2g(M88 *c) {
  const int n = c->n_grid;
  for (int64_t p = 0; p < c->N; p++) {
    int base[2];
    float fx[2], w[3][2];
    m88_weights(&c->x[p * 2], c->inv_dx, base, fx, w);
    float stress = -c->dt * c->p_vol * (c->J[p] - 1.0f) * 4.0f * c->inv_dx * c->inv_dx * c->E;
    float A[2][2] = {{stress + c->p_mass * c->C[p * 4 + 0], c->p_mass * c->C[p * 4 + 1]},
                     {c->p_mass * c->C[p * 4 + 2], stress + c->p_mass * c->C[p * 4 + 3]}};
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        float dpos[2] = {((float)i - fx[0]) * c->dx, ((float)j - fx[1]) * c->dx};
        float weight = w[i][0] * w[j][1];
        int64_t g = (int64_t)(base[0] + i) * n + (base[1] + j);
        for (int d = 0; d < 2; d++)
          c->acc[g * 3 + d] += (double)(weight * (c->p_mass * c->v[p * 2 + d] + (A[d][0] * dpos[0] + A[d][1] * dpos[1])));
        c->acc[g * 3 + 2] += (double)(weight * c->p_mass);
      }
  }
}
static void m88_gridop(M88 *c) {
  const int n = c->n_grid, bound = 3;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      int64_t g = (int64_t)i * n + j;
      c->gv[g * 2] = (float)c->acc[g * 3], c->gv[g * 2 + 1] = (float)c->acc[g * 3 + 1];
      c->gm[g] = (float)c->acc[g * 3 + 2];
      if (c->gm[g] > 0.0f) {
        float inv_m = 1.0f / c->gm[g];
        float *v = &c->gv[g * 2];
        v[0] = inv_m * v[0], v[1] = inv_m * v[1];
        v[1] -= c->dt * 9.8f;
        if (i < bound && v[0] < 0) v[0] = 0;
        if (i > n - bound && v[0] > 0) v[0] = 0;
        if (j < bound && v[1] < 0) v[1] = 0;
        if (j > n - bound && v[1] > 0) v[1] = 0;
      }
    }
}
static void m88_g2p(M88 *c) {
  const int n = c->n_grid;
  for (int64_t p = 0; p < c->N; p++) {
    int base[2];
    float fx[2], w[3][2];
    m88_weights(&c->x[p * 2], c->inv_dx, base, fx, w);
    float nv[2] = {0, 0}, nC[4] = {0, 0, 0, 0};
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        float dpos[2] = {(float)i - fx[0], (float)j - fx[1]};
        const float *gv = &c->gv[((int64_t)(base[0] + i) * n + (base[1] + j)) * 2];
        float weight = w[i][0] * w[j][1];
        for (int a = 0; a < 2; a++) {
          nv[a] += weight * gv[a];
          for (int b = 0; b < 2; b++) nC[a * 2 + b] += 4.0f * weight * (gv[a] * dpos[b]) * c->inv_dx;
        }
      }
    c->v[p * 2] = nv[0], c->v[p * 2 + 1] = nv[1];
    c->x[p * 2] += c->dt * nv[0], c->x[p * 2 + 1] += c->dt * nv[1];
    c->J[p] *= 1.0f + c->dt * (nC[0] + nC[3]);
    memcpy(&c->C[p * 4], nC, sizeof(nC));
  }
}
int orc_mpm88_substeps(int n_grid, int64_t N, float *x, float *v, float *C, float *J, float dt, float E,
                       int nsteps, float *grid_v, float *grid_m) {
  M88 c;
  int64_t cells = (int64_t)n_grid * n_grid;
  c.n_grid = n_grid, c.N = N, c.dt = dt;
  c.dx = 1.0f / (float)n_grid;
  c.inv_dx = 1.0f / c.dx;
  c.p_vol = (c.dx * 0.5f) * (c.dx * 0.5f);
  c.p_mass = c.p_vol * 1.0f;
  c.E = E;
  c.x = x, c.v = v, c.C = C, c.J = J, c.gv = grid_v, c.gm = grid_m;
  c.acc = (double *)malloc(sizeof(double) * 3 * (size_t)cells);
  for (int s = 0; s < nsteps; s++) {
    memset(c.acc, 0, sizeof(double) * 3 * (size_t)cells);
    m88_p2g(&c);
    m88_gridop(&c);
    m88_g2p(&c);
  }
  free(c.acc);
  return 0;
}

/* ====================== 2-D F/SVD model (mpm128.py) ====================== */
static void polar2d(const float A[4], float U[4], float P[4]) { /* _funcs.py:53-98 */
  U[0] = 1, U[1] = 0, U[2] = 0, U[3] = 1;
  memcpy(P, A, 16);
  if (A[0] == 0.0f && A[1] == 0.0f && A[2] == 0.0f && A[3] == 0.0f) return;
  float detA = A[0] * A[3] - A[2] * A[1];
  float adetA = fabsf(detA);
  float B[4] = {A[0] + A[3], A[1] - A[2], A[2] - A[1], A[3] + A[0]};
  if (detA < 0.0f) {
    B[0] = A[0] - A[3], B[1] = A[1] + A[2], B[2] = A[2] + A[1], B[3] = A[3] - A[0];
  }
  float adetB = fabsf(B[0] * B[3] - B[2] * B[1]);
  float k = 1.0f / sqrtf(adetB);
  for (int i = 0; i < 4; i++) U[i] = B[i] * k;
  /* P = (A^T A + |det A| I) k */
  P[0] = (A[0] * A[0] + A[2] * A[2] + adetA) * k;
  P[1] = (A[0] * A[1] + A[2] * A[3]) * k;
  P[2] = (A[1] * A[0] + A[3] * A[2]) * k;
  P[3] = (A[1] * A[1] + A[3] * A[3] + adetA) * k;
}
static void svd2d(const float A[4], float U[4], float sig[2], float V[4]) { /* _funcs.py:117-157 */
  float R[4], S[4];
  polar2d(A, R, S);
  float c, s, s1, s2;
  if (fabsf(S[1]) < 1e-5f) {
    c = 1, s = 0, s1 = S[0], s2 = S[3];
  } else {
    float tao = 0.5f * (S[0] - S[3]);
    float w = sqrtf(tao * tao + S[1] * S[1]);
    float t = tao > 0 ? S[1] / (tao + w) : S[1] / (tao - w);
    c = 1.0f / sqrtf(t * t + 1.0f);
    s = -t * c;
    s1 = c * c * S[0] - 2 * c * s * S[1] + s * s * S[3];
    s2 = s * s * S[0] + 2 * c * s * S[1] + c * c * S[3];
  }
  if (s1 < s2) {
    float tmp = s1;
    s1 = s2, s2 = tmp;
    V[0] = -s, V[1] = c, V[2] = -c, V[3] = -s;
  } else {
    V[0] = c, V[1] = s, V[2] = -s, V[3] = c;
  }
  U[0] = R[0] * V[0] + R[1] * V[2], U[1] = R[0] * V[1] + R[1] * V[3];
  U[2] = R[2] * V[0] + R[3] * V[2], U[3] = R[2] * V[1] + R[3] * V[3];
  sig[0] = s1, sig[1] = s2;
}
void orc_svd2d(const float *A, float *U, float *sig, float *V) { svd2d(A, U, sig, V); }

static void mm2(const float a[4], const float b[4], float o[4]) {
  o[0] = a[0] * b[0] + a[1] * b[2], o[1] = a[0] * b[1] + a[1] * b[3];
  o[2] = a[2] * b[0] + a[3] * b[2], o[3] = a[2] * b[1] + a[3] * b[3];
}

/* mpm128.py substep(), gravity = (gx,gy) (the script multiplies by 30), attractor off unless strength != 0 */
int orc_mpm128_substeps(int n_grid, int64_t N, float *x, float *v, float *C, float *F, int32_t *material,
                        float *Jp, float dt, float E, float nu, float gx, float gy, float ax, float ay,
                        float astrength, int nsteps, float *grid_v, float *grid_m) {
  const int n = n_grid;
  const int64_t cells = (int64_t)n * n;
  const float dx = 1.0f / (float)n, inv_dx = (float)n;
  const float p_vol = (dx * 0.5f) * (dx * 0.5f), p_mass = p_vol * 1.0f;
  const float mu_0 = E / (2 * (1 + nu)), lambda_0 = E * nu / ((1 + nu) * (1 - 2 * nu));
  double *acc = (double *)malloc(sizeof(double) * 3 * (size_t)cells);
  for (int st = 0; st < nsteps; st++) {
    memset(acc, 0, sizeof(double) * 3 * (size_t)cells);
    for (int64_t p = 0; p < N; p++) {
      int base[2];
      float fx[2], w[3][2];
      m88_weights(&x[p * 2], inv_dx, base, fx, w);
      float *Fp = &F[p * 4], *Cp = &C[p * 4];
      float M[4] = {1.0f + dt * Cp[0], dt * Cp[1], dt * Cp[2], 1.0f + dt * Cp[3]}, Fn[4];
      mm2(M, Fp, Fn);
      memcpy(Fp, Fn, 16);
      float h = fmaxf(0.1f, fminf(5.0f, expf(10.0f * (1.0f - Jp[p]))));
      if (material[p] == 1) h = 0.3f;
      float mu = mu_0 * h, la = lambda_0 * h;
      if (material[p] == 0) mu = 0.0f;
      float U[4], sig[2], V[4];
      svd2d(Fp, U, sig, V);
      float J = 1.0f;
      for (int d = 0; d < 2; d++) {
        float new_sig = sig[d];
        if (material[p] == 2) new_sig = fminf(fmaxf(sig[d], 1 - 2.5e-2f), 1 + 4.5e-3f);
        Jp[p] *= sig[d] / new_sig;
        sig[d] = new_sig;
        J *= new_sig;
      }
      float Vt[4] = {V[0], V[2], V[1], V[3]};
      if (material[p] == 0) {
        float sq = sqrtf(J);
        Fp[0] = sq, Fp[1] = 0, Fp[2] = 0, Fp[3] = sq;
      } else if (material[p] == 2) {
        float Us[4] = {U[0] * sig[0], U[1] * sig[1], U[2] * sig[0], U[3] * sig[1]};
        mm2(Us, Vt, Fp);
      }
      float UVt[4], D[4], Ft[4] = {Fp[0], Fp[2], Fp[1], Fp[3]}, DFt[4];
      mm2(U, Vt, UVt);
      for (int i = 0; i < 4; i++) D[i] = Fp[i] - UVt[i];
      mm2(D, Ft, DFt);
      float stress[4];
      for (int i = 0; i < 4; i++) stress[i] = 2 * mu * DFt[i] + ((i == 0 || i == 3) ? la * J * (J - 1) : 0.0f);
      float k = -dt * p_vol * 4 * inv_dx * inv_dx;
      float A[4];
      for (int i = 0; i < 4; i++) A[i] = k * stress[i] + p_mass * Cp[i];
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
          float dpos[2] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx};
          float weight = w[i][0] * w[j][1];
          int64_t g = (int64_t)(base[0] + i) * n + (base[1] + j);
          acc[g * 3 + 0] += (double)(weight * (p_mass * v[p * 2] + (A[0] * dpos[0] + A[1] * dpos[1])));
          acc[g * 3 + 1] += (double)(weight * (p_mass * v[p * 2 + 1] + (A[2] * dpos[0] + A[3] * dpos[1])));
          acc[g * 3 + 2] += (double)(weight * p_mass);
        }
    }
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++) {
        int64_t g = (int64_t)i * n + j;
        float *gv = &grid_v[g * 2];
        gv[0] = (float)acc[g * 3], gv[1] = (float)acc[g * 3 + 1];
        grid_m[g] = (float)acc[g * 3 + 2];
        if (grid_m[g] > 0) {
          float im = 1 / grid_m[g];
          gv[0] = im * gv[0], gv[1] = im * gv[1];
          gv[0] += dt * gx * 30, gv[1] += dt * gy * 30;
          float dist[2] = {ax - dx * (float)i, ay - dx * (float)j};
          float nrm = sqrtf(dist[0] * dist[0] + dist[1] * dist[1]);
          gv[0] += dist[0] / (0.01f + nrm) * astrength * dt * 100;
          gv[1] += dist[1] / (0.01f + nrm) * astrength * dt * 100;
          if (i < 3 && gv[0] < 0) gv[0] = 0;
          if (i > n - 3 && gv[0] > 0) gv[0] = 0;
          if (j < 3 && gv[1] < 0) gv[1] = 0;
          if (j > n - 3 && gv[1] > 0) gv[1] = 0;
        }
      }
    for (int64_t p = 0; p < N; p++) {
      int base[2];
      float fx[2], w[3][2];
      m88_weights(&x[p * 2], inv_dx, base, fx, w);
      float nv[2] = {0, 0}, nC[4] = {0, 0, 0, 0};
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
          float dpos[2] = {(float)i - fx[0], (float)j - fx[1]};
          const float *gv = &grid_v[((int64_t)(base[0] + i) * n + (base[1] + j)) * 2];
          float weight = w[i][0] * w[j][1];
          for (int a = 0; a < 2; a++) {
            nv[a] += weight * gv[a];
            for (int b = 0; b < 2; b++) nC[a * 2 + b] += 4 * inv_dx * weight * (gv[a] * dpos[b]);
          }
        }
      v[p * 2] = nv[0], v[p * 2 + 1] = nv[1];
      memcpy(&C[p * 4], nC, 16);
      x[p * 2] += dt * nv[0], x[p * 2 + 1] += dt * nv[1];
    }
  }
  free(acc);
  return 0;
}
