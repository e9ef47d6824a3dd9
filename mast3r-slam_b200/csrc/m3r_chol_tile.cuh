// m3r_chol_tile.cuh — tile-level building blocks of the fp64 Cholesky solve shared by the
// 32x32 dataflow kernel (m3r_gn_solve.cu, small systems) and the 64x64 macro-tile kernel
// (m3r_gn_chol64.cu, large systems): flag primitives, the 4-warp POTRF / TRSM teams, the
// diagnostics hooks and the final dx / retraction / convergence step.
#pragma once
#include "m3r_gn.cuh"
#include "m3r_sim3.cuh"

namespace m3r {

constexpr int NB = kGnNB;

// ---------------------------------------------------------------------------------
// Tile-dataflow Cholesky (left-looking), NB = 32, one 64-thread CTA per tile task.
// ---------------------------------------------------------------------------------
constexpr int kCholThreads = 128;
constexpr int kTY = kCholThreads / 8;  // thread grid kTY x 8 over a 32x32 tile
constexpr int kRX = NB / kTY;          // micro-tile: rows ty + kTY*x (x < kRX), cols tx + 8*y (y < 4)

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// All threads of the CTA return once *f == epoch (written by another CTA of this launch;
// the cooperative launch guarantees that CTA is resident, so the spin terminates).
__device__ __forceinline__ void wait_flag(const int* f, int epoch) {
  if (threadIdx.x == 0) {
    while (ld_acquire(f) != epoch) __nanosleep(32);
  }
  __syncthreads();
}
// Publish: every thread's global writes so far become visible to whoever acquires *f (the CTA barrier orders
// them before thread 0, whose st.release.gpu is cumulative: no separate __threadfence, which only added a
// second MEMBAR).
__device__ __forceinline__ void post_flag(int* f, int epoch) {
  __syncthreads();
  if (threadIdx.x == 0) { st_release(f, epoch); }
}

// Optional timeline of the dataflow kernel (diagnostics; set by m3r_debug_chol_trace): per
// task 4 x u64 = {kind<<32 | i<<16 | j, t_begin, t_inputs_ready, t_end} in globaltimer ns.
// Process-wide diagnostics switches; atomics so that concurrent solves (one host thread per
// device) read a consistent value each — they are test hooks, not per-solve state.
extern std::atomic<unsigned long long*> g_chol_trace;
extern std::atomic<int> g_chol_trace_cap;
extern std::atomic<int> g_chol_delay_us, g_chol_delay_which;
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t;
}

struct CholArgs {
  double* A; int Dp, D, nb;
  int* flags;      // [nb*nb] tile (i,j) ready, then [nb] x block ready
  double* x;       // [Dp] solution of the normal equations (before the sign flip), then [Dp] 1/diag(L)
  float* Twc; float* dx; int N; float delta_thresh;
  GnHeader* hdr;
  unsigned long long* trace; int trace_cap;
  int delay_us, delay_which;  // diagnostics (m3r_debug_chol_delay): stall some tasks before they start
};

__device__ __forceinline__ void trace_put(const CholArgs& a, int slot, int kind, int i, int j,
                                          unsigned long long t0, unsigned long long t1) {
  if (a.trace && threadIdx.x == 0 && slot < a.trace_cap) {
    unsigned long long* p = a.trace + 4 * (size_t)slot;
    p[0] = ((unsigned long long)kind << 32) | ((unsigned long long)i << 16) | (unsigned long long)j;
    p[1] = t0; p[2] = t1; p[3] = gtimer();
  }
}

// global tile (bi,bj) -> shared [32][33], L2-coherent loads (tiles are written by other SMs)
__device__ __forceinline__ void load_tile(const double* __restrict__ A, int Dp, int bi, int bj,
                                          double (*s)[NB + 1]) {
  const double* src = A + (size_t)bi * NB * Dp + (size_t)bj * NB;
  for (int idx = threadIdx.x; idx < NB * NB / 2; idx += kCholThreads) {
    const int r = idx / (NB / 2), c2 = (idx % (NB / 2)) * 2;
    const double2 v = __ldcg(reinterpret_cast<const double2*>(src + (size_t)r * Dp + c2));
    s[r][c2] = v.x; s[r][c2 + 1] = v.y;
  }
}

// Tile accessor: the [32][33] staging layout (a warp reading one column, lane = row, hits 32 banks).
struct TilePad {
  double* p;
  __device__ __forceinline__ double& operator()(int r, int c) const { return p[r * (NB + 1) + c]; }
};
// 4x4 Cholesky of a diagonal block held (lower) in b, redundantly by every lane: the only
// rsqrt chain of the factorisation.  Rows whose global index is >= D (augmented rhs row,
// padding) and non-positive pivots take pivot 1.
__device__ __forceinline__ bool factor4(double (&b)[4][4], double (&rr)[4], int g, int D) {
  bool fail = false;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    double d = b[j][j];
#pragma unroll
    for (int k = 0; k < j; k++) d = fma(-b[j][k], b[j][k], d);
    const bool real = g + j < D;
    const bool bad = !(d > 0.0);
    fail |= real && bad;
    const bool unit = !real || bad;
    rr[j] = unit ? 1.0 : rsqrt(d);
    b[j][j] = unit ? 1.0 : d * rr[j];
#pragma unroll
    for (int i = j + 1; i < 4; i++) {
      double v = b[i][j];
#pragma unroll
      for (int k = 0; k < j; k++) v = fma(-b[i][k], b[j][k], v);
      b[i][j] = v * rr[j];
    }
  }
  return fail;
}

// Rank-4 update of the own-row entries at columns jj0..jj0+3 with the finished panel at column
// p0 (x = own row of that panel).  All loads first, then the arithmetic, then the stores: the
// shared-memory stores of one entry must not serialise the loads of the next (may-alias).
template <class Tile>
__device__ __forceinline__ void rank4_rows(const Tile s, int r, int p0, int jj0, const double (&x)[4],
                                           double (&t)[4]) {
  double v[4][4];
#pragma unroll
  for (int u = 0; u < 4; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) v[u][k] = s(jj0 + u, p0 + k);
    t[u] = s(r, jj0 + u);
  }
#pragma unroll
  for (int u = 0; u < 4; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) t[u] = fma(-x[k], v[u][k], t[u]);
  }
#pragma unroll
  for (int u = 0; u < 4; u++)
    if (jj0 + u <= r) s(r, jj0 + u) = t[u];
}

// In-place Cholesky of one 32x32 tile by a team of 4 warps (`sync` = barrier over exactly these
// 128 threads; w = warp within the team).  Lane r owns row r, which stays in shared memory.
// Columns are taken four at a time.  Warp 0 is the critical path: it factors the 4x4 diagonal
// block (every lane redundantly, in registers), solves its own row against it, and after the
// barrier applies the rank-4 update to the NEXT panel's four columns only (look-ahead).  Warps
// 1-3 apply the same update to all columns further right while warp 0 is already factoring the
// next block.  Compact loops on purpose: a fully unrolled register-resident form (40 KB of
// code) spent more time in cold instruction fetch than in arithmetic, and the earlier 128-thread
// 2-D form needed a barrier per column (8.0 us per tile; this one 3.7 us,
// profiles/micro/warp_tile.cu).  1/pivot goes to `dinv` for the TRSM and back-substitution.
template <class Tile, class Sync>
__device__ __forceinline__ void potrf_team(const Tile s, int w, int g0, int D, double* __restrict__ dinv,
                                           GnHeader* hdr, Sync sync) {
  const int r = threadIdx.x & 31;
  double x[4] = {0.0, 0.0, 0.0, 0.0};  // warp 0: own row of the current panel
  double cur[4] = {0.0, 0.0, 0.0, 0.0};  // warp 0: own row at the current panel's columns, up to date
  if (w == 0) {
#pragma unroll
    for (int j = 0; j < 4; j++) cur[j] = s(r, j);
  }
#pragma unroll 1
  for (int c0 = 0; c0 < NB; c0 += 4) {
    if (w == 0) {
      double b[4][4], rr[4];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) b[i][j] = s(c0 + i, c0 + j);
      const bool fail = factor4(b, rr, g0 + c0, D);
      if (r == 0) {
        if (fail) hdr->chol_fail = 1;
#pragma unroll
        for (int j = 0; j < 4; j++) dinv[c0 + j] = rr[j];
      }
      const int i = r - c0;  // own row against the block: same operation order as the block itself
#pragma unroll
      for (int j = 0; j < 4; j++) {
        double v = cur[j];
#pragma unroll
        for (int k = 0; k < j; k++) v = fma(-x[k], b[j][k], v);
        x[j] = v * rr[j];
        if (i == j) x[j] = b[j][j];  // the pivot itself (1 for unit rows)
        if (i < j) x[j] = 0.0;       // above the diagonal (rows above the panel: unused)
      }
      if (i >= 0) {
#pragma unroll
        for (int j = 0; j < 4; j++) s(r, c0 + j) = x[j];
      }
    }
    sync();  // panel c0 is final; warps 1-3 have finished the previous panel's updates
    if (w == 0) {
      if (c0 + 4 < NB) rank4_rows(s, r, c0, c0 + 4, x, cur);
      __syncwarp();
    } else {
      double xw[4], t[4];
#pragma unroll
      for (int k = 0; k < 4; k++) xw[k] = s(r, c0 + k);
#pragma unroll 1
      for (int jj = c0 + 8 + 4 * (w - 1); jj < NB; jj += 12) rank4_rows(s, r, c0, jj, xw, t);
    }
  }
}

// X <- X * Ljj^-T for the 32x32 tile X by the 4-warp team (rows are independent), right-looking in
// blocks of four columns: lane r of every warp holds row r; warp w owns column blocks w and w+4
// (8 columns) in registers.  Per block: its owner solves the 4x4 triangle and writes the four
// finished columns, one barrier, then every warp applies the rank-4 update to the blocks it still
// owns.  Each element sees its fma's in the order of a plain forward substitution (finished
// columns ascending, then the triangle), so the bits equal those of the earlier one-warp,
// barrier-free form — whose chain of 112 x (5 shared-memory loads + 4 fma) per lane cost 4.2 us
// of every block column's critical path; this one takes 1.2 us.
template <class Tile, class Sync>
__device__ __forceinline__ void trsm_team(const Tile sx, const Tile sl, const double* __restrict__ dinv, int w,
                                          Sync sync) {
  const int r = threadIdx.x & 31;
  double a[2][4];
#pragma unroll
  for (int q = 0; q < 2; q++)
#pragma unroll
    for (int j = 0; j < 4; j++) a[q][j] = sx(r, 4 * (w + 4 * q) + j);
#pragma unroll
  for (int b = 0; b < NB / 4; b++) {
    const int c0 = 4 * b;
    if (w == (b & 3)) {
      double l[4][4];
#pragma unroll
      for (int j = 1; j < 4; j++)
#pragma unroll
        for (int k = 0; k < j; k++) l[j][k] = sl(c0 + j, c0 + k);
#pragma unroll
      for (int j = 0; j < 4; j++) {
#pragma unroll
        for (int k = 0; k < j; k++) a[b >> 2][j] = fma(-a[b >> 2][k], l[j][k], a[b >> 2][j]);
        a[b >> 2][j] *= dinv[c0 + j];
      }
#pragma unroll
      for (int j = 0; j < 4; j++) sx(r, c0 + j) = a[b >> 2][j];
    }
    sync();
    if (b + 1 < NB / 4) {
      double x[4];
#pragma unroll
      for (int k = 0; k < 4; k++) x[k] = sx(r, c0 + k);
#pragma unroll
      for (int q = 0; q < 2; q++) {
        const int bb = w + 4 * q;
        if (bb > b) {
#pragma unroll
          for (int k = 0; k < 4; k++)
#pragma unroll
            for (int j = 0; j < 4; j++) a[q][j] = fma(-x[k], sl(4 * bb + j, c0 + k), a[q][j]);
        }
      }
    }
  }
}

// Protocol test (m3r_debug_chol_delay): the result must not depend on when a CTA gets to a task
// or how long it takes.  which = 1 / 2: diagonal (j > 0) / off-diagonal owners wait delay_us
// before they start; which >= 3: every task waits a pseudo-random 0..delay_us (a hash of the
// tile, the iteration, the stall point and `which` as the seed) before it starts and again
// before it stores its tile.
__device__ __forceinline__ void chol_debug_stall(const CholArgs& a, int i, int j, int epoch, int point) {
  if (a.delay_us <= 0) return;
  unsigned long long ns = 1000ull * (unsigned long long)a.delay_us;
  if (a.delay_which >= 3) {
    unsigned h = (unsigned)(i * 73856093) ^ (unsigned)(j * 19349663) ^ (unsigned)(epoch * 83492791) ^
                 (unsigned)((2 * a.delay_which + point) * 2654435761u);
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    ns = ns * (h & 1023u) / 1023u;
  } else if (point != 0 || ((i == j && j > 0) ? 1 : (i != j ? 2 : 0)) != a.delay_which) {
    return;
  }
  const unsigned long long until = gtimer() + ns;
  while (gtimer() < until) __nanosleep(200);
}

// dx = -x (zeros on failure), retract, |dx| test  (gn_kernels.cu:1209-1222) — one team of
// `nthr` threads (`sync` = barrier over exactly those threads).
template <class Sync>
__device__ __forceinline__ void gn_update_block(float* __restrict__ Twc, const double* __restrict__ x,
                                                int N, int D, float delta_thresh,
                                                float* __restrict__ dx, GnHeader* hdr, float* red,
                                                int nthr, int tid, Sync sync) {
  const bool timed_out = *reinterpret_cast<volatile int*>(&hdr->peer_timeout) != 0;
  const bool fail = *reinterpret_cast<volatile int*>(&hdr->chol_fail) != 0 || timed_out;
  float ss = 0.f;
  for (int i = tid; i < D; i += nthr) {
    const float v = fail ? 0.f : (float)(-__ldcg(x + i));
    dx[i] = v;
    ss += v * v;
  }
  sync();  // dx visible to the retraction below (same team)
  for (int p = 1 + tid; p < N; p += nthr) {
    float pose[8], xi[7];
#pragma unroll
    for (int m = 0; m < 8; m++) pose[m] = Twc[p * 8 + m];
#pragma unroll
    for (int m = 0; m < 7; m++) xi[m] = dx[(p - 1) * 7 + m];
    retr_sim3_f(xi, pose);
#pragma unroll
    for (int m = 0; m < 8; m++) Twc[p * 8 + m] = pose[m];
  }
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  sync();
  if (tid == 0) {
    float tot = 0.f;
    for (int w = 0; w < (nthr + 31) / 32; w++) tot += red[w];
    hdr->iters += 1;
    // a step zeroed by a peer timeout is not convergence: `done` stays 0, the abort is latched
    // (peer_timeout = 2 turns every later launch of this solve into a no-op) and reported
    if (timed_out) hdr->peer_timeout = 2;
    else if (sqrtf(tot) < delta_thresh) hdr->done = 1;
  }
}

}  // namespace m3r
