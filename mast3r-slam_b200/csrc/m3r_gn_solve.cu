// m3r_gn_solve.cu — device-side assembly, fp64 Cholesky solve, Sim3 retraction and the
// convergence test of one Gauss-Newton iteration (sm_100a).
//
// Replaces the host part of the reference loop (gn_kernels.cu:1198-1222): SparseBlock
// update_lhs/update_rhs (:71-113, D2H copy + Eigen triplets), SparseBlock::solve (:132-153,
// Eigen::SimplicialLLT in fp64 on the CPU), the H2D copy of dx, pose_retr_kernel (:415-453)
// and linalg_norm(dx).item() (3 host syncs per iteration).  Here everything stays on the
// stream: the `break` becomes a device flag that turns later launches into no-ops.
//
// The normal equations are assembled as the dense augmented matrix  [ H  . ; b^T 1 ]  (fp64,
// lower triangle), so the Cholesky factorisation also performs the forward substitution
// (row D of L is y^T = (L^-1 b)^T); back substitution gives x and dx = -x (the reference's
// `dx = -A.solve()`).  A non-positive pivot sets chol_fail and the step becomes zero, as
// SimplicialLLT failure does in the reference (:142-150).
//
// Factorisation, back substitution, retraction and the convergence test are ONE persistent
// cooperative kernel (chol_dataflow_kernel): the matrix is cut into 32x32 tiles, every tile
// (i,j) has a statically assigned owner CTA, and CTAs synchronise through per-tile "ready"
// flags in global memory (release/acquire) instead of kernel boundaries or grid barriers —
// a left-looking tile Cholesky whose critical path is POTRF(k) -> TRSM(k+1,k) -> last SYRK
// term -> POTRF(k+1).  The first version of this file used 2 launches per panel (23 launches
// and 0.44 ms per iteration at N = 32, more than the accumulation pass it follows).
#include "m3r_gn.cuh"
#include "m3r_sim3.cuh"

namespace m3r {

constexpr int NB = kGnNB;

// ---------------------------------------------------------------------------------
// Assembly: one CTA per free pose row; incident edges visited in CSR (= fixed) order.
//   H[ii,ii] += S, H[ii,jj] -= S, H[jj,ii] -= S, H[jj,jj] += S ; b[ii] -= v, b[jj] += v
// (update_lhs/update_rhs with the reference's Hs/gs layout, gn_kernels.cu:1201-1206).
// ---------------------------------------------------------------------------------
constexpr int kAsmThreads = 128;
constexpr int kAsmChunk = 64;               // incident edges whose indices are staged per round
constexpr int kAsmSmemMax = 96 * 1024;      // row buffer: 7 x (7r+7) doubles (pose rows up to r = 249)

__global__ void __launch_bounds__(kAsmThreads)
gn_assemble_kernel(const double* __restrict__ blocks, const int* __restrict__ ie,
                   const int* __restrict__ je, const int* __restrict__ row_ptr,
                   const int* __restrict__ adj, int M, int N, int D, int Dp, double* __restrict__ A,
                   GnHeader* hdr, const int* wait_flags, int wait_world, int wait_epoch,
                   int smem_bytes) {
  extern __shared__ __align__(16) double rowbuf[];  // [7][7r+7]: this pose's rows of the lower triangle
  __shared__ int s_e[kAsmChunk], s_other[kAsmChunk];
  __shared__ unsigned char s_role[kAsmChunk];
  if (*reinterpret_cast<volatile int*>(&hdr->peer_timeout)) return;  // aborted solve: nothing runs any more
  const bool done = *reinterpret_cast<volatile int*>(&hdr->done) != 0;
  // A converged solve still WAITS for the peers' flags of this epoch (one CTA, no work after it):
  // the exchange buffers alternate by epoch parity, and "epoch e+2 may overwrite the buffer of
  // epoch e" holds only if every process has passed the wait of e+1, no-op iteration or not.
  if (done && !(wait_flags && blockIdx.x == 0)) return;
  if (wait_flags) {  // peer-memory exchange: all ranks' blocks of this iteration must have landed
    if (threadIdx.x == 0) {
      unsigned long long t0;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
      for (int r = 0; r < wait_world; r++) {
        for (;;) {
          int v;
          asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(wait_flags + r) : "memory");
          if (v - wait_epoch >= 0) break;
          unsigned long long t1;
          asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
          if (t1 - t0 > 10000000000ull) {  // 10 s: give up — zero step, solve aborted (m3r_gn_status)
            if (!done) hdr->peer_timeout = 1;
            break;
          }
          __nanosleep(64);
        }
      }
    }
    __syncthreads();
  }
  if (done) return;
  const int t = threadIdx.x;
  const int r = blockIdx.x;  // free index; pose row r+1
  if (r == N - 1) {          // augmented row diagonal + identity padding
    if (t == 0) hdr->chol_fail = 0;
    for (int row = D + t; row < Dp; row += kAsmThreads) {
      if (row > D)
        for (int c = 0; c < row; c++) A[(size_t)row * Dp + c] = 0.0;
      A[(size_t)row * Dp + row] = 1.0;
    }
    return;
  }
  // The off-diagonal blocks are read-modify-write targets (an edge and its reverse hit the same
  // block).  They are accumulated in shared memory — a global RMW is an L2 round trip per edge,
  // which made this kernel 17 us at N = 32 — and written out once, coalesced.  Rows too long
  // for the buffer (r >= 250) fall back to accumulating in global memory.
  const int ncols = 7 * r + 7;
  const bool in_smem = (size_t)7 * ncols * sizeof(double) <= (size_t)smem_bytes;
  if (in_smem) {
    for (int c = t; c < 7 * ncols; c += kAsmThreads) rowbuf[c] = 0.0;
  } else {
    for (int k = 0; k < 7; k++)
      for (int c = t; c < ncols; c += kAsmThreads) A[(size_t)(7 * r + k) * Dp + c] = 0.0;
  }
  __syncthreads();
  const int s0 = row_ptr[r + 1], s1 = row_ptr[r + 2];
  double acc = 0.0;  // t<49: diagonal block entry; 49<=t<56: b entry
  const int k = t / 7, l = t % 7;
  for (int base = s0; base < s1; base += kAsmChunk) {
    const int cnt = min(kAsmChunk, s1 - base);
    if (base > s0) __syncthreads();  // previous round consumed
    if (t < cnt) {
      const int slot = adj[base + t];
      const bool role_i = slot < M;
      const int e = role_i ? slot : slot - M;
      s_e[t] = e;
      s_role[t] = role_i ? 1 : 0;
      s_other[t] = (role_i ? je[e] : ie[e]) - 1;  // free index of the other end (-1: fixed)
    }
    __syncthreads();
    if (t < 56) {
      for (int q0 = 0; q0 < cnt; q0 += 4) {
        double v[4];
#pragma unroll
        for (int q = 0; q < 4; q++)
          v[q] = (q0 + q < cnt) ? __ldcg(blocks + (size_t)s_e[q0 + q] * kGnBlock + t) : 0.0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          if (q0 + q >= cnt) break;
          const int other = s_other[q0 + q];
          if (t < 49) {
            acc += v[q];
            if (other == r) acc -= v[q];  // self edge
            else if (other >= 0 && other < r) {
              if (in_smem) rowbuf[k * ncols + 7 * other + l] -= v[q];
              else A[(size_t)(7 * r + k) * Dp + 7 * other + l] -= v[q];
            }
          } else {
            acc += s_role[q0 + q] ? -v[q] : v[q];
          }
        }
      }
    }
  }
  if (t < 49) {
    if (l <= k) {
      if (in_smem) rowbuf[k * ncols + 7 * r + l] = acc;
      else A[(size_t)(7 * r + k) * Dp + 7 * r + l] = acc;
    }
  } else if (t < 56) {
    A[(size_t)D * Dp + 7 * r + (t - 49)] = acc;
  }
  if (in_smem) {
    __syncthreads();
    for (int kk = 0; kk < 7; kk++)
      for (int c = t; c < ncols; c += kAsmThreads) A[(size_t)(7 * r + kk) * Dp + c] = rowbuf[kk * ncols + c];
  }
}


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
// Publish: every thread's global writes so far become visible to whoever acquires *f.
__device__ __forceinline__ void post_flag(int* f, int epoch) {
  __syncthreads();
  if (threadIdx.x == 0) { __threadfence(); st_release(f, epoch); }
}

// Optional timeline of the dataflow kernel (diagnostics; set by m3r_debug_chol_trace): per
// task 4 x u64 = {kind<<32 | i<<16 | j, t_begin, t_inputs_ready, t_end} in globaltimer ns.
// Process-wide diagnostics switches; atomics so that concurrent solves (one host thread per
// device) read a consistent value each — they are test hooks, not per-solve state.
static std::atomic<unsigned long long*> g_chol_trace{nullptr};
static std::atomic<int> g_chol_trace_cap{0};
static std::atomic<int> g_chol_delay_us{0}, g_chol_delay_which{0};
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

// Tile (i,j), i >= j:  A_ij <- (A_ij - sum_{k<j} L_ik L_jk^T) [chol | * L_jj^-T], then flag it.
//
// The critical path of the factorisation is POTRF(j-1) -> TRSM(j,j-1) -> last SYRK term ->
// POTRF(j).  With the TRSM in another CTA that path crossed two flag hops and moved the 8 KB tile
// L(j,j-1) through L2 twice (store + fence + flag, acquire + reload: ~2.4 us per block column).
// The owner of a DIAGONAL tile therefore builds L(j,j-1) itself: it pre-sums the sub-diagonal
// tile's earlier terms while waiting (they are final one block column earlier), and once
// POTRF(j-1) is flagged it runs the same TRSM on its private copy and feeds the result
// straight from shared memory into the last SYRK term.  Same operations in the same order as the
// (j,j-1) task, which still publishes the tile for everybody else: identical bits.
//
// That task overwrites H(j,j-1) with L(j,j-1) IN PLACE, and the diagonal owner needs the
// original: the owner therefore raises flags[(j-1)*nb + j] (the unused upper-triangle mirror
// of the tile's own flag) once its copy is in shared memory, and the (j,j-1) task waits for it
// before it stores.  Without it the order was only a matter of timing (the owner reads ~6 us
// before the store in an undisturbed run; a CTA delayed by 15 us read the factored tile and
// solved it a second time).  Needs one task per CTA: a CTA owning both tiles would wait for
// itself.
constexpr int kOwnTrsmMaxNb = 8;
__device__ __forceinline__ bool chol_own_trsm(const CholArgs& a) {
  return a.nb <= kOwnTrsmMaxNb && a.nb * (a.nb + 1) / 2 <= (int)gridDim.x;
}
__device__ __forceinline__ void chol_tile_task(const CholArgs& a, int i, int j, int epoch,
                                               double (*sA)[NB + 1], double (*sB)[NB + 1],
                                               double (*sC)[NB + 1], double* sv, double* sred, int slot) {
  const int tid = threadIdx.x;
  chol_debug_stall(a, i, j, epoch, 0);
  const unsigned long long t_begin = a.trace ? gtimer() : 0ull;
  const int ty = tid >> 3, tx = tid & 7;  // micro-tile: rows ty+kTY*x, cols tx+8y
  double acc[kRX][4];
  // Left-looking sum, software-pipelined: the two tiles of term k+1 are fetched from L2 into
  // registers while term k is multiplied out of shared memory (N = 512: 3.31 -> 2.77 ms).
  constexpr int kLd = NB * NB / 2 / kCholThreads;  // double2 words per thread and tile
  double2 pa[kLd], pb[kLd];
  // acc = sum_{k < nterms} L(ri,k) L(rj,k)^T
  auto sum_terms = [&](int ri, int rj, int nterms) {
#pragma unroll
    for (int x = 0; x < kRX; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) acc[x][y] = 0.0;
    auto wait2 = [&](int k) {
      if (threadIdx.x == 0) {
        while (ld_acquire(a.flags + ri * a.nb + k) != epoch) __nanosleep(32);
        if (ri != rj) while (ld_acquire(a.flags + rj * a.nb + k) != epoch) __nanosleep(32);
      }
      __syncthreads();
    };
    auto fetch = [&](int k) {
      const double* srcA = a.A + (size_t)ri * NB * a.Dp + (size_t)k * NB;
      const double* srcB = a.A + (size_t)rj * NB * a.Dp + (size_t)k * NB;
#pragma unroll
      for (int q = 0; q < kLd; q++) {
        const int idx = tid + q * kCholThreads;
        const size_t off = (size_t)(idx / (NB / 2)) * a.Dp + (idx % (NB / 2)) * 2;
        pa[q] = __ldcg(reinterpret_cast<const double2*>(srcA + off));
        if (ri != rj) pb[q] = __ldcg(reinterpret_cast<const double2*>(srcB + off));
      }
    };
    if (nterms > 0) { wait2(0); fetch(0); }
    for (int k = 0; k < nterms; k++) {
#pragma unroll
      for (int q = 0; q < kLd; q++) {
        const int idx = tid + q * kCholThreads;
        const int r = idx / (NB / 2), c2 = (idx % (NB / 2)) * 2;
        sA[r][c2] = pa[q].x; sA[r][c2 + 1] = pa[q].y;
        if (ri != rj) { sB[r][c2] = pb[q].x; sB[r][c2 + 1] = pb[q].y; }
      }
      __syncthreads();
      if (k + 1 < nterms) { wait2(k + 1); fetch(k + 1); }
      const double (*sR)[NB + 1] = (ri != rj) ? sB : sA;
#pragma unroll 8
      for (int m = 0; m < NB; m++) {
        double li[kRX], lj[4];
#pragma unroll
        for (int x = 0; x < kRX; x++) li[x] = sA[ty + kTY * x][m];
#pragma unroll
        for (int y = 0; y < 4; y++) lj[y] = sR[tx + 8 * y][m];
#pragma unroll
        for (int x = 0; x < kRX; x++)
#pragma unroll
          for (int y = 0; y < 4; y++) acc[x][y] = fma(li[x], lj[y], acc[x][y]);
      }
      __syncthreads();
    }
  };
  const TilePad tA{&sA[0][0]}, tB{&sB[0][0]}, tC{&sC[0][0]};
  // diagonal owner builds L(j,j-1) privately — small systems only: the private pre-sum has j-1
  // terms, which the diagonal CTA runs back to back with its own j-1 SYRK terms; at nb = 14 that
  // already cancels the gain (calib N = 64: 1121 -> 1112 iters/s) and at nb = 112 it doubles the
  // longest tasks (N = 512: 14.6 -> 17.8 ms per iteration)
  const bool own_trsm = (i == j) && j > 0 && chol_own_trsm(a);
  if (own_trsm) {
    sum_terms(j, j - 1, j - 1);  // earlier terms of the sub-diagonal tile (j, j-1)
    const double* src = a.A + (size_t)j * NB * a.Dp + (size_t)(j - 1) * NB;
#pragma unroll
    for (int x = 0; x < kRX; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) {
        const int r = ty + kTY * x, c = tx + 8 * y;
        sC[r][c] = __ldcg(src + (size_t)r * a.Dp + c) - acc[x][y];
      }
    post_flag(a.flags + (j - 1) * a.nb + j, epoch);  // H(j,j-1) consumed: its owner may overwrite it
  }
  sum_terms(i, j, own_trsm ? j - 1 : j);
  unsigned long long t_ready = a.trace ? gtimer() : 0ull;
  if (own_trsm) {
    wait_flag(a.flags + (j - 1) * a.nb + (j - 1), epoch);  // POTRF(j-1)
    if (a.trace) t_ready = gtimer();
    load_tile(a.A, a.Dp, j - 1, j - 1, sB);
    if (tid < NB) sv[tid] = __ldcg(a.x + a.Dp + (j - 1) * NB + tid);
    __syncthreads();
    trsm_team(tC, tB, sv, tid >> 5, [] { __syncthreads(); });
#pragma unroll 8
    for (int m = 0; m < NB; m++) {  // the last SYRK term, L(j,j-1) L(j,j-1)^T, from shared memory
      double li[kRX], lj[4];
#pragma unroll
      for (int x = 0; x < kRX; x++) li[x] = sC[ty + kTY * x][m];
#pragma unroll
      for (int y = 0; y < 4; y++) lj[y] = sC[tx + 8 * y][m];
#pragma unroll
      for (int x = 0; x < kRX; x++)
#pragma unroll
        for (int y = 0; y < 4; y++) acc[x][y] = fma(li[x], lj[y], acc[x][y]);
    }
  }
  // own tile (assembled by the previous launch; nobody else touches it before the flag)
  double* dst = a.A + (size_t)i * NB * a.Dp + (size_t)j * NB;
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const int r = ty + kTY * x, c = tx + 8 * y;
      sA[r][c] = __ldcg(dst + (size_t)r * a.Dp + c) - acc[x][y];
    }
  __syncthreads();
  if (!own_trsm && a.trace) t_ready = gtimer();
  if (i == j) {
    potrf_team(tA, tid >> 5, j * NB, a.D, a.x + a.Dp + j * NB, a.hdr, [] { __syncthreads(); });
  } else {
    wait_flag(a.flags + j * a.nb + j, epoch);
    if (a.trace) t_ready = gtimer();
    load_tile(a.A, a.Dp, j, j, sB);
    if (tid < NB) sv[tid] = __ldcg(a.x + a.Dp + j * NB + tid);
    __syncthreads();
    trsm_team(tA, tB, sv, tid >> 5, [] { __syncthreads(); });
    if (i == j + 1 && chol_own_trsm(a)) wait_flag(a.flags + j * a.nb + i, epoch);  // see kOwnTrsmMaxNb
  }
  __syncthreads();
  chol_debug_stall(a, i, j, epoch, 1);
  for (int idx = tid; idx < NB * NB; idx += kCholThreads) {
    const int r = idx / NB, c = idx % NB;
    dst[(size_t)r * a.Dp + c] = sA[r][c];
  }
  post_flag(a.flags + i * a.nb + j, epoch);
  trace_put(a, slot, i == j ? 0 : 1, i, j, t_begin, t_ready);
}

// Back substitution for block column c (L^T x = y, y^T = row D of L):
//   x_c = L_cc^-T ( y_c - sum_{m>c} L_mc^T x_m ),  entries with global index >= D are 0.
__device__ __forceinline__ void chol_back_task(const CholArgs& a, int c, int epoch,
                                               double (*sA)[NB + 1], double* sx, double* sred,
                                               int slot) {
  constexpr int kParts = kCholThreads / 32, kRows = NB / kParts;
  const int tid = threadIdx.x, t = tid & 31, part = tid >> 5;
  const unsigned long long t_begin = a.trace ? gtimer() : 0ull;
  int* xflags = a.flags + a.nb * a.nb;
  wait_flag(a.flags + (a.nb - 1) * a.nb + c, epoch);  // row D of L (and, transitively, column c)
  double s = 0.0;
  for (int m = a.nb - 1; m > c; m--) {
    wait_flag(xflags + m, epoch);
    if (tid < NB) sx[tid] = __ldcg(a.x + m * NB + tid);
    __syncthreads();
    const double* Lmc = a.A + (size_t)m * NB * a.Dp + (size_t)c * NB;
#pragma unroll 4
    for (int r = part * kRows; r < part * kRows + kRows; r++)
      s = fma(__ldcg(Lmc + (size_t)r * a.Dp + t), sx[r], s);
    __syncthreads();
  }
  const unsigned long long t_ready = a.trace ? gtimer() : 0ull;
  if (part > 0) sred[(part - 1) * NB + t] = s;
  load_tile(a.A, a.Dp, c, c, sA);
  __syncthreads();
  if (tid < NB) {
    const int g = c * NB + t;
    double tot = s;
#pragma unroll
    for (int pp = 1; pp < kParts; pp++) tot += sred[(pp - 1) * NB + t];
    double rhs = (g < a.D) ? __ldcg(a.A + (size_t)a.D * a.Dp + g) - tot : 0.0;
    const double dinv = __ldcg(a.x + a.Dp + g);
    double xo = 0.0;
#pragma unroll
    for (int k = NB - 1; k >= 0; k--) {
      double xk = __shfl_sync(0xffffffffu, rhs * dinv, k);
      if (c * NB + k >= a.D) xk = 0.0;
      if (t == k) xo = xk;
      if (t < k) rhs = fma(-sA[k][t], xk, rhs);
    }
    a.x[g] = xo;
  }
  post_flag(xflags + c, epoch);
  trace_put(a, slot, 2, c, c, t_begin, t_ready);
}

// Back substitution of a small system (nb <= kBackSingleMax) by ONE CTA — the one that factored
// the last diagonal tile, so every tile is final (the caller waits for the one exception) and
// nothing is waited for inside.  The distributed
// form above costs a flag hop, an L2 round trip for x_m and a fence per block column (3.7 us
// each); here x stays in shared memory and the L tiles of a block column are fetched four at a
// time (32 independent loads per thread in flight).
constexpr int kBackSingleMax = 16;

struct BackLoads {  // global loads of one block column, issued before they are needed
  double v[4][NB / (kCholThreads / 32)];  // rows part*kRows.. of up to four L_mc tiles, column t
  double2 d[NB * NB / 2 / kCholThreads];  // this thread's share of the diagonal tile L_cc
  double rhs, dinv;
};

__device__ __forceinline__ void back_issue(const CholArgs& a, int c, int m0, bool with_diag, BackLoads& L) {
  constexpr int kRows = NB / (kCholThreads / 32);
  const int tid = threadIdx.x, t = tid & 31, part = tid >> 5;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    if (m0 + q < a.nb) {
      const double* Lmc = a.A + (size_t)(m0 + q) * NB * a.Dp + (size_t)c * NB;
#pragma unroll
      for (int rr = 0; rr < kRows; rr++) L.v[q][rr] = __ldcg(Lmc + (size_t)(part * kRows + rr) * a.Dp + t);
    }
  }
  if (with_diag) {
    const double* src = a.A + (size_t)c * NB * a.Dp + (size_t)c * NB;
#pragma unroll
    for (int q = 0; q < NB * NB / 2 / kCholThreads; q++) {
      const int idx = tid + q * kCholThreads;
      L.d[q] = __ldcg(reinterpret_cast<const double2*>(src + (size_t)(idx / (NB / 2)) * a.Dp + (idx % (NB / 2)) * 2));
    }
    const int g = c * NB + t;
    L.rhs = (tid < NB && g < a.D) ? __ldcg(a.A + (size_t)a.D * a.Dp + g) : 0.0;
    L.dinv = (tid < NB) ? __ldcg(a.x + a.Dp + g) : 0.0;
  }
}

__device__ __forceinline__ void chol_back_all(const CholArgs& a, double (*sA)[NB + 1], double* xs,
                                              double* sred, int slot) {
  constexpr int kParts = kCholThreads / 32, kRows = NB / kParts;
  const int tid = threadIdx.x, t = tid & 31, part = tid >> 5;
  BackLoads L;
  back_issue(a, a.nb - 1, a.nb, true, L);
  for (int c = a.nb - 1; c >= 0; c--) {
    const unsigned long long t_top = a.trace ? gtimer() : 0ull;
#pragma unroll
    for (int q = 0; q < NB * NB / 2 / kCholThreads; q++) {
      const int idx = tid + q * kCholThreads;
      const int r = idx / (NB / 2), c2 = (idx % (NB / 2)) * 2;
      sA[r][c2] = L.d[q].x; sA[r][c2 + 1] = L.d[q].y;
    }
    const double rhs0 = L.rhs, dinv = L.dinv;
    double s = 0.0;
    for (int m0 = c + 1; m0 < a.nb; m0 += 4) {
      if (m0 > c + 1) back_issue(a, c, m0, false, L);  // further batches of a long column
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (m0 + q < a.nb) {
#pragma unroll
          for (int rr = 0; rr < kRows; rr++) s = fma(L.v[q][rr], xs[(m0 + q) * NB + part * kRows + rr], s);
        }
      }
    }
    if (part > 0) sred[(part - 1) * NB + t] = s;
    __syncthreads();  // partial sums and the diagonal tile are in shared memory
    const unsigned long long t_mid = a.trace ? gtimer() : 0ull;
    if (c > 0) back_issue(a, c - 1, c, true, L);  // next column's loads fly during the triangular solve
    if (tid < NB) {
      const int g = c * NB + t;
      double tot = s;
#pragma unroll
      for (int pp = 1; pp < kParts; pp++) tot += sred[(pp - 1) * NB + t];
      double rhs = (g < a.D) ? rhs0 - tot : 0.0;
      double xo = 0.0;
#pragma unroll
      for (int k = NB - 1; k >= 0; k--) {
        double xk = __shfl_sync(0xffffffffu, rhs * dinv, k);
        if (c * NB + k >= a.D) xk = 0.0;
        if (t == k) xo = xk;
        if (t < k) rhs = fma(-sA[k][t], xk, rhs);
      }
      xs[g] = xo;
      a.x[g] = xo;
    }
    __syncthreads();  // x_c visible; sA / sred free for the next block column
    trace_put(a, slot + c, 2, c, c, t_top, t_mid);
  }
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

__global__ void __launch_bounds__(kCholThreads, 4)
chol_dataflow_kernel(const CholArgs a) {
  if (*reinterpret_cast<volatile int*>(&a.hdr->done)) return;
  // a peer timeout raised by THIS iteration's assembly still runs the step (as a zero step, so
  // that dx is defined and the iteration counter — the flag epoch — advances); later ones do not
  if (*reinterpret_cast<volatile int*>(&a.hdr->peer_timeout) > 1) return;
  __shared__ double sA[NB][NB + 1];
  __shared__ double sB[NB][NB + 1];
  __shared__ double sC[NB][NB + 1];
  __shared__ double sv[2 * NB], sred[(kCholThreads / 32) * NB];
  __shared__ float red[kCholThreads / 32];
  const int epoch = a.hdr->iters + 1;
  const int G = gridDim.x, cta = blockIdx.x;
  // tiles in column-major order; every dependency of a tile precedes it in this order and
  // every CTA walks its tiles in this order, so the earliest unfinished tile can always run
  int t = 0;
  for (int j = 0; j < a.nb; j++)
    for (int i = j; i < a.nb; i++, t++)
      if (t % G == cta) chol_tile_task(a, i, j, epoch, sA, sB, sC, sv, sred, t);
  if (a.nb <= kBackSingleMax) {  // small system: the owner of the last diagonal tile finishes alone
    if ((t - 1) % G == cta) {
      // every tile is final once the last diagonal tile is — except, with the private TRSM, the
      // tile left of it, which this CTA rebuilt for itself instead of waiting for its owner
      if (a.nb > 1 && chol_own_trsm(a)) wait_flag(a.flags + (a.nb - 1) * a.nb + (a.nb - 2), epoch);
      chol_back_all(a, sA, &sB[0][0], sred, t);
      gn_update_block(a.Twc, a.x, a.N, a.D, a.delta_thresh, a.dx, a.hdr, red, kCholThreads, threadIdx.x,
                      [] { __syncthreads(); });
    }
    return;
  }
  // block columns of the back substitution, last to first (all after every tile)
  for (int c = a.nb - 1; c >= 0; c--)
    if (c % G == cta) {
      chol_back_task(a, c, epoch, sA, sv, sred, t + c);
      if (c == 0) {  // x complete: this CTA finishes the iteration
        __syncthreads();
        gn_update_block(a.Twc, a.x, a.N, a.D, a.delta_thresh, a.dx, a.hdr, red, kCholThreads, threadIdx.x,
                        [] { __syncthreads(); });
      }
    }
}

// standalone retraction (m3r_pose_retr)
__global__ void pose_retr_kernel(float* __restrict__ poses, const float* __restrict__ dx,
                                 int num_poses, int num_fix) {
  const int p = num_fix + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= num_poses) return;
  float pose[8], xi[7];
#pragma unroll
  for (int m = 0; m < 8; m++) pose[m] = poses[p * 8 + m];
#pragma unroll
  for (int m = 0; m < 7; m++) xi[m] = dx[(p - num_fix) * 7 + m];
  retr_sim3_f(xi, pose);
#pragma unroll
  for (int m = 0; m < 8; m++) poses[p * 8 + m] = pose[m];
}

// Co-resident CTA capacity of the device for the dataflow kernel (its CTAs spin on each
// other's flags, so the grid must never exceed it).
static int chol_max_ctas() {
  static std::atomic<int> cache[kMaxDevices];  // per DEVICE: one process may drive several GPUs
  int dev = 0, sms = kNumSMs, per_sm = 1;
  cudaGetDevice(&dev);
  std::atomic<int>* slot = (dev >= 0 && dev < kMaxDevices) ? &cache[dev] : nullptr;
  if (slot && slot->load()) return slot->load();
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_dataflow_kernel, kCholThreads, 0);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  if (slot) slot->store(sms * per_sm);
  return sms * per_sm;
}

int gn_solve_launch(float* Twc, const double* blocks, int N, int M, float delta_thresh, float* dx,
                    void* ws, const GnLayout& L, cudaStream_t st, const int* wait_flags,
                    int wait_world, int wait_epoch) {
  GnHeader* hdr = ws_ptr<GnHeader>(ws, L.hdr);
  double* A = ws_ptr<double>(ws, L.A);
  const int D = L.D, Dp = L.Dp, nb = Dp / NB;
  size_t asm_smem = (size_t)7 * (7 * (size_t)N) * sizeof(double);  // longest row: r = N-2
  if (asm_smem > (size_t)kAsmSmemMax) asm_smem = kAsmSmemMax;
  // the > 48 KB opt-in is a per-DEVICE attribute: set it on every launch (as the accumulate
  // kernels do) rather than once per process, which broke the second GPU of a process at N > 125
  if (asm_smem > 48 * 1024)
    M3R_TRY(check_cuda(cudaFuncSetAttribute(gn_assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kAsmSmemMax)));
  gn_assemble_kernel<<<N, kAsmThreads, asm_smem, st>>>(
      blocks, ws_ptr<int>(ws, L.ie), ws_ptr<int>(ws, L.je), ws_ptr<int>(ws, L.row_ptr), ws_ptr<int>(ws, L.adj),
      M, N, D, Dp, A, hdr, wait_flags, wait_world, wait_epoch, (int)asm_smem);
  M3R_TRY(after_launch());
  CholArgs a;
  a.A = A; a.Dp = Dp; a.D = D; a.nb = nb;
  a.flags = ws_ptr<int>(ws, L.flags);
  a.x = ws_ptr<double>(ws, L.x);
  a.Twc = Twc; a.dx = dx; a.N = N; a.delta_thresh = delta_thresh; a.hdr = hdr;
  a.trace = g_chol_trace.load(); a.trace_cap = a.trace ? g_chol_trace_cap.load() : 0;
  a.delay_us = g_chol_delay_us.load(); a.delay_which = g_chol_delay_which.load();
  const int tiles = nb * (nb + 1) / 2;
  int grid = chol_max_ctas();
  if (grid > tiles) grid = tiles;
  void* args[] = {&a};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)chol_dataflow_kernel, dim3(grid),
                                                    dim3(kCholThreads), args, 0, st);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return check_cuda(e);
}

}  // namespace m3r

using namespace m3r;

extern "C" void m3r_debug_chol_trace(unsigned long long* device_buf, int capacity_tasks) {
  g_chol_trace_cap.store(device_buf ? capacity_tasks : 0);
  g_chol_trace.store(device_buf);
}

extern "C" void m3r_debug_chol_delay(int delay_us, int which) {
  g_chol_delay_us.store(delay_us > 0 ? delay_us : 0);
  g_chol_delay_which.store(which);
}

extern "C" int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix,
                             m3r_stream_t stream) {
  if (num_poses < 0 || num_fix < 0 || !poses || !dx) return M3R_EINVAL;
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<(cnt + 127) / 128, 128, 0, as_stream(stream)>>>(poses, dx, num_poses, num_fix);
  return after_launch();
}
