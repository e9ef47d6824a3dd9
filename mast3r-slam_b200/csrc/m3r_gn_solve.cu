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
#include "m3r_chol_tile.cuh"

namespace m3r {

std::atomic<unsigned long long*> g_chol_trace{nullptr};
std::atomic<int> g_chol_trace_cap{0};
std::atomic<int> g_chol_delay_us{0}, g_chol_delay_which{0};


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
                   int smem_bytes, int* task_queue) {
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
    if (t == 0) {
      hdr->chol_fail = 0;
      if (task_queue) *task_queue = 0;  // macro-tile factorisation: its CTAs draw tasks from 0 again
    }
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
      M, N, D, Dp, A, hdr, wait_flags, wait_world, wait_epoch, (int)asm_smem,
      L.nbm ? ws_ptr<int>(ws, L.flags) + 2 * L.nbm * L.nbm + 2 * L.nbm : nullptr);
  M3R_TRY(after_launch());
  if (L.nbm) return gn_chol64_launch(Twc, N, delta_thresh, dx, ws, L, st);  // large system: 64x64 macro tiles
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
  M3R_NVTX("m3r_pose_retr");
  if (num_poses < 0 || num_fix < 0 || !poses || !dx) return M3R_EINVAL;
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<(cnt + 127) / 128, 128, 0, as_stream(stream)>>>(poses, dx, num_poses, num_fix);
  return after_launch();
}
