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
__global__ void __launch_bounds__(64)
gn_assemble_kernel(const double* __restrict__ blocks, const int* __restrict__ ie,
                   const int* __restrict__ je, const int* __restrict__ row_ptr,
                   const int* __restrict__ adj, int M, int N, int D, int Dp, double* __restrict__ A,
                   GnHeader* hdr, const int* wait_flags, int wait_world, int wait_epoch) {
  if (hdr->done) return;
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
          if (t1 - t0 > 10000000000ull) { hdr->peer_timeout = 1; break; }  // 10 s: give up, zero step
          __nanosleep(64);
        }
      }
    }
    __syncthreads();
  }
  const int t = threadIdx.x;
  const int r = blockIdx.x;  // free index; pose row r+1
  if (r == N - 1) {          // augmented row diagonal + identity padding
    if (t == 0) hdr->chol_fail = 0;
    for (int row = D + t; row < Dp; row += 64) {
      if (row > D)
        for (int c = 0; c < row; c++) A[(size_t)row * Dp + c] = 0.0;
      A[(size_t)row * Dp + row] = 1.0;
    }
    return;
  }
  // zero the lower-triangular part of this pose's 7 rows
  const int ncols = 7 * r + 7;
  for (int k = 0; k < 7; k++)
    for (int c = t; c < ncols; c += 64) A[(size_t)(7 * r + k) * Dp + c] = 0.0;
  __syncthreads();
  const int s0 = row_ptr[r + 1], s1 = row_ptr[r + 2];
  double acc = 0.0;  // t<49: diagonal block entry; 49<=t<56: b entry
  const int k = t / 7, l = t % 7;
  for (int s = s0; s < s1; s++) {
    const int slot = adj[s];
    const bool role_i = slot < M;
    const int e = role_i ? slot : slot - M;
    const int other = (role_i ? je[e] : ie[e]) - 1;  // free index of the other end (-1: fixed)
    if (t < 49) {
      const double v = __ldcg(blocks + (size_t)e * kGnBlock + t);
      acc += v;
      if (other == r) acc -= v;  // self edge
      else if (other >= 0 && other < r) A[(size_t)(7 * r + k) * Dp + 7 * other + l] -= v;
    } else if (t < 56) {
      const double v = __ldcg(blocks + (size_t)e * kGnBlock + t);
      acc += role_i ? -v : v;
    }
  }
  if (t < 49) {
    if (l <= k) A[(size_t)(7 * r + k) * Dp + 7 * r + l] = acc;
  } else if (t < 56) {
    A[(size_t)D * Dp + 7 * r + (t - 49)] = acc;
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
static unsigned long long* g_chol_trace = nullptr;
static int g_chol_trace_cap = 0;
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

// In-place Cholesky of the 32x32 tile in `s` by the whole 64-thread CTA: thread (ty,tx) of an
// 8x8 grid keeps the 4x4 micro-tile {rows ty+8x, cols tx+8y} in registers; per column step the
// raw column goes through a double-buffered shared vector (ONE __syncthreads per step), every
// thread scales the 4+4 entries it needs by rsqrt(pivot) itself, and updates its micro-tile.
// The per-step dependent chain is sts -> bar -> lds -> rsqrt -> dmul -> dfma (~170 cycles,
// profiles/micro/fp64_lat.cu) and the loop body is the same code for every column, so nothing
// is fetched cold: the earlier one-warp, fully unrolled form was 40 KB of straight-line code
// and cost 11 us per tile inside the kernel (4.7 us warm), profiles/r1j_chol_notes.md.
// Rows whose global index is >= D (augmented rhs row, padding) take pivot 1; 1/pivot is kept in
// `dinv_g` for the TRSM and the back-substitution.
// One column step of potrf_tile for a column c = 8*YC + cc whose micro-tile index YC is a
// compile-time constant: register indices are static and whole micro-tile rows/columns that lie
// left of / above the pivot drop out at compile time.
template <int YC>
__device__ __forceinline__ void potrf_step(double (&a)[kRX][4], double (*s)[NB + 1], double* colbuf,
                                           int cc, int ty, int tx, int g0, int D,
                                           double* __restrict__ dinv_g, GnHeader* hdr) {
  const int c = 8 * YC + cc;
  double* col = colbuf + (c & 1) * NB;
  if (tx == cc) {  // owners of column c publish its raw entries (rows < c: unused)
#pragma unroll
    for (int x = 0; x < kRX; x++) col[ty + kTY * x] = a[x][YC];
  }
  __syncthreads();
  const double d = col[c];
  const bool bad = !(d > 0.0);
  const bool unit = (g0 + c >= D) || bad;
  if (threadIdx.x == 0) {
    if (bad && g0 + c < D) hdr->chol_fail = 1;
  }
  const double yv = unit ? 1.0 : rsqrt(d);
  if (threadIdx.x == 0) dinv_g[g0 + c] = yv;
  double lr[kRX], lc[4];
#pragma unroll
  for (int x = 0; x < kRX; x++) lr[x] = (ty + kTY * x > c) ? col[ty + kTY * x] * yv : 0.0;
#pragma unroll
  for (int y = YC; y < 4; y++) lc[y] = col[tx + 8 * y] * yv;
  lc[YC] = (tx > cc) ? lc[YC] : 0.0;  // columns <= c of the pivot's micro-column are final
  if (tx == cc) {  // column c of L: 0 above the diagonal, the pivot, then the scaled entries
#pragma unroll
    for (int x = 0; x < kRX; x++) {
      const int r = ty + kTY * x;
      s[r][c] = (r == c) ? (unit ? 1.0 : d * yv) : lr[x];
    }
  }
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = YC; y < 4; y++) a[x][y] = fma(-lr[x], lc[y], a[x][y]);
}

// In-place Cholesky of the 32x32 tile in `s` by the whole CTA (see above).  The factor is
// written to `s` column by column; the registers only carry the trailing matrix.
__device__ __forceinline__ void potrf_tile(double (*s)[NB + 1], double* colbuf, int g0, int D,
                                           double* __restrict__ dinv_g, GnHeader* hdr) {
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  double a[kRX][4];  // the strict upper triangle is zeroed: it takes part in the (symmetric) updates
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = 0; y < 4; y++)
      a[x][y] = (tx + 8 * y <= ty + kTY * x) ? s[ty + kTY * x][tx + 8 * y] : 0.0;
  __syncthreads();  // everyone has its micro-tile before columns of `s` are overwritten
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) potrf_step<0>(a, s, colbuf, cc, ty, tx, g0, D, dinv_g, hdr);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) potrf_step<1>(a, s, colbuf, cc, ty, tx, g0, D, dinv_g, hdr);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) potrf_step<2>(a, s, colbuf, cc, ty, tx, g0, D, dinv_g, hdr);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) potrf_step<3>(a, s, colbuf, cc, ty, tx, g0, D, dinv_g, hdr);
}

template <int YC>
__device__ __forceinline__ void trsm_step(double (&a)[kRX][4], double (*sx)[NB + 1],
                                          const double (*sl)[NB + 1], double* colbuf,
                                          const double* dinv, int cc, int ty, int tx) {
  const int c = 8 * YC + cc;
  double* col = colbuf + (c & 1) * NB;
  if (tx == cc) {  // column c of X is final after scaling by 1/L[c][c]
    const double di = dinv[c];
#pragma unroll
    for (int x = 0; x < kRX; x++) {
      const double v = a[x][YC] * di;
      col[ty + kTY * x] = v;
      sx[ty + kTY * x][c] = v;
    }
  }
  __syncthreads();
  double xr[kRX], lc[4];
#pragma unroll
  for (int x = 0; x < kRX; x++) xr[x] = col[ty + kTY * x];
#pragma unroll
  for (int y = YC; y < 4; y++) lc[y] = sl[tx + 8 * y][c];
  lc[YC] = (tx > cc) ? lc[YC] : 0.0;
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = YC; y < 4; y++) a[x][y] = fma(-xr[x], lc[y], a[x][y]);
}

// X <- X * Ljj^-T for the 32x32 tile X in `sx`, Ljj in `sl`: same thread grid; per column step
// the finished column of X is broadcast through the double-buffered shared vector.
__device__ __forceinline__ void trsm_tile(double (*sx)[NB + 1], const double (*sl)[NB + 1],
                                          double* colbuf, double* dinv,
                                          const double* __restrict__ dinv_g) {
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  if (tid < NB) dinv[tid] = __ldcg(dinv_g + tid);
  double a[kRX][4];
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) a[x][y] = sx[ty + kTY * x][tx + 8 * y];
  __syncthreads();
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) trsm_step<0>(a, sx, sl, colbuf, dinv, cc, ty, tx);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) trsm_step<1>(a, sx, sl, colbuf, dinv, cc, ty, tx);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) trsm_step<2>(a, sx, sl, colbuf, dinv, cc, ty, tx);
#pragma unroll 1
  for (int cc = 0; cc < 8; cc++) trsm_step<3>(a, sx, sl, colbuf, dinv, cc, ty, tx);
}

// Tile (i,j), i >= j:  A_ij <- (A_ij - sum_{k<j} L_ik L_jk^T) [chol | * L_jj^-T], then flag it.
__device__ __forceinline__ void chol_tile_task(const CholArgs& a, int i, int j, int epoch,
                                               double (*sA)[NB + 1], double (*sB)[NB + 1],
                                               double* sv, double* sred, int slot) {
  const int tid = threadIdx.x;
  const unsigned long long t_begin = a.trace ? gtimer() : 0ull;
  const int ty = tid >> 3, tx = tid & 7;  // micro-tile: rows ty+kTY*x, cols tx+8y
  double acc[kRX][4];
#pragma unroll
  for (int x = 0; x < kRX; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) acc[x][y] = 0.0;
  for (int k = 0; k < j; k++) {
    wait_flag(a.flags + i * a.nb + k, epoch);
    if (i != j) wait_flag(a.flags + j * a.nb + k, epoch);
    load_tile(a.A, a.Dp, i, k, sA);
    if (i != j) load_tile(a.A, a.Dp, j, k, sB);
    __syncthreads();
    const double (*sR)[NB + 1] = (i != j) ? sB : sA;
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
  unsigned long long t_ready = a.trace ? gtimer() : 0ull;
  if (i == j) {
    potrf_tile(sA, sv, j * NB, a.D, a.x + a.Dp, a.hdr);
  } else {
    wait_flag(a.flags + j * a.nb + j, epoch);
    if (a.trace) t_ready = gtimer();
    load_tile(a.A, a.Dp, j, j, sB);
    __syncthreads();
    trsm_tile(sA, sB, sv, sred, a.x + a.Dp + j * NB);
  }
  __syncthreads();
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

// dx = -x (zeros on failure), retract, |dx| test  (gn_kernels.cu:1209-1222) — one CTA.
__device__ __forceinline__ void gn_update_block(float* __restrict__ Twc, const double* __restrict__ x,
                                                int N, int D, float delta_thresh,
                                                float* __restrict__ dx, GnHeader* hdr, float* red) {
  const int nthr = blockDim.x, tid = threadIdx.x;
  const bool fail = *reinterpret_cast<volatile int*>(&hdr->chol_fail) != 0 ||
                    *reinterpret_cast<volatile int*>(&hdr->peer_timeout) != 0;
  float ss = 0.f;
  for (int i = tid; i < D; i += nthr) {
    const float v = fail ? 0.f : (float)(-__ldcg(x + i));
    dx[i] = v;
    ss += v * v;
  }
  __syncthreads();  // dx visible to the retraction below (same CTA)
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
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
    for (int w = 0; w < (nthr + 31) / 32; w++) tot += red[w];
    hdr->iters += 1;
    if (sqrtf(tot) < delta_thresh) hdr->done = 1;
  }
}

__global__ void __launch_bounds__(kCholThreads)
chol_dataflow_kernel(const CholArgs a) {
  if (a.hdr->done) return;
  __shared__ double sA[NB][NB + 1];
  __shared__ double sB[NB][NB + 1];
  __shared__ double sv[2 * NB], sred[(kCholThreads / 32) * NB];
  __shared__ float red[kCholThreads / 32];
  const int epoch = a.hdr->iters + 1;
  const int G = gridDim.x, cta = blockIdx.x;
  // tiles in column-major order; every dependency of a tile precedes it in this order and
  // every CTA walks its tiles in this order, so the earliest unfinished tile can always run
  int t = 0;
  for (int j = 0; j < a.nb; j++)
    for (int i = j; i < a.nb; i++, t++)
      if (t % G == cta) chol_tile_task(a, i, j, epoch, sA, sB, sv, sred, t);
  // block columns of the back substitution, last to first (all after every tile)
  for (int c = a.nb - 1; c >= 0; c--)
    if (c % G == cta) {
      chol_back_task(a, c, epoch, sA, sv, sred, t + c);
      if (c == 0) {  // x complete: this CTA finishes the iteration
        __syncthreads();
        gn_update_block(a.Twc, a.x, a.N, a.D, a.delta_thresh, a.dx, a.hdr, red);
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
  static int cached = 0;
  if (cached) return cached;
  int dev = 0, sms = kNumSMs, per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_dataflow_kernel, kCholThreads, 0);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  cached = sms * per_sm;
  return cached;
}

int gn_solve_launch(float* Twc, const double* blocks, int N, int M, float delta_thresh, float* dx,
                    void* ws, const GnLayout& L, cudaStream_t st, const int* wait_flags,
                    int wait_world, int wait_epoch) {
  GnHeader* hdr = ws_ptr<GnHeader>(ws, L.hdr);
  double* A = ws_ptr<double>(ws, L.A);
  const int D = L.D, Dp = L.Dp, nb = Dp / NB;
  gn_assemble_kernel<<<N, 64, 0, st>>>(blocks, ws_ptr<int>(ws, L.ie), ws_ptr<int>(ws, L.je),
                                       ws_ptr<int>(ws, L.row_ptr), ws_ptr<int>(ws, L.adj), M, N, D,
                                       Dp, A, hdr, wait_flags, wait_world, wait_epoch);
  M3R_TRY(after_launch());
  CholArgs a;
  a.A = A; a.Dp = Dp; a.D = D; a.nb = nb;
  a.flags = ws_ptr<int>(ws, L.flags);
  a.x = ws_ptr<double>(ws, L.x);
  a.Twc = Twc; a.dx = dx; a.N = N; a.delta_thresh = delta_thresh; a.hdr = hdr;
  a.trace = g_chol_trace; a.trace_cap = g_chol_trace_cap;
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
  g_chol_trace = device_buf;
  g_chol_trace_cap = device_buf ? capacity_tasks : 0;
}

extern "C" int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix,
                             m3r_stream_t stream) {
  if (num_poses < 0 || num_fix < 0 || !poses || !dx) return M3R_EINVAL;
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<(cnt + 127) / 128, 128, 0, as_stream(stream)>>>(poses, dx, num_poses, num_fix);
  return after_launch();
}
