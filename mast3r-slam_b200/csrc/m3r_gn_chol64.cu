// m3r_gn_chol64.cu — fp64 Cholesky solve of normal-equation systems of more than 5 panels
// of 32 (more than 23 keyframes: BASELINE configs 3, 4 and 5 — N = 32, 64, 512), sm_100a.
//
// Replaces, like m3r_gn_solve.cu for small systems, SparseBlock::solve of the reference
// (gn_kernels.cu:132-153: Eigen::SimplicialLLT in fp64 on the host) plus the retraction and the
// convergence test that follow it (gn_kernels.cu:1209-1222).  Same augmented matrix [H b; b^T 1],
// same failure semantics (non-positive pivot => zero step).
//
// Why a second kernel.  The 32x32 dataflow kernel (m3r_gn_solve.cu, still used up to 5 panels, where
// it is the faster one) ran N = 512 at 5.4 TFLOP/s: (i) every tile
// product pulled two 8 KB tiles from L2 for 32 K fma (3.8 GB per factorisation) and multiplied
// them with DFMA from shared memory at 6 loads per 8 fma; (ii) a left-looking tile is summed by
// ONE CTA, so the diagonal tile of a late block column (110 products) arrived long after the
// column before it was finished; (iii) per 32 columns the critical path crossed two flag hops.
// Here:
//   * tiles are 64x64, products run on the fp64 tensor pipe (mma.sync.m8n8k4.f64, DMMA): a CTA
//     of 8 warps (2x4, 32x16 outputs each) streams 64x32 operand panels through a cp.async
//     ring of three products (two 32-column stages each); 6 shared-memory loads feed 8 DMMA (64 fma per thread) and a product moves
//     64 KB from L2 for 262 K fma (4x less traffic per flop);
//   * the sum of a tile over block columns k is cut into CHUNKS (a tail of 2-4 block columns
//     and one chunk of 4 near the chain, chunks of 16 far from it; layout in m3r_gn.cuh).  A chunk
//     is a task of its own: any CTA computes it and subtracts it from the tile in place, in
//     chunk order (a per-tile progress counter), so the bits do not depend on who ran what.
//     Tasks sit in one statically sorted list; CTAs take the next entry with an atomic counter.
//     Sort key = max(last block column the chunk needs + 1, J - window): a tile's old chunks are
//     deferred until the chain is `window` block columns away, new chunks run as soon as the
//     chain has produced their inputs, and every dependency of a task precedes it in the list —
//     so blocking waits inside tasks cannot deadlock (the earliest unfinished task can always run);
//   * the critical path — POTRF(J-1) -> L(J,J-1) = S L^-T -> last SYRK term -> POTRF(J) — is
//     ONE dedicated CTA (block 0) that never leaves its SM: L(J-1,J-1) stays in shared memory,
//     the TRSM of the sub-diagonal tile is private, and inside a 64-wide diagonal block the two
//     32-wide POTRFs, the TRSM and the update between them need no flag at all.  The last terms
//     of the two tiles the chain needs next are formed by two SIDEKICK CTAs (blocks 1, 2) on SMs
//     of their own; the back substitution is block 0 again, with one helper CTA per block column.
// Measured (profiles/r3/): N = 512 solve step 2.75 -> 1.17 ms, N = 64 228 -> 150 us, N = 32 105 -> 91 us.
// Summation order differs from the 32x32 kernel (and from Eigen): results agree to fp64
// rounding, not bit for bit; they ARE bit-identical from run to run and between a single-GPU and
// an edge-sharded solve (same kernel, same inputs, fixed chunk order).
#include "m3r_chol_tile.cuh"

namespace m3r {
namespace c64 {

constexpr int MB = kGnMB;          // macro tile
constexpr int kThreads = 256;      // 8 warps: 2 (rows of 32) x 4 (columns of 16)
constexpr int LDP = 34;            // row stride of a 64x32 operand panel (16-B aligned rows for cp.async,
                                   // == 2 mod 16 for the conflict-free fragment pattern below)
constexpr int LDC = 65;            // row stride of a 64x64 chain tile (odd: lane-per-row access is conflict free)
constexpr int kStages = 6;  // three products of two 32-column stages each
constexpr int kPanel = MB * LDP;
constexpr int kStageD = 2 * kPanel;
constexpr int kTileC = MB * LDC;
constexpr int kWindow = 8;         // block columns ahead of the chain whose tiles are kept up to date
constexpr int kMaxSlots = 4096;    // (nbm + 1) * (kWindow + 1) sort groups of the task list
constexpr int kChainTiles = 5;  // T, X, Lp, Tn, Xn of the chain CTA
constexpr size_t kSmemDoubles =
    (size_t)kStages * kStageD > (size_t)kChainTiles * kTileC ? (size_t)kStages * kStageD : (size_t)kChainTiles * kTileC;
constexpr size_t kSmemBytesFact = kSmemDoubles * sizeof(double);  // factorisation; the back substitution may need more (below)

enum { kTaskUpdate = 0, kTaskFinal = 1 };

struct Args {
  CholArgs c;          // A, Dp, D, nb (= Dp / 32), flags, x, Twc, dx, N, delta_thresh, hdr, diagnostics
  const int4* tasks;   // {I | J << 16, k0 | k1 << 16, kind, chunk index}
  int ntasks, nbm;
};

// flag array (GnLayout::flags, (2 nbm)^2 + 2 nbm ints): tile ready | chunk progress | x ready | queue
__device__ __forceinline__ int* tile_flag(const Args& a, int I, int J) { return a.c.flags + I * a.nbm + J; }
__device__ __forceinline__ int* prog_flag(const Args& a, int I, int J) { return a.c.flags + a.nbm * a.nbm + I * a.nbm + J; }
__device__ __forceinline__ int* x_flag(const Args& a, int c32) { return a.c.flags + 2 * a.nbm * a.nbm + c32; }
__host__ __device__ inline int queue_index(int nbm) { return 2 * nbm * nbm + 2 * nbm; }

// ------------------------------------------------------------------------------------------
// Task list (once per solve).  Group (v, J): the tasks of tile column J whose key is v, i.e.
// chunks ending at block column k1 == v, plus — when the column enters the window, J - W == v —
// all its older chunks (k1 < v).  Groups ascend in (v, J); inside a group (k1, I) ascend.
// ------------------------------------------------------------------------------------------
template <class F>
__host__ __device__ __forceinline__ void tasks_ending_at(int nbm, int J, int e, F&& f) {
  if (e > J) return;
  // chain tiles (their last terms are the chain's own)
  if (gn_chunk_end(J - 2, kGnTailChain, e))
    f(J, gn_chunk_k0(J - 2, kGnTailChain, e), e, (int)kTaskUpdate, gn_chunk_index(J - 2, kGnTailChain, e));
  if (J + 1 < nbm && gn_chunk_end(J - 1, kGnTailChain, e))
    f(J + 1, gn_chunk_k0(J - 1, kGnTailChain, e), e, (int)kTaskUpdate, gn_chunk_index(J - 1, kGnTailChain, e));
  // the other tiles of the column
  if (gn_chunk_end(J - 1, kGnKC, e)) {
    const int k0 = gn_chunk_k0(J - 1, kGnKC, e), kc = gn_chunk_index(J - 1, kGnKC, e);
    for (int I = J + 2; I < nbm; I++) f(I, k0, e, (int)kTaskUpdate, kc);
  }
  if (e == J) {  // the last term and the TRSM
    const int kc = gn_nch(J - 1, kGnKC), k0 = J > 0 ? J - 1 : 0;
    for (int I = J + 2; I < nbm; I++) f(I, k0, J, (int)kTaskFinal, kc);
  }
}

template <class F>
__host__ __device__ __forceinline__ void tasks_of_group(int nbm, int v, int J, F&& f) {
  if (J - kWindow == v) {
    for (int e = 0; e <= v; e++) tasks_ending_at(nbm, J, e, f);
  } else {
    tasks_ending_at(nbm, J, v, f);
  }
}

struct TaskCount {
  int n;
  __host__ __device__ void operator()(int, int, int, int, int) { n++; }
};
struct TaskEmit {
  int4* tasks; int ntasks, J, o;
  __host__ __device__ void operator()(int I, int k0, int k1, int kind, int kc) {
    if (o < ntasks) tasks[o] = make_int4(I | (J << 16), k0 | (k1 << 16), kind, kc);
    o++;
  }
};

__global__ void __launch_bounds__(1024)
chol64_tasklist_kernel(int nbm, int ntasks, int4* __restrict__ tasks) {
  __shared__ int offs[kMaxSlots];
  const int nslots = (nbm + 1) * (kWindow + 1);
  for (int s = threadIdx.x; s < nslots; s += blockDim.x) {
    const int v = s / (kWindow + 1), J = v + s % (kWindow + 1);
    TaskCount cnt{0};
    if (J < nbm) tasks_of_group(nbm, v, J, cnt);
    offs[s] = cnt.n;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int s = 0; s < nslots; s++) { const int c = offs[s]; offs[s] = run; run += c; }
    if (run != ntasks) asm volatile("trap;");  // host and device disagree on the task count
  }
  __syncthreads();
  for (int s = threadIdx.x; s < nslots; s += blockDim.x) {
    const int v = s / (kWindow + 1), J = v + s % (kWindow + 1);
    if (J >= nbm) continue;
    TaskEmit emit{tasks, ntasks, J, offs[s]};
    tasks_of_group(nbm, v, J, emit);
  }
}

// ------------------------------------------------------------------------------------------
// fp64 tensor-pipe products from shared memory
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// Which of the 16 k's of a block the lane with quad index q feeds into the kk-th of its four
// DMMAs.  The sum over k does not care about the order, the banks do: a fragment load reads, per
// half warp, rows g = 0..3 at k(q), q = 0..3.  Odd row stride: k = 4q + kk puts the 16 words on 16
// different bank pairs; stride == 2 (mod 16): k = 2kk + (q & 1) + 8 (q >> 1) does.
template <int LD>
__device__ __forceinline__ int frag_k(int q, int kk) {
  if constexpr (LD % 2 == 1) return 4 * q + kk;
  else return 2 * kk + (q & 1) + 8 * (q >> 1);
}

// acc[mi][ni] += A[mi*8 + g][k] * B[ni*8 + g][k] summed over 16 * k16 values of k (C = A B^T):
// acc[mi][ni][e] is C[mi*8 + g][ni*8 + 2q + e], g = lane >> 2, q = lane & 3.
template <int MI, int NI, int LD>
__device__ __forceinline__ void warp_mma(const double* __restrict__ A, const double* __restrict__ B, int k16,
                                         double (&acc)[MI][NI][2]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const double* pa = A + g * LD;
  const double* pb = B + g * LD;
  for (int kb = 0; kb < k16; kb++) {
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
      const int k = kb * 16 + frag_k<LD>(q, kk);
      double av[MI], bv[NI];
#pragma unroll
      for (int mi = 0; mi < MI; mi++) av[mi] = pa[mi * 8 * LD + k];
#pragma unroll
      for (int ni = 0; ni < NI; ni++) bv[ni] = pb[ni * 8 * LD + k];
#pragma unroll
      for (int mi = 0; mi < MI; mi++)
#pragma unroll
        for (int ni = 0; ni < NI; ni++) dmma(acc[mi][ni], av[mi], bv[ni]);
    }
  }
}

template <int MI, int NI>
__device__ __forceinline__ void zero_acc(double (&acc)[MI][NI][2]) {
#pragma unroll
  for (int mi = 0; mi < MI; mi++)
#pragma unroll
    for (int ni = 0; ni < NI; ni++) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
}

// C (shared, stride LDC) -= acc for the warp's MI x NI fragments starting at (r0, c0)
template <int MI, int NI>
__device__ __forceinline__ void sub_acc_smem(double* __restrict__ C, int r0, int c0, const double (&acc)[MI][NI][2]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
#pragma unroll
  for (int mi = 0; mi < MI; mi++)
#pragma unroll
    for (int ni = 0; ni < NI; ni++) {
      double* p = C + (r0 + mi * 8 + g) * LDC + c0 + ni * 8 + 2 * q;
      p[0] -= acc[mi][ni][0];
      p[1] -= acc[mi][ni][1];
    }
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// team barrier: exactly the 128 threads of team `t` (warps 4t..4t+3)
__device__ __forceinline__ void team_sync(int t) { asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory"); }
// one-way barrier over all 256 threads: the arriving team does not wait for the waiting one
__device__ __forceinline__ void cta_arrive(int id) { asm volatile("bar.arrive %0, 256;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void cta_wait(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }

template <int LD>
struct TileLd {
  double* p;
  __device__ __forceinline__ double& operator()(int r, int c) const { return p[r * LD + c]; }
};
using TileC = TileLd<LDC>;

// global 64x64 tile (bi, bj) <-> shared [64][LDC]; L2-coherent (tiles are written by other SMs)
__device__ __forceinline__ void load_tile64(const double* __restrict__ A, int Dp, int bi, int bj, double* s) {
  const double* src = A + (size_t)bi * MB * Dp + (size_t)bj * MB;
  for (int idx = threadIdx.x; idx < MB * MB / 2; idx += kThreads) {
    const int r = idx >> 5, c2 = (idx & 31) * 2;
    const double2 v = __ldcg(reinterpret_cast<const double2*>(src + (size_t)r * Dp + c2));
    s[r * LDC + c2] = v.x; s[r * LDC + c2 + 1] = v.y;
  }
}
__device__ __forceinline__ void store_tile64(double* __restrict__ A, int Dp, int bi, int bj, const double* s) {
  double* dst = A + (size_t)bi * MB * Dp + (size_t)bj * MB;
  for (int idx = threadIdx.x; idx < MB * MB; idx += kThreads) {
    const int r = idx >> 6, c = idx & 63;
    __stcg(dst + (size_t)r * Dp + c, s[r * LDC + c]);
  }
}

// X <- X L^-T for the 64x64 tiles X, L (lower) in shared memory, dinv = 1 / diag(L):
//   [Xa Xb] [L00^T L10^T; 0 L11^T] = [Sa Sb]:  Xa = Sa L00^-T ; Xb = (Sb - Xa L10^T) L11^-T.
// Two 4-warp teams, 32 rows each, run the 32-wide solves side by side; the middle product is DMMA.
__device__ __forceinline__ void trsm64(double* X, double* L, const double* dinv) {
  const int warp = threadIdx.x >> 5, team = warp >> 2, w = warp & 3;
  trsm_team(TileC{X + team * 32 * LDC}, TileC{L}, dinv, w, [team] { team_sync(team); });
  __syncthreads();
  {
    const int wr = warp >> 2, wc = warp & 3;  // outputs: rows wr*32.., columns 32 + wc*8..
    double acc[4][1][2];
    zero_acc(acc);
    warp_mma<4, 1, LDC>(X + wr * 32 * LDC, L + (32 + wc * 8) * LDC, 2, acc);
    sub_acc_smem(X, wr * 32, 32 + wc * 8, acc);
  }
  __syncthreads();
  trsm_team(TileC{X + team * 32 * LDC + 32}, TileC{L + 32 * LDC + 32}, dinv + 32, w, [team] { team_sync(team); });
  __syncthreads();
}

// acc[mi][ni] += A[mi*8 + g][k] * Bt[k][ni*8 + g]: as warp_mma with the second operand stored transposed
// (rows = k).  Conflict free for odd row strides: a half warp reads k = 4q + kk, column g.
template <int MI, int NI, int LD>
__device__ __forceinline__ void warp_mma_bt(const double* __restrict__ A, const double* __restrict__ Bt, int k16,
                                            double (&acc)[MI][NI][2]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const double* pa = A + g * LD;
  for (int kb = 0; kb < k16; kb++) {
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
      const int k = kb * 16 + frag_k<LD>(q, kk);
      double av[MI], bv[NI];
#pragma unroll
      for (int mi = 0; mi < MI; mi++) av[mi] = pa[mi * 8 * LD + k];
#pragma unroll
      for (int ni = 0; ni < NI; ni++) bv[ni] = Bt[k * LD + ni * 8 + g];
#pragma unroll
      for (int mi = 0; mi < MI; mi++)
#pragma unroll
        for (int ni = 0; ni < NI; ni++) dmma(acc[mi][ni], av[mi], bv[ni]);
    }
  }
}

// trsm64 with the first half through the INVERSE of L00: U = L00^-T (upper triangular, [32][LDC], prepared by
// the idle team during the previous chain step), so  Xa = Sa U  is a DMMA product (0.3 us) instead of a 32-step
// dependent solve (1.3 us); the second half keeps the solve, L11 being only just finished.
__device__ __forceinline__ void trsm64w(double* X, double* L, const double* dinv, const double* U) {
  const int warp = threadIdx.x >> 5, team = warp >> 2, w = warp & 3, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const int wr = warp >> 2, wc = warp & 3;
  {
    double acc[4][1][2];
    zero_acc(acc);
    warp_mma_bt<4, 1, LDC>(X + wr * 32 * LDC, U + wc * 8, 2, acc);  // rows wr*32.., columns wc*8..
    __syncthreads();  // every warp has read Sa
#pragma unroll
    for (int mi = 0; mi < 4; mi++) {
      double* p = X + (wr * 32 + mi * 8 + g) * LDC + wc * 8 + 2 * q;
      p[0] = acc[mi][0][0]; p[1] = acc[mi][0][1];
    }
  }
  __syncthreads();
  {
    double acc[4][1][2];
    zero_acc(acc);
    warp_mma<4, 1, LDC>(X + wr * 32 * LDC, L + (32 + wc * 8) * LDC, 2, acc);
    sub_acc_smem(X, wr * 32, 32 + wc * 8, acc);
  }
  __syncthreads();
  trsm_team(TileC{X + team * 32 * LDC + 32}, TileC{L + 32 * LDC + 32}, dinv + 32, w, [team] { team_sync(team); });
  __syncthreads();
}

// POTRF of the 32x32 block at the top of a 64x32 panel AND the TRSM of the 32 rows below it, in
// one pass of potrf_team's loop: lane r owns rows r and 32 + r.  Row 32 + r goes through the same
// 4-column steps as a row below the current diagonal block (solve against the 4x4 block, rank-4
// update of the columns to the right), so the panel below is finished when the factor is — the
// separate T10 <- T10 L00^-T (1.4 us per block column on the critical path) disappears.  Same
// operation order per element as potrf_team / trsm_team.
template <class Tile>
__device__ __forceinline__ void rank4_rows2(const Tile s, int r, int p0, int jj0, const double (&x)[4],
                                            const double (&x2)[4], double (&t)[4], double (&t2)[4]) {
  double v[4][4];
#pragma unroll
  for (int u = 0; u < 4; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) v[u][k] = s(jj0 + u, p0 + k);
    t[u] = s(r, jj0 + u);
    t2[u] = s(32 + r, jj0 + u);
  }
#pragma unroll
  for (int u = 0; u < 4; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) { t[u] = fma(-x[k], v[u][k], t[u]); t2[u] = fma(-x2[k], v[u][k], t2[u]); }
  }
#pragma unroll
  for (int u = 0; u < 4; u++) {
    if (jj0 + u <= r) s(r, jj0 + u) = t[u];
    s(32 + r, jj0 + u) = t2[u];
  }
}

template <class Tile, class Sync>
__device__ __forceinline__ void potrf_tall_team(const Tile s, int w, int g0, int D, double* __restrict__ dinv,
                                                GnHeader* hdr, Sync sync) {
  const int r = threadIdx.x & 31;
  double x[4] = {0.0, 0.0, 0.0, 0.0}, cur[4] = {0.0, 0.0, 0.0, 0.0};      // row r
  double x2[4] = {0.0, 0.0, 0.0, 0.0}, cur2[4] = {0.0, 0.0, 0.0, 0.0};    // row 32 + r
  if (w == 0) {
#pragma unroll
    for (int j = 0; j < 4; j++) { cur[j] = s(r, j); cur2[j] = s(32 + r, j); }
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
      const int i = r - c0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        double v = cur[j], v2 = cur2[j];
#pragma unroll
        for (int k = 0; k < j; k++) { v = fma(-x[k], b[j][k], v); v2 = fma(-x2[k], b[j][k], v2); }
        x[j] = v * rr[j];
        x2[j] = v2 * rr[j];
        if (i == j) x[j] = b[j][j];
        if (i < j) x[j] = 0.0;
      }
      if (i >= 0) {
#pragma unroll
        for (int j = 0; j < 4; j++) s(r, c0 + j) = x[j];
      }
#pragma unroll
      for (int j = 0; j < 4; j++) s(32 + r, c0 + j) = x2[j];
    }
    sync();
    if (w == 0) {
      if (c0 + 4 < NB) rank4_rows2(s, r, c0, c0 + 4, x, x2, cur, cur2);
      __syncwarp();
    } else {
      double xw[4], xw2[4], t[4], t2[4];
#pragma unroll
      for (int k = 0; k < 4; k++) { xw[k] = s(r, c0 + k); xw2[k] = s(32 + r, c0 + k); }
#pragma unroll 1
      for (int jj = c0 + 8 + 4 * (w - 1); jj < NB; jj += 12) rank4_rows2(s, r, c0, jj, xw, xw2, t, t2);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Bulk tasks: one chunk of a tile's left-looking sum, and (kTaskFinal) the tile's TRSM
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void issue_stage(const Args& a, int I, int J, int kcol, double* stage) {
  const double* srcA = a.c.A + (size_t)I * MB * a.c.Dp + kcol;
  const double* srcB = a.c.A + (size_t)J * MB * a.c.Dp + kcol;
  for (int ch = threadIdx.x; ch < MB * 16; ch += kThreads) {
    const int r = ch >> 4, seg = (ch & 15) * 2;
    cp_async16(stage + r * LDP + seg, srcA + (size_t)r * a.c.Dp + seg);
    if (I != J) cp_async16(stage + kPanel + r * LDP + seg, srcB + (size_t)r * a.c.Dp + seg);
  }
}

// publish by the LAST warp: its ~1 us fence then overlaps the next task's queue pop and flag polls,
// which are warp 0's
__device__ __forceinline__ void post_flag_last(int* f, int value) {
  __syncthreads();
  if (threadIdx.x == kThreads - 1) { st_release(f, value); }
}

__device__ __forceinline__ void run_task(const Args& a, const int4 tk, int epoch, double* smem, double* s_dinv,
                                         int slot, int* s_task, int* s_early) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  const int wr = warp >> 2, wc = warp & 3;
  const int I = tk.x & 0xffff, J = tk.x >> 16, k0 = tk.y & 0xffff, k1 = tk.y >> 16, kind = tk.z, kc = tk.w;
  chol_debug_stall(a.c, I, J, epoch, 0);
  const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
  // (Drawing the NEXT list entry here, to hide the atomic's round trip, was tried: an urgent entry
  // then sits behind its holder's current task, up to 48 us — chain stalls of 50-70 us.)
  // inputs: L(I,k) and L(J,k) for every block column of the chunk
  if (tid < 2 * (k1 - k0)) {
    const int* f = tile_flag(a, (tid & 1) ? J : I, k0 + (tid >> 1));
    while (ld_acquire(f) != epoch) __nanosleep(32);
  }
  // Is it already this chunk's turn at the tile (the earlier chunks subtracted)?  Usually yes: then the tile's
  // current value is fetched NOW and rides along in registers instead of costing an L2 round trip at the end.
  if (tid == 64) *s_early = (kc == 0 || ld_acquire(prog_flag(a, I, J)) == ((epoch << 12) | kc)) ? 1 : 0;
  __syncthreads();
  const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
  const bool early = *s_early != 0;
  double* tile = a.c.A + (size_t)I * MB * a.c.Dp + (size_t)J * MB;
  double2 base[4][2];
  auto load_base = [&]() {
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int ni = 0; ni < 2; ni++)
        base[mi][ni] = __ldcg(reinterpret_cast<const double2*>(tile + (size_t)(wr * 32 + mi * 8 + g) * a.c.Dp + wc * 16 + ni * 8 + 2 * q));
  };
  if (early) load_base();
  double acc[4][2][2];
  zero_acc(acc);
  // One product (64 block columns = two 32-column stages) per pipeline step, three steps in the ring: one
  // CTA barrier and one round of copy issue per 262 K fma instead of two.
  const int npairs = k1 - k0;
  auto issue_pair = [&](int pq) {
    double* slot = smem + (pq % 3) * 2 * kStageD;
    issue_stage(a, I, J, (k0 + pq) * MB, slot);
    issue_stage(a, I, J, (k0 + pq) * MB + 32, slot + kStageD);
  };
  for (int pq = 0; pq < 2; pq++) {
    if (pq < npairs) issue_pair(pq);
    cp_async_commit();
  }
  for (int pq = 0; pq < npairs; pq++) {
    cp_async_wait<1>();
    __syncthreads();  // product pq has landed for everybody; the slot of product pq-1 is no longer read
    if (pq + 2 < npairs) issue_pair(pq + 2);
    cp_async_commit();
    const double* st = smem + (pq % 3) * 2 * kStageD;
    warp_mma<4, 2, LDP>(st + wr * 32 * LDP, (I != J ? st + kPanel : st) + wc * 16 * LDP, 2, acc);
    warp_mma<4, 2, LDP>(st + kStageD + wr * 32 * LDP, (I != J ? st + kStageD + kPanel : st + kStageD) + wc * 16 * LDP, 2, acc);
  }
  cp_async_wait<0>();
  if (!early) {
    wait_flag(prog_flag(a, I, J), (epoch << 12) | kc);  // earlier chunks are in (fixed order)
    load_base();
  } else {
    __syncthreads();
  }
  if (kind == kTaskUpdate) {
    if (!(I == J && wc * 16 > wr * 32 + 31)) {  // fragments strictly above the diagonal are never read
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++) {
          double2* p = reinterpret_cast<double2*>(tile + (size_t)(wr * 32 + mi * 8 + g) * a.c.Dp + wc * 16 + ni * 8 + 2 * q);
          double2 v = base[mi][ni];
          v.x -= acc[mi][ni][0]; v.y -= acc[mi][ni][1];
          __stcg(p, v);
        }
    }
    chol_debug_stall(a.c, I, J, epoch, 1);
    if (tid == 0) *s_task = atomicAdd(a.c.flags + queue_index(a.nbm), 1);
    post_flag_last(prog_flag(a, I, J), (epoch << 12) | (kc + 1));
    trace_put(a.c, slot, 3, I, J, t_begin, t_ready);
    return;
  }
  // final: T = tile - last chunk, then L(I,J) = T L(J,J)^-T
  double* X = smem;
  double* Ls = smem + kTileC;
#pragma unroll
  for (int mi = 0; mi < 4; mi++)
#pragma unroll
    for (int ni = 0; ni < 2; ni++) {
      const int r = wr * 32 + mi * 8 + g, c = wc * 16 + ni * 8 + 2 * q;
      const double2 v = base[mi][ni];
      X[r * LDC + c] = v.x - acc[mi][ni][0];
      X[r * LDC + c + 1] = v.y - acc[mi][ni][1];
    }
  wait_flag(tile_flag(a, J, J), epoch);
  const unsigned long long t_diag = a.c.trace ? gtimer() : 0ull;
  load_tile64(a.c.A, a.c.Dp, J, J, Ls);
  if (tid < MB) s_dinv[tid] = __ldcg(a.c.x + a.c.Dp + J * MB + tid);
  __syncthreads();
  trsm64(X, Ls, s_dinv);
  chol_debug_stall(a.c, I, J, epoch, 1);
  store_tile64(a.c.A, a.c.Dp, I, J, X);
  if (tid == 0) *s_task = atomicAdd(a.c.flags + queue_index(a.nbm), 1);
  post_flag_last(tile_flag(a, I, J), epoch);
  trace_put(a.c, slot, 1, I, J, t_begin, t_diag);
}

// ------------------------------------------------------------------------------------------
// Sidekicks: the LAST terms of the two tiles the chain needs next, each on an SM of its own.
// While the chain CTA is in step J (factorising tile (J,J)), block 1 forms
//   tile (J+1,J)   -= L(J+1,J-1) L(J,J-1)^T      (k = J-1, the sub-diagonal tile's last term)
// and block 2
//   tile (J+1,J+1) -= L(J+1,J-1) L(J+1,J-1)^T    (k = J-1 of the next diagonal tile; k = J is the chain's own)
// in place, as one more "chunk" of the tile's progress counter.  Both products need L(J+1,J-1) (a list
// task's TRSM, finished early in step J), block 1 also L(J,J-1), which the chain publishes at the start of
// step J.  These two products used to run on the chain CTA's idle team; their DMMA streams shared the SM's
// four fp64 pipes with the POTRFs (4.3 -> 5.1 us each, and the products themselves took 7.5 us on three
// warps); as list tasks they sat behind the queue.  Here they take ~2 us on an otherwise idle SM.
// ------------------------------------------------------------------------------------------
constexpr int LDS2 = 66;  // row stride of the sidekicks' 64x64 operands: 16-byte rows, == 2 (mod 16)
__device__ __forceinline__ void run_sidekick(const Args& a, int which, int epoch, double* smem) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  double* Ys = smem;
  double* Xs = which == 0 ? smem + MB * LDS2 : Ys;
  // one 32x16 block of outputs per warp: all 8 for tile (J+1,J); the 6 lower ones for the diagonal tile
  int r0, c0;
  bool active = true;
  if (which == 0) { r0 = (warp >> 2) * 32; c0 = (warp & 3) * 16; }
  else { r0 = warp < 2 ? 0 : 32; c0 = warp < 2 ? warp * 16 : (warp - 2) * 16; active = warp < 6; }
  for (int J = 1; J + 1 < a.nbm; J++) {
    const int bi = J + 1, bj = which == 0 ? J : J + 1;
    const int nbulk = which == 0 ? gn_nch_sub(J) : gn_nch_diag(J + 1);
    chol_debug_stall(a.c, bi, bj, epoch, 0);
    if (tid == 0) while (ld_acquire(tile_flag(a, J + 1, J - 1)) != epoch) __nanosleep(20);
    if (tid == 32 && nbulk) while (ld_acquire(prog_flag(a, bi, bj)) != ((epoch << 12) | nbulk)) __nanosleep(20);
    if (tid == 64 && which == 0) while (ld_acquire(tile_flag(a, J, J - 1)) != epoch) { }
    __syncthreads();
    for (int ch = tid; ch < MB * 32; ch += kThreads) {
      const int r = ch >> 5, seg = (ch & 31) * 2;
      cp_async16(Ys + r * LDS2 + seg, a.c.A + (size_t)((J + 1) * MB + r) * a.c.Dp + (size_t)(J - 1) * MB + seg);
      if (which == 0) cp_async16(Xs + r * LDS2 + seg, a.c.A + (size_t)(J * MB + r) * a.c.Dp + (size_t)(J - 1) * MB + seg);
    }
    cp_async_commit();
    double* tile = a.c.A + (size_t)bi * MB * a.c.Dp + (size_t)bj * MB;
    double2 base[4][2];
    if (active) {
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
          base[mi][ni] = __ldcg(reinterpret_cast<const double2*>(tile + (size_t)(r0 + mi * 8 + g) * a.c.Dp + c0 + ni * 8 + 2 * q));
    }
    cp_async_wait<0>();
    __syncthreads();
    if (active) {
      double acc[4][2][2];
      zero_acc(acc);
      warp_mma<4, 2, LDS2>(Ys + r0 * LDS2, Xs + c0 * LDS2, 4, acc);
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++) {
          double2 v = base[mi][ni];
          v.x -= acc[mi][ni][0]; v.y -= acc[mi][ni][1];
          __stcg(reinterpret_cast<double2*>(tile + (size_t)(r0 + mi * 8 + g) * a.c.Dp + c0 + ni * 8 + 2 * q), v);
        }
    }
    chol_debug_stall(a.c, bi, bj, epoch, 1);
    post_flag_last(prog_flag(a, bi, bj), (epoch << 12) | (nbulk + 1));  // (its barrier also frees the operand buffers)
  }
}

// ------------------------------------------------------------------------------------------
// The chain: diagonal tiles and the sub-diagonal tile of every block column, one CTA.
//
// One step, J >= 1.  In shared memory: T = tile (J,J) with all terms but k = J-1 applied, X = tile (J,J-1)
// with ALL its terms applied (list chunks plus the sidekicks' last term), Lp = L(J-1,J-1) from the previous step.
//   all 8 warps   X <- X Lp^-T                      (trsm64: X is L(J,J-1) now), stored to the factor
//   warps 0-6     T00 -= X0 X0^T, T10 -= X1 X0^T    | warp 7  fence + flag of L(J,J-1)
//   team 0 (warps 0-3)                               | team 1 (warps 4-7)
//     tall POTRF: L00 and T10 <- T10 L00^-T          |   T11 -= X1 X1^T
//     ---- one-way barrier: waits for team 1 ----    |   ---- arrives, does not wait ----
//     T11 -= T10 T10^T                               |   waits for the sidekicks' chunks of tiles (J+1,J+1), (J+1,J)
//     POTRF(T11), publish L(J,J)                     |   fetches them (cp.async) as Tn, Xn for the next step
// Publishing (32 KB store, fence, flag) and the loads of the next step's tiles run beside the two POTRFs, which
// keep only four warps busy; the DMMA products for the next tiles' last terms run on two other SMs (above).
// ------------------------------------------------------------------------------------------
// the calling team stores the shared tile to global tile (bi,bj) and raises `flag`
__device__ __forceinline__ void team_publish_tile64(double* __restrict__ A, int Dp, int bi, int bj, const double* s,
                                                    int* flag, int epoch, int team) {
  double* dst = A + (size_t)bi * MB * Dp + (size_t)bj * MB;
  for (int idx = threadIdx.x & 127; idx < MB * MB; idx += 128) __stcg(dst + (size_t)(idx >> 6) * Dp + (idx & 63), s[(idx >> 6) * LDC + (idx & 63)]);
  team_sync(team);
  if ((threadIdx.x & 127) == 0) { st_release(flag, epoch); }
}

// diagnostics: a trace entry written by the calling thread (team 1's timeline inside a chain step)
__device__ __forceinline__ void trace_put3(const CholArgs& c, int slot, int kind, int i, int j,
                                           unsigned long long t0, unsigned long long t1, unsigned long long t2) {
  if (c.trace && slot < c.trace_cap) {
    unsigned long long* p = c.trace + 4 * (size_t)slot;
    p[0] = ((unsigned long long)kind << 32) | ((unsigned long long)i << 16) | (unsigned long long)j;
    p[1] = t0; p[2] = t1; p[3] = t2;
  }
}

__device__ __forceinline__ void run_chain(const Args& a, int epoch, double* smem, double* s_dinv) {
  const int tid = threadIdx.x, warp = tid >> 5, team = warp >> 2, w = warp & 3;
  double* T = smem;
  double* X = smem + kTileC;
  double* Lp = smem + 2 * kTileC;
  double* Tn = smem + 3 * kTileC;
  double* Xn = smem + 4 * kTileC;
  double* Ub = smem + 5 * kTileC;  // [32][LDC]: L00^-T of the tile factorised in the previous step
  double* dT = s_dinv;
  double* dP = s_dinv + MB;
  load_tile64(a.c.A, a.c.Dp, 0, 0, T);
  __syncthreads();
  for (int J = 0; J < a.nbm; J++) {
    chol_debug_stall(a.c, J, J, epoch, 0);
    const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
    if (J > 0) {
      trsm64w(X, Lp, dP, Ub);
      // L(J,J-1) goes out at once: sidekick 1 and the column's list tasks wait for it.  Warp 7 fences and flags
      // while warps 0-6 form  T00 -= X0 X0^T (3 lower 16x16 blocks)  and  T10 -= X1 X0^T (4 blocks).
      store_tile64(a.c.A, a.c.Dp, J, J - 1, X);
      __syncthreads();
      if (warp == 7) {
        if ((tid & 31) == 0) { st_release(tile_flag(a, J, J - 1), epoch); }
      } else {
        const int r0 = warp < 3 ? (warp == 0 ? 0 : 16) : 32 + ((warp - 3) >> 1) * 16;
        const int c0 = warp < 3 ? (warp == 2 ? 16 : 0) : ((warp - 3) & 1) * 16;
        double acc[2][2][2];
        zero_acc(acc);
        warp_mma<2, 2, LDC>(X + r0 * LDC, X + c0 * LDC, 4, acc);
        sub_acc_smem(T, r0, c0, acc);
      }
      __syncthreads();
    }
    const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
    if (team == 0) {
      potrf_tall_team(TileC{T}, w, J * MB, a.c.D, dT, a.c.hdr, [] { team_sync(0); });  // L00 and T10 <- T10 L00^-T
      team_sync(0);
      asm volatile("bar.arrive 6, 256;" ::: "memory");  // L00 and its 1/diag are final: team 1 may invert it
      const unsigned long long p1 = a.c.trace ? gtimer() : 0ull;
      const unsigned long long p2 = p1;
      cta_wait(3);  // (1) team 1 has finished T11 -= X1 X1^T (it arrived and went on)
      {                 // T11 -= T10 T10^T, 16x16 outputs per warp
        const int r0 = 32 + (w >> 1) * 16, c0 = 32 + (w & 1) * 16;
        double acc[2][2][2];
        zero_acc(acc);
        warp_mma<2, 2, LDC>(T + r0 * LDC, T + c0 * LDC, 2, acc);
        sub_acc_smem(T, r0, c0, acc);
      }
      team_sync(0);
      const unsigned long long p3 = a.c.trace ? gtimer() : 0ull;
      potrf_team(TileC{T + 32 * LDC + 32}, w, J * MB + 32, a.c.D, dT + 32, a.c.hdr, [] { team_sync(0); });
      team_sync(0);
      if (tid == 0) {  // team 0's timeline: POTRF(T00) done, TRSM(T10) done, T11 updated | .., POTRF(T11) done
        const int base = a.nbm + a.ntasks + 4 * a.nbm;
        trace_put3(a.c, base + 2 * J, 6, J, J, p1, p2, p3);
        trace_put3(a.c, base + 2 * J + 1, 7, J, J, p3, gtimer(), gtimer());
      }
      chol_debug_stall(a.c, J, J, epoch, 1);
      if (tid < MB) a.c.x[a.c.Dp + J * MB + tid] = dT[tid];
      // L(J,J) and 1/diag.  The fence stalls its warp for ~1 us: by now team 0 only waits for team 1.
      team_publish_tile64(a.c.A, a.c.Dp, J, J, T, tile_flag(a, J, J), epoch, 0);
    } else {
      // team 1: T11 -= X1 X1^T (team 0 waits for it), then the next step's two tiles, whose last terms the
      // sidekick CTAs have added by now (progress counters one past the list's chunks)
      unsigned long long q0 = 0ull, q1 = 0ull, q2 = 0ull;
      if (J > 0) {
#pragma unroll 1
        for (int blk = w; blk < 8; blk += 4) {  // 16x8 blocks
          const int r0 = 32 + (blk >> 2) * 16, c0 = 32 + (blk & 3) * 8;
          double acc[2][1][2];
          zero_acc(acc);
          warp_mma<2, 1, LDC>(X + r0 * LDC, X + c0 * LDC, 4, acc);
          sub_acc_smem(T, r0, c0, acc);
        }
      }
      cta_arrive(3);  // (1) team 0 may touch T11 now; this team does not wait for it
      if (a.c.trace) q0 = gtimer();
      if (J + 1 < a.nbm) {
        const int extra = J > 0 ? 1 : 0;  // the sidekicks' chunk
        const int nd = gn_nch_diag(J + 1) + extra, ns = gn_nch_sub(J) + extra;
        const int lt = tid & 127;
        if (lt == 0 && nd) while (ld_acquire(prog_flag(a, J + 1, J + 1)) != ((epoch << 12) | nd)) __nanosleep(20);
        if (lt == 32 && ns) while (ld_acquire(prog_flag(a, J + 1, J)) != ((epoch << 12) | ns)) __nanosleep(20);
        team_sync(1);
        if (a.c.trace) q1 = gtimer();
        // 8-byte cp.async: the rows of a chain tile are 8-byte aligned only; each of these lines is read once
        // per launch by this SM, right after an acquire
        for (int idx = lt; idx < MB * MB; idx += 128) {
          const int rr = idx >> 6, cc = idx & 63;
          cp_async8(Tn + rr * LDC + cc, a.c.A + (size_t)((J + 1) * MB + rr) * a.c.Dp + (J + 1) * MB + cc);
          cp_async8(Xn + rr * LDC + cc, a.c.A + (size_t)((J + 1) * MB + rr) * a.c.Dp + J * MB + cc);
        }
        cp_async_commit();
      }
      // L00^-T of THIS step's tile for the next step's trsm64w (the loads above fly meanwhile): X = I, X <- X L00^-T
      asm volatile("bar.sync 6, 256;" ::: "memory");
      for (int idx = tid & 127; idx < 32 * 32; idx += 128) Ub[(idx >> 5) * LDC + (idx & 31)] = (idx >> 5) == (idx & 31) ? 1.0 : 0.0;
      team_sync(1);
      trsm_team(TileC{Ub}, TileC{T}, dT, w, [] { team_sync(1); });
      cp_async_wait<0>();
      if (a.c.trace) q2 = gtimer();
      if (tid == 128) {  // team 1's timeline: T11 term done, next tiles flagged, loaded
        const int base = a.nbm + a.ntasks + 2 * a.nbm;
        trace_put3(a.c, base + 2 * J, 4, J, J, q0, q1, q2);
        trace_put3(a.c, base + 2 * J + 1, 5, J, J, q2, q2, gtimer());
      }
    }
    __syncthreads();  // (2)
    trace_put(a.c, J, 0, J, J, t_begin, t_ready);
    double* t = Lp; Lp = T; T = Tn; Tn = t;
    t = X; X = Xn; Xn = t;
    t = dT; dT = dP; dP = t;
  }
}

// ------------------------------------------------------------------------------------------
// Back substitution L^T x = y (y^T = row D of L), 64 columns per step:
//   x_c = L_cc^-T ( y_c - sum_{m>c} L_mc^T x_m ),  entries with global index >= D are 0.
// The 32-column form dealt round robin took 0.40 ms at N = 512: 112 steps, each a flag hop, an L2
// round trip for x_m and a fence.  Here the dependent chain — the last kBackNear terms of the sum
// and the 64-step triangular solve — stays in ONE CTA (block 0): x lives in its shared memory, the
// tiles it needs next are in flight (registers) while it solves.  The far part of every column's
// sum, m > c + kBackNear, is the work of a helper CTA per column that follows the chain kBackNear
// steps behind, so its flag is normally up when the chain arrives.  Fixed summation order.
// ------------------------------------------------------------------------------------------
constexpr int kBackNear = 2;
__device__ __forceinline__ int* far_flag(const Args& a, int c) { return a.c.flags + 2 * a.nbm * a.nbm + a.nbm + c; }
__device__ __forceinline__ int* winv_flag(const Args& a, int c) { return a.c.flags + queue_index(a.nbm) + 1 + c; }

// Helper of column c.  (i) Inverts the two 32x32 diagonal blocks of L(c,c): W^T = L^-T goes into the
// strictly upper part of the block in the factor (unused otherwise; its diagonal is 1/diag(L), kept
// in x[Dp..]) — with it the chain's 64-step dependent triangular solve becomes three 32x32
// matrix-vector products.  (ii) far_c = sum_{m > c + kBackNear} L(m,c)^T x_m -> x[2 Dp + 64 c ..].
// Then far_flag(c).
__device__ __forceinline__ void back_far(const Args& a, int c, int epoch, double* smem, double* s_dinv, int slot) {
  const int tid = threadIdx.x, j = tid & 63, p = tid >> 6;  // column j of the tile, rows 16p .. 16p+15
  const int m_lo = c + kBackNear + 1;
  const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
  {
    const int team = tid >> 7, lt = tid & 127, w = lt >> 5, off = 32 * team;  // team t inverts block (t,t)
    double* Lt = smem;
    double* Ws = smem + kTileC + team * 32 * LDC;
    load_tile64(a.c.A, a.c.Dp, c, c, Lt);
    if (tid < MB) s_dinv[tid] = __ldcg(a.c.x + a.c.Dp + c * MB + tid);
    for (int idx = lt; idx < 32 * 32; idx += 128) Ws[(idx >> 5) * LDC + (idx & 31)] = (idx >> 5) == (idx & 31) ? 1.0 : 0.0;
    __syncthreads();
    trsm_team(TileC{Ws}, TileC{Lt + off * LDC + off}, s_dinv + off, w, [team] { team_sync(team); });
    team_sync(team);
    double* dst = a.c.A + (size_t)(c * MB + off) * a.c.Dp + (size_t)c * MB + off;
    for (int idx = lt; idx < 32 * 32; idx += 128) {
      const int r = idx >> 5, cc = idx & 31;
      if (cc > r) __stcg(dst + (size_t)r * a.c.Dp + cc, Ws[r * LDC + cc]);
    }
    post_flag(winv_flag(a, c), epoch);
    __syncthreads();
  }
  const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
  if (m_lo <= a.nbm - 1) {
    double* xs = smem;
    double* sred = smem + MB;
    double cur[16], nxt[16], s = 0.0;
    auto fetch = [&](int m, double (&v)[16]) {
      const double* T = a.c.A + (size_t)(m * MB + p * 16) * a.c.Dp + (size_t)c * MB + j;
#pragma unroll
      for (int r = 0; r < 16; r++) v[r] = __ldcg(T + (size_t)r * a.c.Dp);
    };
    fetch(a.nbm - 1, cur);
    for (int m = a.nbm - 1; m >= m_lo; m--) {
      if (m - 1 >= m_lo) fetch(m - 1, nxt);
      if (tid == 0) while (ld_acquire(a.c.flags + 2 * a.nbm * a.nbm + m) != epoch) { }
      __syncthreads();
      if (tid < MB) xs[tid] = __ldcg(a.c.x + m * MB + tid);
      __syncthreads();
#pragma unroll
      for (int r = 0; r < 16; r++) s = fma(cur[r], xs[p * 16 + r], s);
#pragma unroll
      for (int r = 0; r < 16; r++) cur[r] = nxt[r];
    }
    sred[p * MB + j] = s;
    __syncthreads();
    if (tid < MB) a.c.x[2 * a.c.Dp + c * MB + tid] = (sred[tid] + sred[MB + tid]) + (sred[2 * MB + tid] + sred[3 * MB + tid]);
    post_flag(far_flag(a, c), epoch);
  }
  trace_put(a.c, slot, 2, c, 0, t_begin, t_ready);
}

// The chain of the back substitution (block 0).  One step per block column c, descending; roles by warp:
//   warp 0      the solve: rhs = y_c - far_c - near sums, then x_c = L_cc^-T rhs as three 32x32 matrix-vector
//               products with the block inverses; x_c goes to the ring in shared memory and to global memory
//   warps 1-6   the near part of the sum, sum_{m = c+1 .. c+kBackNear} L(m,c)^T x_m, from tiles that a
//               cp.async ring brought into shared memory ONE STEP EARLIER, and the cp.async of the next step
//   warp 7      fence + flag of x_c (it has no loads in flight, so the fence is short); polls the flags the
//               NEXT step needs while the others work, so that no acquire round trip sits on the chain
// Everything a step loads — kBackNear near tiles and the diagonal tile, 96 KB — goes through cp.async into one
// of two shared-memory slots: no registers are tied up by loads in flight (an earlier form kept them in 160
// registers per thread across the solve; at the 255-register cap the solve's own loads were then serialised,
// 2 900 cycles for 96 fma, and the step took 3.7 us), and the issuing warps are not the solving one (the load
// queue of one SM takes ~2 300 cycles to accept a step's loads).
constexpr int kBackSlot = kBackNear * MB * MB + MB * LDS2;  // doubles per slot: near tiles [64][64], diagonal [64][66]
constexpr int kBackScratch = 16 * MB;                       // xr, sred, ex, xb
constexpr size_t kBackSmemBytes = (size_t)(kBackScratch + 2 * kBackSlot) * sizeof(double);
constexpr size_t kSmemBytes = kSmemBytesFact > kBackSmemBytes ? kSmemBytesFact : kBackSmemBytes;  // 206 848 B
static_assert(kSmemBytes >= (size_t)kStages * kStageD * sizeof(double), "pipeline ring");
static_assert(kSmemBytes <= 227 * 1024, "dynamic shared memory of chol64_kernel exceeds the 227 KB of sm_100a");

__device__ __forceinline__ void back_chain(const Args& a, int epoch, double* smem, int slot0) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nbm = a.nbm, Dp = a.c.Dp, D = a.c.D;
  // warp 4 shares its scheduler (and its queue towards the load/store unit) with warp 0: it is the one that
  // stays quiet; the six loaders are warps 1-3 and 5-7
  const bool loader = warp != 0 && warp != 4;
  const int part = warp < 4 ? warp - 1 : warp - 2;  // 0..5 among the loaders
  double* xr = smem;                       // ring of the last 4 x blocks
  double* sred = smem + 4 * MB;            // [6][64] partial near sums
  double* xb = smem + 10 * MB;             // [64] exchange inside warp 0
  double* farv = smem + 11 * MB;           // [2][64] far sums of this and the next step (fetched by warp 4)
  double* slots = smem + kBackScratch;
  // loads of step c into slot c & 1 (warps 1-6, 192 threads): near tiles L(c+1+i, c) and the diagonal tile
  auto issue = [&](int c) {
    double* sl = slots + (c & 1) * kBackSlot;
    const int t6 = part * 32 + lane;
#pragma unroll
    for (int i = 0; i < kBackNear; i++) {
      const int m = c + 1 + i;
      if (m < nbm) {
        const double* T = a.c.A + (size_t)m * MB * Dp + (size_t)c * MB;
        for (int ch = t6; ch < MB * 32; ch += 192) cp_async16(sl + i * MB * MB + (ch >> 5) * MB + (ch & 31) * 2, T + (size_t)(ch >> 5) * Dp + (ch & 31) * 2);
      }
    }
    const double* T = a.c.A + (size_t)c * MB * Dp + (size_t)c * MB;
    double* dg = sl + kBackNear * MB * MB;
    for (int ch = t6; ch < MB * 32; ch += 192) cp_async16(dg + (ch >> 5) * LDS2 + (ch & 31) * 2, T + (size_t)(ch >> 5) * Dp + (ch & 31) * 2);
    cp_async_commit();
  };
  // winv_flag(c): the block inverses inside the diagonal tile of column c are written (needed before the loads of
  // step c are issued); far_flag(c): the far sum of column c (needed by step c's rhs)
  auto poll = [&](int c) {  // what step c itself and the issue inside step c need
    if (c > 0) while (ld_acquire(winv_flag(a, c - 1)) != epoch) { }
    if (c + kBackNear + 1 <= nbm - 1) while (ld_acquire(far_flag(a, c)) != epoch) { }
  };
  double yA = 0.0, yB = 0.0, iA = 0.0, iB = 0.0;  // warp 0: y_c and 1/diag(L_cc), entries lane and 32 + lane
  auto load_y = [&](int c) {
    const int gA = c * MB + lane, gB = gA + 32;
    yA = gA < D ? __ldcg(a.c.A + (size_t)D * Dp + gA) : 0.0;
    yB = gB < D ? __ldcg(a.c.A + (size_t)D * Dp + gB) : 0.0;
    iA = __ldcg(a.c.x + Dp + gA);
    iB = __ldcg(a.c.x + Dp + gB);
  };
  if (tid == 0) { while (ld_acquire(winv_flag(a, nbm - 1)) != epoch) { } poll(nbm - 1); }
  __syncthreads();
  if (loader) issue(nbm - 1);
  if (warp == 0) load_y(nbm - 1);
  for (int c = nbm - 1; c >= 0; c--) {
    const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
    const double* sl = slots + (c & 1) * kBackSlot;
    if (loader) cp_async_wait<0>();
    __syncthreads();  // (T) slot c & 1 has landed; x_{c+1} is in the ring
    if (loader) {
      // rows [r_lo, r_hi) of every near tile, columns 2 lane and 2 lane + 1
      const int r_lo = part * MB / 6, r_hi = (part + 1) * MB / 6;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int i = 0; i < kBackNear; i++) {
        const int m = c + 1 + i;
        if (m < nbm) {
          const double* tile = sl + i * MB * MB + 2 * lane;
          const double* xm = xr + (m & 3) * MB;
          for (int r = r_lo; r < r_hi; r++) {
            const double2 v = *reinterpret_cast<const double2*>(tile + r * MB);
            s0 = fma(v.x, xm[r], s0); s1 = fma(v.y, xm[r], s1);
          }
        }
      }
      sred[part * MB + 2 * lane] = s0; sred[part * MB + 2 * lane + 1] = s1;
    }
    __syncthreads();  // (A) near sums are in; the other slot is free
    const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
    if (loader) {
      if (c > 0) issue(c - 1);
    } else if (warp == 4) {
      if (c > 0) {  // for the next step, off the chain: its flags, then its far sum into shared memory
        if (lane == 0) poll(c - 1);
        __syncwarp();
        if (c - 1 + kBackNear + 1 <= nbm - 1) {
          farv[((c - 1) & 1) * MB + lane] = __ldcg(a.c.x + 2 * Dp + (c - 1) * MB + lane);
          farv[((c - 1) & 1) * MB + 32 + lane] = __ldcg(a.c.x + 2 * Dp + (c - 1) * MB + 32 + lane);
        }
      }
    } else {
      // x = L_cc^-T rhs with the inverses W^T = L^-T of the two 32x32 diagonal blocks (strictly upper part
      // of the block, diagonal = 1/diag(L)):  x_B = W11^T r_B ; r_A -= L10^T x_B ; x_A = W00^T r_A
      const double* Ld = sl + kBackNear * MB * MB;
      const bool has_far = c + kBackNear + 1 <= nbm - 1;
      double rA = yA, rB = yB;
#pragma unroll
      for (int pp = 0; pp < 6; pp++) { rA -= sred[pp * MB + lane]; rB -= sred[pp * MB + 32 + lane]; }
      if (has_far) { rA -= farv[(c & 1) * MB + lane]; rB -= farv[(c & 1) * MB + 32 + lane]; }
      if (c * MB + lane >= D) rA = 0.0;
      if (c * MB + 32 + lane >= D) rB = 0.0;
      const double dA = iA, dB = iB;
      if (c > 0) load_y(c - 1);  // next step's y and 1/diag, in flight during the solve
      xb[32 + lane] = rB;
      __syncwarp();
      double p0 = dB * rB, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
      for (int k = 0; k < 32; k += 4) {
        const double l0 = k + 0 > lane ? Ld[(32 + lane) * LDS2 + 32 + k + 0] : 0.0;
        const double l1 = k + 1 > lane ? Ld[(32 + lane) * LDS2 + 32 + k + 1] : 0.0;
        const double l2 = k + 2 > lane ? Ld[(32 + lane) * LDS2 + 32 + k + 2] : 0.0;
        const double l3 = k + 3 > lane ? Ld[(32 + lane) * LDS2 + 32 + k + 3] : 0.0;
        p0 = fma(l0, xb[32 + k + 0], p0); p1 = fma(l1, xb[32 + k + 1], p1);
        p2 = fma(l2, xb[32 + k + 2], p2); p3 = fma(l3, xb[32 + k + 3], p3);
      }
      const double xB = (p0 + p1) + (p2 + p3);
      __syncwarp();
      xb[lane] = xB;
      __syncwarp();
      p0 = p1 = p2 = p3 = 0.0;
#pragma unroll
      for (int k = 0; k < 32; k += 4) {
        p0 = fma(Ld[(32 + k + 0) * LDS2 + lane], xb[k + 0], p0);
        p1 = fma(Ld[(32 + k + 1) * LDS2 + lane], xb[k + 1], p1);
        p2 = fma(Ld[(32 + k + 2) * LDS2 + lane], xb[k + 2], p2);
        p3 = fma(Ld[(32 + k + 3) * LDS2 + lane], xb[k + 3], p3);
      }
      rA -= (p0 + p1) + (p2 + p3);
      __syncwarp();
      xb[lane] = rA;
      __syncwarp();
      p0 = dA * rA; p1 = p2 = p3 = 0.0;
#pragma unroll
      for (int k = 0; k < 32; k += 4) {
        const double l0 = k + 0 > lane ? Ld[lane * LDS2 + k + 0] : 0.0;
        const double l1 = k + 1 > lane ? Ld[lane * LDS2 + k + 1] : 0.0;
        const double l2 = k + 2 > lane ? Ld[lane * LDS2 + k + 2] : 0.0;
        const double l3 = k + 3 > lane ? Ld[lane * LDS2 + k + 3] : 0.0;
        p0 = fma(l0, xb[k + 0], p0); p1 = fma(l1, xb[k + 1], p1);
        p2 = fma(l2, xb[k + 2], p2); p3 = fma(l3, xb[k + 3], p3);
      }
      const double xA = (p0 + p1) + (p2 + p3);
      xr[(c & 3) * MB + lane] = xA; xr[(c & 3) * MB + 32 + lane] = xB;
      a.c.x[c * MB + lane] = xA; a.c.x[c * MB + 32 + lane] = xB;
    }
    __syncthreads();  // (C) x_c is written
    if (tid == 4 * 32) { st_release(a.c.flags + 2 * nbm * nbm + c, epoch); }  // x_c for the helpers
    trace_put(a.c, slot0 + c, 2, c, c, t_begin, t_ready);
  }
}

__global__ void __launch_bounds__(kThreads, 1)
chol64_kernel(const Args a) {
  if (*reinterpret_cast<volatile int*>(&a.c.hdr->done)) return;
  if (*reinterpret_cast<volatile int*>(&a.c.hdr->peer_timeout) > 1) return;  // see chol_dataflow_kernel
  extern __shared__ __align__(16) double smem[];
  __shared__ double s_dinv[2 * MB];
  __shared__ float red[kThreads / 32];
  __shared__ int s_task, s_early;
  const int epoch = a.c.hdr->iters + 1;
  const int G = gridDim.x, cta = blockIdx.x;
  if (cta == 0) {
    run_chain(a, epoch, smem, s_dinv);
  } else if (cta <= 2) {
    run_sidekick(a, cta - 1, epoch, smem);
  } else {
    int* queue = a.c.flags + queue_index(a.nbm);
    if (threadIdx.x == 0) s_task = atomicAdd(queue, 1);
    for (;;) {
      __syncthreads();
      const int t = s_task;
      __syncthreads();
      if (t >= a.ntasks) break;
      run_task(a, __ldg(a.tasks + t), epoch, smem, s_dinv, a.nbm + t, &s_task, &s_early);  // leaves the next entry in s_task
    }
  }
  // back substitution: block 0 runs the chain, blocks 1.. the far sums of the columns (descending)
  if (cta == 0) {
    __syncthreads();
    back_chain(a, epoch, smem, a.nbm + a.ntasks + a.nbm);
    __syncthreads();  // x complete
    gn_update_block(a.c.Twc, a.c.x, a.c.N, a.c.D, a.c.delta_thresh, a.c.dx, a.c.hdr, red, kThreads, threadIdx.x,
                    [] { __syncthreads(); });
  } else {
    // the last diagonal tile depends, transitively, on every tile of the factor
    wait_flag(tile_flag(a, a.nbm - 1, a.nbm - 1), epoch);
    for (int c = a.nbm - 1 - (cta - 1); c >= 0; c -= G - 1) back_far(a, c, epoch, smem, s_dinv, a.nbm + a.ntasks + c);
  }
}

static int max_ctas() {
  static std::atomic<int> cache[kMaxDevices];
  int dev = 0, sms = kNumSMs, per_sm = 0;
  cudaGetDevice(&dev);
  std::atomic<int>* slot = (dev >= 0 && dev < kMaxDevices) ? &cache[dev] : nullptr;
  if (slot && slot->load()) return slot->load();
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol64_kernel, kThreads, kSmemBytes);
  if (per_sm < 1) return 0;
  if (slot) slot->store(sms);  // one CTA per SM: the chain CTA must have its SM to itself
  return sms;
}

}  // namespace c64

int gn_chol64_prepare(void* ws, const GnLayout& L, cudaStream_t st) {
  if ((L.nbm + 1) * (c64::kWindow + 1) > c64::kMaxSlots) return M3R_EINVAL;
  c64::chol64_tasklist_kernel<<<1, 1024, 0, st>>>(L.nbm, L.ntasks, ws_ptr<int4>(ws, L.tasks));
  return after_launch();
}

int gn_chol64_launch(float* Twc, int N, float delta_thresh, float* dx, void* ws, const GnLayout& L,
                     cudaStream_t st) {
  c64::Args a;
  a.c.A = ws_ptr<double>(ws, L.A); a.c.Dp = L.Dp; a.c.D = L.D; a.c.nb = L.Dp / NB;
  a.c.flags = ws_ptr<int>(ws, L.flags);
  a.c.x = ws_ptr<double>(ws, L.x);
  a.c.Twc = Twc; a.c.dx = dx; a.c.N = N; a.c.delta_thresh = delta_thresh;
  a.c.hdr = ws_ptr<GnHeader>(ws, L.hdr);
  a.c.trace = g_chol_trace.load(); a.c.trace_cap = a.c.trace ? g_chol_trace_cap.load() : 0;
  a.c.delay_us = g_chol_delay_us.load(); a.c.delay_which = g_chol_delay_which.load();
  a.tasks = ws_ptr<int4>(ws, L.tasks); a.ntasks = L.ntasks; a.nbm = L.nbm;
  // > 48 KB of dynamic shared memory is a per-device opt-in: set it on every launch
  M3R_TRY(check_cuda(cudaFuncSetAttribute(c64::chol64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)c64::kSmemBytes)));
  int grid = c64::max_ctas();
  if (grid < 4) return M3R_ELAUNCH;  // the chain CTA, two sidekicks and at least one worker must be co-resident
  if (grid > 3 + L.ntasks) grid = 3 + (L.ntasks > 0 ? L.ntasks : 1);
  void* args[] = {&a};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)c64::chol64_kernel, dim3(grid),
                                                    dim3(c64::kThreads), args, c64::kSmemBytes, st);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return check_cuda(e);
}

}  // namespace m3r

// Diagnostics / tests: the task list of the macro-tile factorisation for nbm block columns, built
// on the HOST by the same enumeration the device kernel runs (4 ints per task:
// I | J << 16, k0 | k1 << 16, kind, chunk index).  Returns the number of tasks; writes at most
// `capacity` of them.
namespace {
struct HostTaskSink {
  int* out; int capacity; int J; int o;
  __host__ __device__ void operator()(int I, int k0, int k1, int kind, int kc) {
    if (out && o < capacity) {
      out[4 * o] = I | (J << 16); out[4 * o + 1] = k0 | (k1 << 16); out[4 * o + 2] = kind; out[4 * o + 3] = kc;
    }
    o++;
  }
};
}  // namespace

extern "C" int m3r_debug_chol64_tasklist(int nbm, int* out, int capacity) {
  using namespace m3r;
  HostTaskSink sink{out, capacity, 0, 0};
  for (int v = 0; v <= nbm; v++)
    for (int J = v; J <= v + c64::kWindow && J < nbm; J++) {
      sink.J = J;
      c64::tasks_of_group(nbm, v, J, sink);
    }
  return sink.o == gn_chol64_ntasks(nbm) ? sink.o : -1;
}
