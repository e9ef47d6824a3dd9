// m3r_gn.cuh — workspace layout shared by the Gauss-Newton kernels (m3r_gn_*.cu).
//
// HBM layout of one solve (all inside the caller-provided workspace, 256-B aligned):
//   hdr      : GnHeader (flags, counters)
//   ie, je   : int32[M]      pose row of each directed edge's i / j keyframe
//   keys     : int64[P]      sort scratch (P = pow2 >= 2M)
//   rank     : int32[P]      rank of each sorted key among the unique ids
//   row_ptr  : int32[N+1]    CSR over pose rows -> incident (edge, role) slots
//   adj      : int32[2M]     slot = e (pose is the i end) or M+e (pose is the j end)
//   xc       : float4[N*n]   (X, C) packed: one 16-B aligned gather per correspondence
//   part     : float[Ml*S*36] per-CTA partial sums of the LOCAL 28+7 normal-equation terms
//   blocks   : double[M*56]  per-edge S (7x7) and v (7) after the adjoint transform
//   A        : double[Dp*Dp] augmented normal matrix [H b; b^T 1], padded to 32
//   x        : double[3*Dp]  the solution x, 1/diag(L), scratch sink
//   flags    : int32[nb*nb + nb] tile-ready / x-block-ready epochs of the dataflow Cholesky
//   tasks    : int4[ntasks]  large systems only (Dp > 5 panels of 32): the statically ordered task list of
//                            the 64x64 macro-tile Cholesky (m3r_gn_chol64.cu), built once per solve
//   tickets  : int32[M+2]    per-edge count of finished point-splits (the last one to arrive runs
//                            the edge's fp64 epilogue), [M] = edges finished in this pass
#pragma once
#include <cstdlib>

#include "m3r_common.cuh"

namespace m3r {

constexpr int kGnNB = 32;        // Cholesky panel width
constexpr int kGnPartStride = 36;  // 28 + 7 (+1 pad) floats per partial
constexpr int kGnBlock = 56;     // doubles per edge block: S(49) + v(7)

struct GnHeader {
  int done;        // 1 once |dx| < delta_thresh (the reference's `break`)
  int chol_fail;   // set by the factorisation of the current iteration
  int iters;       // iterations actually executed
  int peer_timeout;  // a peer's blocks did not arrive in time (m3r_gn_solve_step_p2p): the step is
                     // zero, `done` stays 0 and every later launch of the solve is a no-op
};

// true once the loop has ended: converged (`done`) or aborted by a peer timeout
__device__ __forceinline__ bool gn_stopped(const GnHeader* hdr) {
  return (*reinterpret_cast<const volatile int*>(&hdr->done) |
          *reinterpret_cast<const volatile int*>(&hdr->peer_timeout)) != 0;
}

// ---- large systems: 64x64 macro tiles (m3r_gn_chol64.cu) ----------------------------------
// A system of more than kGnSmallMaxNb 32-wide panels is padded to a multiple of 64 and
// factorised by the macro-tile kernel (measured, profiles/r3/r3_gn_threshold.txt: the macro kernel is
// level with the 32x32 dataflow kernel at 2 and 4 panels, loses at 3 and 5 — an odd panel count pays for
// the padding to 64 — and wins from 6 panels on: 97 against 105 us at 7 panels (N = 32), 163 against 228 us
// at 14 (N = 64), 394 against 583 us at 33 (N = 150)).  The left-looking sum of tile (I,J) over block columns
// k is cut into chunks; every chunk is one task.  A tile's list tasks cover k in [0, R) and the
// chunks are laid out FROM THE END: a tail of `t` block columns (the chunk released last, one
// or two chain steps before the tile is needed, is short), then chunks of kGnKC going down;
// the partial chunk is the first.  Because R depends on the tile column, the chunk boundaries of
// neighbouring columns are staggered and the chain releases work evenly (with boundaries at
// multiples of 4 everywhere, every fourth chain step released ~400 tasks ahead of the urgent ones).
//   diagonal tile (J,J):       R = J-2, t = 2; the chain CTA adds k = J-2 and k = J-1 itself
//   sub-diagonal tile (J+1,J): R = J-1, t = 2; the chain CTA adds k = J-1 itself
//   any other tile (I >= J+2): R = J-1, t = 4; k = J-1 is alone in the tile's final (TRSM) task
// Shared by the host (workspace size) and the device (task list, progress counters).
constexpr int kGnSmallMaxNb = 5;
constexpr int kGnMB = 64;
constexpr int kGnKC = 4;
constexpr int kGnTailChain = 2;
__host__ __device__ inline int gn_cdiv(int a, int b) { return (a + b - 1) / b; }
// Chunk ends of [0, R), walking down from R: the tail t, then kGnRecent / kGnKC chunks of kGnKC,
// then chunks of kGnKCOld.  The recent chunks are released one at a time as the chain advances;
// the old ones are all available when the tile's column enters the task list's window and are
// released TOGETHER — their subtractions from the tile are serialised by the progress counter, so
// they are few and long (with chunks of 4 everywhere a column's tiles spent 30 % of their CTA time
// waiting for their turn).  Swept at N = 512 (profiles/r3/r3_sweep_chunks.txt): recent region 4 / 8 / 12
// block columns -> solve step 1.24 / 1.36 / 1.44 ms; old chunks of 8 / 16 / 24 / 32 within 2 %.
constexpr int kGnRecent = 4;
constexpr int kGnKCOld = 16;
__host__ __device__ inline int gn_chunk_step(int d) { return d < kGnRecent ? kGnKC : kGnKCOld; }  // d = R - t - end
__host__ __device__ inline int gn_nch(int R, int t) {
  if (R <= 0) return 0;
  int n = 1;
  for (int e = R - t, d = 0; e > 0; n++) { const int st = gn_chunk_step(d); e -= st; d += st; }
  return n;
}
// is e the end of a chunk of a tile with range [0, R) and tail t; its start and index if so
__host__ __device__ inline bool gn_chunk_end(int R, int t, int e) {
  if (R <= 0 || e <= 0 || e > R) return false;
  if (e == R) return true;
  for (int c = R - t, d = 0; c > 0;) {
    if (c == e) return true;
    if (c < e) return false;
    const int st = gn_chunk_step(d); c -= st; d += st;
  }
  return false;
}
__host__ __device__ inline int gn_chunk_k0(int R, int t, int e) {
  const int k0 = e == R ? R - t : e - gn_chunk_step(R - t - e);
  return k0 > 0 ? k0 : 0;
}
__host__ __device__ inline int gn_chunk_index(int R, int t, int e) {
  int above = 0;  // chunk ends above e
  if (e != R) {
    above = 1;
    for (int c = R - t, d = 0; c > e;) { const int st = gn_chunk_step(d); c -= st; d += st; above++; }
  }
  return gn_nch(R, t) - 1 - above;
}
__host__ __device__ inline int gn_nch_diag(int J) { return gn_nch(J - 2, kGnTailChain); }
__host__ __device__ inline int gn_nch_sub(int J) { return gn_nch(J - 1, kGnTailChain); }
__host__ __device__ inline int gn_ntask_other(int J) { return gn_nch(J - 1, kGnKC) + 1; }
inline int gn_chol64_ntasks(int nbm) {
  long long t = 0;
  for (int J = 0; J < nbm; J++) {
    t += gn_nch_diag(J);
    if (J + 1 < nbm) t += gn_nch_sub(J);
    if (J + 2 < nbm) t += (long long)(nbm - J - 2) * gn_ntask_other(J);
  }
  return (int)t;
}

struct GnLayout {
  size_t hdr, ie, je, keys, rank, row_ptr, adj, xc, part, blocks, A, x, flags, tickets, tasks, total;
  int nbm;     // 64-wide block columns of the macro-tile factorisation, 0 for small systems
  int ntasks;  // entries of `tasks`
  int P;       // sort size
  int S;       // point splits per edge
  int chunk;   // points per split
  int D;       // 7*(N-1)
  int Dp;      // padded dimension of the augmented system
  int nflags;  // entries of `flags`
};

inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

// Splits per edge depend only on (n, M) — never on how many edges THIS process owns —
// so an edge's block is bit-identical whether the graph is solved on 1 GPU or sharded.
inline void gn_splits(int n, int M, int& S, int& chunk) {
  const int unit = 1024;  // 256 threads x 4 points
  if (n <= 0) { S = 1; chunk = unit; return; }  // layout queries that only need (N, M) offsets
  int want = (8 * 2 * kNumSMs + (M > 0 ? M : 1) - 1) / (M > 0 ? M : 1);  // ~8 waves of 2 CTA/SM
  if (want < 1) want = 1;
  if (want > 32) want = 32;
  int max_s = (n + unit - 1) / unit;
  if (max_s < 1) max_s = 1;
  if (want > max_s) want = max_s;
  chunk = ((n + want - 1) / want + unit - 1) / unit * unit;
  S = (n + chunk - 1) / chunk;
  if (S < 1) S = 1;
}

// Largest system (in 32-wide panels) that the 32x32 dataflow kernel takes; M3R_GN_SMALL_MAX_NB overrides the
// default for experiments (profiles/time_gn_threshold.py), read once per process.
inline int gn_small_max_nb() {
  static const int v = [] {
    const char* e = getenv("M3R_GN_SMALL_MAX_NB");
    const int x = e ? atoi(e) : kGnSmallMaxNb;
    return x < 1 ? 1 : x;
  }();
  return v;
}

inline GnLayout gn_layout(int N, int M, int Ml, int n) {
  GnLayout L;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  L.P = next_pow2(2 * (M > 0 ? M : 1));
  gn_splits(n, M, L.S, L.chunk);
  L.D = 7 * (N - 1);
  L.Dp = (L.D + 1 + kGnNB - 1) / kGnNB * kGnNB;
  L.nbm = 0; L.ntasks = 0;
  if (L.Dp / kGnNB > gn_small_max_nb()) {  // large system: macro tiles
    L.Dp = (L.D + 1 + kGnMB - 1) / kGnMB * kGnMB;
    L.nbm = L.Dp / kGnMB;
    L.ntasks = gn_chol64_ntasks(L.nbm);
  }
  L.hdr = take(sizeof(GnHeader));
  L.ie = take(sizeof(int) * (size_t)M);
  L.je = take(sizeof(int) * (size_t)M);
  L.keys = take(sizeof(int64_t) * (size_t)L.P);
  L.rank = take(sizeof(int) * (size_t)L.P);
  L.row_ptr = take(sizeof(int) * (size_t)(N + 1));
  L.adj = take(sizeof(int) * (size_t)2 * M);
  L.blocks = take(sizeof(double) * (size_t)M * kGnBlock);
  L.A = take(sizeof(double) * (size_t)L.Dp * L.Dp);
  L.x = take(sizeof(double) * (size_t)3 * L.Dp);
  L.nflags = (L.Dp / kGnNB) * (L.Dp / kGnNB) + (L.Dp / kGnNB);
  L.flags = take(sizeof(int) * (size_t)L.nflags);
  L.tickets = take(sizeof(int) * ((size_t)M + 2));
  L.tasks = take(sizeof(int4) * (size_t)L.ntasks);
  // n- and Ml-dependent arrays last: every offset above depends on (N, M) only, `xc` on
  // (N, M, n), and only the TOTAL depends on how many edges this process accumulates.
  L.xc = take(sizeof(float4) * (size_t)N * n);
  L.part = take(sizeof(float) * (size_t)Ml * L.S * kGnPartStride);
  L.total = off;
  return L;
}

template <typename T>
inline T* ws_ptr(void* ws, size_t off) { return reinterpret_cast<T*>(static_cast<char*>(ws) + off); }

// m3r_gn_solve.cu.  wait_flags != nullptr: before assembling, wait on the device until all
// `wait_world` entries of wait_flags have reached `wait_epoch` (peer-memory block exchange).
// m3r_gn_chol64.cu: the task list (once per solve) and the factorise / solve / retract launch.
int gn_chol64_prepare(void* ws, const GnLayout& L, cudaStream_t st);
int gn_chol64_launch(float* Twc, int N, float delta_thresh, float* dx, void* ws, const GnLayout& L,
                     cudaStream_t st);

int gn_solve_launch(float* Twc, const double* blocks, int N, int M, float delta_thresh, float* dx,
                    void* ws, const GnLayout& L, cudaStream_t st, const int* wait_flags = nullptr,
                    int wait_world = 0, int wait_epoch = 0);

}  // namespace m3r
