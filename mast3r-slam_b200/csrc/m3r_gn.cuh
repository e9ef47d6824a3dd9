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
//   tasks    : int4[ntasks]  large systems only (Dp > 16 panels): the statically ordered task list of
//                            the 64x64 macro-tile Cholesky (m3r_gn_chol64.cu), built once per solve
//   tickets  : int32[M+2]    per-edge count of finished point-splits (the last one to arrive runs
//                            the edge's fp64 epilogue), [M] = edges finished in this pass
#pragma once
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
// factorised by the macro-tile kernel.  The left-looking sum of tile (I,J) over block columns
// k is cut into chunks of kGnKC block columns; every chunk is one task.  Chunk counts per tile,
// shared by the host (workspace size) and the device (task list, progress counters):
constexpr int kGnSmallMaxNb = 16;
constexpr int kGnMB = 64;
constexpr int kGnKC = 4;
__host__ __device__ inline int gn_cdiv(int a, int b) { return (a + b - 1) / b; }
// diagonal tile (J,J): bulk tasks sum k in [0, J-1); the chain adds k = J-1 privately
__host__ __device__ inline int gn_nch_diag(int J) { return J >= 2 ? gn_cdiv(J - 1, kGnKC) : 0; }
// sub-diagonal tile (J+1,J): chunks of [0, J-1) and the single k = J-1 on its own (the chain
// waits for that one, so it must not drag three other products with it)
__host__ __device__ inline int gn_nch_sub(int J) { return (J >= 1 ? 1 : 0) + gn_nch_diag(J); }
// any other tile (I >= J+2): chunks of [0, J); the last one is merged into the tile's TRSM task
__host__ __device__ inline int gn_ntask_other(int J) { return J >= 1 ? gn_cdiv(J, kGnKC) : 1; }
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

inline GnLayout gn_layout(int N, int M, int Ml, int n) {
  GnLayout L;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  L.P = next_pow2(2 * (M > 0 ? M : 1));
  gn_splits(n, M, L.S, L.chunk);
  L.D = 7 * (N - 1);
  L.Dp = (L.D + 1 + kGnNB - 1) / kGnNB * kGnNB;
  L.nbm = 0; L.ntasks = 0;
  if (L.Dp / kGnNB > kGnSmallMaxNb) {  // large system: macro tiles
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
