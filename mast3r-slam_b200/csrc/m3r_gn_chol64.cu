// m3r_gn_chol64.cu — fp64 Cholesky solve of LARGE normal-equation systems (more than 16 panels
// of 32, i.e. more than ~74 keyframes; BASELINE config 5 has N = 512, D = 3577), sm_100a.
//
// Replaces, like m3r_gn_solve.cu for small systems, SparseBlock::solve of the reference
// (gn_kernels.cu:132-153: Eigen::SimplicialLLT in fp64 on the host) plus the retraction and the
// convergence test that follow it (gn_kernels.cu:1209-1222).  Same augmented matrix [H b; b^T 1],
// same failure semantics (non-positive pivot => zero step).
//
// Why a second kernel.  The 32x32 dataflow kernel ran N = 512 at 5.4 TFLOP/s: (i) every tile
// product pulled two 8 KB tiles from L2 for 32 K fma (3.8 GB per factorisation) and multiplied
// them with DFMA from shared memory at 6 loads per 8 fma; (ii) a left-looking tile is summed by
// ONE CTA, so the diagonal tile of a late block column (110 products) arrived long after the
// column before it was finished; (iii) per 32 columns the critical path crossed two flag hops.
// Here:
//   * tiles are 64x64, products run on the fp64 tensor pipe (mma.sync.m8n8k4.f64, DMMA): a CTA
//     of 8 warps (2x4, 32x16 outputs each) streams 64x32 operand panels through a 4-stage
//     cp.async ring; 2 shared-memory loads feed 8 fma per thread and a product moves 64 KB from
//     L2 for 262 K fma (4x less traffic per flop);
//   * the sum of a tile over block columns k is cut into CHUNKS of 4 block columns.  A chunk
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
//     32-wide POTRFs, the TRSM and the update between them need no flag at all.
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
constexpr int kStages = 4;
constexpr int kPanel = MB * LDP;
constexpr int kStageD = 2 * kPanel;
constexpr int kTileC = MB * LDC;
constexpr int kWindow = 8;         // block columns ahead of the chain whose tiles are kept up to date
constexpr int kMaxSlots = 4096;    // (nbm + 1) * (kWindow + 1) sort groups of the task list
constexpr int kChainTiles = 3;
constexpr size_t kSmemDoubles =
    (size_t)kStages * kStageD > (size_t)kChainTiles * kTileC ? (size_t)kStages * kStageD : (size_t)kChainTiles * kTileC;
constexpr size_t kSmemBytes = kSmemDoubles * sizeof(double);

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
  if (e == 0) {
    if (J == 0)
      for (int I = 2; I < nbm; I++) f(I, 0, 0, (int)kTaskFinal, 0);
    return;
  }
  const bool reg = (e % kGnKC) == 0;
  if (e <= J - 1 && (reg || e == J - 1)) {  // chunks of [0, J-1): diagonal and sub-diagonal tile
    const int kc = (e - 1) / kGnKC, k0 = kc * kGnKC;
    f(J, k0, e, (int)kTaskUpdate, kc);
    if (J + 1 < nbm) f(J + 1, k0, e, (int)kTaskUpdate, kc);
  } else if (e == J && J + 1 < nbm) {       // the sub-diagonal tile's last term, alone
    f(J + 1, J - 1, J, (int)kTaskUpdate, gn_nch_diag(J));
  }
  if (e < J && reg) {
    for (int I = J + 2; I < nbm; I++) f(I, e - kGnKC, e, (int)kTaskUpdate, e / kGnKC - 1);
  } else if (e == J) {
    const int kc = (J - 1) / kGnKC;
    for (int I = J + 2; I < nbm; I++) f(I, kc * kGnKC, J, (int)kTaskFinal, kc);
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

__global__ void __launch_bounds__(1024)
chol64_tasklist_kernel(int nbm, int ntasks, int4* __restrict__ tasks) {
  __shared__ int offs[kMaxSlots];
  const int nslots = (nbm + 1) * (kWindow + 1);
  for (int s = threadIdx.x; s < nslots; s += blockDim.x) {
    const int v = s / (kWindow + 1), J = v + s % (kWindow + 1);
    int cnt = 0;
    if (J < nbm) tasks_of_group(nbm, v, J, [&](int, int, int, int, int) { cnt++; });
    offs[s] = cnt;
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
    int o = offs[s];
    tasks_of_group(nbm, v, J, [&](int I, int k0, int k1, int kind, int kc) {
      if (o < ntasks) tasks[o] = make_int4(I | (J << 16), k0 | (k1 << 16), kind, kc);
      o++;
    });
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
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// team barrier: exactly the 128 threads of team `t` (warps 4t..4t+3)
__device__ __forceinline__ void team_sync(int t) { asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory"); }

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

// In-place Cholesky of the 64x64 tile T (lower) in shared memory: POTRF(T00), T10 <- T10 L00^-T,
// T11 -= T10 T10^T, POTRF(T11) — the four steps of one block column of the 32-wide algorithm,
// without a flag between them.  dinv[64] receives 1 / diag(L).
__device__ __forceinline__ void potrf64(double* T, int g0, int D, double* dinv, GnHeader* hdr) {
  const int warp = threadIdx.x >> 5, team = warp >> 2, w = warp & 3;
  if (team == 0) potrf_team(TileC{T}, w, g0, D, dinv, hdr, [] { team_sync(0); });
  __syncthreads();
  if (team == 1) trsm_team(TileC{T + 32 * LDC}, TileC{T}, dinv, w, [] { team_sync(1); });
  __syncthreads();
  {
    const int wr = warp >> 2, wc = warp & 3;  // outputs: rows 32 + wr*16.., columns 32 + wc*8..
    double acc[2][1][2];
    zero_acc(acc);
    warp_mma<2, 1, LDC>(T + (32 + wr * 16) * LDC, T + (32 + wc * 8) * LDC, 2, acc);
    sub_acc_smem(T, 32 + wr * 16, 32 + wc * 8, acc);
  }
  __syncthreads();
  if (team == 0) potrf_team(TileC{T + 32 * LDC + 32}, w, g0 + 32, D, dinv + 32, hdr, [] { team_sync(0); });
  __syncthreads();
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

__device__ __forceinline__ void run_task(const Args& a, const int4 tk, int epoch, double* smem, double* s_dinv,
                                         int slot) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  const int wr = warp >> 2, wc = warp & 3;
  const int I = tk.x & 0xffff, J = tk.x >> 16, k0 = tk.y & 0xffff, k1 = tk.y >> 16, kind = tk.z, kc = tk.w;
  chol_debug_stall(a.c, I, J, epoch, 0);
  const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
  // inputs: L(I,k) and L(J,k) for every block column of the chunk
  if (tid < 2 * (k1 - k0)) {
    const int* f = tile_flag(a, (tid & 1) ? J : I, k0 + (tid >> 1));
    while (ld_acquire(f) != epoch) __nanosleep(32);
  }
  __syncthreads();
  const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
  double acc[4][2][2];
  zero_acc(acc);
  const int nsteps = 2 * (k1 - k0);  // 32 columns per step
  for (int s = 0; s < kStages - 1; s++) {
    if (s < nsteps) issue_stage(a, I, J, k0 * MB + s * 32, smem + s * kStageD);
    cp_async_commit();
  }
  for (int s = 0; s < nsteps; s++) {
    cp_async_wait<kStages - 2>();
    __syncthreads();  // stage s has landed for everybody; stage s-1 is no longer read
    const int nx = s + kStages - 1;
    if (nx < nsteps) issue_stage(a, I, J, k0 * MB + nx * 32, smem + (nx % kStages) * kStageD);
    cp_async_commit();
    const double* st = smem + (s % kStages) * kStageD;
    warp_mma<4, 2, LDP>(st + wr * 32 * LDP, (I != J ? st + kPanel : st) + wc * 16 * LDP, 2, acc);
  }
  cp_async_wait<0>();
  if (kc > 0) wait_flag(prog_flag(a, I, J), (epoch << 12) | kc);  // earlier chunks are in (fixed order)
  else __syncthreads();
  double* tile = a.c.A + (size_t)I * MB * a.c.Dp + (size_t)J * MB;
  if (kind == kTaskUpdate) {
    if (!(I == J && wc * 16 > wr * 32 + 31)) {  // fragments strictly above the diagonal are never read
#pragma unroll
      for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++) {
          double2* p = reinterpret_cast<double2*>(tile + (size_t)(wr * 32 + mi * 8 + g) * a.c.Dp + wc * 16 + ni * 8 + 2 * q);
          double2 v = __ldcg(p);
          v.x -= acc[mi][ni][0]; v.y -= acc[mi][ni][1];
          __stcg(p, v);
        }
    }
    chol_debug_stall(a.c, I, J, epoch, 1);
    post_flag(prog_flag(a, I, J), (epoch << 12) | (kc + 1));
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
      const double2 v = __ldcg(reinterpret_cast<const double2*>(tile + (size_t)r * a.c.Dp + c));
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
  post_flag(tile_flag(a, I, J), epoch);
  trace_put(a.c, slot, 1, I, J, t_begin, t_diag);
}

// ------------------------------------------------------------------------------------------
// The chain: diagonal tiles and the sub-diagonal tile of every block column, one CTA
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void run_chain(const Args& a, int epoch, double* smem, double* s_dinv) {
  const int warp = threadIdx.x >> 5, wr = warp >> 2, wc = warp & 3;
  double* T = smem;
  double* X = smem + kTileC;
  double* Lp = smem + 2 * kTileC;
  double* dT = s_dinv;
  double* dP = s_dinv + MB;
  for (int J = 0; J < a.nbm; J++) {
    chol_debug_stall(a.c, J, J, epoch, 0);
    const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
    const int nd = gn_nch_diag(J);
    if (nd) wait_flag(prog_flag(a, J, J), (epoch << 12) | nd);
    load_tile64(a.c.A, a.c.Dp, J, J, T);
    unsigned long long t_ready = t_begin;
    if (J > 0) {
      const int ns = gn_nch_sub(J - 1);
      if (ns) wait_flag(prog_flag(a, J, J - 1), (epoch << 12) | ns);
      if (a.c.trace) t_ready = gtimer();
      load_tile64(a.c.A, a.c.Dp, J, J - 1, X);
      __syncthreads();
      trsm64(X, Lp, dP);
      store_tile64(a.c.A, a.c.Dp, J, J - 1, X);
      post_flag(tile_flag(a, J, J - 1), epoch);  // L(J,J-1) for the rest of block column J
      double acc[4][2][2];
      zero_acc(acc);
      if (!(wc * 16 > wr * 32 + 31)) {  // T -= X X^T, lower fragments only
        warp_mma<4, 2, LDC>(X + wr * 32 * LDC, X + wc * 16 * LDC, 4, acc);
        sub_acc_smem(T, wr * 32, wc * 16, acc);
      }
    }
    __syncthreads();
    potrf64(T, J * MB, a.c.D, dT, a.c.hdr);
    chol_debug_stall(a.c, J, J, epoch, 1);
    store_tile64(a.c.A, a.c.Dp, J, J, T);
    if (threadIdx.x < MB) a.c.x[a.c.Dp + J * MB + threadIdx.x] = dT[threadIdx.x];
    post_flag(tile_flag(a, J, J), epoch);
    trace_put(a.c, J, 0, J, J, t_begin, t_ready);
    double* t = T; T = Lp; Lp = t;
    t = dT; dT = dP; dP = t;
  }
}

// ------------------------------------------------------------------------------------------
// Back substitution L^T x = y (y^T = row D of L), 32 columns per task, tasks dealt round robin:
//   x_c = L_cc^-T ( y_c - sum_{m>c} L_mc^T x_m ),  entries with global index >= D are 0.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void back_task(const Args& a, int c, int epoch, double* smem, int slot) {
  constexpr int kParts = kThreads / 32, kRows = NB / kParts;
  const int tid = threadIdx.x, t = tid & 31, part = tid >> 5;
  const int nb = a.c.nb;
  double (*sA)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(smem);
  double* sx = smem + NB * (NB + 1);
  double* sred = sx + NB;
  const unsigned long long t_begin = a.c.trace ? gtimer() : 0ull;
  // the last diagonal tile depends, transitively, on every tile of the factor
  wait_flag(tile_flag(a, a.nbm - 1, a.nbm - 1), epoch);
  double s = 0.0;
  for (int m = nb - 1; m > c; m--) {
    wait_flag(x_flag(a, m), epoch);
    if (tid < NB) sx[tid] = __ldcg(a.c.x + m * NB + tid);
    __syncthreads();
    const double* Lmc = a.c.A + (size_t)m * NB * a.c.Dp + (size_t)c * NB;
#pragma unroll
    for (int r = part * kRows; r < part * kRows + kRows; r++)
      s = fma(__ldcg(Lmc + (size_t)r * a.c.Dp + t), sx[r], s);
    __syncthreads();
  }
  const unsigned long long t_ready = a.c.trace ? gtimer() : 0ull;
  if (part > 0) sred[(part - 1) * NB + t] = s;
  {
    const double* src = a.c.A + (size_t)c * NB * a.c.Dp + (size_t)c * NB;
    for (int idx = tid; idx < NB * NB; idx += kThreads) sA[idx >> 5][idx & 31] = __ldcg(src + (size_t)(idx >> 5) * a.c.Dp + (idx & 31));
  }
  __syncthreads();
  if (tid < NB) {
    const int gi = c * NB + t;
    double tot = s;
#pragma unroll
    for (int pp = 1; pp < kParts; pp++) tot += sred[(pp - 1) * NB + t];
    double rhs = (gi < a.c.D) ? __ldcg(a.c.A + (size_t)a.c.D * a.c.Dp + gi) - tot : 0.0;
    const double dinv = __ldcg(a.c.x + a.c.Dp + gi);
    double xo = 0.0;
#pragma unroll
    for (int k = NB - 1; k >= 0; k--) {
      double xk = __shfl_sync(0xffffffffu, rhs * dinv, k);
      if (c * NB + k >= a.c.D) xk = 0.0;
      if (t == k) xo = xk;
      if (t < k) rhs = fma(-sA[k][t], xk, rhs);
    }
    a.c.x[gi] = xo;
  }
  post_flag(x_flag(a, c), epoch);
  trace_put(a.c, slot, 2, c, c, t_begin, t_ready);
}

__global__ void __launch_bounds__(kThreads, 1)
chol64_kernel(const Args a) {
  if (*reinterpret_cast<volatile int*>(&a.c.hdr->done)) return;
  if (*reinterpret_cast<volatile int*>(&a.c.hdr->peer_timeout) > 1) return;  // see chol_dataflow_kernel
  extern __shared__ __align__(16) double smem[];
  __shared__ double s_dinv[2 * MB];
  __shared__ float red[kThreads / 32];
  __shared__ int s_task;
  const int epoch = a.c.hdr->iters + 1;
  const int G = gridDim.x, cta = blockIdx.x;
  if (cta == 0) {
    run_chain(a, epoch, smem, s_dinv);
  } else {
    int* queue = a.c.flags + queue_index(a.nbm);
    for (;;) {
      if (threadIdx.x == 0) s_task = atomicAdd(queue, 1);
      __syncthreads();
      const int t = s_task;
      __syncthreads();
      if (t >= a.ntasks) break;
      run_task(a, __ldg(a.tasks + t), epoch, smem, s_dinv, a.nbm + t);
    }
  }
  for (int c = a.c.nb - 1; c >= 0; c--)
    if (c % G == cta) {
      back_task(a, c, epoch, smem, a.nbm + a.ntasks + c);
      if (c == 0) {  // x complete: this CTA finishes the iteration
        __syncthreads();
        gn_update_block(a.c.Twc, a.c.x, a.c.N, a.c.D, a.c.delta_thresh, a.c.dx, a.c.hdr, red, kThreads,
                        threadIdx.x, [] { __syncthreads(); });
      }
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
  if (grid < 2) return M3R_ELAUNCH;  // the chain CTA and at least one worker must be co-resident
  if (grid > 1 + L.ntasks) grid = 1 + (L.ntasks > 0 ? L.ntasks : 1);
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
extern "C" int m3r_debug_chol64_tasklist(int nbm, int* out, int capacity) {
  using namespace m3r;
  int o = 0;
  for (int v = 0; v <= nbm; v++)
    for (int J = v; J <= v + c64::kWindow && J < nbm; J++)
      c64::tasks_of_group(nbm, v, J, [&](int I, int k0, int k1, int kind, int kc) {
        if (out && o < capacity) {
          out[4 * o] = I | (J << 16); out[4 * o + 1] = k0 | (k1 << 16); out[4 * o + 2] = kind; out[4 * o + 3] = kc;
        }
        o++;
      });
  return o == gn_chol64_ntasks(nbm) ? o : -1;
}
