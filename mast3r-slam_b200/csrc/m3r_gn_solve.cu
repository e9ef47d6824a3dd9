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
                   GnHeader* hdr) {
  if (hdr->done) return;
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
      const double v = blocks[(size_t)e * kGnBlock + t];
      acc += v;
      if (other == r) acc -= v;  // self edge
      else if (other >= 0 && other < r) A[(size_t)(7 * r + k) * Dp + 7 * other + l] -= v;
    } else if (t < 56) {
      const double v = blocks[(size_t)e * kGnBlock + t];
      acc += role_i ? -v : v;
    }
  }
  if (t < 49) {
    if (l <= k) A[(size_t)(7 * r + k) * Dp + 7 * r + l] = acc;
  } else if (t < 56) {
    A[(size_t)D * Dp + 7 * r + (t - 49)] = acc;
  }
}

