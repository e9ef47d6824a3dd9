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
// lower triangle), so the blocked right-looking Cholesky also performs the forward
// substitution (row D of L is y^T = (L^-1 b)^T); a blocked back substitution gives x and
// dx = -x (the reference's `dx = -A.solve()`).  A non-positive pivot sets chol_fail and the
// step becomes zero, as SimplicialLLT failure does in the reference (:142-150).
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

// ---------------------------------------------------------------------------------
// Blocked right-looking Cholesky, panel width NB = 32.
// Panel kernel: every CTA factors the diagonal block (redundantly, in shared memory),
// CTA 0 stores it, CTA c >= 1 solves block row k+c:  L_ik = A_ik L_kk^-T.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
chol_panel_kernel(double* __restrict__ A, int Dp, int D, int k, GnHeader* hdr) {
  if (hdr->done) return;
  __shared__ double a[NB][NB + 1];
  __shared__ double bmat[NB][NB + 1];
  __shared__ double piv_s;
  const int tid = threadIdx.x;
  const int k0 = k * NB;
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int i = idx / NB, j = idx % NB;
    a[i][j] = (j <= i) ? A[(size_t)(k0 + i) * Dp + k0 + j] : 0.0;
  }
  const int c = blockIdx.x;
  if (c > 0) {
    const int i0 = (k + c) * NB;
    for (int idx = tid; idx < NB * NB; idx += 256) {
      const int i = idx / NB, j = idx % NB;
      bmat[i][j] = A[(size_t)(i0 + i) * Dp + k0 + j];
    }
  }
  __syncthreads();
  for (int j = 0; j < NB; j++) {
    if (tid == 0) {
      const double d = a[j][j];
      double p;
      if (k0 + j >= D) p = 1.0;           // augmented row / padding: pivot not used
      else if (d > 0.0) p = sqrt(d);
      else { p = 1.0; if (c == 0) hdr->chol_fail = 1; }
      a[j][j] = p;
      piv_s = p;
    }
    __syncthreads();
    if (tid > j && tid < NB) a[tid][j] /= piv_s;
    __syncthreads();
    for (int idx = tid; idx < NB * NB; idx += 256) {
      const int i = idx / NB, cc = idx % NB;
      if (cc > j && cc <= i) a[i][cc] -= a[i][j] * a[cc][j];
    }
    __syncthreads();
  }
  if (c == 0) {
    for (int idx = tid; idx < NB * NB; idx += 256) {
      const int i = idx / NB, j = idx % NB;
      if (j <= i) A[(size_t)(k0 + i) * Dp + k0 + j] = a[i][j];
    }
  } else {
    // row-wise forward substitution against L_kk^T: one thread per row of the block
    if (tid < NB) {
      const int i = tid;
      for (int j = 0; j < NB; j++) {
        double s = bmat[i][j];
        for (int m = 0; m < j; m++) s -= bmat[i][m] * a[j][m];
        bmat[i][j] = s / a[j][j];
      }
    }
    __syncthreads();
    const int i0 = (k + c) * NB;
    for (int idx = tid; idx < NB * NB; idx += 256) {
      const int i = idx / NB, j = idx % NB;
      A[(size_t)(i0 + i) * Dp + k0 + j] = bmat[i][j];
    }
  }
}

// Trailing update A_ij -= L_ik L_jk^T over 64x64 tiles (lower tiles only), 4x4 per thread.
__global__ void __launch_bounds__(256)
chol_syrk_kernel(double* __restrict__ A, int Dp, int k, const GnHeader* hdr) {
  if (hdr->done) return;
  __shared__ double Li[64][NB + 1];
  __shared__ double Lj[64][NB + 1];
  // linear tile index -> (ti >= tj)
  int ti = (int)((sqrtf(8.f * (float)blockIdx.x + 1.f) - 1.f) * 0.5f);
  while ((ti + 1) * (ti + 2) / 2 <= (int)blockIdx.x) ti++;
  while (ti * (ti + 1) / 2 > (int)blockIdx.x) ti--;
  const int tj = blockIdx.x - ti * (ti + 1) / 2;
  const int base = (k + 1) * NB;
  const int i0 = base + ti * 64, j0 = base + tj * 64;
  const int k0 = k * NB;
  const int tid = threadIdx.x;
  for (int idx = tid; idx < 64 * NB; idx += 256) {
    const int r = idx / NB, cidx = idx % NB;
    Li[r][cidx] = (i0 + r < Dp) ? A[(size_t)(i0 + r) * Dp + k0 + cidx] : 0.0;
    Lj[r][cidx] = (j0 + r < Dp) ? A[(size_t)(j0 + r) * Dp + k0 + cidx] : 0.0;
  }
  __syncthreads();
  const int tr = tid / 16, tc = tid % 16;  // rows tr+16*a, cols tc+16*b
  double acc[4][4];
#pragma unroll
  for (int x = 0; x < 4; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) acc[x][y] = 0.0;
#pragma unroll 8
  for (int m = 0; m < NB; m++) {
    double li[4], lj[4];
#pragma unroll
    for (int x = 0; x < 4; x++) { li[x] = Li[tr + 16 * x][m]; lj[x] = Lj[tc + 16 * x][m]; }
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) acc[x][y] = fma(li[x], lj[y], acc[x][y]);
  }
#pragma unroll
  for (int x = 0; x < 4; x++) {
    const int row = i0 + tr + 16 * x;
    if (row >= Dp) continue;
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const int col = j0 + tc + 16 * y;
      if (col <= row) A[(size_t)row * Dp + col] -= acc[x][y];
    }
  }
}

// y <- row D of L (the forward-substituted right-hand side), padded with zeros.
__global__ void chol_fetch_y_kernel(const double* __restrict__ A, int Dp, int D,
                                    double* __restrict__ y, const GnHeader* hdr) {
  if (hdr->done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Dp) y[i] = i < D ? A[(size_t)D * Dp + i] : 0.0;
}

// Back substitution block k: x_k = L_kk^-T y_k (every CTA, redundantly); CTA c < k then
// updates y_c -= L[k-block rows, c-block cols]^T x_k; CTA k stores x_k.
__global__ void __launch_bounds__(NB * 8)
chol_back_kernel(const double* __restrict__ A, int Dp, int D, int k, double* __restrict__ y,
                 double* __restrict__ xsol, const GnHeader* hdr) {
  if (hdr->done) return;
  __shared__ double l[NB][NB + 1];
  __shared__ double xk[NB];
  __shared__ double red[8][NB];
  const int tid = threadIdx.x;
  const int k0 = k * NB;
  for (int idx = tid; idx < NB * NB; idx += NB * 8) {
    const int i = idx / NB, j = idx % NB;
    l[i][j] = (j <= i) ? A[(size_t)(k0 + i) * Dp + k0 + j] : 0.0;
  }
  if (tid < NB) xk[tid] = (k0 + tid < D) ? y[k0 + tid] : 0.0;
  __syncthreads();
  if (tid == 0) {
    for (int i = NB - 1; i >= 0; i--) {
      if (k0 + i >= D) { xk[i] = 0.0; continue; }
      double s = xk[i];
      for (int m = i + 1; m < NB; m++) s -= l[m][i] * xk[m];
      xk[i] = s / l[i][i];
    }
  }
  __syncthreads();
  const int c = blockIdx.x;
  if (c == k) {
    if (tid < NB) xsol[k0 + tid] = xk[tid];  // separate array: other CTAs still read y_k
    return;
  }
  // y_c[m] -= sum_r L[k0+r][c*NB+m] * x_k[r] : thread (part, m) sums rows r = part, part+8, ...
  const int m = tid % NB, part = tid / NB;
  double s = 0.0;
  for (int r = part; r < NB; r += 8) s += A[(size_t)(k0 + r) * Dp + c * NB + m] * xk[r];
  red[part][m] = s;
  __syncthreads();
  if (tid < NB) {
    double tot = 0.0;
    for (int p = 0; p < 8; p++) tot += red[p][tid];
    y[c * NB + tid] -= tot;
  }
}

// dx = -x (zeros on failure), retract, |dx| test  (gn_kernels.cu:1209-1222)
__global__ void __launch_bounds__(256)
gn_update_kernel(float* __restrict__ Twc, const double* __restrict__ x, int N, int D,
                 float delta_thresh, float* __restrict__ dx, GnHeader* hdr) {
  if (hdr->done) return;
  __shared__ float red[8];
  const bool fail = hdr->chol_fail != 0;
  float ss = 0.f;
  for (int i = threadIdx.x; i < D; i += 256) {
    const float v = fail ? 0.f : (float)(-x[i]);
    dx[i] = v;
    ss += v * v;
  }
  __syncthreads();  // dx visible to the retraction below (same CTA)
  for (int p = 1 + threadIdx.x; p < N; p += 256) {
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
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int w = 0; w < 8; w++) tot += red[w];
    hdr->iters += 1;
    if (sqrtf(tot) < delta_thresh) hdr->done = 1;
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

int gn_solve_launch(float* Twc, const double* blocks, int N, int M, float delta_thresh, float* dx,
                    void* ws, const GnLayout& L, cudaStream_t st) {
  GnHeader* hdr = ws_ptr<GnHeader>(ws, L.hdr);
  double* A = ws_ptr<double>(ws, L.A);
  double* x = ws_ptr<double>(ws, L.x);
  const int D = L.D, Dp = L.Dp, nb = Dp / NB;
  gn_assemble_kernel<<<N, 64, 0, st>>>(blocks, ws_ptr<int>(ws, L.ie), ws_ptr<int>(ws, L.je),
                                       ws_ptr<int>(ws, L.row_ptr), ws_ptr<int>(ws, L.adj), M, N, D,
                                       Dp, A, hdr);
  M3R_TRY(after_launch());
  for (int k = 0; k < nb; k++) {
    chol_panel_kernel<<<nb - k, 256, 0, st>>>(A, Dp, D, k, hdr);
    M3R_TRY(after_launch());
    const int rem = nb - k - 1;
    if (rem > 0) {
      const int T = (rem + 1) / 2;
      chol_syrk_kernel<<<T * (T + 1) / 2, 256, 0, st>>>(A, Dp, k, hdr);
      M3R_TRY(after_launch());
    }
  }
  chol_fetch_y_kernel<<<(Dp + 255) / 256, 256, 0, st>>>(A, Dp, D, x, hdr);
  M3R_TRY(after_launch());
  const int kb = (D - 1) / NB;  // last block holding unknowns
  for (int k = kb; k >= 0; k--) {
    chol_back_kernel<<<k + 1, NB * 8, 0, st>>>(A, Dp, D, k, x, x + Dp, hdr);
    M3R_TRY(after_launch());
  }
  gn_update_kernel<<<1, 256, 0, st>>>(Twc, x + Dp, N, D, delta_thresh, dx, hdr);
  return after_launch();
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix,
                             m3r_stream_t stream) {
  if (num_poses < 0 || num_fix < 0 || !poses || !dx) return M3R_EINVAL;
  const int cnt = num_poses - num_fix;
  if (cnt <= 0) return M3R_OK;
  pose_retr_kernel<<<(cnt + 127) / 128, 128, 0, as_stream(stream)>>>(poses, dx, num_poses, num_fix);
  return after_launch();
}
