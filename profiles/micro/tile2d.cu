// Isolated timing of the 64-thread 2-D tile POTRF / TRSM of m3r_gn_solve.cu (same code, pasted by
// profiles/micro/gen_tile2d.py) with clock64 — and with parts disabled to see what dominates.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NB = 32;
struct GnHeader { int done, chol_fail, iters, pad; };
#ifndef VARIANT
#define VARIANT 0
#endif
#if VARIANT == 1   // no barrier (wrong results; timing only)
#define __syncthreads() __syncwarp()
#endif
#if VARIANT == 2   // no rsqrt
#define rsqrt(x) (x)
#endif
template <int YC>
__device__ __forceinline__ void potrf_step(double (&a)[4][4], double (*s)[NB + 1], double* colbuf,
                                           int cc, int ty, int tx, int g0, int D,
                                           double* __restrict__ dinv_g, GnHeader* hdr) {
  const int c = 8 * YC + cc;
  double* col = colbuf + (c & 1) * NB;
  if (tx == cc) {  // owners of column c publish its raw entries (rows >= 8*YC are enough)
#pragma unroll
    for (int x = YC; x < 4; x++) col[ty + 8 * x] = a[x][YC];
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
  double lr[4], lc[4];
#pragma unroll
  for (int x = YC; x < 4; x++) lr[x] = col[ty + 8 * x] * yv;
#pragma unroll
  for (int y = YC; y < 4; y++) lc[y] = col[tx + 8 * y] * yv;
  lr[YC] = (ty > cc) ? lr[YC] : 0.0;  // rows <= c of the pivot's micro-row do not move
  lc[YC] = (tx > cc) ? lc[YC] : 0.0;  // columns <= c of the pivot's micro-column are final
  if (tx == cc) {  // column c of L: 0 above the diagonal, the pivot, then the scaled entries
#pragma unroll
    for (int x = 0; x < 4; x++) {
      const int r = ty + 8 * x;
      double v = 0.0;
      if (x > YC) v = lr[x];
      else if (x == YC) v = (ty > cc) ? lr[YC] : (ty == cc ? (unit ? 1.0 : d * yv) : 0.0);
      s[r][c] = v;
    }
  }
#pragma unroll
  for (int x = YC; x < 4; x++)
#pragma unroll
    for (int y = YC; y < 4; y++) a[x][y] = fma(-lr[x], lc[y], a[x][y]);
}

// In-place Cholesky of the 32x32 tile in `s` by the whole 64-thread CTA (see above).  The factor
// is written to `s` column by column; the registers only carry the trailing matrix.
__device__ __forceinline__ void potrf_tile(double (*s)[NB + 1], double* colbuf, int g0, int D,
                                           double* __restrict__ dinv_g, GnHeader* hdr) {
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  double a[4][4];  // the strict upper triangle is zeroed: it takes part in the (symmetric) updates
#pragma unroll
  for (int x = 0; x < 4; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) a[x][y] = (tx + 8 * y <= ty + 8 * x) ? s[ty + 8 * x][tx + 8 * y] : 0.0;
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
__device__ __forceinline__ void trsm_step(double (&a)[4][4], double (*sx)[NB + 1],
                                          const double (*sl)[NB + 1], double* colbuf,
                                          const double* dinv, int cc, int ty, int tx) {
  const int c = 8 * YC + cc;
  double* col = colbuf + (c & 1) * NB;
  if (tx == cc) {  // column c of X is final after scaling by 1/L[c][c]
    const double di = dinv[c];
#pragma unroll
    for (int x = 0; x < 4; x++) {
      const double v = a[x][YC] * di;
      col[ty + 8 * x] = v;
      sx[ty + 8 * x][c] = v;
    }
  }
  __syncthreads();
  double xr[4], lc[4];
#pragma unroll
  for (int x = 0; x < 4; x++) xr[x] = col[ty + 8 * x];
#pragma unroll
  for (int y = YC; y < 4; y++) lc[y] = sl[tx + 8 * y][c];
  lc[YC] = (tx > cc) ? lc[YC] : 0.0;
#pragma unroll
  for (int x = 0; x < 4; x++)
#pragma unroll
    for (int y = YC; y < 4; y++) a[x][y] = fma(-xr[x], lc[y], a[x][y]);
}

// X <- X * Ljj^-T for the 32x32 tile X in `sx`, Ljj in `sl`: same 8x8 thread grid; per column
// step the finished column of X is broadcast through the double-buffered shared vector.
__device__ __forceinline__ void trsm_tile(double (*sx)[NB + 1], const double (*sl)[NB + 1],
                                          double* colbuf, double* dinv,
                                          const double* __restrict__ dinv_g) {
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  if (tid < NB) dinv[tid] = __ldcg(dinv_g + tid);
  double a[4][4];
#pragma unroll
  for (int x = 0; x < 4; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) a[x][y] = sx[ty + 8 * x][tx + 8 * y];
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


__global__ void run(double* A, double* dinv_g, GnHeader* hdr, long long* out) {
  __shared__ double sA[NB][NB + 1], sB[NB][NB + 1];
  __shared__ double sv[2 * NB], sred[NB];
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 64) { sA[i / NB][i % NB] = A[i]; sB[i / NB][i % NB] = A[i] * 0.5; }
  __syncthreads();
  long long c0 = clock64();
  potrf_tile(sA, sv, 0, 1000, dinv_g, hdr);
  __syncthreads();
  long long c1 = clock64();
  trsm_tile(sB, sA, sv, sred, dinv_g);
  __syncthreads();
  long long c2 = clock64();
  if (tid == 0) { out[0] = c1 - c0; out[1] = c2 - c1; }
  for (int i = tid; i < NB * NB; i += 64) A[i] = sA[i / NB][i % NB] + sB[i / NB][i % NB];
}
int main() {
  double h[NB * NB];
  for (int i = 0; i < NB; i++) for (int j = 0; j < NB; j++) h[i * NB + j] = (i == j ? 40.0 : 0.0) + 1.0 / (1 + abs(i - j));
  double *A, *dinv; GnHeader* hdr; long long* out;
  cudaMalloc(&A, sizeof(h)); cudaMalloc(&dinv, NB * 8); cudaMalloc(&hdr, 16); cudaMalloc(&out, 16);
  for (int rep = 0; rep < 3; rep++) {
    cudaMemcpy(A, h, sizeof(h), cudaMemcpyHostToDevice);
    run<<<1, 64>>>(A, dinv, hdr, out);
    long long o[2]; cudaMemcpy(o, out, 16, cudaMemcpyDeviceToHost);
    printf("variant %d rep %d: potrf %lld cycles, trsm %lld cycles\n", VARIANT, rep, o[0], o[1]);
  }
  return 0;
}
