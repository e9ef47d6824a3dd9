// Phase timing (clock64) of the one-warp tile POTRF / TRSM of m3r_gn_solve.cu in isolation.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o warp_tile warp_tile.cu && ./warp_tile
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
constexpr int NB = 32;
struct GnHeader { int done, chol_fail, iters, pad; };
struct TilePad {
  double* p;
  __device__ __forceinline__ double& operator()(int r, int c) const { return p[r * (NB + 1) + c]; }
};
__device__ long long g_t[8];

// 4x4 Cholesky of the diagonal block held (lower) in b, redundantly by every lane.
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

// rank-4 update of own-row entries t[0..CNT) at columns jj0.. with the finished panel at column p0:
// all loads first, then the arithmetic, then (optionally) the stores, so that nothing serialises on
// shared-memory aliasing.
template <class Tile, int CNT>
__device__ __forceinline__ void rank4_rows(const Tile s, int r, int p0, int jj0, const double (&x)[4], double (&t)[CNT],
                                           bool store) {
  double v[CNT][4];
#pragma unroll
  for (int u = 0; u < CNT; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) v[u][k] = s(jj0 + u, p0 + k);
    t[u] = s(r, jj0 + u);
  }
#pragma unroll
  for (int u = 0; u < CNT; u++) {
#pragma unroll
    for (int k = 0; k < 4; k++) t[u] = fma(-x[k], v[u][k], t[u]);
  }
  if (store) {
#pragma unroll
    for (int u = 0; u < CNT; u++)
      if (jj0 + u <= r) s(r, jj0 + u) = t[u];
  }
}

template <class Tile>
__device__ __forceinline__ void potrf_warp(const Tile s, int g0, int D, double* __restrict__ dinv_g, GnHeader* hdr,
                                           long long* tphase) {
  const int r = threadIdx.x & 31;
  double x[4] = {0.0, 0.0, 0.0, 0.0};  // own row of the previous panel
  double cur[4];                       // own row, columns of the current panel (up to date)
#pragma unroll
  for (int j = 0; j < 4; j++) cur[j] = s(r, j);
#pragma unroll 1
  for (int c0 = 0; c0 < NB; c0 += 4) {
    long long t0 = clock64();
    double b[4][4], rr[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) b[i][j] = s(c0 + i, c0 + j);
    // deferred: previous panel's update of the columns right of this panel (independent of the
    // factor chain below; ptxas interleaves the two)
    if (c0 >= 4) {
#pragma unroll 1
      for (int jj = c0 + 4; jj < NB; jj += 4) {
        double t[4];
        rank4_rows<Tile, 4>(s, r, c0 - 4, jj, x, t, true);
      }
    }
    const bool fail = factor4(b, rr, g0 + c0, D);
    if (r == 0) {
      if (fail) hdr->chol_fail = 1;
#pragma unroll
      for (int j = 0; j < 4; j++) dinv_g[g0 + c0 + j] = rr[j];
    }
    long long t1 = clock64();
    {  // own row against the block: same operation order as the block itself
      const int i = r - c0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        double v = cur[j];
#pragma unroll
        for (int k = 0; k < j; k++) v = fma(-x[k], b[j][k], v);
        x[j] = v * rr[j];
        if (i == j) x[j] = b[j][j];  // the pivot itself (1 for unit rows)
        if (i < j) x[j] = 0.0;       // above the diagonal (and rows above the panel: unused)
      }
      if (i >= 0) {
#pragma unroll
        for (int j = 0; j < 4; j++) s(r, c0 + j) = x[j];
      }
    }
    __syncwarp();
    long long t2 = clock64();
    // look-ahead: the next panel's columns of the own row, kept in registers and stored
    if (c0 + 4 < NB) {
      rank4_rows<Tile, 4>(s, r, c0, c0 + 4, x, cur, true);
    }
    __syncwarp();
    long long t3 = clock64();
    if (r == 0) { tphase[0] += t1 - t0; tphase[1] += t2 - t1; tphase[2] += t3 - t2; }
  }
}

// X <- X * Ljj^-T, lane r solves row r; four columns per trip: the part against finished
// columns is a load-only loop (pipelined), the 4x4 triangle is solved in registers.
template <class Tile>
__device__ __forceinline__ void trsm_warp(const Tile sx, const Tile sl, const double* __restrict__ dinv) {
  const int r = threadIdx.x & 31;
#pragma unroll 1
  for (int c0 = 0; c0 < NB; c0 += 4) {
    double a[4];
#pragma unroll
    for (int j = 0; j < 4; j++) a[j] = sx(r, c0 + j);
#pragma unroll 4
    for (int m = 0; m < c0; m++) {
      const double xm = sx(r, m);
#pragma unroll
      for (int j = 0; j < 4; j++) a[j] = fma(-xm, sl(c0 + j, m), a[j]);
    }
    double l[4][4];
#pragma unroll
    for (int j = 1; j < 4; j++)
#pragma unroll
      for (int k = 0; k < j; k++) l[j][k] = sl(c0 + j, c0 + k);
#pragma unroll
    for (int j = 0; j < 4; j++) {
#pragma unroll
      for (int k = 0; k < j; k++) a[j] = fma(-a[k], l[j][k], a[j]);
      a[j] *= dinv[c0 + j];
    }
#pragma unroll
    for (int j = 0; j < 4; j++) sx(r, c0 + j) = a[j];
  }
}

__global__ void __launch_bounds__(128) k(const double* A, double* out, double* dinv_g, GnHeader* hdr) {
  __shared__ double sA[NB][NB + 1], sB[NB][NB + 1], sv[NB];
  for (int i = threadIdx.x; i < NB * NB; i += 128) { sA[i / NB][i % NB] = A[i]; sB[i / NB][i % NB] = A[NB * NB + i]; }
  __syncthreads();
  long long tp[3] = {0, 0, 0};
  const TilePad tA{&sA[0][0]}, tB{&sB[0][0]};
  long long t0 = clock64();
  if (threadIdx.x < 32) potrf_warp(tA, 0, 1000, dinv_g, hdr, tp);
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x < NB) sv[threadIdx.x] = dinv_g[threadIdx.x];
  __syncthreads();
  long long t2 = clock64();
  if (threadIdx.x < 32) trsm_warp(tB, tA, sv);
  __syncthreads();
  long long t3 = clock64();
  if (threadIdx.x == 0) { g_t[0] = t1 - t0; g_t[1] = t3 - t2; g_t[2] = tp[0]; g_t[3] = tp[1]; g_t[4] = tp[2]; }
  for (int i = threadIdx.x; i < NB * NB; i += 128) { out[i] = sA[i / NB][i % NB]; out[NB * NB + i] = sB[i / NB][i % NB]; }
}

int main() {
  static double hA[2 * NB * NB], hO[2 * NB * NB];
  for (int r = 0; r < NB; r++)
    for (int c = 0; c < NB; c++) {
      hA[r * NB + c] = (r == c ? 40.0 : 0.0) + 1.0 / (1 + abs(r - c)) + 0.01 * ((r * 7 + c * 3) % 5 + (c * 7 + r * 3) % 5);
      hA[NB * NB + r * NB + c] = sin(0.1 * r + 0.37 * c);
    }
  double *dA, *dO, *dD; GnHeader* dH;
  cudaMalloc(&dA, sizeof(hA)); cudaMalloc(&dO, sizeof(hA)); cudaMalloc(&dD, 8 * NB); cudaMalloc(&dH, sizeof(GnHeader));
  cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice); cudaMemset(dH, 0, sizeof(GnHeader));
  for (int it = 0; it < 3; it++) k<<<1, 128>>>(dA, dO, dD, dH);
  cudaDeviceSynchronize();
  long long t[8]; cudaMemcpyFromSymbol(t, g_t, sizeof(t));
  cudaMemcpy(hO, dO, sizeof(hO), cudaMemcpyDeviceToHost);
  // check: L L^T == A (lower), X L^T == B
  double e1 = 0, e2 = 0;
  for (int r = 0; r < NB; r++)
    for (int c = 0; c <= r; c++) {
      double s = 0; for (int m = 0; m <= c; m++) s += hO[r * NB + m] * hO[c * NB + m];
      e1 = fmax(e1, fabs(s - hA[r * NB + c]));
    }
  for (int r = 0; r < NB; r++)
    for (int c = 0; c < NB; c++) {
      double s = 0; for (int m = 0; m <= c; m++) s += hO[NB * NB + r * NB + m] * hO[c * NB + m];
      e2 = fmax(e2, fabs(s - hA[NB * NB + r * NB + c]));
    }
  printf("potrf %lld cycles (factor4x4 %lld, own-row %lld, trailing %lld)  trsm %lld cycles   err %.2e %.2e  (%s)\n", t[0], t[2],
         t[3], t[4], t[1], e1, e2, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
