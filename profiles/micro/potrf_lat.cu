// One-warp 32x32 fp64 tile Cholesky variants, timed with clock64 and globaltimer (B200).
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NB = 32;
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t;
}
// A: fully unrolled, lane r owns row r (the shipped formulation)
__device__ void potrf_unrolled(double (*s)[NB + 1], double* col) {
  const int r = threadIdx.x;
  double a[NB];
#pragma unroll
  for (int c = 0; c < NB; c++) a[c] = s[r][c];
#pragma unroll
  for (int c = 0; c < NB; c++) {
    const double d = __shfl_sync(0xffffffffu, a[c], c);
    const double y = rsqrt(d);
    const double l = (r == c) ? d * y : a[c] * y;
    a[c] = l; col[r] = l;
    __syncwarp();
#pragma unroll
    for (int c2 = c + 1; c2 < NB; c2++) if (r >= c2) a[c2] = fma(-l, col[c2], a[c2]);
    __syncwarp();
  }
#pragma unroll
  for (int c = 0; c < NB; c++) s[r][c] = (c <= r) ? a[c] : 0.0;
}
// B: lane c owns COLUMN c; the scaled column is broadcast by shuffles from the owner lane
// (no shared memory in the loop): step c: lane c scales its column; every lane j>c needs
// l[j] (from lane c's register j -> dynamic register index is impossible), so instead
// row-owner + shuffle: l_c2 = shfl(l, c2) gives lane c2's l, which is exactly col[c2].
__device__ void potrf_shfl(double (*s)[NB + 1]) {
  const int r = threadIdx.x;
  double a[NB];
#pragma unroll
  for (int c = 0; c < NB; c++) a[c] = s[r][c];
#pragma unroll
  for (int c = 0; c < NB; c++) {
    const double d = __shfl_sync(0xffffffffu, a[c], c);
    const double y = rsqrt(d);
    const double l = (r == c) ? d * y : a[c] * y;
    a[c] = l;
#pragma unroll
    for (int c2 = c + 1; c2 < NB; c2++) {
      const double lc = __shfl_sync(0xffffffffu, l, c2);
      if (r >= c2) a[c2] = fma(-l, lc, a[c2]);
    }
  }
#pragma unroll
  for (int c = 0; c < NB; c++) s[r][c] = (c <= r) ? a[c] : 0.0;
}
__global__ void run(double* A, long long* out, int variant) {
  __shared__ double s[NB][NB + 1];
  __shared__ double col[NB];
  const int r = threadIdx.x;
  for (int c = 0; c < NB; c++) s[r][c] = A[r * NB + c];
  __syncwarp();
  const long long c0 = clock64(); const unsigned long long g0 = gtimer();
  if (variant == 0) potrf_unrolled(s, col); else potrf_shfl(s);
  __syncwarp();
  const long long c1 = clock64(); const unsigned long long g1 = gtimer();
  if (r == 0) { out[0] = c1 - c0; out[1] = (long long)(g1 - g0); }
  for (int c = 0; c < NB; c++) A[r * NB + c] = s[r][c];
}
int main() {
  double h[NB * NB];
  for (int i = 0; i < NB; i++) for (int j = 0; j < NB; j++) h[i * NB + j] = (i == j ? 40.0 : 0.0) + 1.0 / (1 + abs(i - j));
  double* A; long long* out; cudaMalloc(&A, sizeof(h)); cudaMalloc(&out, 16);
  for (int v = 0; v < 2; v++)
    for (int rep = 0; rep < 3; rep++) {
      cudaMemcpy(A, h, sizeof(h), cudaMemcpyHostToDevice);
      run<<<1, 32>>>(A, out, v);
      long long o[2]; cudaMemcpy(o, out, 16, cudaMemcpyDeviceToHost);
      printf("variant %d rep %d: %lld cycles, %lld ns\n", v, rep, o[0], o[1]);
    }
  return 0;
}
