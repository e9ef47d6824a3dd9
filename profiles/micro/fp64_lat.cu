// Single-warp fp64 latency probes (B200): what one potrf column step is made of.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_lat fp64_lat.cu && ./fp64_lat
#include <cstdio>
#include <cuda_runtime.h>
__global__ void probe(double* out, long long* cyc, double seed) {
  __shared__ double col[64];
  const int r = threadIdx.x;
  double x = seed + r * 1e-3;
  long long t0, t1;
  // 1. dependent DFMA chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) x = fma(x, 1.0000001, 1e-9);
  t1 = clock64(); if (r == 0) cyc[0] = (t1 - t0) / 64;
  // 2. dependent rsqrt chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; i++) x = rsqrt(x + 2.0);
  t1 = clock64(); if (r == 0) cyc[1] = (t1 - t0) / 32;
  // 3. dependent sqrt + div chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; i++) x = 1.0 / sqrt(x + 2.0);
  t1 = clock64(); if (r == 0) cyc[2] = (t1 - t0) / 32;
  // 4. dependent shfl(double) chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; i++) x = __shfl_sync(0xffffffffu, x, (i * 7) & 31) + 1.0;
  t1 = clock64(); if (r == 0) cyc[3] = (t1 - t0) / 32;
  // 5. sts + syncwarp + lds round trip (dependent)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; i++) { col[r] = x; __syncwarp(); x = col[(r + 1) & 31] + 1.0; __syncwarp(); }
  t1 = clock64(); if (r == 0) cyc[4] = (t1 - t0) / 32;
  // 6. 32 independent (LDS broadcast + DFMA) pairs
  double a[32];
#pragma unroll
  for (int k = 0; k < 32; k++) a[k] = x + k;
  col[r] = x; col[r + 32] = x * 0.5; __syncwarp();
  t0 = clock64();
#pragma unroll
  for (int rep = 0; rep < 8; rep++)
#pragma unroll
    for (int k = 0; k < 32; k++) a[k] = fma(-x, col[(k + rep) & 63], a[k]);
  t1 = clock64(); if (r == 0) cyc[5] = (t1 - t0) / 8;
  // 7. 32 independent DFMAs from registers
  t0 = clock64();
#pragma unroll
  for (int rep = 0; rep < 8; rep++)
#pragma unroll
    for (int k = 0; k < 32; k++) a[k] = fma(-x, 1.0000001, a[k]);
  t1 = clock64(); if (r == 0) cyc[6] = (t1 - t0) / 8;
  // 8. float rsqrt seed + 2 Newton steps in double
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; i++) {
    const double d = x + 2.0;
    double y = (double)rsqrtf((float)d);
    y = y * fma(-0.5 * d * y, y, 1.5);
    y = y * fma(-0.5 * d * y, y, 1.5);
    x = y;
  }
  t1 = clock64(); if (r == 0) cyc[7] = (t1 - t0) / 32;
  double s = x;
#pragma unroll
  for (int k = 0; k < 32; k++) s += a[k];
  out[r] = s;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 8 * 8);
  for (int i = 0; i < 2; i++) probe<<<1, 32>>>(out, cyc, 1.5);
  long long h[8];
  cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
  const char* n[8] = {"dfma dependent", "rsqrt(double) dependent", "1/sqrt dependent", "shfl(double)+dadd",
                      "sts+syncwarp+lds+dadd+syncwarp", "32x (lds bcast + dfma)", "32x dfma regs",
                      "rsqrtf seed + 2 NR"};
  for (int i = 0; i < 8; i++) printf("%-34s %lld cycles\n", n[i], h[i]);
  return 0;
}
