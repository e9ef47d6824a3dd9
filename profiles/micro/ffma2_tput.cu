// FFMA vs FFMA2 (fma.rn.f32x2, sm_100) issue/throughput probe: 148x8 CTAs of 256 threads, 16 independent
// accumulator chains per thread.  Prints GFMA/s (scalar FMAs per second) for both forms.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_ffma(float* out, float a, float b, int iters) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x + i;
  for (int it = 0; it < iters; it++)
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = fmaf(acc[i], a, b);
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b, int iters) {
  float2 acc[8];
#pragma unroll
  for (int i = 0; i < 8; i++) acc[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.999f);
  for (int it = 0; it < iters; it++)
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = __ffma2_rn(acc[i], a2, b2);
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int v = 0; v < 2; v++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(e0);
      if (v == 0) k_ffma<<<148 * 8, 256>>>(out, 1.0001f, 0.5f, iters);
      else k_ffma2<<<148 * 8, 256>>>(out, 1.0001f, 0.5f, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      const double fmas = 148.0 * 8 * 256 * 16.0 * iters;
      if (rep) printf("%s: %.3f ms, %.1f TFMA/s (x2 = TFLOP/s %.1f)\n", v ? "FFMA2" : "FFMA ", ms, fmas / ms / 1e9, 2 * fmas / ms / 1e9);
    }
  }
  return 0;
}
