// fp64 throughput on B200: mma.sync.m8n8k4.f64 (DMMA) vs DFMA, register operands only.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma_tput dmma_tput.cu && ./dmma_tput
// Prints FMA per clock per SM for 1 / 4 / 8 / 16 / 32 warps per SM (1 CTA per SM), and the
// dependent-issue latency of one DMMA.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int ACC>
__global__ void k_dmma(double* out, int iters, long long* cyc) {
  double c[ACC][2];
#pragma unroll
  for (int i = 0; i < ACC; i++) { c[i][0] = threadIdx.x * 1e-3; c[i][1] = i; }
  double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ACC; i++) dmma(c[i][0], c[i][1], a, b);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ACC; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int ACC>
__global__ void k_dfma(double* out, int iters, long long* cyc) {
  double c[ACC];
#pragma unroll
  for (int i = 0; i < ACC; i++) c[i] = threadIdx.x * 1e-3 + i;
  double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ACC; i++) c[i] = fma(c[i], a, b);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ACC; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * sms * 1024);
  cudaMallocManaged(&cyc, sizeof(long long));
  const int iters = 4096;
  for (int warps : {1, 4, 8, 16, 32}) {
    k_dmma<8><<<sms, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    k_dmma<8><<<sms, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    const double fma_dmma = (double)warps * iters * 8 * 256 / (double)cyc[0];
    k_dfma<16><<<sms, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    k_dfma<16><<<sms, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    const double fma_dfma = (double)warps * iters * 16 * 32 / (double)cyc[0];
    printf("%2d warps/SM: DMMA %.1f FMA/clk/SM   DFMA %.1f FMA/clk/SM\n", warps, fma_dmma, fma_dfma);
  }
  k_dmma<1><<<1, 32>>>(out, iters, cyc); cudaDeviceSynchronize();
  printf("dependent DMMA latency: %.1f cycles\n", (double)cyc[0] / iters);
  k_dmma<2><<<1, 32>>>(out, iters, cyc); cudaDeviceSynchronize();
  printf("2 independent DMMA per warp: %.1f cycles per pair\n", (double)cyc[0] / iters);
  k_dmma<8><<<1, 32>>>(out, iters, cyc); cudaDeviceSynchronize();
  printf("8 independent DMMA per warp: %.1f cycles per 8\n", (double)cyc[0] / iters);
  cudaError_t e = cudaGetLastError();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
