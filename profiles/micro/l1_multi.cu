// Cost of the 3 x LD.128 that fetch one 48-byte descriptor per lane (lane stride 48 B), by issue pattern (B200).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_multi l1_multi.cu && ./l1_multi
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ unsigned long long g_cycles;

__device__ __forceinline__ uint32_t ld16(const char* p) {
  uint32_t a, b, c, d;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  return a ^ b ^ c ^ d;
}

// MODE 0: one candidate per step, K chunks (+16 each).  MODE 1: two candidates (rows 24 KB apart), chunks interleaved
// A0 B0 A1 B1 A2 B2.  MODE 2: two candidates, A0 A1 A2 B0 B1 B2.  WARPS: resident warps per SM.
template <int MODE, int K, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k(const char* __restrict__ buf, int steps, uint32_t* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t acc = 0;
  unsigned off = (unsigned)warp * 1536u;
  const long long t0 = clock64();
  for (int s = 0; s < steps; s++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
      off = (off + 240u) & 16383u;
      const char* pa = buf + (off & ~15u) + lane * 48;
      const char* pb = pa + 24576;
      if (MODE == 0) {
#pragma unroll
        for (int w = 0; w < K; w++) acc ^= ld16(pa + 16 * w);
      } else if (MODE == 1) {
#pragma unroll
        for (int w = 0; w < K; w++) { acc ^= ld16(pa + 16 * w); acc ^= ld16(pb + 16 * w); }
      } else {
#pragma unroll
        for (int w = 0; w < K; w++) acc ^= ld16(pa + 16 * w);
#pragma unroll
        for (int w = 0; w < K; w++) acc ^= ld16(pb + 16 * w);
      }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) atomicMax(&g_cycles, (unsigned long long)(t1 - t0));
}

template <int MODE, int K, int THREADS>
void run(const char* buf, uint32_t* out) {
  const int steps = 10000;
  unsigned long long zero = 0;
  k<MODE, K, THREADS><<<148, THREADS>>>(buf, 100, out);
  cudaMemcpyToSymbol(g_cycles, &zero, sizeof(zero));
  k<MODE, K, THREADS><<<148, THREADS>>>(buf, steps, out);
  cudaDeviceSynchronize();
  unsigned long long cc; cudaMemcpyFromSymbol(&cc, g_cycles, sizeof(cc));
  const double n = (double)steps * 4 * (THREADS / 32) * K * (MODE == 0 ? 1 : 2);
  printf("mode %d, %d chunks per candidate, %2d warps/SM: %.2f clk per warp LD.128 per SM (%s)\n", MODE, K, THREADS / 32,
         cc / n, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  char* buf; uint32_t* out;
  cudaMalloc(&buf, 1 << 17); cudaMemset(buf, 1, 1 << 17);
  cudaMalloc(&out, 148 * 1024 * 4);
  run<0, 1, 768>(buf, out); run<0, 2, 768>(buf, out); run<0, 3, 768>(buf, out);
  run<1, 3, 768>(buf, out); run<2, 3, 768>(buf, out);
  run<0, 3, 256>(buf, out); run<0, 3, 512>(buf, out); run<0, 3, 1024>(buf, out);
  run<1, 3, 256>(buf, out); run<1, 3, 1024>(buf, out);
  return 0;
}
