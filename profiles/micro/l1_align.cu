// What sets the cost of a 128-bit warp load that hits L1 (B200)?  Alignment of the warp's base, lane stride,
// and how many lines the warp touches.  1 CTA of 768 threads per SM, clock64 timing.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_align l1_align.cu && ./l1_align
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ unsigned long long g_cycles;

__device__ __forceinline__ uint32_t ld16(const char* p) {
  uint32_t a, b, c, d;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  return a ^ b ^ c ^ d;
}

// base of the warp = (off & ~(ALIGN-1)) + SKEW; lane address = base + lane*STRIDE
template <int ALIGN, int SKEW, int STRIDE>
__global__ void __launch_bounds__(768, 1) k(const char* __restrict__ buf, int steps, uint32_t* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t acc = 0;
  unsigned off = (unsigned)warp * 2048u;
  const long long t0 = clock64();
  for (int s = 0; s < steps; s++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      off = (off + 1024u + 128u) & 32767u;
      acc ^= ld16(buf + (off & ~(unsigned)(ALIGN - 1)) + SKEW + lane * STRIDE);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) atomicMax(&g_cycles, (unsigned long long)(t1 - t0));
}

template <int ALIGN, int SKEW, int STRIDE>
void run(const char* buf, uint32_t* out) {
  const int steps = 20000;
  unsigned long long zero = 0;
  k<ALIGN, SKEW, STRIDE><<<148, 768>>>(buf, 100, out);
  cudaMemcpyToSymbol(g_cycles, &zero, sizeof(zero));
  k<ALIGN, SKEW, STRIDE><<<148, 768>>>(buf, steps, out);
  cudaDeviceSynchronize();
  unsigned long long cc; cudaMemcpyFromSymbol(&cc, g_cycles, sizeof(cc));
  const double n = (double)steps * 8 * 24;
  printf("base %% %3d == %3d, lane stride %2d B: %.2f clk per warp LD.128 per SM (%s)\n", ALIGN, SKEW, STRIDE, cc / n,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  char* buf; uint32_t* out;
  cudaMalloc(&buf, 1 << 17); cudaMemset(buf, 1, 1 << 17);
  cudaMalloc(&out, 148 * 768 * 4);
  run<128, 0, 16>(buf, out); run<128, 16, 16>(buf, out); run<128, 32, 16>(buf, out); run<128, 64, 16>(buf, out); run<128, 112, 16>(buf, out);
  run<128, 0, 48>(buf, out); run<128, 16, 48>(buf, out); run<128, 32, 48>(buf, out); run<128, 48, 48>(buf, out); run<128, 96, 48>(buf, out);
  run<128, 0, 32>(buf, out); run<128, 0, 64>(buf, out); run<128, 0, 80>(buf, out); run<128, 0, 96>(buf, out); run<128, 0, 128>(buf, out);
  return 0;
}
