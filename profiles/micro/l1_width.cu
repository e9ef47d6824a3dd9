// L1-hit bytes/clk/SM of warp loads by access width (B200), contiguous lanes, 1 CTA of 768 threads per SM.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_width l1_width.cu && ./l1_width
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ unsigned long long g_cycles;

template <int WIDTH, bool SHARED>
__device__ __forceinline__ uint32_t ld(const char* p) {
  uint32_t r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (SHARED) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    if (WIDTH == 4) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r[0]) : "r"(a));
    if (WIDTH == 8) asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(a));
    if (WIDTH == 16) asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
  } else {
    if (WIDTH == 4) asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(r[0]) : "l"(p));
    if (WIDTH == 8) asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "l"(p));
    if (WIDTH == 16) asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(p));
    if (WIDTH == 32) asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                  : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
  }
  return r[0] ^ r[1] ^ r[2] ^ r[3] ^ r[4] ^ r[5] ^ r[6] ^ r[7];
}

template <int WIDTH, bool SHARED>
__global__ void __launch_bounds__(768, 1) k(const char* __restrict__ buf, int steps, uint32_t* out) {
  extern __shared__ char sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (SHARED) for (int i = threadIdx.x; i < 16384; i += 768) reinterpret_cast<uint32_t*>(sm)[i] = i;
  __syncthreads();
  const char* base = SHARED ? sm : buf;
  uint32_t acc = 0;
  unsigned off = (unsigned)warp * 2048u;
  const long long t0 = clock64();
  for (int s = 0; s < steps; s++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      off = (off + 1024u + 32u) & 32767u;
      acc ^= ld<WIDTH, SHARED>(base + (off & ~31u) + lane * WIDTH);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) atomicMax(&g_cycles, (unsigned long long)(t1 - t0));
}

template <int WIDTH, bool SHARED>
void run(const char* buf, uint32_t* out) {
  const int steps = 20000;
  unsigned long long zero = 0;
  cudaFuncSetAttribute(k<WIDTH, SHARED>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  k<WIDTH, SHARED><<<148, 768, SHARED ? 65536 : 0>>>(buf, 100, out);
  cudaMemcpyToSymbol(g_cycles, &zero, sizeof(zero));
  k<WIDTH, SHARED><<<148, 768, SHARED ? 65536 : 0>>>(buf, steps, out);
  cudaDeviceSynchronize();
  unsigned long long cc; cudaMemcpyFromSymbol(&cc, g_cycles, sizeof(cc));
  const double n = (double)steps * 8 * 24;
  printf("%s %3d-bit per lane: %.2f clk per warp load per SM = %.1f B/clk/SM (%s)\n", SHARED ? "shared" : "global",
         WIDTH * 8, cc / n, 32.0 * WIDTH * n / cc, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  char* buf; uint32_t* out;
  cudaMalloc(&buf, 1 << 17); cudaMemset(buf, 1, 1 << 17);
  cudaMalloc(&out, 148 * 768 * 4);
  run<4, false>(buf, out); run<8, false>(buf, out); run<16, false>(buf, out); run<32, false>(buf, out);
  run<4, true>(buf, out); run<8, true>(buf, out); run<16, true>(buf, out);
  return 0;
}
