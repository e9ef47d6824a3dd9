// L1-hit throughput of 128-bit global loads by instruction flavour (B200): .nc (LDG.CONSTANT) vs plain
// vs shared memory, contiguous 16 B per lane.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_ldkinds l1_ldkinds.cu && ./l1_ldkinds
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ long long g_cycles;
template <int KIND>
__device__ __forceinline__ uint4 ld(const char* p) {
  uint4 r;
  if (KIND == 0) asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  if (KIND == 1) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  if (KIND == 2) asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  if (KIND == 3) asm volatile("ld.global.nc.L1::evict_last.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <int KIND, int STRIDE>
__global__ void __launch_bounds__(768, 1) k(const char* __restrict__ buf, int region, int steps, uint32_t* out) {
  extern __shared__ char sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t acc = 0;
  unsigned off = (unsigned)warp * 2048u;
  const long long t0 = clock64();
  if (KIND == 4) for (int i = threadIdx.x; i < region / 4; i += 768) reinterpret_cast<uint32_t*>(sm)[i] = i;
  __syncthreads();
  for (int s = 0; s < steps; s++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {  // 4 independent candidates per trip: 12 loads in flight per thread
      off = (off + 48u * 5u) & (unsigned)(region / 2 - 1);
      const unsigned o = (off & ~15u) + lane * STRIDE;
      uint4 a, b, c;
      if (KIND == 4) {
        a = *reinterpret_cast<const uint4*>(sm + o);
        b = *reinterpret_cast<const uint4*>(sm + o + 16 * (STRIDE == 16 ? 40 : 1));
        c = *reinterpret_cast<const uint4*>(sm + o + 32 * (STRIDE == 16 ? 40 : 1));
      } else {
        a = ld<KIND>(buf + o);
        b = ld<KIND>(buf + o + 16 * (STRIDE == 16 ? 40 : 1));
        c = ld<KIND>(buf + o + 32 * (STRIDE == 16 ? 40 : 1));
      }
      acc ^= a.x ^ a.w ^ b.y ^ c.z;
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) atomicMax((unsigned long long*)&g_cycles, (unsigned long long)(t1 - t0));
}

template <int KIND, int STRIDE>
void run(const char* name, const char* buf, int region, uint32_t* out) {
  const int steps = 10000, grid = 148;
  long long zero = 0; cudaMemcpyToSymbol(g_cycles, &zero, sizeof(zero));
  cudaEvent_t a, z;
  cudaEventCreate(&a); cudaEventCreate(&z);
  const int smem = KIND == 4 ? region : 0;
  cudaFuncSetAttribute(k<KIND, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k<KIND, STRIDE><<<grid, 768, smem>>>(buf, region, 100, out);
  cudaMemcpyToSymbol(g_cycles, &zero, sizeof(zero));
  cudaEventRecord(a);
  k<KIND, STRIDE><<<grid, 768, smem>>>(buf, region, steps, out);
  cudaEventRecord(z);
  cudaEventSynchronize(z);
  float ms;
  cudaEventElapsedTime(&ms, a, z);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double cyc = ms * 1e-3 * clk * 1e3;
  const double instr_per_sm = (double)steps * 8 * 3 * 3 * 4;
  long long cc; cudaMemcpyFromSymbol(&cc, g_cycles, sizeof(cc));
  printf("%-40s lane stride %2d: %.3f ms  %.2f clk (event x nominal clock) / %.2f clk (clock64) per warp LD.128 per SM; SM clock %.0f MHz (%s)\n",
         name, STRIDE, ms, cyc / instr_per_sm, (double)cc / instr_per_sm, cc / (ms * 1e3), cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int region = 64 * 1024;
  char* buf; uint32_t* out;
  cudaMalloc(&buf, 4 * region); cudaMemset(buf, 1, 4 * region);
  cudaMalloc(&out, 148 * 768 * 4);
  run<0, 16>("ld.global.nc (LDG.CONSTANT)", buf, region, out);
  run<1, 16>("ld.global (plain)", buf, region, out);
  run<2, 16>("ld.global.ca", buf, region, out);
  run<3, 16>("ld.global.nc.L1::evict_last", buf, region, out);
  run<4, 16>("ld.shared", buf, region, out);
  run<0, 48>("ld.global.nc (LDG.CONSTANT)", buf, region, out);
  run<1, 48>("ld.global (plain)", buf, region, out);
  run<4, 48>("ld.shared", buf, region, out);
  return 0;
}
