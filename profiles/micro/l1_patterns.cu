// L1 data-pipe cost of the access patterns refine_matches could use (all hits; B200).
// Each warp reads, per step, one "candidate" per lane: BYTES bytes per lane starting at
// base + lane*STRIDE (+ plane offsets for the planar forms), as 128- or 256-bit loads.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_patterns l1_patterns.cu && ./l1_patterns
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint4 ld128(const char* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ld256(const char* p) {
  uint32_t r[8];
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
  return r[0] ^ r[1] ^ r[2] ^ r[3] ^ r[4] ^ r[5] ^ r[6] ^ r[7];
}

// MODE 0: AoS 48 B, 3 x LDG.128 at lane*48          (shipped layout)
// MODE 1: planar 3 x 16 B, 3 x LDG.128 at lane*16 in three planes
// MODE 2: two planes 32 B + 16 B: LDG.256 at lane*32, LDG.128 at lane*16
// MODE 3: AoS padded to 64 B: 2 x LDG.256 at lane*64 (second half-used)
// MODE 4: AoS 48 B, LDG.256 + LDG.128 at lane*48 (32-B alignment holds for even pixels only -> use 16-B aligned? no: 96-B pairs)
template <int MODE>
__global__ void __launch_bounds__(256) k(const char* __restrict__ buf, int region, int steps, int shift, uint32_t* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t acc = 0;
  const int plane = region;  // planes are `region` bytes apart
  unsigned off = (unsigned)warp * 4096u;
  for (int s = 0; s < steps; s++) {
    off = (off + (unsigned)shift) % (unsigned)(region - 4096);
    if (MODE == 0) {
      const char* p = buf + off + lane * 48;
      uint4 a = ld128(p), b = ld128(p + 16), c = ld128(p + 32);
      acc ^= a.x ^ a.w ^ b.y ^ c.z;
    } else if (MODE == 1) {
      const char* p = buf + (off & ~15u) + lane * 16;
      uint4 a = ld128(p), b = ld128(p + plane), c = ld128(p + 2 * plane);
      acc ^= a.x ^ a.w ^ b.y ^ c.z;
    } else if (MODE == 2) {
      const char* p = buf + (off & ~31u) + lane * 32;
      const char* p2 = buf + plane + ((off >> 1) & ~15u) + lane * 16;
      acc ^= ld256(p);
      uint4 c = ld128(p2);
      acc ^= c.x ^ c.w;
    } else if (MODE == 3) {
      const char* p = buf + (off & ~63u) + lane * 64;
      acc ^= ld256(p) ^ ld256(p + 32);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char* name, const char* buf, int region, int shift, uint32_t* out, int bytes_per_lane) {
  const int steps = 20000, grid = 148 * 3;
  cudaEvent_t a, z;
  cudaEventCreate(&a); cudaEventCreate(&z);
  k<MODE><<<grid, 256>>>(buf, region, 100, shift, out);
  cudaEventRecord(a);
  k<MODE><<<grid, 256>>>(buf, region, steps, shift, out);
  cudaEventRecord(z);
  cudaEventSynchronize(z);
  float ms;
  cudaEventElapsedTime(&ms, a, z);
  const double warp_steps_per_sm = (double)steps * 8 * 3;  // warps resident per SM x steps
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double cyc = ms * 1e-3 * clk * 1e3;
  printf("%-44s shift %5d: %.3f ms  %.2f clk per warp-candidate per SM  (%.1f B/clk/SM useful)\n", name, shift, ms,
         cyc / warp_steps_per_sm, warp_steps_per_sm * 32 * bytes_per_lane / cyc);
}

int main() {
  const int region = 48 * 1024;
  char* buf; uint32_t* out;
  cudaMalloc(&buf, 4 * region); cudaMemset(buf, 1, 4 * region);
  cudaMalloc(&out, 148 * 3 * 256 * 4);
  for (int shift : {48, 240, 48 * 512}) {
    run<0>("AoS 48B: 3 x LDG.128 stride 48", buf, region, shift, out, 48);
    run<1>("planar 3x16B: 3 x LDG.128 contiguous", buf, region, shift, out, 48);
    run<2>("2-plane 32B+16B: LDG.256 + LDG.128 contiguous", buf, region, shift, out, 48);
    run<3>("AoS 64B padded: 2 x LDG.256 stride 64", buf, region, shift, out, 48);
  }
  return 0;
}
