// m3r_common.cuh — shared host/device plumbing for libm3r_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <atomic>
#include <nvtx3/nvToolsExt.h>

#include "../../include/m3r_b200.h"

namespace m3r {

extern thread_local int g_last_cuda_error;
extern std::atomic<unsigned long long> g_launch_count;

inline cudaStream_t as_stream(m3r_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// Record a kernel launch and convert a launch failure into M3R_ELAUNCH.
inline int after_launch() {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_last_cuda_error = (int)e; return M3R_ELAUNCH; }
  return M3R_OK;
}

inline int check_cuda(cudaError_t e) {
  if (e != cudaSuccess) { g_last_cuda_error = (int)e; return M3R_ELAUNCH; }
  return M3R_OK;
}

// NVTX range around a C-ABI entry point: visible as a named span on the host timeline of
// nsys / ncu --nvtx; a few nanoseconds when no tool is attached (nvtx3 is header-only and lazy).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
#define M3R_NVTX(name) ::m3r::NvtxRange m3r_nvtx_range_(name)

#define M3R_TRY(expr) do { int _rc = (expr); if (_rc != M3R_OK) return _rc; } while (0)

constexpr int kNumSMs = 148;  // B200
constexpr int kMaxDevices = 64;  // per-device caches (occupancy, SM count) are indexed by ordinal

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// 128-bit read-only streaming loads (data touched once per pass: keep it out of L1).
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld_stream_f4(const void* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f(const void* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace m3r
