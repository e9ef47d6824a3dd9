// m3r_abi.cu — library-wide state and the small C-ABI entry points.
#include "m3r_common.cuh"

namespace m3r {
thread_local int g_last_cuda_error = 0;
std::atomic<unsigned long long> g_launch_count{0};
}  // namespace m3r

extern "C" int m3r_version(void) { return 100; }  // 0.1.0

extern "C" const char* m3r_error_string(int code) {
  switch (code) {
    case M3R_OK: return "ok";
    case M3R_EINVAL: return "invalid argument (shape, null pointer or unsupported parameter)";
    case M3R_ELAUNCH: return "CUDA launch/runtime error";
    case M3R_EWORKSPACE: return "workspace too small";
    case M3R_ETOOLARGE: return "size beyond the fixed maximum of a no-allocation entry point (see include/m3r_b200.h)";
    default: return "unknown error";
  }
}

extern "C" int m3r_last_cuda_error(void) { return m3r::g_last_cuda_error; }

extern "C" unsigned long long m3r_launch_count(void) {
  return m3r::g_launch_count.load(std::memory_order_relaxed);
}
