#include <cuda_fp16.h>
// What does "products on the fp32 pipe" cost?  fp16 pair -> two fp32, FMUL, pack back with RN.
__global__ void k(const unsigned* __restrict__ c, const float2* __restrict__ q, unsigned* o) {
  const unsigned w = c[threadIdx.x];
  const __half2 h = *reinterpret_cast<const __half2*>(&w);
  const float2 f = __half22float2(h);
  const float2 qq = q[threadIdx.x];
  const __half2 p = __floats2half2_rn(f.x * qq.x, f.y * qq.y);
  o[threadIdx.x] = *reinterpret_cast<const unsigned*>(&p);
}
