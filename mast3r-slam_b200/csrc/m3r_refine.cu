// m3r_refine.cu — dilated-window descriptor argmax (sm_100a).
//
// Replaces refine_matches_kernel / refine_matches_cuda (reference
// matching_kernels.cu:25-116).  The output is an integer pixel, so the reference's
// score arithmetic is reproduced bit-for-bit:
//   score_k = rn16( float(score_{k-1}) + float( rn16( float(q_k) * float(c_k) ) ) ),  k = 0..F-1
// (c10::Half operators: every product and every partial sum is formed in fp32 and rounded
// to fp16 — SURVEY H1).  fp32 has >= 2*11+2 significand bits, so the double rounding
// fp32->fp16 is innocuous for one + or * of fp16 operands: rn16(float(a) op float(b)) ==
// the correctly rounded fp16 result == HADD2.RN / HMUL2.RN per lane.  Two candidates are
// therefore scored at once in the two lanes of half2 registers (mul.rn/add.rn, which
// ptxas may not fuse into an fma), 2 instructions per 2 MACs instead of the reference's
// FMUL + F2F + FADD + F2F per MAC.  Scan order (u outer, v inner, dilation_max..1),
// strict '>' and the fp16-zero initial maximum are kept.
#include "m3r_common.cuh"

namespace m3r {

constexpr int kRefineThreads = 128;

__device__ __forceinline__ bool inside_image(int u, int v, int W, int H) {
  return v >= 0 && v < H && u >= 0 && u < W;
}

// Score two candidates (A in the low lane, B in the high lane) against the query.
// q2[k] holds (q_k, q_k).  F8 = F/8 uint4 words per descriptor.
template <int F8>
__device__ __forceinline__ __half2 score_pair(const uint4* __restrict__ ca,
                                              const uint4* __restrict__ cb,
                                              const __half2* q2) {
  __half2 acc = __float2half2_rn(0.f);
#pragma unroll
  for (int w = 0; w < F8; w++) {
    const uint4 a = __ldg(ca + w);
    const uint4 b = __ldg(cb + w);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
    const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const uint32_t lo = __byte_perm(aw[t], bw[t], 0x5410);  // (A_2t  , B_2t  )
      const uint32_t hi = __byte_perm(aw[t], bw[t], 0x7632);  // (A_2t+1, B_2t+1)
      const int k = w * 8 + t * 2;
      acc = __hadd2_rn(acc, __hmul2_rn(*reinterpret_cast<const __half2*>(&lo), q2[k]));
      acc = __hadd2_rn(acc, __hmul2_rn(*reinterpret_cast<const __half2*>(&hi), q2[k + 1]));
    }
  }
  return acc;
}

// Core search for one query.  D11b: descriptor image of this batch element (H,W,F) fp16.
template <int F8>
__device__ __forceinline__ void refine_query(const uint16_t* __restrict__ D11b,
                                             const __half2* q2, int H, int W, int radius,
                                             int dilation_max, long long& u0, long long& v0) {
  constexpr int F = F8 * 8;
  float max_score = 0.f;  // c10::Half() == +0
  long long u_new = u0, v_new = v0;
  const int side = 2 * radius + 1;
  const int cnt = side * side;
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    for (int c = 0; c < cnt; c += 2) {
      // candidate c -> (i = c / side outer = u, j = c % side inner = v)
      const int ia = c / side, ja = c - ia * side;
      const int cbi = c + 1;
      const int ib = cbi / side, jb = cbi - ib * side;
      const long long ua = u0 - rd + (long long)ia * d, va = v0 - rd + (long long)ja * d;
      const long long ub = u0 - rd + (long long)ib * d, vb = v0 - rd + (long long)jb * d;
      const bool oka = inside_image((int)ua, (int)va, W, H);
      const bool okb = (cbi < cnt) && inside_image((int)ub, (int)vb, W, H);
      if (!oka && !okb) continue;
      const size_t offa = oka ? ((size_t)(int)va * W + (int)ua) * F : 0;
      const size_t offb = okb ? ((size_t)(int)vb * W + (int)ub) * F : offa;
      const __half2 s = score_pair<F8>(reinterpret_cast<const uint4*>(D11b + offa),
                                       reinterpret_cast<const uint4*>(D11b + offb), q2);
      const float sa = __low2float(s), sb = __high2float(s);
      if (oka && sa > max_score) { max_score = sa; u_new = ua; v_new = va; }
      if (okb && sb > max_score) { max_score = sb; u_new = ub; v_new = vb; }
    }
    u0 = u_new; v0 = v_new;
  }
}

// Plain form: D21 already fp16, output (u,v) pairs.
template <int F8>
__global__ void __launch_bounds__(kRefineThreads)
refine_f16_kernel(const uint16_t* __restrict__ D11, const uint16_t* __restrict__ D21,
                  const int64_t* __restrict__ p1, int H, int W, int n, int radius,
                  int dilation_max, int64_t* __restrict__ p1_new) {
  constexpr int F = F8 * 8;
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  __half2 q2[F];
  const uint4* qp = reinterpret_cast<const uint4*>(D21 + bk * F);
#pragma unroll
  for (int w = 0; w < F8; w++) {
    const uint4 qv = ld_stream_u4(qp + w);
    const uint32_t qw[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const __half2 h = *reinterpret_cast<const __half2*>(&qw[t]);
      q2[w * 8 + t * 2] = __low2half2(h);
      q2[w * 8 + t * 2 + 1] = __high2half2(h);
    }
  }
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  long long u0 = p.x, v0 = p.y;
  refine_query<F8>(D11 + (size_t)blockIdx.y * H * W * F, q2, H, W, radius, dilation_max, u0, v0);
  longlong2 o; o.x = u0; o.y = v0;
  reinterpret_cast<longlong2*>(p1_new)[bk] = o;
}

// Fused form: D21 fp32 rounded to fp16 on the fly (== .half()), output linear index.
template <int F8>
__global__ void __launch_bounds__(kRefineThreads)
refine_fused_kernel(const uint16_t* __restrict__ D11h, const float* __restrict__ D21,
                    const int64_t* __restrict__ p1, int H, int W, int radius,
                    int dilation_max, int64_t* __restrict__ idx_out) {
  constexpr int F = F8 * 8;
  const int n = H * W;
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  __half2 q2[F];
  const float4* qp = reinterpret_cast<const float4*>(D21 + bk * F);
#pragma unroll
  for (int w = 0; w < F / 4; w++) {
    const float4 qv = ld_stream_f4(qp + w);
    q2[w * 4 + 0] = __half2half2(__float2half_rn(qv.x));
    q2[w * 4 + 1] = __half2half2(__float2half_rn(qv.y));
    q2[w * 4 + 2] = __half2half2(__float2half_rn(qv.z));
    q2[w * 4 + 3] = __half2half2(__float2half_rn(qv.w));
  }
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  long long u0 = p.x, v0 = p.y;
  if (radius > 0)
    refine_query<F8>(D11h + (size_t)blockIdx.y * n * F, q2, H, W, radius, dilation_max, u0, v0);
  idx_out[bk] = u0 + (long long)W * v0;  // pixel_to_lin, matching.py:13-15
}

// <float> instantiation of the reference: sequential fp32 accumulate (FFMA), max from FLT_MIN.
__global__ void __launch_bounds__(kRefineThreads)
refine_f32_kernel(const float* __restrict__ D11, const float* __restrict__ D21,
                  const int64_t* __restrict__ p1, int H, int W, int n, int F, int radius,
                  int dilation_max, int64_t* __restrict__ p1_new) {
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  const float* q = D21 + bk * F;
  const float* D11b = D11 + (size_t)blockIdx.y * H * W * F;
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  long long u0 = p.x, v0 = p.y, u_new = p.x, v_new = p.y;
  float max_score = 1.17549435e-38f;  // numeric_limits<float>::min()
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i, v = v0 - rd + j;
        if (inside_image((int)u, (int)v, W, H)) {
          const float* c = D11b + ((size_t)(int)v * W + (int)u) * F;
          float score = 0.f;
          for (int t = 0; t < F; t++) score = fmaf(__ldg(q + t), __ldg(c + t), score);
          if (score > max_score) { max_score = score; u_new = u; v_new = v; }
        }
      }
    }
    u0 = u_new; v0 = v_new;
  }
  longlong2 o; o.x = u_new; o.y = v_new;
  reinterpret_cast<longlong2*>(p1_new)[bk] = o;
}

}  // namespace m3r

using namespace m3r;

#define M3R_DISPATCH_F8(F8v, CALL)                   \
  switch (F8v) {                                     \
    case 1: { constexpr int F8 = 1; CALL; } break;   \
    case 2: { constexpr int F8 = 2; CALL; } break;   \
    case 3: { constexpr int F8 = 3; CALL; } break;   \
    case 4: { constexpr int F8 = 4; CALL; } break;   \
    default: return M3R_EINVAL;                      \
  }

extern "C" int m3r_refine_matches_f16(const uint16_t* D11, const uint16_t* D21,
                                      const int64_t* p1, int b, int H, int W, int n, int F,
                                      int radius, int dilation_max, int64_t* p1_new,
                                      m3r_stream_t stream) {
  if (b < 0 || n < 0 || H <= 0 || W <= 0 || F <= 0 || (F % 8) != 0 || F > 32 || radius < 0 ||
      b > 65535)
    return M3R_EINVAL;
  if (b == 0 || n == 0) return M3R_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_EINVAL;
  dim3 grid((n + kRefineThreads - 1) / kRefineThreads, b);
  M3R_DISPATCH_F8(F / 8, (refine_f16_kernel<F8><<<grid, kRefineThreads, 0, as_stream(stream)>>>(
                             D11, D21, p1, H, W, n, radius, dilation_max, p1_new)));
  return after_launch();
}

extern "C" int m3r_match_refine_fused(const uint16_t* D11h, const float* D21,
                                      const int64_t* p1, int b, int H, int W, int F, int radius,
                                      int dilation_max, int64_t* idx_1_to_2,
                                      m3r_stream_t stream) {
  if (b < 0 || H <= 0 || W <= 0 || F <= 0 || (F % 8) != 0 || F > 32 || radius < 0 || b > 65535)
    return M3R_EINVAL;
  if (b == 0) return M3R_OK;
  if (!D11h || !D21 || !p1 || !idx_1_to_2) return M3R_EINVAL;
  const int n = H * W;
  dim3 grid((n + kRefineThreads - 1) / kRefineThreads, b);
  M3R_DISPATCH_F8(F / 8, (refine_fused_kernel<F8><<<grid, kRefineThreads, 0, as_stream(stream)>>>(
                             D11h, D21, p1, H, W, radius, dilation_max, idx_1_to_2)));
  return after_launch();
}

extern "C" int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                                      int b, int H, int W, int n, int F, int radius,
                                      int dilation_max, int64_t* p1_new, m3r_stream_t stream) {
  if (b < 0 || n < 0 || H <= 0 || W <= 0 || F <= 0 || radius < 0 || b > 65535) return M3R_EINVAL;
  if (b == 0 || n == 0) return M3R_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_EINVAL;
  dim3 grid((n + kRefineThreads - 1) / kRefineThreads, b);
  refine_f32_kernel<<<grid, kRefineThreads, 0, as_stream(stream)>>>(D11, D21, p1, H, W, n, F,
                                                                    radius, dilation_max, p1_new);
  return after_launch();
}
