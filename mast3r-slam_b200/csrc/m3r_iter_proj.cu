// m3r_iter_proj.cu — per-pixel Levenberg-Marquardt projective ray matching (sm_100a).
//
// Replaces iter_proj_kernel / iter_proj_cuda (reference matching_kernels.cu:119-315).
// One thread per query pixel, 256-thread CTAs (the reference launches 16-thread CTAs).
// The arithmetic of one LM iteration is kept expression-for-expression, including the
// float->double promotions of the reference's un-suffixed literals, so the
// iterates agree to rounding.  Two exact shortcuts remove most of the gathers:
//   (1) after an ACCEPTED step the next iteration's ray sample is the candidate's ray
//       sample (same position, same expression) — only the 6 gradient channels are read;
//   (2) after a REJECTED step nothing but lambda changes — the sampled values, J^T J and
//       J^T r are reused; and when the clamped candidate equals the current position
//       bit-for-bit its cost equals the current cost, so it is not re-sampled.
// The reference re-reads 4x9 + 4x3 floats in every one of its max_iter iterations.
#include "m3r_common.cuh"

namespace m3r {

__device__ __forceinline__ void clampf(float& x, const float lo, const float hi) {
  x = fminf(fmaxf(x, lo), hi);
}

// CH = 9: the reference's (b,H,W,9) image, scalar gathers.  CH = 12: the private layout of the
// fused pipeline, three 16-B words per pixel [ray,0 | d/du,0 | d/dv,0]: a bilinear sample of one
// group is four LDG.128 instead of twelve scalar loads (same values, same expression).
template <int CH>
struct Bilin {
  const float* r11; const float* r12; const float* r21; const float* r22;
  float w11, w12, w21, w22;
};

// Weights and texel pointers for a sample at (u,v); matching_kernels.cu:155-170.
template <int CH>
__device__ __forceinline__ Bilin<CH> bilin_setup(const float* __restrict__ img, int W, float u, float v) {
  Bilin<CH> s;
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = u - static_cast<float>(u11);
  const float dv = v - static_cast<float>(v11);
  s.w11 = du * dv;
  s.w12 = (1.0 - du) * dv;          // double product, rounded once (as the reference)
  s.w21 = du * (1.0 - dv);
  s.w22 = (1.0 - du) * (1.0 - dv);
  const float* base = img + ((size_t)v11 * W + u11) * CH;
  s.r22 = base;                      // (v11  , u11  )
  s.r21 = base + CH;                 // (v11  , u11+1)
  s.r12 = base + (size_t)W * CH;     // (v11+1, u11  )
  s.r11 = s.r12 + CH;                // (v11+1, u11+1)
  return s;
}

// channels 3g .. 3g+2 (g = 0 ray, 1 d/du, 2 d/dv) of the sample
template <int CH>
__device__ __forceinline__ void bilin_group(const Bilin<CH>& s, int g, float* out) {
  if (CH == 12) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(s.r11) + g);
    const float4 b = __ldg(reinterpret_cast<const float4*>(s.r12) + g);
    const float4 c = __ldg(reinterpret_cast<const float4*>(s.r21) + g);
    const float4 d = __ldg(reinterpret_cast<const float4*>(s.r22) + g);
    out[0] = s.w11 * a.x + s.w12 * b.x + s.w21 * c.x + s.w22 * d.x;
    out[1] = s.w11 * a.y + s.w12 * b.y + s.w21 * c.y + s.w22 * d.y;
    out[2] = s.w11 * a.z + s.w12 * b.z + s.w21 * c.z + s.w22 * d.z;
  } else {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int ch = 3 * g + j;
      out[j] = s.w11 * __ldg(s.r11 + ch) + s.w12 * __ldg(s.r12 + ch) + s.w21 * __ldg(s.r21 + ch) +
               s.w22 * __ldg(s.r22 + ch);
    }
  }
}

__device__ __forceinline__ void normalize3(float* r) {  // matching_kernels.cu:186-191
  const float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  const float r_norm_inv = 1.0 / r_norm;
  r[0] *= r_norm_inv; r[1] *= r_norm_inv; r[2] *= r_norm_inv;
}

// The LM loop for one pixel.  Returns the final (u,v) and the `converged` flag.
template <int CH>
__device__ __forceinline__ void lm_pixel(const float* __restrict__ img, int H, int W,
                                         const float t0, const float t1, const float t2,
                                         float& u, float& v, int max_iter, float lambda,
                                         const float cost_thresh, bool& conv) {
  clampf(u, 1, W - 2);
  clampf(v, 1, H - 2);
  conv = false;
  float gx[3], gy[3], err[3];
  float cost = 0.f, A00 = 0.f, A01 = 0.f, A11 = 0.f, b0 = 0.f, b1 = 0.f;
  bool moved = true;   // position changed since J^T J was formed
  bool have_ray = false;  // err/cost already valid for (u,v)
  for (int i = 0; i < max_iter; i++) {
    if (moved) {
      const Bilin<CH> s = bilin_setup<CH>(img, W, u, v);
      if (!have_ray) {
        float r[3];
        bilin_group<CH>(s, 0, r);
        normalize3(r);
        err[0] = r[0] - t0; err[1] = r[1] - t1; err[2] = r[2] - t2;
        cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
      }
      bilin_group<CH>(s, 1, gx);
      bilin_group<CH>(s, 2, gy);
      A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2];
      A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
      A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
      b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
      b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
      moved = false;
    }
    // LM diagonal and closed-form 2x2 solve, matching_kernels.cu:209-215
    const float a00 = A00 + lambda;
    const float a11 = A11 + lambda;
    const float det_inv = 1.0 / (a00 * a11 - A01 * A01);
    const float delta_u = det_inv * (a11 * b0 - A01 * b1);
    const float delta_v = det_inv * (-A01 * b0 + a00 * b1);
    float u_new = u + delta_u;
    float v_new = v + delta_v;
    clampf(u_new, 1, W - 2);
    clampf(v_new, 1, H - 2);

    float e2[3];
    float new_cost;
    if (u_new == u && v_new == v) {
      new_cost = cost;  // same position, same expression => same bits; step is rejected
    } else {
      const Bilin<CH> s = bilin_setup<CH>(img, W, u_new, v_new);
      float r[3];
      bilin_group<CH>(s, 0, r);
      normalize3(r);
      e2[0] = r[0] - t0; e2[1] = r[1] - t1; e2[2] = r[2] - t2;
      new_cost = e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2];
    }
    if (new_cost < cost) {  // matching_kernels.cu:259-268
      u = u_new; v = v_new;
      lambda *= 0.1;
      conv = new_cost < cost_thresh;
      err[0] = e2[0]; err[1] = e2[1]; err[2] = e2[2];
      cost = new_cost;
      moved = true; have_ray = true;
    } else {
      lambda *= 10.0;
      conv = cost < cost_thresh;
    }
  }
}

constexpr int kIterProjThreads = 256;

__global__ void __launch_bounds__(kIterProjThreads)
iter_proj_kernel(const float* __restrict__ rays_img, const float* __restrict__ pts,
                 const float* __restrict__ p_init, int H, int W, int n, int max_iter,
                 float lambda_init, float cost_thresh, float* __restrict__ p_new,
                 uint8_t* __restrict__ converged) {
  const int k = blockIdx.x * kIterProjThreads + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  const float* img = rays_img + (size_t)blockIdx.y * H * W * 9;
  const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + bk);
  float u = p0.x, v = p0.y;
  const float t0 = __ldg(pts + bk * 3 + 0), t1 = __ldg(pts + bk * 3 + 1), t2 = __ldg(pts + bk * 3 + 2);
  bool conv;
  lm_pixel<9>(img, H, W, t0, t1, t2, u, v, max_iter, lambda_init, cost_thresh, conv);
  reinterpret_cast<float2*>(p_new)[bk] = make_float2(u, v);
  converged[bk] = conv ? 1 : 0;
}

// Fused form (matching.py:25-76): target = F.normalize(X21) on the fly, p_init from the
// linear index (identity when idx_init == nullptr), then p1 = trunc(p) and the occlusion
// test |X11[p1] - X21| < dist_thresh folded into `valid`.
template <int CH>
__global__ void __launch_bounds__(kIterProjThreads)
iter_proj_fused_kernel(const float* __restrict__ rays_img, const float* __restrict__ X11,
                       const float* __restrict__ X21, const int64_t* __restrict__ idx_init,
                       int H, int W, int max_iter, float lambda_init, float cost_thresh,
                       float dist_thresh, int64_t* __restrict__ p1, uint8_t* __restrict__ valid,
                       float* __restrict__ p_new) {
  const int n = H * W;
  const int k = blockIdx.x * kIterProjThreads + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  const float* img = rays_img + (size_t)blockIdx.y * n * CH;
  const float x = __ldg(X21 + bk * 3 + 0), y = __ldg(X21 + bk * 3 + 1), z = __ldg(X21 + bk * 3 + 2);
  // F.normalize(dim=-1, eps=1e-12): x / max(|x|_2, eps)
  const float nrm = fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
  const float t0 = x / nrm, t1 = y / nrm, t2 = z / nrm;
  const int64_t lin = idx_init ? __ldg(idx_init + bk) : (int64_t)k;
  // lin_to_pixel().float(), matching.py:18-22,47-48; 32-bit division whenever the index fits
  // (always, for indices into the image) instead of the 64-bit division subroutine
  float u, v;
  if ((unsigned long long)lin < 0x80000000ULL) {
    const unsigned l32 = (unsigned)lin, q = l32 / (unsigned)W;
    u = (float)(l32 - q * (unsigned)W); v = (float)q;
  } else {
    u = (float)(lin % W); v = (float)(lin / W);
  }
  bool conv;
  lm_pixel<CH>(img, H, W, t0, t1, t2, u, v, max_iter, lambda_init, cost_thresh, conv);
  if (p_new) reinterpret_cast<float2*>(p_new)[bk] = make_float2(u, v);
  const int ui = (int)u, vi = (int)v;  // .long(): truncation; u,v are within [1,W-2]x[1,H-2]
  const float* Xm = X11 + ((size_t)blockIdx.y * n + (size_t)vi * W + ui) * 3;
  const float dx = __ldg(Xm + 0) - x, dy = __ldg(Xm + 1) - y, dz = __ldg(Xm + 2) - z;
  const float dist = sqrtf(dx * dx + dy * dy + dz * dz);  // torch.linalg.norm(..., dim=-1)
  longlong2 out; out.x = ui; out.y = vi;
  reinterpret_cast<longlong2*>(p1)[bk] = out;
  valid[bk] = (conv && dist < dist_thresh) ? 1 : 0;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_iter_proj(const float* rays_img_with_grad, const float* pts_3d_norm,
                             const float* p_init, int b, int H, int W, int n, int max_iter,
                             float lambda_init, float cost_thresh, float* p_new,
                             uint8_t* converged, m3r_stream_t stream) {
  M3R_NVTX("m3r_iter_proj");
  if (b < 0 || n < 0 || H < 3 || W < 3 || max_iter < 0) return M3R_EINVAL;
  if (b == 0 || n == 0) return M3R_OK;
  if (!rays_img_with_grad || !pts_3d_norm || !p_init || !p_new || !converged) return M3R_EINVAL;
  if (b > 65535) return M3R_EINVAL;
  dim3 grid((n + kIterProjThreads - 1) / kIterProjThreads, b);
  iter_proj_kernel<<<grid, kIterProjThreads, 0, as_stream(stream)>>>(
      rays_img_with_grad, pts_3d_norm, p_init, H, W, n, max_iter, lambda_init, cost_thresh,
      p_new, converged);
  return after_launch();
}

static int iter_proj_fused_launch(int ch, const float* rays, const float* X11, const float* X21,
                                  const int64_t* idx_init, int b, int H, int W, int max_iter,
                                  float lambda_init, float cost_thresh, float dist_thresh, int64_t* p1,
                                  uint8_t* valid, float* p_new, m3r_stream_t stream) {
  if (b < 0 || H < 3 || W < 3 || max_iter < 0 || b > 65535) return M3R_EINVAL;
  if (b == 0) return M3R_OK;
  if (!rays || !X11 || !X21 || !p1 || !valid) return M3R_EINVAL;
  if (ch == 12 && (reinterpret_cast<uintptr_t>(rays) & 15) != 0) return M3R_EINVAL;
  const int n = H * W;
  dim3 grid((n + kIterProjThreads - 1) / kIterProjThreads, b);
  if (ch == 12)
    iter_proj_fused_kernel<12><<<grid, kIterProjThreads, 0, as_stream(stream)>>>(
        rays, X11, X21, idx_init, H, W, max_iter, lambda_init, cost_thresh, dist_thresh, p1, valid, p_new);
  else
    iter_proj_fused_kernel<9><<<grid, kIterProjThreads, 0, as_stream(stream)>>>(
        rays, X11, X21, idx_init, H, W, max_iter, lambda_init, cost_thresh, dist_thresh, p1, valid, p_new);
  return after_launch();
}

extern "C" int m3r_match_iter_proj_fused(const float* rays_img_with_grad, const float* X11,
                                         const float* X21, const int64_t* idx_init, int b,
                                         int H, int W, int max_iter, float lambda_init,
                                         float cost_thresh, float dist_thresh, int64_t* p1,
                                         uint8_t* valid, float* p_new, m3r_stream_t stream) {
  M3R_NVTX("m3r_match_iter_proj_fused");
  return iter_proj_fused_launch(9, rays_img_with_grad, X11, X21, idx_init, b, H, W, max_iter, lambda_init,
                                cost_thresh, dist_thresh, p1, valid, p_new, stream);
}

extern "C" int m3r_match_iter_proj_packed(const float* rays_packed, const float* X11,
                                          const float* X21, const int64_t* idx_init, int b,
                                          int H, int W, int max_iter, float lambda_init,
                                          float cost_thresh, float dist_thresh, int64_t* p1,
                                          uint8_t* valid, float* p_new, m3r_stream_t stream) {
  M3R_NVTX("m3r_match_iter_proj_packed");
  return iter_proj_fused_launch(12, rays_packed, X11, X21, idx_init, b, H, W, max_iter, lambda_init,
                                cost_thresh, dist_thresh, p1, valid, p_new, stream);
}
