// m3r_prep.cu — fused prologue of match_iterative_proj (sm_100a).
//
// Replaces the PyTorch glue of matching.prep_for_iter_proj (reference matching.py:25-50) and
// image.img_gradient (image.py:5-38): F.normalize -> permute -> 2x (reflect pad + depthwise
// 3x3 conv) -> cat -> permute -> contiguous (six kernels, each a round trip through HBM) become
// one pass: X11 (b,H,W,3) -> rays_img_with_grad (b,H,W,9).  The fp32 -> fp16 cast of the
// descriptor image (matching.py:80) is the second kernel here.
#include "m3r_common.cuh"

namespace m3r {

constexpr int kPrepTX = 32, kPrepTY = 8;

__device__ __forceinline__ int reflect(int i, int n) {  // F.pad(mode="reflect"), pad = 1
  if (i < 0) return -i;
  if (i >= n) return 2 * n - 2 - i;
  return i;
}

// One 32x8 pixel tile (bx, by) of image bz; 256 threads, tid = ty*32 + tx.
// CH = 9: the reference's (b,H,W,9) layout.  CH = 12: three 16-B words per pixel
// [ray,0 | d/du,0 | d/dv,0], the layout m3r_match_iter_proj_packed gathers with LDG.128.
template <int CH>
__device__ __forceinline__ void prep_rays_tile(const float* __restrict__ X11, int H, int W,
                                               float* __restrict__ out, int bx, int by, int bz) {
  __shared__ float tile[kPrepTY + 2][kPrepTX + 2][3];
  // output staging: one tile row is kPrepTX*9 contiguous floats in the (b,H,W,9) image, written
  // back as coalesced 16-B words (9 scalar 36-B-strided stores per thread kept the LSU 69 % busy)
  __shared__ __align__(16) float stage[kPrepTY][kPrepTX * CH];
  const int tid = threadIdx.x;
  const int thx = tid % kPrepTX, thy = tid / kPrepTX;
  const int x0 = bx * kPrepTX, y0 = by * kPrepTY;
  const float* Xb = X11 + (size_t)bz * H * W * 3;
  for (int idx = tid; idx < (kPrepTY + 2) * (kPrepTX + 2); idx += kPrepTX * kPrepTY) {
    const int ty = idx / (kPrepTX + 2), tx = idx % (kPrepTX + 2);
    const int y = reflect(min(y0 + ty - 1, H), H), x = reflect(min(x0 + tx - 1, W), W);
    const float* p = Xb + ((size_t)min(y, H - 1) * W + min(x, W - 1)) * 3;
    const float a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    // F.normalize(dim=-1, eps=1e-12): v / max(|v|_2, eps)
    const float nrm = fmaxf(sqrtf(a * a + b * b + c * c), 1e-12f);
    tile[ty][tx][0] = a / nrm; tile[ty][tx][1] = b / nrm; tile[ty][tx][2] = c / nrm;
  }
  __syncthreads();
  {
    const int tx = thx + 1, ty = thy + 1;
    float* o = &stage[thy][thx * CH];
    constexpr int G = CH == 12 ? 4 : 3;  // stride between the ray / d-du / d-dv groups
    if (CH == 12) { o[3] = 0.f; o[7] = 0.f; o[11] = 0.f; }
    const float k3 = 3.0f / 32.0f, k10 = 10.0f / 32.0f;  // (1/32)*[-3,0,3;-10,0,10;-3,0,3]
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float tl = tile[ty - 1][tx - 1][c], tc = tile[ty - 1][tx][c], tr = tile[ty - 1][tx + 1][c];
      const float ml = tile[ty][tx - 1][c], mr = tile[ty][tx + 1][c];
      const float bl = tile[ty + 1][tx - 1][c], bc = tile[ty + 1][tx][c], br = tile[ty + 1][tx + 1][c];
      o[c] = tile[ty][tx][c];
      o[G + c] = -k3 * tl + k3 * tr - k10 * ml + k10 * mr - k3 * bl + k3 * br;   // gx
      o[2 * G + c] = -k3 * tl - k10 * tc - k3 * tr + k3 * bl + k10 * bc + k3 * br;   // gy
    }
  }
  __syncthreads();
  const int cols = min(kPrepTX, W - x0);   // valid pixels of this tile row
  const int rows = min(kPrepTY, H - y0);
  const int nfl = cols * CH;               // valid floats per row
  float* dst0 = out + (((size_t)bz * H + y0) * W + x0) * CH;
  const size_t row_stride = (size_t)W * CH;
  if ((nfl & 3) == 0 && (row_stride & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst0) & 15) == 0)) {
    const int per_row = nfl / 4;
    for (int i = tid; i < rows * per_row; i += kPrepTX * kPrepTY) {
      const int r = i / per_row, k = i - r * per_row;
      reinterpret_cast<float4*>(dst0 + r * row_stride)[k] = reinterpret_cast<const float4*>(stage[r])[k];
    }
  } else {
    for (int i = tid; i < rows * nfl; i += kPrepTX * kPrepTY) {
      const int r = i / nfl, k = i - r * nfl;
      dst0[r * row_stride + k] = stage[r][k];
    }
  }
}

// fp32 -> fp16 (round to nearest even == Tensor.half()), 8 elements per thread; block `blk` of `nblk`.
__device__ __forceinline__ void cast_f16_block(const float* __restrict__ src, size_t count8, size_t count,
                                               uint16_t* __restrict__ dst, int blk, int nblk) {
  const size_t stride = (size_t)nblk * 256;
  for (size_t i = (size_t)blk * 256 + threadIdx.x; i < count8; i += stride) {
    const float4 a = ld_stream_f4(src + i * 8), b = ld_stream_f4(src + i * 8 + 4);
    __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w);
    __half2 h2 = __floats2half2_rn(b.x, b.y), h3 = __floats2half2_rn(b.z, b.w);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&h0); o.y = *reinterpret_cast<uint32_t*>(&h1);
    o.z = *reinterpret_cast<uint32_t*>(&h2); o.w = *reinterpret_cast<uint32_t*>(&h3);
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
  if (blk == 0) {
    for (size_t i = count8 * 8 + threadIdx.x; i < count; i += 256)
      dst[i] = __half_as_ushort(__float2half_rn(src[i]));
  }
}

template <int CH>
__global__ void __launch_bounds__(256)
prep_rays_kernel(const float* __restrict__ X11, int H, int W, int tiles_x, int tiles_y, float* __restrict__ out) {
  const int t = blockIdx.x, per = tiles_x * tiles_y;
  prep_rays_tile<CH>(X11, H, W, out, (t % per) % tiles_x, (t % per) / tiles_x, t / per);
}

__global__ void __launch_bounds__(256)
cast_f16_kernel(const float* __restrict__ src, size_t count8, size_t count, uint16_t* __restrict__ dst) {
  cast_f16_block(src, count8, count, dst, blockIdx.x, gridDim.x);
}

// Both jobs of m3r_match_prep in ONE launch: they are independent, one is bound by L1/shared
// and instruction issue, the other by HBM, so their CTAs overlap on the SMs instead of running
// back to back with a launch gap in between.  Cast blocks are interleaved 1:2 with ray tiles.
template <int CH>
__global__ void __launch_bounds__(256)
match_prep_kernel(const float* __restrict__ X11, int H, int W, int tiles_x, int tiles_y, int ray_blocks,
                  float* __restrict__ rays, const float* __restrict__ D11, size_t count8, size_t count,
                  int cast_blocks, uint16_t* __restrict__ D11h) {
  const int t = blockIdx.x;
  const int trio = t / 3, pos = t % 3;
  // blocks 3k+2 (while cast blocks last) are cast blocks; everything else is a ray tile
  const bool is_cast = pos == 2 && trio < cast_blocks;
  if (is_cast) {
    cast_f16_block(D11, count8, count, D11h, trio, cast_blocks);
  } else {
    const int casts_before = min(trio + (pos == 2 ? 1 : 0), cast_blocks);
    const int r = t - casts_before;
    if (r < ray_blocks) {
      const int per = tiles_x * tiles_y;
      prep_rays_tile<CH>(X11, H, W, rays, (r % per) % tiles_x, (r % per) / tiles_x, r / per);
    }
  }
}

// image.img_gradient (image.py:5-38) as a function of its own: (planes, H, W) f32 -> gx, gy of the
// same shape, planes = b*c (the convolution is depthwise, so batch and channel are one index).
// Same taps and the same order of additions as prep_rays_tile.  One thread per pixel, rows of 32
// pixels per warp: the nine taps of a warp are three 128-B rows shifted by one element, so L1
// serves the overlap and HBM sees each element once (12 B per element: 4 in, 8 out).
__global__ void __launch_bounds__(256)
img_gradient_kernel(const float* __restrict__ img, int H, int W, float* __restrict__ gx, float* __restrict__ gy) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= W || y >= H) return;
  const float* pl = img + (size_t)blockIdx.z * H * W;
  const int xm = reflect(x - 1, W), xp = reflect(x + 1, W), ym = reflect(y - 1, H), yp = reflect(y + 1, H);
  const float* r0 = pl + (size_t)ym * W; const float* r1 = pl + (size_t)y * W; const float* r2 = pl + (size_t)yp * W;
  const float tl = __ldg(r0 + xm), tc = __ldg(r0 + x), tr = __ldg(r0 + xp);
  const float ml = __ldg(r1 + xm), mr = __ldg(r1 + xp);
  const float bl = __ldg(r2 + xm), bc = __ldg(r2 + x), br = __ldg(r2 + xp);
  const float k3 = 3.0f / 32.0f, k10 = 10.0f / 32.0f;
  const size_t o = (size_t)blockIdx.z * H * W + (size_t)y * W + x;
  gx[o] = -k3 * tl + k3 * tr - k10 * ml + k10 * mr - k3 * bl + k3 * br;
  gy[o] = -k3 * tl - k10 * tc - k3 * tr + k3 * bl + k10 * bc + k3 * br;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_img_gradient(const float* img, int planes, int H, int W, float* gx, float* gy,
                                m3r_stream_t stream) {
  M3R_NVTX("m3r_img_gradient");
  if (planes < 0 || H < 2 || W < 2 || planes > 65535) return M3R_EINVAL;  // reflect padding needs 2 rows / columns
  if (planes == 0) return M3R_OK;
  if (!img || !gx || !gy) return M3R_EINVAL;
  const dim3 grid((W + 31) / 32, (H + 7) / 8, planes);
  if (grid.y > 65535) return M3R_EINVAL;
  img_gradient_kernel<<<grid, 256, 0, as_stream(stream)>>>(img, H, W, gx, gy);
  return after_launch();
}

template <int CH>
static int match_prep_launch(const float* X11, const float* D11, int b, int H, int W, int F,
                             float* rays, uint16_t* D11h, m3r_stream_t stream) {
  if (b < 0 || H < 2 || W < 2 || b > 65535) return M3R_EINVAL;
  if (b == 0) return M3R_OK;
  cudaStream_t st = as_stream(stream);
  if (rays && !X11) return M3R_EINVAL;
  if (D11h && (!D11 || F <= 0)) return M3R_EINVAL;
  if (CH == 12 && rays && (reinterpret_cast<uintptr_t>(rays) & 15) != 0) return M3R_EINVAL;
  const int tiles_x = (W + kPrepTX - 1) / kPrepTX, tiles_y = (H + kPrepTY - 1) / kPrepTY;
  const long long ray_blocks_ll = (long long)tiles_x * tiles_y * b;
  const size_t count = D11h ? (size_t)b * H * W * F : 0;
  const size_t count8 = count / 8;
  size_t cast_blocks = (count8 + 255) / 256;
  if (cast_blocks > (size_t)kNumSMs * 16) cast_blocks = (size_t)kNumSMs * 16;
  if (cast_blocks < 1) cast_blocks = 1;
  if (ray_blocks_ll + (long long)cast_blocks > 0x7fffffffLL) return M3R_EINVAL;
  const int ray_blocks = (int)ray_blocks_ll;
  if (rays && D11h && ray_blocks >= 2 * (int)cast_blocks) {
    match_prep_kernel<CH><<<ray_blocks + (int)cast_blocks, 256, 0, st>>>(
        X11, H, W, tiles_x, tiles_y, ray_blocks, rays, D11, count8, count, (int)cast_blocks, D11h);
    return after_launch();
  }
  if (rays) {
    prep_rays_kernel<CH><<<ray_blocks, 256, 0, st>>>(X11, H, W, tiles_x, tiles_y, rays);
    M3R_TRY(after_launch());
  }
  if (D11h) {
    cast_f16_kernel<<<(int)cast_blocks, 256, 0, st>>>(D11, count8, count, D11h);
    M3R_TRY(after_launch());
  }
  return M3R_OK;
}

extern "C" int m3r_match_prep(const float* X11, const float* D11, int b, int H, int W, int F,
                              float* rays_img_with_grad, uint16_t* D11h, m3r_stream_t stream) {
  M3R_NVTX("m3r_match_prep");
  return match_prep_launch<9>(X11, D11, b, H, W, F, rays_img_with_grad, D11h, stream);
}

extern "C" int m3r_match_prep_packed(const float* X11, const float* D11, int b, int H, int W, int F,
                                     float* rays_packed, uint16_t* D11h, m3r_stream_t stream) {
  M3R_NVTX("m3r_match_prep_packed");
  return match_prep_launch<12>(X11, D11, b, H, W, F, rays_packed, D11h, stream);
}
