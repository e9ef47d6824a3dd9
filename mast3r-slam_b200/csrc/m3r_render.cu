// m3r_render.cu — headless splat of the viewer's pointmap shaders into a device buffer (sm_100a).
//
// Restates, as CUDA kernels, the arithmetic of the reference viewer's two GLSL programs
//   resources/programs/surfelmap.glsl   (geometry :28-90, fragment :106-135)
//   resources/programs/trianglemap.glsl (geometry :41-93, fragment :118-143)
// as driven by Window.render_pointmap (visualization.py:334-356: model = T_WC.matrix(),
// conf_threshold, use_img, depth_bias; DEPTH_TEST + CULL_FACE, :100-103).  There is no graphics
// API here: the "window" is a caller-owned device buffer of 64-bit words
//     fb[y * width + x] = (float_bits(window depth) << 32) | RGBA8        (y up, GL window origin)
// and the depth test GL_LESS is a single 64-bit atomicMin (window depths lie in [0,1), where the
// float bit pattern is monotone).  One thread = one pointmap pixel = one geometry-shader
// invocation: it emits its quad, rasterises the two strip triangles itself (primitives are a
// few pixels wide), runs the fragment shader per covered pixel and commits with the atomic.
//
// Stated deviations from a GL rasteriser (SURVEY 8f last row; parity is "unpinned": no GL here):
//   * vertices are snapped to 1/256 pixel and coverage is decided by exact 64-bit edge functions
//     at pixel centres with a top-left tie rule (GL leaves the tie rule to the implementation);
//   * a primitive with a vertex at w <= 0, a non-finite vertex, or a vertex beyond +-2^20 px is
//     dropped instead of clipped; near/far clipping is per fragment (0 <= depth < 1);
//   * equal depths are resolved by the smaller colour word, not by draw order (deterministic);
//   * pow(x, spec) is evaluated by repeated squaring when spec is a small integer (the
//     defaults 32, 2, 3), so that every operation is an IEEE single-rounding one: the CPU
//     restatement (oracle/m3r_render_oracle.c) then reproduces the image bit for bit.
// Every float operation is an explicit round-to-nearest intrinsic (no FMA contraction except the
// fmaf written here), mirrored one to one by the oracle.
#include "m3r_common.cuh"

namespace m3r {

struct RenderParams {
  float camera[16], proj[16];  // row-major: clip = proj * camera * model * [p, 1]
  int H, W;                    // pointmap size
  int fbw, fbh;                // target size
  float conf_threshold, radius, slant_threshold, depth_bias;
  float light[3], phong[3], spec, base_color[3];
  int use_img, show_normal, cull;
};

__device__ __forceinline__ float mulf(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float addf(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float subf(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float divf(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fmaf_(float a, float b, float c) { return __fmaf_rn(a, b, c); }

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 sub3(V3 a, V3 b) { return v3(subf(a.x, b.x), subf(a.y, b.y), subf(a.z, b.z)); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return fmaf_(a.x, b.x, fmaf_(a.y, b.y, mulf(a.z, b.z))); }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
  return v3(fmaf_(a.y, b.z, -mulf(a.z, b.y)), fmaf_(a.z, b.x, -mulf(a.x, b.z)),
            fmaf_(a.x, b.y, -mulf(a.y, b.x)));
}
__device__ __forceinline__ V3 normalize3v(V3 a) {
  const float len = __fsqrt_rn(dot3(a, a));
  return v3(divf(a.x, len), divf(a.y, len), divf(a.z, len));
}
// rows 0..2 of a row-major 4x4 applied to [p, 1]
__device__ __forceinline__ float row_apply(const float* m, V3 p) {
  return fmaf_(m[0], p.x, fmaf_(m[1], p.y, fmaf_(m[2], p.z, m[3])));
}
__device__ __forceinline__ V3 ld3(const float* p) { return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2)); }

// (a * b) for row-major 4x4 float matrices: exact double products, fixed-order double sums,
// one rounding to float — independent of FMA contraction, reproduced by the oracle.
__device__ void mat4_mul(const float* a, const float* b, float* out) {
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0.0;
      for (int k = 0; k < 4; k++) s += (double)a[i * 4 + k] * (double)b[k * 4 + j];
      out[i * 4 + j] = (float)s;
    }
}

struct Vtx {
  float cx, cy, cz, cw;  // clip position
  V3 pos;                // view-space position (mv * p)
  V3 col;                // vertex colour (triangle map)
  float tu, tv;          // fragTexCoord (surfel map)
};

struct Shade {
  V3 normal;   // view-space normal, constant over the primitive
  V3 col;      // constant colour (surfel map)
  bool disk;   // surfel: discard |texcoord| > 1, colour constant; triangle: colour interpolated
  float spec_power_noimg;
};

__device__ __forceinline__ float pow_spec(float x, float p) {
  const int n = (int)p;
  if ((float)n == p && n >= 1 && n <= 64) {  // square-and-multiply, fixed order
    float r = 1.f, b = x;
    bool have = false;
    for (int e = n; e > 0; e >>= 1) {
      if (e & 1) { r = have ? mulf(r, b) : b; have = true; }
      b = mulf(b, b);
    }
    return r;
  }
  return powf(x, p);
}

__device__ __forceinline__ uint32_t to_unorm8(float c) {
  if (!(c > 0.f)) return 0u;  // NaN and negatives
  if (c > 1.f) c = 1.f;
  return (uint32_t)(int)addf(mulf(c, 255.f), 0.5f);
}

// Fragment shader (surfelmap.glsl:106-135, trianglemap.glsl:118-143).
__device__ __forceinline__ uint32_t shade(const RenderParams& P, const Shade& S, V3 position, V3 color) {
  const bool ui = P.use_img != 0;
  const float kA = ui ? P.phong[0] : 0.1f, kD = ui ? P.phong[1] : 0.2f, kS = ui ? P.phong[2] : 1.3f;
  const V3 fc = ui ? color : v3(P.base_color[0], P.base_color[1], P.base_color[2]);
  const float sp = ui ? P.spec : S.spec_power_noimg;
  const V3 L = normalize3v(sub3(v3(P.light[0], P.light[1], P.light[2]), position));
  V3 N = normalize3v(S.normal);
  const float lambertian = fmaxf(dot3(N, L), 0.f);
  float specular = 0.f;
  if (lambertian > 0.f) {
    // R = 2 * (L * N) * N - L with a COMPONENT-WISE product, as written in the shader
    const V3 R = v3(subf(mulf(mulf(2.f, mulf(L.x, N.x)), N.x), L.x),
                    subf(mulf(mulf(2.f, mulf(L.y, N.y)), N.y), L.y),
                    subf(mulf(mulf(2.f, mulf(L.z, N.z)), N.z), L.z));
    const V3 V = normalize3v(v3(-position.x, -position.y, -position.z));
    specular = pow_spec(fmaxf(dot3(R, V), 0.f), sp);
  }
  const float diff = addf(kA, mulf(lambertian, kD)), sadd = mulf(kS, specular);
  V3 out = v3(addf(mulf(fc.x, diff), sadd), addf(mulf(fc.y, diff), sadd), addf(mulf(fc.z, diff), sadd));
  if (P.show_normal) {  // N = (N.x, -N.y, -N.z); out = -N * 0.5 + 0.5
    out = v3(addf(mulf(-N.x, 0.5f), 0.5f), addf(mulf(N.y, 0.5f), 0.5f), addf(mulf(N.z, 0.5f), 0.5f));
  }
  return to_unorm8(out.x) | (to_unorm8(out.y) << 8) | (to_unorm8(out.z) << 16) | 0xFF000000u;
}

struct Snap { long long x, y; float zndc, w; bool ok; };

__device__ __forceinline__ Snap snap_vertex(const RenderParams& P, const Vtx& v) {
  Snap s;
  s.w = v.cw;
  s.ok = v.cw > 0.f;
  const float nx = divf(v.cx, v.cw), ny = divf(v.cy, v.cw);
  s.zndc = divf(v.cz, v.cw);
  const float wx = mulf(fmaf_(nx, 0.5f, 0.5f), (float)P.fbw), wy = mulf(fmaf_(ny, 0.5f, 0.5f), (float)P.fbh);
  const float lim = 1048576.f;
  s.ok = s.ok && (fabsf(wx) <= lim) && (fabsf(wy) <= lim) && (s.zndc == s.zndc) && (fabsf(s.zndc) <= 3.0e38f);
  s.x = s.ok ? (long long)__float2int_rn(mulf(wx, 256.f)) : 0;
  s.y = s.ok ? (long long)__float2int_rn(mulf(wy, 256.f)) : 0;
  return s;
}

__device__ __forceinline__ long long edge_fn(long long ax, long long ay, long long bx, long long by,
                                             long long px, long long py) {
  return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}
// pixel centres exactly on an edge belong to the triangle iff the edge is a left or a top edge
__device__ __forceinline__ bool owns_edge(long long dx, long long dy) { return dy < 0 || (dy == 0 && dx < 0); }

__device__ __forceinline__ long long floor_div256(long long v) { return v >> 8; }  // arithmetic shift

__device__ void raster_triangle(const RenderParams& P, const Shade& S, unsigned long long* __restrict__ fb,
                                Vtx a, Vtx b, Vtx c) {
  Snap sa = snap_vertex(P, a), sb = snap_vertex(P, b), sc = snap_vertex(P, c);
  if (!(sa.ok && sb.ok && sc.ok)) return;
  long long area2 = edge_fn(sa.x, sa.y, sb.x, sb.y, sc.x, sc.y);
  if (area2 == 0) return;
  if (area2 < 0) {
    if (P.cull) return;  // back face (CCW is front, GL default)
    const Vtx tv = b; b = c; c = tv;
    const Snap ts = sb; sb = sc; sc = ts;
    area2 = -area2;
  }
  const long long minx = min(sa.x, min(sb.x, sc.x)), maxx = max(sa.x, max(sb.x, sc.x));
  const long long miny = min(sa.y, min(sb.y, sc.y)), maxy = max(sa.y, max(sb.y, sc.y));
  // pixel (px, py) has its centre at (256 px + 128, 256 py + 128)
  const int x0 = (int)max(0LL, floor_div256(minx - 128 + 255)), x1 = (int)min((long long)P.fbw - 1, floor_div256(maxx - 128));
  const int y0 = (int)max(0LL, floor_div256(miny - 128 + 255)), y1 = (int)min((long long)P.fbh - 1, floor_div256(maxy - 128));
  if (x0 > x1 || y0 > y1) return;
  const bool own_a = owns_edge(sc.x - sb.x, sc.y - sb.y);  // edge b->c, opposite a
  const bool own_b = owns_edge(sa.x - sc.x, sa.y - sc.y);
  const bool own_c = owns_edge(sb.x - sa.x, sb.y - sa.y);
  const float fa2 = (float)area2;
  for (int py = y0; py <= y1; py++) {
    for (int px = x0; px <= x1; px++) {
      const long long cx = 256LL * px + 128, cy = 256LL * py + 128;
      const long long wa = edge_fn(sb.x, sb.y, sc.x, sc.y, cx, cy);
      const long long wb = edge_fn(sc.x, sc.y, sa.x, sa.y, cx, cy);
      const long long wc = edge_fn(sa.x, sa.y, sb.x, sb.y, cx, cy);
      if (wa < 0 || wb < 0 || wc < 0) continue;
      if ((wa == 0 && !own_a) || (wb == 0 && !own_b) || (wc == 0 && !own_c)) continue;
      const float la = divf((float)wa, fa2), lb = divf((float)wb, fa2), lc = divf((float)wc, fa2);
      const float zn = fmaf_(lc, sc.zndc, fmaf_(lb, sb.zndc, mulf(la, sa.zndc)));
      const float zw = fmaf_(zn, 0.5f, 0.5f);
      if (!(zw >= 0.f && zw < 1.f)) continue;
      unsigned long long* dst = fb + (size_t)py * P.fbw + px;
      const uint32_t zbits = __float_as_uint(zw);
      // early depth test: the stored word only ever decreases, so a stale read is conservative
      if ((uint32_t)(*reinterpret_cast<volatile unsigned long long*>(dst) >> 32) < zbits) continue;
      // perspective-correct weights
      const float qa = divf(la, sa.w), qb = divf(lb, sb.w), qc = divf(lc, sc.w);
      const float qs = addf(addf(qa, qb), qc);
#define M3R_INTERP(fa_, fb_, fc_) divf(fmaf_(qc, fc_, fmaf_(qb, fb_, mulf(qa, fa_))), qs)
      if (S.disk) {  // surfelmap.glsl:107-111
        const float tu = M3R_INTERP(a.tu, b.tu, c.tu), tv = M3R_INTERP(a.tv, b.tv, c.tv);
        const float dist = __fsqrt_rn(fmaf_(tu, tu, mulf(tv, tv)));
        if (dist > 1.0f) continue;
      }
      const V3 pos = v3(M3R_INTERP(a.pos.x, b.pos.x, c.pos.x), M3R_INTERP(a.pos.y, b.pos.y, c.pos.y),
                        M3R_INTERP(a.pos.z, b.pos.z, c.pos.z));
      V3 col = S.col;
      if (!S.disk)
        col = v3(M3R_INTERP(a.col.x, b.col.x, c.col.x), M3R_INTERP(a.col.y, b.col.y, c.col.y),
                 M3R_INTERP(a.col.z, b.col.z, c.col.z));
#undef M3R_INTERP
      const uint32_t rgba = shade(P, S, pos, col);
      atomicMin(dst, ((unsigned long long)zbits << 32) | rgba);
    }
  }
}

// per-CTA: mv = camera * model, mvp = proj * mv for keyframe blockIdx.y
__device__ __forceinline__ void block_matrices(const RenderParams& P, const float* __restrict__ models,
                                               float* mv, float* mvp) {
  if (threadIdx.x == 0) {
    float model[16];
    for (int i = 0; i < 16; i++) model[i] = models[(size_t)blockIdx.y * 16 + i];
    mat4_mul(P.camera, model, mv);
    mat4_mul(P.proj, mv, mvp);
  }
  __syncthreads();
}

__device__ __forceinline__ void transform_vertex(const float* mv, const float* mvp, V3 p, Vtx& v) {
  v.pos = v3(row_apply(mv, p), row_apply(mv + 4, p), row_apply(mv + 8, p));
  v.cx = row_apply(mvp, p); v.cy = row_apply(mvp + 4, p); v.cz = row_apply(mvp + 8, p);
  v.cw = row_apply(mvp + 12, p);
}
__device__ __forceinline__ V3 mat3_apply(const float* mv, V3 n) {
  return v3(fmaf_(mv[0], n.x, fmaf_(mv[1], n.y, mulf(mv[2], n.z))),
            fmaf_(mv[4], n.x, fmaf_(mv[5], n.y, mulf(mv[6], n.z))),
            fmaf_(mv[8], n.x, fmaf_(mv[9], n.y, mulf(mv[10], n.z))));
}

constexpr int kRenderThreads = 128;

// surfelmap.glsl geometry shader (:28-90): a camera-facing-independent square of half-side
// `radius` in the tangent plane of the surface, emitted as the strip p1 p2 p3 p4.
__global__ void __launch_bounds__(kRenderThreads)
render_surfel_kernel(const RenderParams P, const float* __restrict__ X, const float* __restrict__ Cf,
                     const float* __restrict__ img, const float* __restrict__ models,
                     unsigned long long* __restrict__ fb) {
  __shared__ float mv[16], mvp[16];
  block_matrices(P, models, mv, mvp);
  const int id = blockIdx.x * kRenderThreads + threadIdx.x;
  if (id >= P.H * P.W) return;
  const int y = id / P.W, x = id - y * P.W;
  if (x < 1 || x >= P.W - 1 || y < 1 || y >= P.H - 1) return;
  const size_t base = (size_t)blockIdx.y * P.H * P.W;
  if (__ldg(Cf + base + id) < P.conf_threshold) return;
  const float* Xk = X + base * 3;
  const V3 p00 = ld3(Xk + (size_t)id * 3), p01 = ld3(Xk + (size_t)(id + P.W) * 3), p10 = ld3(Xk + (size_t)(id + 1) * 3);
  const V3 right = sub3(p10, p00), down = sub3(p01, p00);
  const V3 N = normalize3v(cross3(down, right));
  const V3 tangent = normalize3v(cross3(N, v3(0.f, -1.f, 0.f)));  // "assume y up"
  const V3 bitangent = cross3(N, tangent);
  Shade S;
  S.normal = mat3_apply(mv, N);
  S.col = ld3(img + (base + id) * 3);
  S.disk = true;
  S.spec_power_noimg = 2.f;
  Vtx v[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float st = (k & 1) ? 1.f : -1.f, sb = (k & 2) ? 1.f : -1.f;  // p1..p4: (-,-) (+,-) (-,+) (+,+)
    // xyz00 + (+-tangent +- bitangent) * radius, as written (surfelmap.glsl:55-58)
    const V3 p = v3(addf(p00.x, mulf(addf(mulf(st, tangent.x), mulf(sb, bitangent.x)), P.radius)),
                    addf(p00.y, mulf(addf(mulf(st, tangent.y), mulf(sb, bitangent.y)), P.radius)),
                    addf(p00.z, mulf(addf(mulf(st, tangent.z), mulf(sb, bitangent.z)), P.radius)));
    transform_vertex(mv, mvp, p, v[k]);
    v[k].tu = st; v[k].tv = sb;
    v[k].col = S.col;
  }
  raster_triangle(P, S, fb, v[0], v[1], v[2]);  // strip: (0,1,2), (2,1,3)
  raster_triangle(P, S, fb, v[2], v[1], v[3]);
}

// trianglemap.glsl geometry shader (:41-93): the pixel quad TL BL TR BR of the pointmap itself.
__global__ void __launch_bounds__(kRenderThreads)
render_triangle_kernel(const RenderParams P, const float* __restrict__ X, const float* __restrict__ Cf,
                       const float* __restrict__ img, const float* __restrict__ models,
                       unsigned long long* __restrict__ fb) {
  __shared__ float mv[16], mvp[16];
  block_matrices(P, models, mv, mvp);
  const int id = blockIdx.x * kRenderThreads + threadIdx.x;
  if (id >= P.H * P.W) return;
  const int y = id / P.W, x = id - y * P.W;
  if (x < 10 || x >= P.W - 10 || y < 10 || y >= P.H - 10) return;
  const size_t base = (size_t)blockIdx.y * P.H * P.W;
  const float* Xk = X + base * 3;
  const float* Ik = img + base * 3;
  const int ids[4] = {id, id + 1, id + P.W, id + P.W + 1};  // TL TR BL BR
  V3 pos[4], col[4];
#pragma unroll
  for (int k = 0; k < 4; k++) { pos[k] = ld3(Xk + (size_t)ids[k] * 3); col[k] = ld3(Ik + (size_t)ids[k] * 3); }
  const V3 n1 = normalize3v(cross3(sub3(pos[2], pos[0]), sub3(pos[1], pos[0])));  // normal(TL, BL, TR)
  const V3 n2 = normalize3v(cross3(sub3(pos[2], pos[1]), sub3(pos[3], pos[1])));  // normal(TR, BL, BR)
  const V3 ray1 = normalize3v(pos[0]), ray2 = normalize3v(pos[1]);
  if (fabsf(dot3(n1, ray1)) < P.slant_threshold) return;
  if (fabsf(dot3(n2, ray2)) < P.slant_threshold) return;
  if (__ldg(Cf + base + id) < P.conf_threshold) return;  // the shader reads conf at (x, y) for all four
  Shade S;
  S.normal = mat3_apply(mv, v3(divf(addf(n1.x, n2.x), 2.f), divf(addf(n1.y, n2.y), 2.f), divf(addf(n1.z, n2.z), 2.f)));
  S.col = v3(0.f, 0.f, 0.f);
  S.disk = false;
  S.spec_power_noimg = 3.f;
  Vtx v[4];
  const int order[4] = {0, 2, 1, 3};  // TL, BL, TR, BR ("CCW")
#pragma unroll
  for (int k = 0; k < 4; k++) {
    transform_vertex(mv, mvp, pos[order[k]], v[k]);
    v[k].cz = subf(v[k].cz, P.depth_bias);
    v[k].col = col[order[k]];
    v[k].tu = 0.f; v[k].tv = 0.f;
  }
  raster_triangle(P, S, fb, v[0], v[1], v[2]);
  raster_triangle(P, S, fb, v[2], v[1], v[3]);
}

__global__ void __launch_bounds__(256)
render_clear_kernel(unsigned long long* __restrict__ fb, size_t count, unsigned long long word) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) fb[i] = word;
}

// fb (y up) -> top-down RGBA8 image and float depth
__global__ void __launch_bounds__(256)
render_resolve_kernel(const unsigned long long* __restrict__ fb, int w, int h, uint32_t* __restrict__ rgba,
                      float* __restrict__ depth) {
  const size_t count = (size_t)w * h;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    const int row = (int)(i / w), col = (int)(i - (size_t)row * w);
    const unsigned long long v = fb[(size_t)(h - 1 - row) * w + col];
    if (rgba) rgba[i] = (uint32_t)v;
    if (depth) depth[i] = __uint_as_float((uint32_t)(v >> 32));
  }
}

static int fill_params(RenderParams& P, const m3r_render_view* view, int H, int W,
                       const m3r_render_opts* o) {
  if (!view || !o || H <= 0 || W <= 0 || view->width <= 0 || view->height <= 0) return M3R_EINVAL;
  if ((long long)view->width * view->height > (1LL << 30) || (long long)H * W > (1LL << 30)) return M3R_EINVAL;
  for (int i = 0; i < 16; i++) { P.camera[i] = view->camera[i]; P.proj[i] = view->proj[i]; }
  P.H = H; P.W = W; P.fbw = view->width; P.fbh = view->height;
  P.conf_threshold = o->conf_threshold; P.radius = o->radius;
  P.slant_threshold = o->slant_threshold; P.depth_bias = o->depth_bias;
  for (int i = 0; i < 3; i++) { P.light[i] = o->lightpos[i]; P.phong[i] = o->phong[i]; P.base_color[i] = o->base_color[i]; }
  P.spec = o->spec;
  P.use_img = o->use_img; P.show_normal = o->show_normal; P.cull = o->cull_face;
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" void m3r_render_default_opts(m3r_render_opts* o) {
  if (!o) return;
  o->conf_threshold = 1.5f;   // WindowMsg.C_conf_threshold, visualization.py:38
  o->radius = 0.003f;         // surfelmap.glsl:17
  o->slant_threshold = 0.1f;  // trianglemap.glsl:17
  o->depth_bias = 0.f;        // trianglemap.glsl:18
  o->lightpos[0] = 0.1f; o->lightpos[1] = 0.1f; o->lightpos[2] = 0.f;  // surfelmap.glsl:99
  o->phong[0] = 0.3f; o->phong[1] = 0.5f; o->phong[2] = 0.4f;
  o->spec = 32.f;
  o->base_color[0] = o->base_color[1] = o->base_color[2] = 1.f;
  o->use_img = 1; o->show_normal = 0; o->cull_face = 1;
}

extern "C" int m3r_render_clear(uint64_t* fb, int width, int height, uint32_t clear_rgba,
                                m3r_stream_t stream) {
  if (!fb || width <= 0 || height <= 0) return M3R_EINVAL;
  const size_t count = (size_t)width * height;
  const unsigned long long word = ((unsigned long long)0x3F800000u << 32) | clear_rgba;  // depth 1.0
  const int grid = (int)((count + 255) / 256 < (size_t)(kNumSMs * 8) ? (count + 255) / 256 : (size_t)(kNumSMs * 8));
  render_clear_kernel<<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<unsigned long long*>(fb), count, word);
  return after_launch();
}

static int render_launch(int program, const float* X, const float* C, const float* img,
                         const float* models, int K, int H, int W, const m3r_render_view* view,
                         const m3r_render_opts* opts, uint64_t* fb, m3r_stream_t stream) {
  if (K < 0 || K > 65535) return M3R_EINVAL;
  RenderParams P;
  M3R_TRY(fill_params(P, view, H, W, opts));
  if (K == 0) return M3R_OK;
  if (!X || !C || !img || !models || !fb) return M3R_EINVAL;
  dim3 grid((H * W + kRenderThreads - 1) / kRenderThreads, K);
  auto* f = reinterpret_cast<unsigned long long*>(fb);
  if (program == 0)
    render_surfel_kernel<<<grid, kRenderThreads, 0, as_stream(stream)>>>(P, X, C, img, models, f);
  else
    render_triangle_kernel<<<grid, kRenderThreads, 0, as_stream(stream)>>>(P, X, C, img, models, f);
  return after_launch();
}

extern "C" int m3r_render_surfels(const float* X, const float* C, const float* img,
                                  const float* models, int num_keyframes, int H, int W,
                                  const m3r_render_view* view, const m3r_render_opts* opts,
                                  uint64_t* fb, m3r_stream_t stream) {
  return render_launch(0, X, C, img, models, num_keyframes, H, W, view, opts, fb, stream);
}

extern "C" int m3r_render_triangles(const float* X, const float* C, const float* img,
                                    const float* models, int num_keyframes, int H, int W,
                                    const m3r_render_view* view, const m3r_render_opts* opts,
                                    uint64_t* fb, m3r_stream_t stream) {
  return render_launch(1, X, C, img, models, num_keyframes, H, W, view, opts, fb, stream);
}

extern "C" int m3r_render_resolve(const uint64_t* fb, int width, int height, uint8_t* rgba,
                                  float* depth, m3r_stream_t stream) {
  if (!fb || width <= 0 || height <= 0) return M3R_EINVAL;
  const size_t count = (size_t)width * height;
  const int grid = (int)((count + 255) / 256 < (size_t)(kNumSMs * 8) ? (count + 255) / 256 : (size_t)(kNumSMs * 8));
  render_resolve_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const unsigned long long*>(fb), width, height, reinterpret_cast<uint32_t*>(rgba), depth);
  return after_launch();
}
