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
//     defaults 32, 2, 3), normalize(v) as v * (1 / sqrt(v.v)), the perspective divide and the
//     attribute interpolation with one reciprocal each, length(t) > 1 as t.t > 1 — so that every
//     operation is an IEEE single-rounding one and divisions are few: the CPU restatement
//     (oracle/m3r_render_oracle.c) then reproduces the image bit for bit;
//   * targets are limited to 2^24 pixels and 8192 per side.
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

// normalize(v) := v * (1 / sqrt(v.v))  — one square root, one division (GLSL leaves the precision open)
__device__ __forceinline__ V3 normalize_rcp(V3 a) {
  const float inv = divf(1.f, __fsqrt_rn(dot3(a, a)));
  return v3(mulf(a.x, inv), mulf(a.y, inv), mulf(a.z, inv));
}

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

// Fragment shader (surfelmap.glsl:106-135, trianglemap.glsl:118-143).  N is the view-space normal,
// already normalised (it is constant over the primitive).
__device__ __forceinline__ uint32_t shade(const RenderParams& P, float spec_noimg, V3 N, V3 position, V3 color) {
  const bool ui = P.use_img != 0;
  const float kA = ui ? P.phong[0] : 0.1f, kD = ui ? P.phong[1] : 0.2f, kS = ui ? P.phong[2] : 1.3f;
  const V3 fc = ui ? color : v3(P.base_color[0], P.base_color[1], P.base_color[2]);
  const float sp = ui ? P.spec : spec_noimg;
  const V3 L = normalize_rcp(sub3(v3(P.light[0], P.light[1], P.light[2]), position));
  const float lambertian = fmaxf(dot3(N, L), 0.f);
  float specular = 0.f;
  if (lambertian > 0.f) {
    // R = 2 * (L * N) * N - L with a COMPONENT-WISE product, as written in the shader
    const V3 R = v3(subf(mulf(mulf(2.f, mulf(L.x, N.x)), N.x), L.x),
                    subf(mulf(mulf(2.f, mulf(L.y, N.y)), N.y), L.y),
                    subf(mulf(mulf(2.f, mulf(L.z, N.z)), N.z), L.z));
    const V3 V = normalize_rcp(v3(-position.x, -position.y, -position.z));
    specular = pow_spec(fmaxf(dot3(R, V), 0.f), sp);
  }
  const float diff = addf(kA, mulf(lambertian, kD)), sadd = mulf(kS, specular);
  V3 out = v3(addf(mulf(fc.x, diff), sadd), addf(mulf(fc.y, diff), sadd), addf(mulf(fc.z, diff), sadd));
  if (P.show_normal) {  // N = (N.x, -N.y, -N.z); out = -N * 0.5 + 0.5
    out = v3(addf(mulf(-N.x, 0.5f), 0.5f), addf(mulf(N.y, 0.5f), 0.5f), addf(mulf(N.z, 0.5f), 0.5f));
  }
  return to_unorm8(out.x) | (to_unorm8(out.y) << 8) | (to_unorm8(out.z) << 16) | 0xFF000000u;
}

constexpr int kRenderThreads = 128;

// Per-primitive record in shared memory, one column per thread (word-major: lanes that read
// the same word of different primitives hit different banks, lanes on the same primitive broadcast).
//   ints  : X[4], Y[4] snapped window coordinates (1/256 px), x0, y0, bw (bounding box), flags
//   floats: zndc[4], rw[4] (1/w_clip), pos[4][3], N[3] (normalised view-space normal),
//           inv_area[2], col[3] (surfel) or col[4][3] (triangle map)
enum { I_X = 0, I_Y = 4, I_X0 = 8, I_Y0 = 9, I_BW = 10, I_FLAGS = 11, kPrimInts = 12 };
enum { F_Z = 0, F_RW = 4, F_POS = 8, F_N = 20, F_IA = 23, F_COL = 25 };
template <int PROGRAM> struct PrimWords { static constexpr int kFloats = PROGRAM == 0 ? 28 : 37; };
// flags: bit 0/1 = triangle 0/1 drawn, bit 2/3 = triangle 0/1 flipped to CCW (culling off)

struct PrimVertex { float cx, cy, cz, cw; V3 pos; };

__device__ __forceinline__ long long edge_fn(int ax, int ay, int bx, int by, int px, int py) {
  return (long long)(bx - ax) * (long long)(py - ay) - (long long)(by - ay) * (long long)(px - ax);
}
// pixel centres exactly on an edge belong to the triangle iff the edge is a left or a top edge
__device__ __forceinline__ bool owns_edge(int dx, int dy) { return dy < 0 || (dy == 0 && dx < 0); }

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

__device__ __forceinline__ void transform_vertex(const float* mv, const float* mvp, V3 p, PrimVertex& v) {
  v.pos = v3(row_apply(mv, p), row_apply(mv + 4, p), row_apply(mv + 8, p));
  v.cx = row_apply(mvp, p); v.cy = row_apply(mvp + 4, p); v.cz = row_apply(mvp + 8, p);
  v.cw = row_apply(mvp + 12, p);
}
__device__ __forceinline__ V3 mat3_apply(const float* mv, V3 n) {
  return v3(fmaf_(mv[0], n.x, fmaf_(mv[1], n.y, mulf(mv[2], n.z))),
            fmaf_(mv[4], n.x, fmaf_(mv[5], n.y, mulf(mv[6], n.z))),
            fmaf_(mv[8], n.x, fmaf_(mv[9], n.y, mulf(mv[10], n.z))));
}

// Viewport transform + snapping of the 4 strip vertices, triangle set-up (orientation, cull,
// 1/area), bounding box of what will be drawn.  Writes the record; returns its pixel count.
template <int PROGRAM>
__device__ __forceinline__ int prim_setup(const RenderParams& P, const PrimVertex* v, int* sI, float* sF) {
  int X[4], Y[4];
  unsigned okm = 0;  // bit k: strip vertex k is drawable
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float rw = divf(1.f, v[k].cw);
    const float nx = mulf(v[k].cx, rw), ny = mulf(v[k].cy, rw), nz = mulf(v[k].cz, rw);
    const float wx = mulf(fmaf_(nx, 0.5f, 0.5f), (float)P.fbw), wy = mulf(fmaf_(ny, 0.5f, 0.5f), (float)P.fbh);
    const bool ok = (v[k].cw > 0.f) && (fabsf(wx) <= 1048576.f) && (fabsf(wy) <= 1048576.f) && (fabsf(nz) <= 3.0e38f);
    okm |= (ok ? 1u : 0u) << k;
    X[k] = ok ? __float2int_rn(mulf(wx, 256.f)) : 0;
    Y[k] = ok ? __float2int_rn(mulf(wy, 256.f)) : 0;
    sI[(I_X + k) * kRenderThreads] = X[k];
    sI[(I_Y + k) * kRenderThreads] = Y[k];
    sF[(F_Z + k) * kRenderThreads] = nz;
    sF[(F_RW + k) * kRenderThreads] = rw;
    sF[(F_POS + 3 * k + 0) * kRenderThreads] = v[k].pos.x;
    sF[(F_POS + 3 * k + 1) * kRenderThreads] = v[k].pos.y;
    sF[(F_POS + 3 * k + 2) * kRenderThreads] = v[k].pos.z;
  }
  int flags = 0, minx = 0x7fffffff, maxx = -0x7fffffff, miny = 0x7fffffff, maxy = -0x7fffffff;
  {
#pragma unroll
    for (int t = 0; t < 2; t++) {  // strip: (0,1,2), (2,1,3)
      const int a = t ? 2 : 0, b = 1, c = t ? 3 : 2;
      long long area2 = edge_fn(X[a], Y[a], X[b], Y[b], X[c], Y[c]);
      bool draw = area2 != 0 && ((okm >> a) & (okm >> b) & (okm >> c) & 1u);
      if (area2 < 0) {
        if (P.cull) draw = false;  // back face (CCW is front, GL default)
        else { flags |= 4 << t; area2 = -area2; }
      }
      if (draw) {
        flags |= 1 << t;
        minx = min(minx, min(X[a], min(X[b], X[c]))); maxx = max(maxx, max(X[a], max(X[b], X[c])));
        miny = min(miny, min(Y[a], min(Y[b], Y[c]))); maxy = max(maxy, max(Y[a], max(Y[b], Y[c])));
      }
      sF[(F_IA + t) * kRenderThreads] = draw ? divf(1.f, (float)area2) : 0.f;
    }
  }
  int count = 0, x0 = 0, y0 = 0, bw = 0;
  if (flags & 3) {  // pixel (px, py) has its centre at (256 px + 128, 256 py + 128)
    x0 = max(0, (minx - 128 + 255) >> 8);
    y0 = max(0, (miny - 128 + 255) >> 8);
    const int x1 = min(P.fbw - 1, (maxx - 128) >> 8), y1 = min(P.fbh - 1, (maxy - 128) >> 8);
    if (x0 <= x1 && y0 <= y1) {
      bw = x1 - x0 + 1;
      const long long c = (long long)bw * (y1 - y0 + 1);
      count = c > (1LL << 24) ? (1 << 24) : (int)c;  // unreachable for fbw*fbh <= 2^24; larger targets clamp
    }
  }
  sI[I_X0 * kRenderThreads] = x0; sI[I_Y0 * kRenderThreads] = y0; sI[I_BW * kRenderThreads] = bw;
  sI[I_FLAGS * kRenderThreads] = flags;
  return count;
}

// A fragment that survived coverage, depth range, the early depth test and (surfel) the disk test,
// queued for shading: which record / triangle / pixel, its depth bits and perspective weights.
struct FragEntry { uint32_t where; uint32_t zbits; float qa, qb, qc, iq; };
constexpr int kFragQueue = 64;  // per warp: at most 31 waiting + 32 pushed

__device__ __forceinline__ void tri_indices(int flags, int t, int& ia, int& ib, int& ic) {
  const bool flip = (flags >> (2 + t)) & 1;
  ia = t ? 2 : 0; ib = flip ? (t ? 3 : 2) : 1; ic = flip ? 1 : (t ? 3 : 2);
}

// Coverage of strip triangle t of record (sI, sF) at pixel (px, py); fills `e` when a fragment results.
template <int PROGRAM>
__device__ __forceinline__ bool coverage_item(const RenderParams& P, const int* sI, const float* sF, int t, int px,
                                              int py, const unsigned long long* __restrict__ fb, FragEntry& e) {
  const int flags = sI[I_FLAGS * kRenderThreads];
  if (!((flags >> t) & 1)) return false;
  int ia, ib, ic;
  tri_indices(flags, t, ia, ib, ic);
  const int cx = 256 * px + 128, cy = 256 * py + 128;
  const int ax = sI[(I_X + ia) * kRenderThreads], ay = sI[(I_Y + ia) * kRenderThreads];
  const int bx = sI[(I_X + ib) * kRenderThreads], by = sI[(I_Y + ib) * kRenderThreads];
  const int ex = sI[(I_X + ic) * kRenderThreads], ey = sI[(I_Y + ic) * kRenderThreads];
  const long long wa = edge_fn(bx, by, ex, ey, cx, cy);
  const long long wb = edge_fn(ex, ey, ax, ay, cx, cy);
  const long long wc = edge_fn(ax, ay, bx, by, cx, cy);
  if (wa < 0 || wb < 0 || wc < 0) return false;
  if ((wa == 0 && !owns_edge(ex - bx, ey - by)) || (wb == 0 && !owns_edge(ax - ex, ay - ey)) ||
      (wc == 0 && !owns_edge(bx - ax, by - ay)))
    return false;
  const float ia2 = sF[(F_IA + t) * kRenderThreads];
  const float la = mulf((float)wa, ia2), lb = mulf((float)wb, ia2), lc = mulf((float)wc, ia2);
  const float zn = fmaf_(lc, sF[(F_Z + ic) * kRenderThreads],
                         fmaf_(lb, sF[(F_Z + ib) * kRenderThreads], mulf(la, sF[(F_Z + ia) * kRenderThreads])));
  const float zw = fmaf_(zn, 0.5f, 0.5f);
  if (!(zw >= 0.f && zw < 1.f)) return false;
  e.zbits = __float_as_uint(zw);
  // early depth test: the stored word only ever decreases, so a stale read is conservative
  const unsigned long long cur = *reinterpret_cast<const volatile unsigned long long*>(fb + (size_t)py * P.fbw + px);
  if ((uint32_t)(cur >> 32) < e.zbits) return false;
  // perspective-correct weights lambda_i / w_i, normalised by one reciprocal
  e.qa = mulf(la, sF[(F_RW + ia) * kRenderThreads]);
  e.qb = mulf(lb, sF[(F_RW + ib) * kRenderThreads]);
  e.qc = mulf(lc, sF[(F_RW + ic) * kRenderThreads]);
  e.iq = divf(1.f, addf(addf(e.qa, e.qb), e.qc));
  if (PROGRAM == 0) {  // surfelmap.glsl:107-111: fragTexCoord of strip vertex k is (+-1, +-1)
    const float tu = mulf(fmaf_(e.qc, (ic & 1) ? 1.f : -1.f, fmaf_(e.qb, (ib & 1) ? 1.f : -1.f, mulf(e.qa, (ia & 1) ? 1.f : -1.f))), e.iq);
    const float tv = mulf(fmaf_(e.qc, (ic & 2) ? 1.f : -1.f, fmaf_(e.qb, (ib & 2) ? 1.f : -1.f, mulf(e.qa, (ia & 2) ? 1.f : -1.f))), e.iq);
    if (fmaf_(tu, tu, mulf(tv, tv)) > 1.0f) return false;  // length(t) > 1  <=>  t.t > 1
  }
  return true;
}

// Fragment shader + commit for one queued fragment of record column `col` (lanes fully converged).
template <int PROGRAM>
__device__ __forceinline__ void shade_item(const RenderParams& P, const int* sIw, const float* sFw,
                                           const FragEntry& e, unsigned long long* __restrict__ fb) {
  const int col = e.where & 31, t = (e.where >> 5) & 1, px = (e.where >> 6) & 0x1FFF, py = e.where >> 19;
  const int* sI = sIw + col;
  const float* sF = sFw + col;
  int ia, ib, ic;
  tri_indices(sI[I_FLAGS * kRenderThreads], t, ia, ib, ic);
#define M3R_INTERP(fa_, fb_, fc_) mulf(fmaf_(e.qc, fc_, fmaf_(e.qb, fb_, mulf(e.qa, fa_))), e.iq)
#define M3R_POS(i_, c_) sF[(F_POS + 3 * (i_) + (c_)) * kRenderThreads]
  const V3 pos = v3(M3R_INTERP(M3R_POS(ia, 0), M3R_POS(ib, 0), M3R_POS(ic, 0)),
                    M3R_INTERP(M3R_POS(ia, 1), M3R_POS(ib, 1), M3R_POS(ic, 1)),
                    M3R_INTERP(M3R_POS(ia, 2), M3R_POS(ib, 2), M3R_POS(ic, 2)));
#undef M3R_POS
  V3 colr;
  if (PROGRAM == 0) {
    colr = v3(sF[(F_COL + 0) * kRenderThreads], sF[(F_COL + 1) * kRenderThreads], sF[(F_COL + 2) * kRenderThreads]);
  } else {
#define M3R_COL(i_, c_) sF[(F_COL + 3 * (i_) + (c_)) * kRenderThreads]
    colr = v3(M3R_INTERP(M3R_COL(ia, 0), M3R_COL(ib, 0), M3R_COL(ic, 0)),
              M3R_INTERP(M3R_COL(ia, 1), M3R_COL(ib, 1), M3R_COL(ic, 1)),
              M3R_INTERP(M3R_COL(ia, 2), M3R_COL(ib, 2), M3R_COL(ic, 2)));
#undef M3R_COL
  }
#undef M3R_INTERP
  const V3 N = v3(sF[(F_N + 0) * kRenderThreads], sF[(F_N + 1) * kRenderThreads], sF[(F_N + 2) * kRenderThreads]);
  const uint32_t rgba = shade(P, PROGRAM == 0 ? 2.f : 3.f, N, pos, colr);
  atomicMin(fb + (size_t)py * P.fbw + px, ((unsigned long long)e.zbits << 32) | rgba);
}

// PROGRAM 0: surfelmap.glsl geometry shader (:28-90) — a square of half-side `radius` in the
// tangent plane of the surface, emitted as the strip p1 p2 p3 p4.
// PROGRAM 1: trianglemap.glsl geometry shader (:41-93) — the pixel quad TL BL TR BR itself.
// Phase 1: one thread = one geometry-shader invocation -> a primitive record in shared memory.
// Phase 2: the warp flattens the (primitive, bounding-box pixel) pairs of its 32 records with
// a prefix sum and walks them 32 at a time, so lanes stay converged whatever the primitives'
// sizes (a lane-per-primitive loop ran at 12.8 active lanes of 32).
template <int PROGRAM>
__global__ void __launch_bounds__(kRenderThreads)
render_kernel(const RenderParams P, const float* __restrict__ X, const float* __restrict__ Cf,
              const float* __restrict__ img, const float* __restrict__ models,
              unsigned long long* __restrict__ fb) {
  __shared__ float mv[16], mvp[16];
  __shared__ int sI[kPrimInts][kRenderThreads];
  __shared__ float sF[PrimWords<PROGRAM>::kFloats][kRenderThreads];
  __shared__ int sPre[kRenderThreads];
  __shared__ FragEntry sQ[kRenderThreads / 32][kFragQueue];
  block_matrices(P, models, mv, mvp);
  const int tid = threadIdx.x, lane = tid & 31, wbase = tid & ~31;
  const int id = blockIdx.x * kRenderThreads + tid;
  int count = 0;
  const int border = PROGRAM == 0 ? 1 : 10;
  const int y = id / P.W, x = id - y * P.W;
  const size_t base = (size_t)blockIdx.y * P.H * P.W;
  bool emit = id < P.H * P.W && !(x < border || x >= P.W - border || y < border || y >= P.H - border);
  if (emit && __ldg(Cf + base + id) < P.conf_threshold) emit = false;  // trianglemap reads conf at (x, y) for all four
  if (emit) {
    const float* Xk = X + base * 3;
    PrimVertex v[4];
    V3 Nv;
    if (PROGRAM == 0) {
      const V3 p00 = ld3(Xk + (size_t)id * 3), p01 = ld3(Xk + (size_t)(id + P.W) * 3), p10 = ld3(Xk + (size_t)(id + 1) * 3);
      const V3 right = sub3(p10, p00), down = sub3(p01, p00);
      const V3 N = normalize_rcp(cross3(down, right));
      const V3 tangent = normalize_rcp(cross3(N, v3(0.f, -1.f, 0.f)));  // "assume y up"
      const V3 bitangent = cross3(N, tangent);
      Nv = mat3_apply(mv, N);
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const float st = (k & 1) ? 1.f : -1.f, sb = (k & 2) ? 1.f : -1.f;  // p1..p4: (-,-) (+,-) (-,+) (+,+)
        // xyz00 + (+-tangent +- bitangent) * radius, as written (surfelmap.glsl:55-58)
        const V3 p = v3(addf(p00.x, mulf(addf(mulf(st, tangent.x), mulf(sb, bitangent.x)), P.radius)),
                        addf(p00.y, mulf(addf(mulf(st, tangent.y), mulf(sb, bitangent.y)), P.radius)),
                        addf(p00.z, mulf(addf(mulf(st, tangent.z), mulf(sb, bitangent.z)), P.radius)));
        transform_vertex(mv, mvp, p, v[k]);
      }
      const V3 col = ld3(img + (base + id) * 3);
      sF[F_COL + 0][tid] = col.x; sF[F_COL + 1][tid] = col.y; sF[F_COL + 2][tid] = col.z;
    } else {
      const float* Ik = img + base * 3;
      const int ids[4] = {id, id + 1, id + P.W, id + P.W + 1};  // TL TR BL BR
      V3 pos[4];
#pragma unroll
      for (int k = 0; k < 4; k++) pos[k] = ld3(Xk + (size_t)ids[k] * 3);
      const V3 n1 = normalize_rcp(cross3(sub3(pos[2], pos[0]), sub3(pos[1], pos[0])));  // normal(TL, BL, TR)
      const V3 n2 = normalize_rcp(cross3(sub3(pos[2], pos[1]), sub3(pos[3], pos[1])));  // normal(TR, BL, BR)
      const V3 ray1 = normalize_rcp(pos[0]), ray2 = normalize_rcp(pos[1]);
      if (fabsf(dot3(n1, ray1)) < P.slant_threshold || fabsf(dot3(n2, ray2)) < P.slant_threshold) emit = false;
      Nv = mat3_apply(mv, v3(mulf(addf(n1.x, n2.x), 0.5f), mulf(addf(n1.y, n2.y), 0.5f), mulf(addf(n1.z, n2.z), 0.5f)));
      const int order[4] = {0, 2, 1, 3};  // TL, BL, TR, BR ("CCW")
#pragma unroll
      for (int k = 0; k < 4; k++) {
        transform_vertex(mv, mvp, pos[order[k]], v[k]);
        v[k].cz = subf(v[k].cz, P.depth_bias);
        const V3 col = ld3(Ik + (size_t)ids[order[k]] * 3);
        sF[F_COL + 3 * k + 0][tid] = col.x; sF[F_COL + 3 * k + 1][tid] = col.y; sF[F_COL + 3 * k + 2][tid] = col.z;
      }
    }
    if (emit) {
      const V3 Nn = normalize_rcp(Nv);
      sF[F_N + 0][tid] = Nn.x; sF[F_N + 1][tid] = Nn.y; sF[F_N + 2][tid] = Nn.z;
      count = prim_setup<PROGRAM>(P, v, &sI[0][tid], &sF[0][tid]);
    }
  }
  // exclusive prefix of the pixel counts over the warp
  int incl = count;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int nb = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += nb;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  sPre[tid] = incl - count;
  __syncwarp();
  // Phase 2a: coverage for 32 (record, pixel) pairs at a time; survivors are queued (ballot
  // compaction) and — Phase 2b — shaded 32 at a time, so the fragment shader also runs converged
  // (a bounding box is ~40 % covered by its disk / triangle).
  FragEntry* q = sQ[tid >> 5];
  int qhead = 0, qcount = 0;  // warp-uniform
  const int* sIw = &sI[0][wbase];
  const float* sFw = &sF[0][wbase];
  auto drain = [&](int n) {  // shade the n (<= 32) oldest entries
    if (lane < n) shade_item<PROGRAM>(P, sIw, sFw, q[(qhead + lane) & (kFragQueue - 1)], fb);
    qhead = (qhead + n) & (kFragQueue - 1);
    qcount -= n;
  };
  for (int tb = 0; tb < total; tb += 32) {
    const int t = tb + lane;
    int colw = 0, px = 0, py = 0;
    const bool have = t < total;
    if (have) {
      int lo = 0;  // owner = last record whose exclusive prefix is <= t
#pragma unroll
      for (int s = 16; s > 0; s >>= 1)
        if (sPre[wbase + lo + s] <= t) lo += s;
      colw = lo;
      const int local = t - sPre[wbase + lo];
      const int bw = sI[I_BW][wbase + lo];
      const int ry = local / bw, rx = local - ry * bw;
      px = sI[I_X0][wbase + lo] + rx; py = sI[I_Y0][wbase + lo] + ry;
    }
#pragma unroll
    for (int tri = 0; tri < 2; tri++) {
      FragEntry e;
      const bool hit = have && coverage_item<PROGRAM>(P, sIw + colw, sFw + colw, tri, px, py, fb, e);
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (hit) {
        e.where = (uint32_t)colw | ((uint32_t)tri << 5) | ((uint32_t)px << 6) | ((uint32_t)py << 19);
        q[(qhead + qcount + __popc(m & ((1u << lane) - 1u))) & (kFragQueue - 1)] = e;
      }
      qcount += __popc(m);
      __syncwarp();
      if (qcount >= 32) { drain(32); __syncwarp(); }
    }
  }
  if (qcount > 0) drain(qcount);
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
  // <= 2^24 target pixels: 32 bounding boxes then sum to < 2^31 in the warp's prefix sum
  if ((long long)view->width * view->height > (1LL << 24) || (long long)H * W > (1LL << 30)) return M3R_EINVAL;
  if (view->width > 8192 || view->height > 8192) return M3R_EINVAL;  // 13-bit pixel fields of the fragment queue
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
  M3R_NVTX("m3r_render_clear");
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
    render_kernel<0><<<grid, kRenderThreads, 0, as_stream(stream)>>>(P, X, C, img, models, f);
  else
    render_kernel<1><<<grid, kRenderThreads, 0, as_stream(stream)>>>(P, X, C, img, models, f);
  return after_launch();
}

extern "C" int m3r_render_surfels(const float* X, const float* C, const float* img,
                                  const float* models, int num_keyframes, int H, int W,
                                  const m3r_render_view* view, const m3r_render_opts* opts,
                                  uint64_t* fb, m3r_stream_t stream) {
  M3R_NVTX("m3r_render_surfels");
  return render_launch(0, X, C, img, models, num_keyframes, H, W, view, opts, fb, stream);
}

extern "C" int m3r_render_triangles(const float* X, const float* C, const float* img,
                                    const float* models, int num_keyframes, int H, int W,
                                    const m3r_render_view* view, const m3r_render_opts* opts,
                                    uint64_t* fb, m3r_stream_t stream) {
  M3R_NVTX("m3r_render_triangles");
  return render_launch(1, X, C, img, models, num_keyframes, H, W, view, opts, fb, stream);
}

extern "C" int m3r_render_resolve(const uint64_t* fb, int width, int height, uint8_t* rgba,
                                  float* depth, m3r_stream_t stream) {
  M3R_NVTX("m3r_render_resolve");
  if (!fb || width <= 0 || height <= 0) return M3R_EINVAL;
  const size_t count = (size_t)width * height;
  const int grid = (int)((count + 255) / 256 < (size_t)(kNumSMs * 8) ? (count + 255) / 256 : (size_t)(kNumSMs * 8));
  render_resolve_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const unsigned long long*>(fb), width, height, reinterpret_cast<uint32_t*>(rgba), depth);
  return after_launch();
}
