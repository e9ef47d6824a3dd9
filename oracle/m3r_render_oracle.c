/* m3r_render_oracle.c — CPU restatement of the viewer's pointmap shaders.  TEST INFRASTRUCTURE ONLY
 * (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; never by the product).
 *
 * Follows, statement by statement, the reference's GLSL programs
 *   /root/reference/resources/programs/surfelmap.glsl    geometry :28-90, fragment :106-135
 *   /root/reference/resources/programs/trianglemap.glsl  geometry :41-93, fragment :118-143
 * under the state Window.render sets (visualization.py:100-103: DEPTH_TEST, CULL_FACE) and the
 * uniforms Window.render_pointmap writes (visualization.py:334-356).
 *
 * PARITY UNPINNED: the shaders only run under OpenGL, which is unavailable on this box and in the
 * build container, and the reference has no golden images.  What the fixed-function stages
 * between the geometry and fragment shader do is therefore DEFINED here (and in DESIGN.md 4.6):
 *   window x = (ndc.x*0.5+0.5)*width (same for y, origin bottom-left), depth = ndc.z*0.5+0.5;
 *   vertices snapped to 1/256 px (round to nearest even); a pixel is covered when its centre is
 *   inside the triangle by exact integer edge functions, centres on an edge belonging to the
 *   triangle for left edges (dy < 0) and top edges (dy == 0, dx < 0) of the CCW-oriented triangle;
 *   triangle strip a0 a1 a2 a3 -> (a0,a1,a2), (a2,a1,a3); back faces (signed area <= 0) culled;
 *   attributes interpolated perspective-correctly (lambda_i / w_i), depth linearly in window space;
 *   fragments outside 0 <= depth < 1 dropped, primitives with a vertex at w <= 0 / non-finite /
 *   beyond 2^20 px dropped (no geometric clipping); depth test LESS, equal depth -> smaller colour
 *   word; RGBA8 = floor(clamp(c,0,1)*255 + 0.5); pow(x, n) for integral 1 <= n <= 64 by
 *   square-and-multiply; normalize(v) = v * (1/sqrt(v.v)); perspective divide, barycentrics and
 *   attribute interpolation through one reciprocal each; length(t) > 1 as t.t > 1.
 * Compiled with -ffp-contract=off: every * and + below is one IEEE rounding; fmaf is explicit. */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef struct { float x, y, z; } v3;

typedef struct {
  float camera[16], proj[16];
  int width, height;
} orc_render_view;

typedef struct {
  float conf_threshold, radius, slant_threshold, depth_bias;
  float lightpos[3], phong[3], spec, base_color[3];
  int use_img, show_normal, cull_face;
} orc_render_opts;

static v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static float dot(v3 a, v3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
static v3 cross(v3 a, v3 b) {
  return V(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
/* normalize(v) := v * (1 / sqrt(v.v)) */
static v3 normalize(v3 a) {
  const float inv = 1.f / sqrtf(dot(a, a));
  return V(a.x * inv, a.y * inv, a.z * inv);
}
static float row4(const float* m, v3 p) { return fmaf(m[0], p.x, fmaf(m[1], p.y, fmaf(m[2], p.z, m[3]))); }
static v3 mat3v(const float* mv, v3 n) {
  return V(fmaf(mv[0], n.x, fmaf(mv[1], n.y, mv[2] * n.z)), fmaf(mv[4], n.x, fmaf(mv[5], n.y, mv[6] * n.z)),
           fmaf(mv[8], n.x, fmaf(mv[9], n.y, mv[10] * n.z)));
}
static void mat4mul(const float* a, const float* b, float* out) {
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0.0;
      for (int k = 0; k < 4; k++) s += (double)a[i * 4 + k] * (double)b[k * 4 + j];
      out[i * 4 + j] = (float)s;
    }
}

typedef struct {
  float clip[4];
  v3 pos, col;
  float tu, tv;
} vtx;

typedef struct {
  const orc_render_view* view;
  const orc_render_opts* o;
  v3 normal, col;
  int disk;
  float spec_noimg;
  uint64_t* fb;
} prim;

static float pow_spec(float x, float p) {
  const int n = (int)p;
  if ((float)n == p && n >= 1 && n <= 64) {
    float r = 1.f, b = x;
    int have = 0;
    for (int e = n; e > 0; e >>= 1) {
      if (e & 1) { r = have ? r * b : b; have = 1; }
      b = b * b;
    }
    return r;
  }
  return powf(x, p);
}

static uint32_t unorm8(float c) {
  if (!(c > 0.f)) return 0u;
  if (c > 1.f) c = 1.f;
  return (uint32_t)(int)(c * 255.f + 0.5f);
}

/* fragment shader main(): surfelmap.glsl:112-134 == trianglemap.glsl:119-142 (spec_power differs) */
static uint32_t fragment(const prim* P, v3 position, v3 color) {
  const orc_render_opts* o = P->o;
  const float kA = o->use_img ? o->phong[0] : 0.1f;
  const float kD = o->use_img ? o->phong[1] : 0.2f;
  const float kS = o->use_img ? o->phong[2] : 1.3f;
  const v3 frag_color = o->use_img ? color : V(o->base_color[0], o->base_color[1], o->base_color[2]);
  const float spec_power = o->use_img ? o->spec : P->spec_noimg;
  const v3 L = normalize(sub(V(o->lightpos[0], o->lightpos[1], o->lightpos[2]), position));
  const v3 N = P->normal; /* normalised once per primitive: it is constant over it */
  const float lambertian = fmaxf(dot(N, L), 0.f);
  float specular = 0.f;
  if (lambertian > 0.f) {
    const v3 R = V(2.f * (L.x * N.x) * N.x - L.x, 2.f * (L.y * N.y) * N.y - L.y, 2.f * (L.z * N.z) * N.z - L.z);
    const v3 Vv = normalize(V(-position.x, -position.y, -position.z));
    specular = pow_spec(fmaxf(dot(R, Vv), 0.f), spec_power);
  }
  const float diff = kA + lambertian * kD, sadd = kS * specular;
  v3 out = V(frag_color.x * diff + sadd, frag_color.y * diff + sadd, frag_color.z * diff + sadd);
  if (o->show_normal) out = V(-N.x * 0.5f + 0.5f, N.y * 0.5f + 0.5f, N.z * 0.5f + 0.5f);
  return unorm8(out.x) | (unorm8(out.y) << 8) | (unorm8(out.z) << 16) | 0xFF000000u;
}

typedef struct { long long x, y; float zndc, rw; int ok; } snapped;

static snapped snap(const orc_render_view* view, const vtx* v) {
  snapped s;
  s.rw = 1.f / v->clip[3];
  const float nx = v->clip[0] * s.rw, ny = v->clip[1] * s.rw;
  s.zndc = v->clip[2] * s.rw;
  const float wx = fmaf(nx, 0.5f, 0.5f) * (float)view->width, wy = fmaf(ny, 0.5f, 0.5f) * (float)view->height;
  s.ok = v->clip[3] > 0.f && fabsf(wx) <= 1048576.f && fabsf(wy) <= 1048576.f && fabsf(s.zndc) <= 3.0e38f;
  s.x = s.ok ? (long long)lrintf(wx * 256.f) : 0;
  s.y = s.ok ? (long long)lrintf(wy * 256.f) : 0;
  return s;
}

static long long edgef(long long ax, long long ay, long long bx, long long by, long long px, long long py) {
  return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}
static int owns(long long dx, long long dy) { return dy < 0 || (dy == 0 && dx < 0); }
static long long fdiv256(long long v) { return (v >= 0) ? v / 256 : -((-v + 255) / 256); }
static long long mn(long long a, long long b) { return a < b ? a : b; }
static long long mx(long long a, long long b) { return a > b ? a : b; }

static void raster(const prim* P, vtx a, vtx b, vtx c) {
  const orc_render_view* view = P->view;
  snapped sa = snap(view, &a), sb = snap(view, &b), sc = snap(view, &c);
  if (!(sa.ok && sb.ok && sc.ok)) return;
  long long area2 = edgef(sa.x, sa.y, sb.x, sb.y, sc.x, sc.y);
  if (area2 == 0) return;
  if (area2 < 0) {
    if (P->o->cull_face) return;
    vtx tv = b; b = c; c = tv;
    snapped ts = sb; sb = sc; sc = ts;
    area2 = -area2;
  }
  const long long minx = mn(sa.x, mn(sb.x, sc.x)), maxx = mx(sa.x, mx(sb.x, sc.x));
  const long long miny = mn(sa.y, mn(sb.y, sc.y)), maxy = mx(sa.y, mx(sb.y, sc.y));
  const long long x0 = mx(0, fdiv256(minx - 128 + 255)), x1 = mn(view->width - 1, fdiv256(maxx - 128));
  const long long y0 = mx(0, fdiv256(miny - 128 + 255)), y1 = mn(view->height - 1, fdiv256(maxy - 128));
  const int own_a = owns(sc.x - sb.x, sc.y - sb.y), own_b = owns(sa.x - sc.x, sa.y - sc.y),
            own_c = owns(sb.x - sa.x, sb.y - sa.y);
  const float inv_area = 1.f / (float)area2;
  for (long long py = y0; py <= y1; py++)
    for (long long px = x0; px <= x1; px++) {
      const long long cx = 256 * px + 128, cy = 256 * py + 128;
      const long long wa = edgef(sb.x, sb.y, sc.x, sc.y, cx, cy), wb = edgef(sc.x, sc.y, sa.x, sa.y, cx, cy),
                      wc = edgef(sa.x, sa.y, sb.x, sb.y, cx, cy);
      if (wa < 0 || wb < 0 || wc < 0) continue;
      if ((wa == 0 && !own_a) || (wb == 0 && !own_b) || (wc == 0 && !own_c)) continue;
      const float la = (float)wa * inv_area, lb = (float)wb * inv_area, lc = (float)wc * inv_area;
      const float zn = fmaf(lc, sc.zndc, fmaf(lb, sb.zndc, la * sa.zndc));
      const float zw = fmaf(zn, 0.5f, 0.5f);
      if (!(zw >= 0.f && zw < 1.f)) continue;
      const float qa = la * sa.rw, qb = lb * sb.rw, qc = lc * sc.rw;
      const float iq = 1.f / ((qa + qb) + qc);
#define INTERP(A, B, C) (fmaf(qc, C, fmaf(qb, B, qa * (A))) * iq)
      if (P->disk) { /* surfelmap.glsl:107-111 */
        const float tu = INTERP(a.tu, b.tu, c.tu), tv = INTERP(a.tv, b.tv, c.tv);
        if (fmaf(tu, tu, tv * tv) > 1.0f) continue; /* length(t) > 1 <=> t.t > 1 */
      }
      const v3 pos = V(INTERP(a.pos.x, b.pos.x, c.pos.x), INTERP(a.pos.y, b.pos.y, c.pos.y),
                       INTERP(a.pos.z, b.pos.z, c.pos.z));
      v3 col = P->col;
      if (!P->disk)
        col = V(INTERP(a.col.x, b.col.x, c.col.x), INTERP(a.col.y, b.col.y, c.col.y),
                INTERP(a.col.z, b.col.z, c.col.z));
#undef INTERP
      uint32_t zbits;
      memcpy(&zbits, &zw, 4);
      const uint64_t word = ((uint64_t)zbits << 32) | fragment(P, pos, col);
      uint64_t* dst = P->fb + (size_t)py * view->width + px;
      if (word < *dst) *dst = word;
    }
}

static void xform(const float* mv, const float* mvp, v3 p, vtx* v) {
  v->pos = V(row4(mv, p), row4(mv + 4, p), row4(mv + 8, p));
  for (int r = 0; r < 4; r++) v->clip[r] = row4(mvp + 4 * r, p);
}
static v3 ld(const float* p) { return V(p[0], p[1], p[2]); }

void orc_render_clear(uint64_t* fb, int width, int height, uint32_t clear_rgba) {
  for (size_t i = 0; i < (size_t)width * height; i++) fb[i] = ((uint64_t)0x3F800000u << 32) | clear_rgba;
}

/* program 0: surfelmap.glsl, 1: trianglemap.glsl.  X (K,H,W,3), C (K,H,W), img (K,H,W,3), models (K,16). */
void orc_render(int program, const float* X, const float* Cf, const float* img, const float* models, int K,
                int H, int W, const orc_render_view* view, const orc_render_opts* o, uint64_t* fb) {
  for (int kf = 0; kf < K; kf++) {
    float mv[16], mvp[16];
    mat4mul(view->camera, models + 16 * (size_t)kf, mv); /* mv = m_camera * m_model */
    mat4mul(view->proj, mv, mvp);                        /* mvp = m_proj * mv */
    const float* Xk = X + (size_t)kf * H * W * 3;
    const float* Ik = img + (size_t)kf * H * W * 3;
    const float* Ck = Cf + (size_t)kf * H * W;
    prim P;
    P.view = view; P.o = o; P.fb = fb;
    for (int id = 0; id < H * W; id++) {
      const int y = id / W, x = id - y * W;
      vtx v[4];
      if (program == 0) {
        if (x < 1 || x >= W - 1 || y < 1 || y >= H - 1) continue; /* :32-34 */
        if (Ck[id] < o->conf_threshold) continue;                 /* :36-39 */
        const v3 xyz00 = ld(Xk + 3 * (size_t)id), xyz01 = ld(Xk + 3 * (size_t)(id + W)),
                 xyz10 = ld(Xk + 3 * (size_t)(id + 1));
        const v3 right = sub(xyz10, xyz00), down = sub(xyz01, xyz00);
        const v3 N = normalize(cross(down, right));
        const v3 tangent = normalize(cross(N, V(0.f, -1.f, 0.f)));
        const v3 bitangent = cross(N, tangent);
        P.normal = normalize(mat3v(mv, N));
        P.col = ld(Ik + 3 * (size_t)id);
        P.disk = 1; P.spec_noimg = 2.f;
        for (int k = 0; k < 4; k++) { /* p1..p4, :55-58, texcoords :64,71,78,85 */
          const float st = (k & 1) ? 1.f : -1.f, sb = (k & 2) ? 1.f : -1.f;
          const v3 p = V(xyz00.x + (st * tangent.x + sb * bitangent.x) * o->radius,
                         xyz00.y + (st * tangent.y + sb * bitangent.y) * o->radius,
                         xyz00.z + (st * tangent.z + sb * bitangent.z) * o->radius);
          xform(mv, mvp, p, &v[k]);
          v[k].tu = st; v[k].tv = sb; v[k].col = P.col;
        }
      } else {
        if (x < 10 || x >= W - 10 || y < 10 || y >= H - 10) continue; /* :44-46 */
        const int ids[4] = {id, id + 1, id + W, id + W + 1};          /* TL TR BL BR, :51 */
        v3 pos[4], col[4];
        for (int k = 0; k < 4; k++) { pos[k] = ld(Xk + 3 * (size_t)ids[k]); col[k] = ld(Ik + 3 * (size_t)ids[k]); }
        const v3 n1 = normalize(cross(sub(pos[2], pos[0]), sub(pos[1], pos[0]))); /* normal(TL,BL,TR) */
        const v3 n2 = normalize(cross(sub(pos[2], pos[1]), sub(pos[3], pos[1]))); /* normal(TR,BL,BR) */
        const v3 ray1 = normalize(pos[0]), ray2 = normalize(pos[1]);
        if (fabsf(dot(n1, ray1)) < o->slant_threshold) continue;
        if (fabsf(dot(n2, ray2)) < o->slant_threshold) continue;
        if (Ck[id] < o->conf_threshold) continue; /* conf fetched at (x,y) for all four, :58 */
        P.normal = normalize(mat3v(mv, V((n1.x + n2.x) / 2.f, (n1.y + n2.y) / 2.f, (n1.z + n2.z) / 2.f)));
        P.col = V(0.f, 0.f, 0.f);
        P.disk = 0; P.spec_noimg = 3.f;
        const int order[4] = {0, 2, 1, 3}; /* TL, BL, TR, BR, :81 */
        for (int k = 0; k < 4; k++) {
          xform(mv, mvp, pos[order[k]], &v[k]);
          v[k].clip[2] = v[k].clip[2] - o->depth_bias; /* :84 */
          v[k].col = col[order[k]];
          v[k].tu = 0.f; v[k].tv = 0.f;
        }
      }
      raster(&P, v[0], v[1], v[2]);
      raster(&P, v[2], v[1], v[3]);
    }
  }
}

void orc_render_resolve(const uint64_t* fb, int width, int height, uint8_t* rgba, float* depth) {
  for (int row = 0; row < height; row++)
    for (int col = 0; col < width; col++) {
      const uint64_t v = fb[(size_t)(height - 1 - row) * width + col];
      const size_t i = (size_t)row * width + col;
      if (rgba) { const uint32_t c = (uint32_t)v; memcpy(rgba + 4 * i, &c, 4); }
      if (depth) { const uint32_t z = (uint32_t)(v >> 32); memcpy(depth + i, &z, 4); }
    }
}
