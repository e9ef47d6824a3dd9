/*
 * m3r_oracle.c — CPU restatement of the MASt3R-SLAM `mast3r_slam_backends` hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (mast3r-slam_b200/) may import,
 * link or call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker / baseline.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/mast3r_slam/backend/src/).  The reference kernels are CUDA-only;
 * this is a plain-C restatement of their arithmetic:
 *   - per-element math keeps the reference's types, including the float->double
 *     promotions caused by un-suffixed literals;
 *   - compiled with -ffp-contract=off, so no FMA contraction happens here, whereas
 *     nvcc contracts a*b+c in the reference: results agree to rounding, not bitwise,
 *     except refine_matches (integer output, fp16 rounding sequence reproduced exactly);
 *   - reductions over points (Hessian blocks) are accumulated in double; the
 *     reference sums in fp32 with a 256-thread strided + tree order that a CPU cannot
 *     reproduce — parity is by tolerance (1e-4 relative per block).
 * Parity pin: checked against outputs of the real reference extension run on a B200
 * (oracle/_ref, fixtures under tests/golden/, generator tests/golden/gen_ref_golden.py).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/build_oracle.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

/* ------------------------------------------------------------------------- */
/* IEEE binary16 <-> binary32, round-to-nearest-even (c10::Half semantics,    */
/* torch/headeronly/util/Half.h: fp16_ieee_from_fp32_value / to_fp32_value).  */
/* ------------------------------------------------------------------------- */
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float orc_h2f(uint16_t h) {
  const uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu;
  uint32_t man = h & 0x3FFu;
  if (exp == 0) {
    if (man == 0) return u2f(sign);
    /* subnormal: value = man * 2^-24 */
    float v = (float)man * 5.9604644775390625e-8f;
    return (sign ? -v : v);
  }
  if (exp == 31) return u2f(sign | 0x7F800000u | (man << 13));
  return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t orc_f2h(float f) {
  const uint32_t x = f2u(f);
  const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
  const uint32_t ax = x & 0x7FFFFFFFu;
  if (ax > 0x7F800000u) return (uint16_t)(sign | 0x7E00u);       /* NaN */
  if (ax >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);      /* >= 65520 -> inf */
  if (ax < 0x33000001u) return sign;                              /* <= 2^-25 -> 0 (ties to even) */
  int e = (int)(ax >> 23) - 127;
  uint32_t m = (ax & 0x7FFFFFu) | 0x800000u; /* 24-bit significand */
  int shift;                                 /* bits to drop */
  uint32_t base;
  if (e < -14) { shift = 13 + (-14 - e); base = 0; }             /* subnormal half */
  else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7FFFFFu; }
  uint32_t q = m >> shift;
  uint32_t rem = m & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  return (uint16_t)(sign | (base + q)); /* carry into exponent is correct by construction */
}

void orc_f2h_array(const float* src, uint16_t* dst, int64_t count) {
  for (int64_t i = 0; i < count; i++) dst[i] = orc_f2h(src[i]);
}
void orc_h2f_array(const uint16_t* src, float* dst, int64_t count) {
  for (int64_t i = 0; i < count; i++) dst[i] = orc_h2f(src[i]);
}

/* ------------------------------------------------------------------------- */
/* refine_matches  (matching_kernels.cu:25-81)                                */
/* ------------------------------------------------------------------------- */
static inline int inside_image(int u, int v, int W, int H) { /* :17-19 */
  return v >= 0 && v < H && u >= 0 && u < W;
}

/* <c10::Half> instantiation.  score and max_score are fp16 values; each product and
 * each partial sum is computed in fp32 and rounded to fp16 (Half.h operator*,+=).
 * max_score starts at c10::Half() == 0 (libcu++ numeric_limits primary template,
 * SURVEY H2), compare is on the float values with strict '>'. */
void orc_refine_matches_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1,
                            int b, int H, int W, int n, int F, int radius,
                            int dilation_max, int64_t* p1_new) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int k = 0; k < n; k++) {
      const uint16_t* q = D21 + ((int64_t)bi * n + k) * F;
      long u0 = p1[((int64_t)bi * n + k) * 2 + 0];
      long v0 = p1[((int64_t)bi * n + k) * 2 + 1];
      uint16_t max_score = 0; /* :47 */
      long u_new = u0, v_new = v0;
      for (int d = dilation_max; d > 0; d--) { /* :51 */
        const int rd = radius * d;
        const int diam = 2 * rd + 1;
        for (int i = 0; i < diam; i += d) {
          for (int j = 0; j < diam; j += d) {
            const long u = u0 - rd + i;
            const long v = v0 - rd + j;
            if (inside_image((int)u, (int)v, W, H)) { /* long -> int truncation, :59 */
              const uint16_t* c = D11 + (((int64_t)bi * H + (int)v) * W + (int)u) * F;
              uint16_t score = 0;
              for (int t = 0; t < F; t++) { /* :61-63 */
                const uint16_t prod = orc_f2h(orc_h2f(q[t]) * orc_h2f(c[t]));
                score = orc_f2h(orc_h2f(score) + orc_h2f(prod));
              }
              if (orc_h2f(score) > orc_h2f(max_score)) { /* :65 */
                max_score = score; u_new = u; v_new = v;
              }
            }
          }
        }
        u0 = u_new; v0 = v_new; /* :75-76 */
      }
      p1_new[((int64_t)bi * n + k) * 2 + 0] = u_new;
      p1_new[((int64_t)bi * n + k) * 2 + 1] = v_new;
    }
  }
}

/* <float> instantiation: fp32 accumulate, max_score starts at FLT_MIN (:47). */
void orc_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                            int b, int H, int W, int n, int F, int radius,
                            int dilation_max, int64_t* p1_new) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int k = 0; k < n; k++) {
      const float* q = D21 + ((int64_t)bi * n + k) * F;
      long u0 = p1[((int64_t)bi * n + k) * 2 + 0];
      long v0 = p1[((int64_t)bi * n + k) * 2 + 1];
      float max_score = FLT_MIN;
      long u_new = u0, v_new = v0;
      for (int d = dilation_max; d > 0; d--) {
        const int rd = radius * d;
        const int diam = 2 * rd + 1;
        for (int i = 0; i < diam; i += d) {
          for (int j = 0; j < diam; j += d) {
            const long u = u0 - rd + i;
            const long v = v0 - rd + j;
            if (inside_image((int)u, (int)v, W, H)) {
              const float* c = D11 + (((int64_t)bi * H + (int)v) * W + (int)u) * F;
              float score = 0.0f;
              /* nvcc contracts score += a*b into FFMA; fmaf reproduces that. */
              for (int t = 0; t < F; t++) score = fmaf(q[t], c[t], score);
              if (score > max_score) { max_score = score; u_new = u; v_new = v; }
            }
          }
        }
        u0 = u_new; v0 = v_new;
      }
      p1_new[((int64_t)bi * n + k) * 2 + 0] = u_new;
      p1_new[((int64_t)bi * n + k) * 2 + 1] = v_new;
    }
  }
}

/* ------------------------------------------------------------------------- */
/* iter_proj  (matching_kernels.cu:119-275)                                   */
/* ------------------------------------------------------------------------- */
static inline float clampf(float x, float lo, float hi) { /* :21-23 fmin(fmax()) */
  return fminf(fmaxf(x, lo), hi);
}

/* Bilinear sample of channels [c0,c1) of the (H,W,9) ray image at (u,v).
 * Weights follow :157-170 exactly: du,dv in float; (1.0-du) promotes to double, the
 * product is formed in double and rounded to float once.  "w11 = du*dv" multiplies
 * texel (v11+1,u11+1) (the reference's naming is transposed but self-consistent). */
static inline void bilinear9(const float* img, int W, float u, float v, int c0, int c1,
                             float* out) {
  const int u11 = (int)floorf(u);
  const int v11 = (int)floorf(v);
  const float du = u - (float)u11;
  const float dv = v - (float)v11;
  const float w11 = du * dv;
  const float w12 = (float)((1.0 - (double)du) * (double)dv);
  const float w21 = (float)((double)du * (1.0 - (double)dv));
  const float w22 = (float)((1.0 - (double)du) * (1.0 - (double)dv));
  const float* r11 = img + ((int64_t)(v11 + 1) * W + (u11 + 1)) * 9;
  const float* r12 = img + ((int64_t)(v11 + 1) * W + u11) * 9;
  const float* r21 = img + ((int64_t)v11 * W + (u11 + 1)) * 9;
  const float* r22 = img + ((int64_t)v11 * W + u11) * 9;
  for (int j = c0; j < c1; j++)
    out[j - c0] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
}

static inline void normalize3_ref(float* r) { /* :186-191 */
  const float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  const float r_norm_inv = (float)(1.0 / (double)r_norm);
  r[0] *= r_norm_inv; r[1] *= r_norm_inv; r[2] *= r_norm_inv;
}

void orc_iter_proj(const float* rays_img, const float* pts_3d_norm, const float* p_init,
                   int b, int H, int W, int n, int max_iter, float lambda_init,
                   float cost_thresh, float* p_new, uint8_t* converged) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int bi = 0; bi < b; bi++) {
    for (int k = 0; k < n; k++) {
      const float* img = rays_img + (int64_t)bi * H * W * 9;
      const float* tgt = pts_3d_norm + ((int64_t)bi * n + k) * 3;
      float u = p_init[((int64_t)bi * n + k) * 2 + 0];
      float v = p_init[((int64_t)bi * n + k) * 2 + 1];
      u = clampf(u, 1.0f, (float)(W - 2)); /* :143-144 */
      v = clampf(v, 1.0f, (float)(H - 2));
      float lambda = lambda_init;
      uint8_t conv = 0; /* torch::zeros, :300 */
      for (int it = 0; it < max_iter; it++) {
        float s[9], r[3], gx[3], gy[3], err[3];
        bilinear9(img, W, u, v, 0, 9, s);
        r[0] = s[0]; r[1] = s[1]; r[2] = s[2];
        gx[0] = s[3]; gx[1] = s[4]; gx[2] = s[5];
        gy[0] = s[6]; gy[1] = s[7]; gy[2] = s[8];
        normalize3_ref(r);
        for (int j = 0; j < 3; j++) err[j] = r[j] - tgt[j];
        const float cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
        float A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2]; /* :202-204 */
        const float A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
        float A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
        const float b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
        const float b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
        A00 += lambda; A11 += lambda;
        const float det = A00 * A11 - A01 * A01;                   /* fp32 inner, :213 */
        const float det_inv = (float)(1.0 / (double)det);
        const float delta_u = det_inv * (A11 * b0 - A01 * b1);
        const float delta_v = det_inv * (-A01 * b0 + A00 * b1);
        float u_new = clampf(u + delta_u, 1.0f, (float)(W - 2));
        float v_new = clampf(v + delta_v, 1.0f, (float)(H - 2));
        float rn[3];
        bilinear9(img, W, u_new, v_new, 0, 3, rn);
        normalize3_ref(rn);
        for (int j = 0; j < 3; j++) err[j] = rn[j] - tgt[j];
        const float new_cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
        if (new_cost < cost) { /* :259-268 */
          u = u_new; v = v_new;
          lambda = (float)((double)lambda * 0.1);
          conv = new_cost < cost_thresh;
        } else {
          lambda = (float)((double)lambda * 10.0);
          conv = cost < cost_thresh;
        }
      }
      p_new[((int64_t)bi * n + k) * 2 + 0] = u;
      p_new[((int64_t)bi * n + k) * 2 + 1] = v;
      converged[(int64_t)bi * n + k] = conv;
    }
  }
}

/* ------------------------------------------------------------------------- */
/* Sim3 device helpers  (gn_kernels.cu:172-413)                               */
/* ------------------------------------------------------------------------- */
static inline float huber_ref(float r) { /* :172-175, double literals */
  const float r_abs = fabsf(r);
  return ((double)r_abs < 1.345) ? 1.0f : (float)(1.345 / (double)r_abs);
}
static void quat_comp(const float* qi, const float* qj, float* out) { /* :178-184 */
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}
static void actSO3(const float* q, const float* X, float* Y) { /* :195-205, Y may alias X */
  float uv[3];
  uv[0] = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  uv[1] = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  uv[2] = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  const float y1 = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  const float y2 = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}
static void actSim3(const float* t, const float* q, const float* s, const float* X,
                    float* Y) { /* :207-219 */
  actSO3(q, X, Y);
  Y[0] *= s[0]; Y[1] *= s[0]; Y[2] *= s[0];
  Y[0] += t[0]; Y[1] += t[1]; Y[2] += t[2];
}
static void relSim3(const float* ti, const float* qi, const float* si, const float* tj,
                    const float* qj, const float* sj, float* tij, float* qij,
                    float* sij) { /* :252-272 */
  const float si_inv = (float)(1.0 / (double)si[0]);
  sij[0] = si_inv * sj[0];
  const float qi_inv[4] = {-qi[0], -qi[1], -qi[2], qi[3]};
  quat_comp(qi_inv, qj, qij);
  tij[0] = tj[0] - ti[0]; tij[1] = tj[1] - ti[1]; tij[2] = tj[2] - ti[2];
  actSO3(qi_inv, tij, tij);
  tij[0] *= si_inv; tij[1] *= si_inv; tij[2] *= si_inv;
}
/* Row vector X (tau, omega, s) times Adj(T)^-1, :277-297 */
static void apply_Sim3_adj_inv(const float* t, const float* q, const float* s,
                               const float* X, float* Y) {
  const float s_inv = (float)(1.0 / (double)s[0]);
  float Ra[3];
  actSO3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0]; Y[1] = s_inv * Ra[1]; Y[2] = s_inv * Ra[2];
  actSO3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]));
}
#define ORC_EPS 1e-6 /* :34, a double literal */
static void expSO3(const float* phi, float* q) { /* :299-321 */
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if ((double)theta_sq < ORC_EPS) {
    const float theta_p4 = theta_sq * theta_sq;
    imag = (float)(0.5 - (1.0 / 48.0) * (double)theta_sq + (1.0 / 3840.0) * (double)theta_p4);
    real = (float)(1.0 - (1.0 / 8.0) * (double)theta_sq + (1.0 / 384.0) * (double)theta_p4);
  } else {
    const float theta = sqrtf(theta_sq);
    imag = sinf((float)(0.5 * (double)theta)) / theta;
    real = cosf((float)(0.5 * (double)theta));
  }
  q[0] = imag * phi[0]; q[1] = imag * phi[1]; q[2] = imag * phi[2]; q[3] = real;
}
static void cross_inplace(const float* a, float* b) { /* :229-240 */
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}
static void expSim3(const float* xi, float* t, float* q, float* s) { /* :323-390 */
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  expSO3(phi, q);
  s[0] = scale;
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float A, B, C;
  const float one = 1.0f, half = 0.5f;
  if ((double)fabsf(sigma) < ORC_EPS) {
    C = one;
    if ((double)fabsf(theta) < ORC_EPS) { A = half; B = (float)(1.0 / 6.0); }
    else {
      A = (one - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if ((double)fabsf(theta) < ORC_EPS) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      B = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float bb = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - bb) * theta) / (theta * c);
      B = (C - ((bb - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0]; t[1] = C * tau[1]; t[2] = C * tau[2];
  cross_inplace(phi, tau);
  t[0] += A * tau[0]; t[1] += A * tau[1]; t[2] += A * tau[2];
  cross_inplace(phi, tau);
  t[0] += B * tau[0]; t[1] += B * tau[1]; t[2] += B * tau[2];
}
static void retrSim3(const float* xi, const float* t, const float* q, const float* s,
                     float* t1, float* q1, float* s1) { /* :392-413 */
  float dt[3] = {0, 0, 0}, dq[4] = {0, 0, 0, 1}, ds[1] = {0};
  expSim3(xi, dt, dq, ds);
  quat_comp(dq, q, q1);
  actSO3(dq, t, t1);
  t1[0] *= ds[0]; t1[1] *= ds[0]; t1[2] *= ds[0];
  t1[0] += dt[0]; t1[1] += dt[1]; t1[2] += dt[2];
  s1[0] = ds[0] * s[0];
}

/* pose_retr_kernel, :415-453 */
void orc_pose_retr(float* poses, const float* dx, int num_poses, int num_fix) {
  for (int k = num_fix; k < num_poses; k++) {
    float* p = poses + (int64_t)k * 8;
    float t1[3], q1[4], s1[1];
    retrSim3(dx + (int64_t)(k - num_fix) * 7, p, p + 3, p + 7, t1, q1, s1);
    p[0] = t1[0]; p[1] = t1[1]; p[2] = t1[2];
    p[3] = q1[0]; p[4] = q1[1]; p[5] = q1[2]; p[6] = q1[3];
    p[7] = s1[0];
  }
}
/* Exposed for unit tests of the group algebra. */
void orc_relSim3(const float* Ti, const float* Tj, float* Tij) {
  relSim3(Ti, Ti + 3, Ti + 7, Tj, Tj + 3, Tj + 7, Tij, Tij + 3, Tij + 7);
}
void orc_actSim3(const float* T, const float* X, float* Y) { actSim3(T, T + 3, T + 7, X, Y); }
void orc_adj_inv_row(const float* T, const float* X, float* Y) {
  apply_Sim3_adj_inv(T, T + 3, T + 7, X, Y);
}

/* ------------------------------------------------------------------------- */
/* Per-edge Hessian/gradient blocks                                           */
/*   rays:   ray_align_kernel   gn_kernels.cu:813-1138                        */
/*   calib:  calib_proj_kernel  gn_kernels.cu:1231-1543                       */
/*   points: point_align_kernel gn_kernels.cu:455-723                         */
/* ------------------------------------------------------------------------- */
typedef struct orc_gn_params {
  int mode; /* 0 rays, 1 calib, 2 points */
  float sigma_a, sigma_b, C_thresh, Q_thresh;
  float fx, fy, cx, cy;
  int height, width, pixel_border;
  float z_eps;
} orc_gn_params;

/* Residual rows of one (edge, point).  Returns the number of rows (4 or 3), fills
 * err[], w[] and the LOCAL Jacobian rows Jl[r][7] (what the reference writes into Ji
 * before apply_Sim3_adj_inv).  valid_base = valid_match & q>Q_thresh & ci>C & cj>C. */
static int residual_rows(const orc_gn_params* prm, const float* Xi, const float* P,
                         int64_t ind_Xi, float q, int valid_base, float* err, float* w,
                         float Jl[4][7]) {
  if (prm->mode == 0) { /* rays :924-1073 */
    const float sigma_ray_inv = (float)(1.0 / (double)prm->sigma_a);
    const float sigma_dist_inv = (float)(1.0 / (double)prm->sigma_b);
    const float norm2_i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
    const float norm1_i = sqrtf(norm2_i);
    const float norm1_i_inv = (float)(1.0 / (double)norm1_i);
    float ri[3], rj[3];
    for (int i = 0; i < 3; i++) ri[i] = norm1_i_inv * Xi[i];
    const float norm2_j = P[0] * P[0] + P[1] * P[1] + P[2] * P[2];
    const float norm1_j = sqrtf(norm2_j);
    const float norm1_j_inv = (float)(1.0 / (double)norm1_j);
    for (int i = 0; i < 3; i++) rj[i] = norm1_j_inv * P[i];
    err[0] = rj[0] - ri[0]; err[1] = rj[1] - ri[1]; err[2] = rj[2] - ri[2];
    err[3] = norm1_j - norm1_i;
    const float sqrt_w_ray = valid_base ? sigma_ray_inv * sqrtf(q) : 0.0f;
    const float sqrt_w_dist = valid_base ? sigma_dist_inv * sqrtf(q) : 0.0f;
    w[0] = huber_ref(sqrt_w_ray * err[0]);
    w[1] = huber_ref(sqrt_w_ray * err[1]);
    w[2] = huber_ref(sqrt_w_ray * err[2]);
    w[3] = huber_ref(sqrt_w_dist * err[3]);
    const float w_const_ray = sqrt_w_ray * sqrt_w_ray;
    const float w_const_dist = sqrt_w_dist * sqrt_w_dist;
    w[0] *= w_const_ray; w[1] *= w_const_ray; w[2] *= w_const_ray; w[3] *= w_const_dist;
    const float norm3_j_inv = norm1_j_inv / norm2_j;
    const float drx_dPx = norm1_j_inv - P[0] * P[0] * norm3_j_inv;
    const float dry_dPy = norm1_j_inv - P[1] * P[1] * norm3_j_inv;
    const float drz_dPz = norm1_j_inv - P[2] * P[2] * norm3_j_inv;
    const float drx_dPy = -P[0] * P[1] * norm3_j_inv;
    const float drx_dPz = -P[0] * P[2] * norm3_j_inv;
    const float dry_dPz = -P[1] * P[2] * norm3_j_inv;
    const float r0[7] = {drx_dPx, drx_dPy, drx_dPz, 0.0f, rj[2], -rj[1], 0.0f};
    const float r1[7] = {drx_dPy, dry_dPy, dry_dPz, -rj[2], 0.0f, rj[0], 0.0f};
    const float r2[7] = {drx_dPz, dry_dPz, drz_dPz, rj[1], -rj[0], 0.0f, 0.0f};
    const float r3[7] = {rj[0], rj[1], rj[2], 0.0f, 0.0f, 0.0f, norm1_j};
    memcpy(Jl[0], r0, sizeof r0); memcpy(Jl[1], r1, sizeof r1);
    memcpy(Jl[2], r2, sizeof r2); memcpy(Jl[3], r3, sizeof r3);
    return 4;
  }
  if (prm->mode == 1) { /* calib :1360-1479 */
    const float fx = prm->fx, fy = prm->fy, cx = prm->cx, cy = prm->cy;
    const int width = prm->width, height = prm->height, border = prm->pixel_border;
    const float sigma_pixel_inv = (float)(1.0 / (double)prm->sigma_a);
    const float sigma_depth_inv = (float)(1.0 / (double)prm->sigma_b);
    const int u_target = (int)(ind_Xi % width);
    const int v_target = (int)(ind_Xi / width);
    const int valid_z = (P[2] > prm->z_eps) && (Xi[2] > prm->z_eps);
    const float zj_inv = valid_z ? (float)(1.0 / (double)P[2]) : 0.0f;
    const float zj_log = valid_z ? logf(P[2]) : 0.0f;
    const float zi_log = valid_z ? logf(Xi[2]) : 0.0f;
    const float x_div_z = P[0] * zj_inv;
    const float y_div_z = P[1] * zj_inv;
    const float u = fx * x_div_z + cx;
    const float v = fy * y_div_z + cy;
    const int valid_u = (u > (float)border) && (u < (float)(width - 1 - border));
    const int valid_v = (v > (float)border) && (v < (float)(height - 1 - border));
    err[0] = u - (float)u_target;
    err[1] = v - (float)v_target;
    err[2] = zj_log - zi_log;
    const int valid = valid_base & valid_u & valid_v & valid_z;
    const float sqrt_w_pixel = valid ? sigma_pixel_inv * sqrtf(q) : 0.0f;
    const float sqrt_w_depth = valid ? sigma_depth_inv * sqrtf(q) : 0.0f;
    w[0] = huber_ref(sqrt_w_pixel * err[0]);
    w[1] = huber_ref(sqrt_w_pixel * err[1]);
    w[2] = huber_ref(sqrt_w_depth * err[2]);
    const float w_const_pixel = sqrt_w_pixel * sqrt_w_pixel;
    const float w_const_depth = sqrt_w_depth * sqrt_w_depth;
    w[0] *= w_const_pixel; w[1] *= w_const_pixel; w[2] *= w_const_depth;
    const float r0[7] = {fx * zj_inv, 0.0f, -fx * x_div_z * zj_inv, -fx * x_div_z * y_div_z,
                         fx * (1 + x_div_z * x_div_z), -fx * y_div_z, 0.0f};
    const float r1[7] = {0.0f, fy * zj_inv, -fy * y_div_z * zj_inv,
                         -fy * (1 + y_div_z * y_div_z), fy * x_div_z * y_div_z,
                         fy * x_div_z, 0.0f};
    const float r2[7] = {0.0f, 0.0f, zj_inv, y_div_z, -x_div_z, 0.0f, 1.0f};
    memcpy(Jl[0], r0, sizeof r0); memcpy(Jl[1], r1, sizeof r1); memcpy(Jl[2], r2, sizeof r2);
    return 3;
  }
  /* points :567-658 */
  const float sigma_point_inv = (float)(1.0 / (double)prm->sigma_a);
  err[0] = P[0] - Xi[0]; err[1] = P[1] - Xi[1]; err[2] = P[2] - Xi[2];
  const float sqrt_w_point = valid_base ? sigma_point_inv * sqrtf(q) : 0.0f;
  w[0] = huber_ref(sqrt_w_point * err[0]);
  w[1] = huber_ref(sqrt_w_point * err[1]);
  w[2] = huber_ref(sqrt_w_point * err[2]);
  const float w_const_point = sqrt_w_point * sqrt_w_point;
  w[0] *= w_const_point; w[1] *= w_const_point; w[2] *= w_const_point;
  const float r0[7] = {1.0f, 0.0f, 0.0f, 0.0f, P[2], -P[1], P[0]};
  const float r1[7] = {0.0f, 1.0f, 0.0f, -P[2], 0.0f, P[0], P[1]};
  const float r2[7] = {0.0f, 0.0f, 1.0f, P[1], -P[0], 0.0f, P[2]};
  memcpy(Jl[0], r0, sizeof r0); memcpy(Jl[1], r1, sizeof r1); memcpy(Jl[2], r2, sizeof r2);
  return 3;
}

/* One accumulation pass: Hs (4,M,7,7) f32 and gs (2,M,7) f32 in the reference layout
 * (:1096-1137).  ii_edge/jj_edge are pose ROWS (already searchsorted, :1162-1164).
 * Optionally also returns the double-precision values (Hs64/gs64, may be NULL). */
void orc_gn_blocks(const orc_gn_params* prm, const float* Twc, const float* Xs,
                   const float* Cs, const int64_t* ii_edge, const int64_t* jj_edge,
                   const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q,
                   int n, int M, float* Hs, float* gs, double* Hs64, double* gs64) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int e = 0; e < M; e++) {
    const int ix = (int)ii_edge[e], jx = (int)jj_edge[e];
    const float* Ti = Twc + (int64_t)ix * 8;
    const float* Tj = Twc + (int64_t)jx * 8;
    float tij[3], qij[4], sij[1];
    relSim3(Ti, Ti + 3, Ti + 7, Tj, Tj + 3, Tj + 7, tij, qij, sij);
    double hij[105], vi[7], vj[7];
    memset(hij, 0, sizeof hij); memset(vi, 0, sizeof vi); memset(vj, 0, sizeof vj);
    for (int k = 0; k < n; k++) {
      const int vm = valid_match[(int64_t)e * n + k] != 0;
      const int64_t ind_Xi = vm ? idx_ii2jj[(int64_t)e * n + k] : 0; /* :913-914 */
      const float* Xi = Xs + ((int64_t)ix * n + ind_Xi) * 3;
      const float* Xj = Xs + ((int64_t)jx * n + k) * 3;
      float P[3];
      actSim3(tij, qij, sij, Xj, P);
      const float q = Q[(int64_t)e * n + k];
      const float ci = Cs[(int64_t)ix * n + ind_Xi];
      const float cj = Cs[(int64_t)jx * n + k];
      const int valid = vm & (q > prm->Q_thresh) & (ci > prm->C_thresh) & (cj > prm->C_thresh);
      float err[4], w[4], Jl[4][7];
      const int rows = residual_rows(prm, Xi, P, ind_Xi, q, valid, err, w, Jl);
      for (int r = 0; r < rows; r++) {
        float Jx[14];
        apply_Sim3_adj_inv(Ti, Ti + 3, Ti + 7, Jl[r], &Jx[7]); /* Jj */
        for (int m = 0; m < 7; m++) Jx[m] = -Jx[7 + m];        /* Ji = -Jj */
        int l = 0;
        for (int a = 0; a < 14; a++)
          for (int c = 0; c <= a; c++) { hij[l] += (double)(w[r] * Jx[a] * Jx[c]); l++; }
        for (int m = 0; m < 7; m++) {
          vi[m] += (double)(w[r] * err[r] * Jx[m]);
          vj[m] += (double)(w[r] * err[r] * Jx[7 + m]);
        }
      }
    }
    for (int m = 0; m < 7; m++) {
      gs[((int64_t)0 * M + e) * 7 + m] = (float)vi[m];
      gs[((int64_t)1 * M + e) * 7 + m] = (float)vj[m];
      if (gs64) { gs64[((int64_t)0 * M + e) * 7 + m] = vi[m]; gs64[((int64_t)1 * M + e) * 7 + m] = vj[m]; }
    }
    int l = 0;
    for (int a = 0; a < 14; a++) {
      for (int c = 0; c <= a; c++) {
        const double hv = hij[l++];
#define HS(blk, r, cc) (((((int64_t)(blk)) * M + e) * 7 + (r)) * 7 + (cc))
        if (a < 7 && c < 7) {
          Hs[HS(0, a, c)] = (float)hv; Hs[HS(0, c, a)] = (float)hv;
          if (Hs64) { Hs64[HS(0, a, c)] = hv; Hs64[HS(0, c, a)] = hv; }
        } else if (a >= 7 && c < 7) {
          Hs[HS(1, c, a - 7)] = (float)hv; Hs[HS(2, a - 7, c)] = (float)hv;
          if (Hs64) { Hs64[HS(1, c, a - 7)] = hv; Hs64[HS(2, a - 7, c)] = hv; }
        } else {
          Hs[HS(3, a - 7, c - 7)] = (float)hv; Hs[HS(3, c - 7, a - 7)] = (float)hv;
          if (Hs64) { Hs64[HS(3, a - 7, c - 7)] = hv; Hs64[HS(3, c - 7, a - 7)] = hv; }
        }
#undef HS
      }
    }
  }
}

/* ------------------------------------------------------------------------- */
/* SparseBlock assembly + solve  (gn_kernels.cu:57-159)                       */
/* Dense fp64 restatement of Eigen::SimplicialLLT on the assembled system:    */
/* setFromTriplets sums duplicates; blocks with a negative index are dropped; */
/* factorisation failure => zeros.  Returns 1 on success, 0 on failure.       */
/* ------------------------------------------------------------------------- */
int orc_sparse_block_solve(const float* Hs, const float* gs, const int64_t* ii_opt,
                           const int64_t* jj_opt, int M, int num_free, double* A_out,
                           double* b_out, float* dx /* (num_free,7), = -x */) {
  const int D = 7 * num_free;
  double* A = (double*)calloc((size_t)D * D, sizeof(double));
  double* bvec = (double*)calloc((size_t)D, sizeof(double));
  /* update_lhs(Hs.reshape(-1,7,7), cat(ii,ii,jj,jj), cat(ii,jj,ii,jj))  :1201-1203 */
  for (int blk = 0; blk < 4; blk++) {
    for (int e = 0; e < M; e++) {
      const int64_t i = (blk < 2) ? ii_opt[e] : jj_opt[e];
      const int64_t j = (blk == 0 || blk == 2) ? ii_opt[e] : jj_opt[e];
      if (i >= 0 && j >= 0)
        for (int k = 0; k < 7; k++)
          for (int l = 0; l < 7; l++)
            A[(size_t)(7 * i + k) * D + (7 * j + l)] +=
                (double)Hs[((((int64_t)blk) * M + e) * 7 + k) * 7 + l];
    }
  }
  /* update_rhs(gs.reshape(-1,7), cat(ii,jj))  :1205-1206 */
  for (int blk = 0; blk < 2; blk++)
    for (int e = 0; e < M; e++) {
      const int64_t i = blk == 0 ? ii_opt[e] : jj_opt[e];
      if (i >= 0)
        for (int k = 0; k < 7; k++) bvec[7 * i + k] += (double)gs[(((int64_t)blk) * M + e) * 7 + k];
    }
  if (A_out) memcpy(A_out, A, (size_t)D * D * sizeof(double));
  if (b_out) memcpy(b_out, bvec, (size_t)D * sizeof(double));
  /* LL^T, lower, in place */
  int ok = 1;
  for (int j = 0; j < D && ok; j++) {
    double d = A[(size_t)j * D + j];
    for (int k = 0; k < j; k++) d -= A[(size_t)j * D + k] * A[(size_t)j * D + k];
    if (!(d > 0.0)) { ok = 0; break; }
    const double ljj = sqrt(d);
    A[(size_t)j * D + j] = ljj;
    for (int i = j + 1; i < D; i++) {
      double s = A[(size_t)i * D + j];
      for (int k = 0; k < j; k++) s -= A[(size_t)i * D + k] * A[(size_t)j * D + k];
      A[(size_t)i * D + j] = s / ljj;
    }
  }
  if (ok) {
    for (int i = 0; i < D; i++) { /* L y = b */
      double s = bvec[i];
      for (int k = 0; k < i; k++) s -= A[(size_t)i * D + k] * bvec[k];
      bvec[i] = s / A[(size_t)i * D + i];
    }
    for (int i = D - 1; i >= 0; i--) { /* L^T x = y */
      double s = bvec[i];
      for (int k = i + 1; k < D; k++) s -= A[(size_t)k * D + i] * bvec[k];
      bvec[i] = s / A[(size_t)i * D + i];
    }
    for (int i = 0; i < D; i++) dx[i] = (float)(-bvec[i]); /* dx = -A.solve(), :1209 */
  } else {
    for (int i = 0; i < D; i++) dx[i] = -0.0f; /* -zeros */
  }
  free(A); free(bvec);
  return ok;
}

static int cmp_i64(const void* a, const void* b) {
  const int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
  return (x > y) - (x < y);
}

/* gauss_newton_{rays,calib,points}_cuda host loop (:1140-1228 and twins).
 * Twc is updated in place.  dx_hist (max_iter, N-1, 7) optional: every iteration's dx.
 * Returns the number of iterations executed. */
int orc_gauss_newton(const orc_gn_params* prm, float* Twc, const float* Xs, const float* Cs,
                     const int64_t* ii, const int64_t* jj, const int64_t* idx_ii2jj,
                     const uint8_t* valid_match, const float* Q, int num_poses, int n,
                     int M, int max_iter, float delta_thresh, float* dx, float* dx_hist) {
  const int num_fix = 1; /* :1157 */
  /* get_unique_kf_idx + create_inds :161-170 */
  int64_t* uniq = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)(M > 0 ? M : 1));
  for (int e = 0; e < M; e++) { uniq[e] = ii[e]; uniq[M + e] = jj[e]; }
  qsort(uniq, 2 * (size_t)M, sizeof(int64_t), cmp_i64);
  int nu = 0;
  for (int t = 0; t < 2 * M; t++) if (t == 0 || uniq[t] != uniq[t - 1]) uniq[nu++] = uniq[t];
  int64_t* ie = (int64_t*)malloc(sizeof(int64_t) * 4 * (size_t)(M > 0 ? M : 1));
  int64_t *je = ie + M, *io = ie + 2 * M, *jo = ie + 3 * M;
  for (int e = 0; e < M; e++) {
    int lo = 0, hi = nu; /* searchsorted, left */
    while (lo < hi) { int mid = (lo + hi) / 2; if (uniq[mid] < ii[e]) lo = mid + 1; else hi = mid; }
    ie[e] = lo; io[e] = lo - num_fix;
    lo = 0; hi = nu;
    while (lo < hi) { int mid = (lo + hi) / 2; if (uniq[mid] < jj[e]) lo = mid + 1; else hi = mid; }
    je[e] = lo; jo[e] = lo - num_fix;
  }
  const int num_free = num_poses - num_fix;
  float* Hs = (float*)calloc((size_t)4 * M * 49 + 1, sizeof(float));
  float* gs = (float*)calloc((size_t)2 * M * 7 + 1, sizeof(float));
  for (int i = 0; i < num_free * 7; i++) dx[i] = 0.0f;
  int itr = 0;
  for (; itr < max_iter; itr++) {
    orc_gn_blocks(prm, Twc, Xs, Cs, ie, je, idx_ii2jj, valid_match, Q, n, M, Hs, gs, NULL, NULL);
    orc_sparse_block_solve(Hs, gs, io, jo, M, num_free, NULL, NULL, dx);
    orc_pose_retr(Twc, dx, num_poses, num_fix);
    if (dx_hist) memcpy(dx_hist + (size_t)itr * num_free * 7, dx, sizeof(float) * num_free * 7);
    double nrm = 0.0;
    for (int i = 0; i < num_free * 7; i++) nrm += (double)dx[i] * (double)dx[i];
    if ((float)sqrt(nrm) < delta_thresh) { itr++; break; } /* :1219-1222 */
  }
  free(Hs); free(gs); free(uniq); free(ie);
  return itr;
}
