// m3r_sim3.cuh — Sim3 group algebra on the device, templated on the scalar type.
//
// Pose layout (lietorch Sim3 / reference Twc rows): [t(3), q = (x,y,z,w), s].
// Tangent layout: [tau(3), phi(3), sigma].
// The formulas restate the reference's device helpers (gn_kernels.cu:172-413):
// quat_comp, actSO3, actSim3, relSim3, apply_Sim3_adj_inv, expSO3, expSim3, retrSim3.
// The fp32 instantiations of exp/retr keep the reference's operation order and its
// double-typed literals so pose updates agree to rounding; the fp64 instantiations are
// used where this library re-associates the reference's sums (per-edge adjoint applied
// once, after the reduction over points).
#pragma once
#include <cuda_runtime.h>

namespace m3r {

template <typename T>
__device__ __forceinline__ void quat_comp(const T* qi, const T* qj, T* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

// Y = R(q) X.  Y may alias X.
template <typename T>
__device__ __forceinline__ void act_so3(const T* q, const T* X, T* Y) {
  const T uv0 = T(2) * (q[1] * X[2] - q[2] * X[1]);
  const T uv1 = T(2) * (q[2] * X[0] - q[0] * X[2]);
  const T uv2 = T(2) * (q[0] * X[1] - q[1] * X[0]);
  const T y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const T y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const T y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}

// T_ij = T_i^-1 * T_j  (relSim3, gn_kernels.cu:252-272)
template <typename T>
__device__ __forceinline__ void rel_sim3(const T* ti, const T* qi, T si, const T* tj,
                                         const T* qj, T sj, T* tij, T* qij, T& sij) {
  const T si_inv = T(1.0 / (double)si);
  sij = si_inv * sj;
  const T qi_inv[4] = {-qi[0], -qi[1], -qi[2], qi[3]};
  quat_comp(qi_inv, qj, qij);
  tij[0] = tj[0] - ti[0]; tij[1] = tj[1] - ti[1]; tij[2] = tj[2] - ti[2];
  act_so3(qi_inv, tij, tij);
  tij[0] *= si_inv; tij[1] *= si_inv; tij[2] *= si_inv;
}

// Row vector X (tau, omega, s) times Adj(T)^-1  (apply_Sim3_adj_inv, gn_kernels.cu:277-297)
template <typename T>
__device__ __forceinline__ void adj_inv_row(const T* t, const T* q, T s, const T* X, T* Y) {
  const T s_inv = T(1.0 / (double)s);
  T Ra[3];
  act_so3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0]; Y[1] = s_inv * Ra[1]; Y[2] = s_inv * Ra[2];
  act_so3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]));
}

#define M3R_SIM3_EPS 1e-6  // gn_kernels.cu:34 (a double literal)

__device__ __forceinline__ void exp_so3_f(const float* phi, float* q) {  // gn_kernels.cu:299-321
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if (theta_sq < M3R_SIM3_EPS) {
    const float theta_p4 = theta_sq * theta_sq;
    imag = 0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * theta_p4;
    real = 1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * theta_p4;
  } else {
    const float theta = sqrtf(theta_sq);
    imag = sinf(0.5 * theta) / theta;
    real = cosf(0.5 * theta);
  }
  q[0] = imag * phi[0]; q[1] = imag * phi[1]; q[2] = imag * phi[2]; q[3] = real;
}

__device__ __forceinline__ void cross_inplace_f(const float* a, float* b) {
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}

__device__ __forceinline__ void exp_sim3_f(const float* xi, float* t, float* q, float& s) {
  float tau[3] = {xi[0], xi[1], xi[2]};  // gn_kernels.cu:323-390
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  exp_so3_f(phi, q);
  s = scale;
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float A, B, C;
  const float one = 1.0f, half = 0.5f;
  if (fabsf(sigma) < M3R_SIM3_EPS) {
    C = one;
    if (fabsf(theta) < M3R_SIM3_EPS) {
      A = half;
      B = 1.0 / 6.0;
    } else {
      A = (one - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if (fabsf(theta) < M3R_SIM3_EPS) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      B = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - b) * theta) / (theta * c);
      B = (C - ((b - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0]; t[1] = C * tau[1]; t[2] = C * tau[2];
  cross_inplace_f(phi, tau);
  t[0] += A * tau[0]; t[1] += A * tau[1]; t[2] += A * tau[2];
  cross_inplace_f(phi, tau);
  t[0] += B * tau[0]; t[1] += B * tau[1]; t[2] += B * tau[2];
}

// pose (8 floats) <- Exp(xi) * pose   (retrSim3, gn_kernels.cu:392-413)
__device__ __forceinline__ void retr_sim3_f(const float* xi, float* pose) {
  float dt[3], dq[4], ds;
  exp_sim3_f(xi, dt, dq, ds);
  const float t[3] = {pose[0], pose[1], pose[2]};
  const float q[4] = {pose[3], pose[4], pose[5], pose[6]};
  float q1[4], t1[3];
  quat_comp(dq, q, q1);
  act_so3(dq, t, t1);
  t1[0] *= ds; t1[1] *= ds; t1[2] *= ds;
  t1[0] += dt[0]; t1[1] += dt[1]; t1[2] += dt[2];
  pose[0] = t1[0]; pose[1] = t1[1]; pose[2] = t1[2];
  pose[3] = q1[0]; pose[4] = q1[1]; pose[5] = q1[2]; pose[6] = q1[3];
  pose[7] = ds * pose[7];
}

}  // namespace m3r
