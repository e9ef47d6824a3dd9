// m3r_track.cu — two-frame Sim3 ray/distance Gauss-Newton (the tracker solve), sm_100a.
//
// Replaces the body of FrameTracker.opt_pose_ray_dist_sim3 + FrameTracker.solve (reference
// tracker.py:156-214) with geometry.point_to_ray_dist / act_Sim3 (geometry.py:17-52) and
// nonlinear_optimizer.check_convergence / huber (:5-33).  The reference runs ~25 PyTorch
// kernels per iteration, materialises an (n,4,7) Jacobian (22 MB at 512x384) and syncs the
// host twice per iteration (`.item()`, norm compare).
//
// Here the WHOLE loop is one persistent cooperative kernel: one CTA per SM slot; per iteration
// each CTA streams its share of the 29 B/point inputs (L2-resident after the first pass),
// accumulates the 28+7+1 normal-equation terms in registers, writes one partial, the grid
// synchronises once, and every CTA redundantly (and identically) reduces the partials in
// fp64, solves the 7x7 system, retracts T and evaluates the stopping test — so no second
// barrier or broadcast is needed.  No host round trip until the final status read.
//
// Residual convention: tracker r = z - h(x) and J = -dh/dT (tracker.py:190-192); with
// e = h - z and Jh = dh/dT this is H = sum w Jh^T Jh, g = -sum w Jh^T e, tau = H^-1 g.
#include <cooperative_groups.h>

#include "m3r_common.cuh"
#include "m3r_sim3.cuh"

namespace cg = cooperative_groups;

namespace m3r {

constexpr int kTrkThreads = 256;
constexpr int kTrkVals = 36;  // 28 H + 7 g + 1 cost
constexpr int kTrkChunks = kTrkThreads / kTrkVals;  // 7

constexpr int kTrkRays = 0, kTrkCalib = 1;

struct TrackArgs {
  const float* Xf; const float* Xk; const float* Qk; const uint8_t* valid;
  // calib mode: Xk is meas_k (n,3) = [u, v, log z_k], valid_meas (n) its validity
  const uint8_t* valid_meas;
  const float* K;  // calib: (3,3) f32 row-major on the device (read once per CTA, no host copy)
  float u_lo, u_hi, v_lo, v_hi, z_eps;
  int n;
  float inv_var_ray, inv_var_dist, huber_k;  // calib: 1/sigma_pixel^2, 1/sigma_depth^2
  int max_iters; float rel_error, delta_norm;
  float* T;      // (8) in/out
  int* status;   // (4)
  float* cost;   // (max_iters) or null
  float* part;   // 2 * grid * kTrkVals
};

__device__ __forceinline__ float huber_weight(float wc, float err, float k) {
  const float s2 = wc * err * err;  // (sqrt_info * r)^2
  const float hub = s2 < k * k ? 1.0f : k * rsqrtf(s2);
  return hub * wc;
}

template <unsigned MASK>
__device__ __forceinline__ void trk_row(float* h, float* g, float& cost, float w, float e,
                                        const float* J) {
  float wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1u) wJ[a] = w * J[a];
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1u)) continue;
#pragma unroll
    for (int c = 0; c <= a; c++)
      if ((MASK >> c) & 1u) h[a * (a + 1) / 2 + c] = fmaf(wJ[a], J[c], h[a * (a + 1) / 2 + c]);
    g[a] = fmaf(wJ[a], e, g[a]);
  }
  cost = fmaf(w * e, e, cost);
}

template <int MODE>
__global__ void __launch_bounds__(kTrkThreads)
track_kernel(const TrackArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float red[kTrkThreads / 32][kTrkVals];
  __shared__ double tot[kTrkVals];
  __shared__ double tot_part[kTrkChunks][kTrkVals];
  __shared__ float Tsh[8];
  __shared__ int stop_sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 8) Tsh[tid] = a.T[tid];
  __syncthreads();
  float T[8];
#pragma unroll
  for (int k = 0; k < 8; k++) T[k] = Tsh[k];
  float kfx = 0.f, kfy = 0.f, kcx = 0.f, kcy = 0.f;
  if (MODE == kTrkCalib) { kfx = __ldg(a.K); kfy = __ldg(a.K + 4); kcx = __ldg(a.K + 2); kcy = __ldg(a.K + 5); }
  float old_cost = __int_as_float(0x7f800000);  // +inf (tracker.py:185)
  int iters = 0, converged = 0, failed = 0;
  const int nthreads = gridDim.x * kTrkThreads;

  for (int step = 0; step < a.max_iters; step++) {
    // affine map of the current estimate: P = s R Xf + t  (lietorch Sim3.act)
    float R[9];
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float ev[3] = {0.f, 0.f, 0.f}, col[3];
      ev[c] = 1.f;
      act_so3<float>(T + 3, ev, col);
      R[0 + c] = T[7] * col[0]; R[3 + c] = T[7] * col[1]; R[6 + c] = T[7] * col[2];
    }
    float h[28], g[7], cost = 0.f;
#pragma unroll
    for (int k = 0; k < 28; k++) h[k] = 0.f;
#pragma unroll
    for (int k = 0; k < 7; k++) g[k] = 0.f;
    for (int p = blockIdx.x * kTrkThreads + tid; p < a.n; p += nthreads) {
      const float fx = __ldg(a.Xf + 3 * p), fy = __ldg(a.Xf + 3 * p + 1), fz = __ldg(a.Xf + 3 * p + 2);
      const float kx = __ldg(a.Xk + 3 * p), ky = __ldg(a.Xk + 3 * p + 1), kz = __ldg(a.Xk + 3 * p + 2);
      const float q = __ldg(a.Qk + p);
      const bool ok = a.valid[p] != 0;
      const float Px = fmaf(R[0], fx, fmaf(R[1], fy, fmaf(R[2], fz, T[0])));
      const float Py = fmaf(R[3], fx, fmaf(R[4], fy, fmaf(R[5], fz, T[1])));
      const float Pz = fmaf(R[6], fx, fmaf(R[7], fy, fmaf(R[8], fz, T[2])));
      if (MODE == kTrkRays) {
        const float n2k = kx * kx + ky * ky + kz * kz;
        const float nk_inv = rsqrtf(n2k);
        const float n2 = Px * Px + Py * Py + Pz * Pz;
        const float n_inv = rsqrtf(n2);
        const float nrm = n2 * n_inv;
        const float rx = Px * n_inv, ry = Py * n_inv, rz = Pz * n_inv;
        // e = h(x) - z
        const float e0 = rx - kx * nk_inv, e1 = ry - ky * nk_inv, e2 = rz - kz * nk_inv;
        const float e3 = nrm - n2k * nk_inv;
        const float wc_ray = ok ? q * a.inv_var_ray : 0.f;
        const float wc_dist = ok ? q * a.inv_var_dist : 0.f;
        const float w0 = huber_weight(wc_ray, e0, a.huber_k), w1 = huber_weight(wc_ray, e1, a.huber_k);
        const float w2 = huber_weight(wc_ray, e2, a.huber_k), w3 = huber_weight(wc_dist, e3, a.huber_k);
        const float dxx = n_inv - rx * rx * n_inv, dyy = n_inv - ry * ry * n_inv;
        const float dzz = n_inv - rz * rz * n_inv;
        const float dxy = -rx * ry * n_inv, dxz = -rx * rz * n_inv, dyz = -ry * rz * n_inv;
        { const float J[7] = {dxx, dxy, dxz, 0.f, rz, -ry, 0.f}; trk_row<0x37u>(h, g, cost, w0, e0, J); }
        { const float J[7] = {dxy, dyy, dyz, -rz, 0.f, rx, 0.f}; trk_row<0x2Fu>(h, g, cost, w1, e1, J); }
        { const float J[7] = {dxz, dyz, dzz, ry, -rx, 0.f, 0.f}; trk_row<0x1Fu>(h, g, cost, w2, e2, J); }
        { const float J[7] = {rx, ry, rz, 0.f, 0.f, 0.f, nrm}; trk_row<0x47u>(h, g, cost, w3, e3, J); }
      } else {
        // project_calib (geometry.py:63-104): pz = [u, v, log z], r = meas - pz  =>  e = pz - meas
        const bool valid_z = Pz > a.z_eps;
        const float u = (kfx * Px + kcx * Pz) / Pz, v = (kfy * Py + kcy * Pz) / Pz;
        const bool in_img = (u > a.u_lo) && (u < a.u_hi) && (v > a.v_lo) && (v < a.v_hi);
        const bool ok2 = ok && valid_z && in_img && (a.valid_meas[p] != 0);
        const float zinv = 1.0f / Pz;
        const float logz = valid_z ? logf(Pz) : 0.f;
        const float xz = Px * zinv, yz = Py * zinv;
        const float e0 = u - kx, e1 = v - ky, e2 = logz - kz;
        const float wc_pix = ok2 ? q * a.inv_var_ray : 0.f;
        const float wc_dep = ok2 ? q * a.inv_var_dist : 0.f;
        if (ok2) {  // weight-0 rows are skipped (0 * inf would poison the sums when z == 0)
          const float w0 = huber_weight(wc_pix, e0, a.huber_k), w1 = huber_weight(wc_pix, e1, a.huber_k);
          const float w2 = huber_weight(wc_dep, e2, a.huber_k);
          { const float J[7] = {kfx * zinv, 0.f, -kfx * xz * zinv, -kfx * xz * yz,
                                kfx * (1.f + xz * xz), -kfx * yz, 0.f};
            trk_row<0x3Du>(h, g, cost, w0, e0, J); }
          { const float J[7] = {0.f, kfy * zinv, -kfy * yz * zinv, -kfy * (1.f + yz * yz),
                                kfy * xz * yz, kfy * xz, 0.f};
            trk_row<0x3Eu>(h, g, cost, w1, e1, J); }
          { const float J[7] = {0.f, 0.f, zinv, yz, -xz, 0.f, 1.f}; trk_row<0x5Cu>(h, g, cost, w2, e2, J); }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 28; k++) { const float v = warp_sum(h[k]); if (lane == 0) red[warp][k] = v; }
#pragma unroll
    for (int k = 0; k < 7; k++) { const float v = warp_sum(g[k]); if (lane == 0) red[warp][28 + k] = v; }
    { const float v = warp_sum(cost); if (lane == 0) red[warp][35] = v; }
    __syncthreads();
    float* part = a.part + (size_t)(step & 1) * gridDim.x * kTrkVals;
    if (tid < kTrkVals) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < kTrkThreads / 32; w++) s += red[w][tid];
      part[(size_t)blockIdx.x * kTrkVals + tid] = s;
    }
    grid.sync();  // (orders the partial stores before every CTA's loads below: no separate fence needed)
    // every CTA: identical fixed-order fp64 reduction + solve
    // 7 x 36 threads: thread (chunk, value) sums every 7th partial (independent L2 loads in
    // flight instead of one 296-long dependent walk per value), then the 7 chunk sums are added
    // in chunk order — the same order in every CTA
    {
      const int val = tid % kTrkVals, chunk = tid / kTrkVals;
      if (chunk < kTrkChunks) {
        double s = 0.0;
#pragma unroll 4
        for (int c = chunk; c < (int)gridDim.x; c += kTrkChunks) s += (double)__ldcg(part + (size_t)c * kTrkVals + val);
        tot_part[chunk][val] = s;
      }
    }
    __syncthreads();
    if (tid < kTrkVals) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < kTrkChunks; k++) s += tot_part[k][tid];
      tot[tid] = s;
    }
    __syncthreads();
    if (tid == 0) {
      // 7x7 Cholesky solve on one thread: it is on the critical path of every iteration, so no fp64 square
      // root or division (~35 of them, ~1.4 us): one rsqrt per pivot, everything else multiplies by it
      double L[7][7], rhs[7], rinv[7];
      for (int r = 0; r < 7; r++) {
        for (int c = 0; c <= r; c++) L[r][c] = tot[r * (r + 1) / 2 + c];
        rhs[r] = -tot[28 + r];
      }
      const float new_cost = (float)(0.5 * tot[35]);
      bool okc = true;
      for (int j = 0; j < 7 && okc; j++) {  // torch.linalg.cholesky(H) (tracker.py:168)
        double d = L[j][j];
        for (int k = 0; k < j; k++) d = fma(-L[j][k], L[j][k], d);
        if (!(d > 0.0)) { okc = false; break; }
        const double rr = rsqrt(d);
        rinv[j] = rr;
        L[j][j] = d * rr;
        for (int i = j + 1; i < 7; i++) {
          double s = L[i][j];
          for (int k = 0; k < j; k++) s = fma(-L[i][k], L[j][k], s);
          L[i][j] = s * rr;
        }
      }
      int stop = 0;
      if (!okc) {
        stop = 2;
      } else {
        for (int i = 0; i < 7; i++) {  // cholesky_solve
          double s = rhs[i];
          for (int k = 0; k < i; k++) s = fma(-L[i][k], rhs[k], s);
          rhs[i] = s * rinv[i];
        }
        for (int i = 6; i >= 0; i--) {
          double s = rhs[i];
          for (int k = i + 1; k < 7; k++) s = fma(-L[k][i], rhs[k], s);
          rhs[i] = s * rinv[i];
        }
        float tau[7], pose[8];
        float nn = 0.f;
        for (int i = 0; i < 7; i++) { tau[i] = (float)rhs[i]; nn += tau[i] * tau[i]; }
        for (int i = 0; i < 8; i++) pose[i] = Tsh[i];
        retr_sim3_f(tau, pose);  // T <- Exp(tau) * T  (lietorch retr, tracker.py:195)
        for (int i = 0; i < 8; i++) Tsh[i] = pose[i];
        // check_convergence (nonlinear_optimizer.py:5-26); first pass: |(inf-c)/inf| = nan
        const float rel_dec = fabsf((old_cost - new_cost) / old_cost);
        if (rel_dec < a.rel_error || sqrtf(nn) < a.delta_norm) stop = 1;
      }
      if (blockIdx.x == 0 && a.cost) a.cost[step] = new_cost;
      red[0][0] = new_cost;
      stop_sh = stop;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; k++) T[k] = Tsh[k];
    old_cost = red[0][0];
    const int stop = stop_sh;
    __syncthreads();
    if (stop == 2) { failed = 1; break; }
    iters = step + 1;
    if (stop == 1) { converged = 1; break; }
  }
  if (blockIdx.x == 0 && tid == 0) {
#pragma unroll
    for (int k = 0; k < 8; k++) a.T[k] = T[k];
    a.status[0] = iters; a.status[1] = converged; a.status[2] = failed; a.status[3] = 0;
  }
}

static int track_grid_size(int mode) {
  static std::atomic<int> cache[kMaxDevices];  // SM count per DEVICE (one process may drive several)
  int dev = 0, sms = kNumSMs, per_sm = 1;
  cudaGetDevice(&dev);
  std::atomic<int>* slot = (dev >= 0 && dev < kMaxDevices) ? &cache[dev] : nullptr;
  if (slot && slot->load()) return slot->load();
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (mode == kTrkRays)
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, track_kernel<kTrkRays>, kTrkThreads, 0);
  else
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, track_kernel<kTrkCalib>, kTrkThreads, 0);
  if (per_sm < 1) per_sm = 1;
  // one CTA per SM: the pass itself is a few microseconds (5.7 MB, L2-resident), what an
  // iteration pays for is the grid barrier and the reduction of the partials, both of which grow
  // with the CTA count (measured 296 / 148 / 111 / 74 CTAs: 59.7 / 52.7 / 53.0 / 56.1 us per solve)
  per_sm = 1;
  if (slot) slot->store(sms * per_sm);
  return sms * per_sm;
}

static int track_launch(int mode, TrackArgs& a, size_t workspace_bytes, cudaStream_t st) {
  int grid = track_grid_size(mode);
  const int need = (a.n + kTrkThreads - 1) / kTrkThreads;
  if (grid > need) grid = need;
  if (workspace_bytes < sizeof(float) * 2 * (size_t)grid * kTrkVals) return M3R_EWORKSPACE;
  void* args[] = {&a};
  const void* fn = mode == kTrkRays ? (const void*)track_kernel<kTrkRays> : (const void*)track_kernel<kTrkCalib>;
  cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kTrkThreads), args, 0, st);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return check_cuda(e);
}

}  // namespace m3r

using namespace m3r;

extern "C" size_t m3r_track_workspace_bytes(int n) {
  (void)n;
  return sizeof(float) * 2 * (size_t)(4 * kNumSMs) * kTrkVals;  // generous: <= 4 CTAs/SM
}

extern "C" int m3r_track_ray_dist(const float* Xf, const float* Xk, const float* Qk,
                                  const uint8_t* valid, int n, float sigma_ray, float sigma_dist,
                                  float huber_k, int max_iters, float rel_error, float delta_norm,
                                  float* T_CkCf, int* status, float* cost, void* workspace,
                                  size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_track_ray_dist");
  if (n <= 0 || max_iters < 0 || !Xf || !Xk || !Qk || !valid || !T_CkCf || !status || !workspace)
    return M3R_EINVAL;
  if (sigma_ray == 0.f || sigma_dist == 0.f) return M3R_EINVAL;
  TrackArgs a;
  a.Xf = Xf; a.Xk = Xk; a.Qk = Qk; a.valid = valid; a.n = n;
  a.inv_var_ray = (float)(1.0 / ((double)sigma_ray * sigma_ray));
  a.inv_var_dist = (float)(1.0 / ((double)sigma_dist * sigma_dist));
  a.huber_k = huber_k; a.max_iters = max_iters; a.rel_error = rel_error; a.delta_norm = delta_norm;
  a.T = T_CkCf; a.status = status; a.cost = cost; a.part = static_cast<float*>(workspace);
  a.valid_meas = nullptr; a.K = nullptr;
  return track_launch(kTrkRays, a, workspace_bytes, as_stream(stream));
}

extern "C" int m3r_track_calib(const float* Xf, const float* meas_k, const float* Qk,
                               const uint8_t* valid, const uint8_t* valid_meas_k, int n,
                               const float* K_dev, int height, int width, int pixel_border, float z_eps, float sigma_pixel, float sigma_depth,
                               float huber_k, int max_iters, float rel_error, float delta_norm,
                               float* T_CkCf, int* status, float* cost, void* workspace,
                               size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_track_calib");
  if (n <= 0 || max_iters < 0 || !Xf || !meas_k || !Qk || !valid || !valid_meas_k || !K_dev || !T_CkCf ||
      !status || !workspace)
    return M3R_EINVAL;
  if (sigma_pixel == 0.f || sigma_depth == 0.f) return M3R_EINVAL;
  TrackArgs a;
  a.Xf = Xf; a.Xk = meas_k; a.Qk = Qk; a.valid = valid; a.valid_meas = valid_meas_k; a.n = n;
  a.K = K_dev;
  a.u_lo = (float)pixel_border; a.u_hi = (float)(width - 1 - pixel_border);
  a.v_lo = (float)pixel_border; a.v_hi = (float)(height - 1 - pixel_border);
  a.z_eps = z_eps;
  a.inv_var_ray = (float)(1.0 / ((double)sigma_pixel * sigma_pixel));
  a.inv_var_dist = (float)(1.0 / ((double)sigma_depth * sigma_depth));
  a.huber_k = huber_k; a.max_iters = max_iters; a.rel_error = rel_error; a.delta_norm = delta_norm;
  a.T = T_CkCf; a.status = status; a.cost = cost; a.part = static_cast<float*>(workspace);
  return track_launch(kTrkCalib, a, workspace_bytes, as_stream(stream));
}

// ---------------------------------------------------------------------------------
// m3r_track_ray_dist_noalloc — the same solve for callers that cannot allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-CTA partial sums, the only scratch of
// the tracker kernel, live in a file-scope device array instead of a caller workspace:
// 2 x kTrkStaticCtas x 36 floats = 85 248 bytes, held by every GPU that loads the library until
// the process exits.  Nothing in it has to start at zero (every partial an iteration reads was
// written by that iteration, before the grid barrier), so the array is never cleared.  All calls
// on a device share it: two calls that overlap on the GPU (different streams) race.
// ---------------------------------------------------------------------------------
namespace m3r {
constexpr int kTrkStaticCtas = 2 * kNumSMs;  // one CTA per SM is launched; room for a 296-SM part
static __device__ __align__(256) float g_trk_static_part[2 * kTrkStaticCtas * kTrkVals];

// Device address of the array on the CURRENT device.  In host code `g_trk_static_part` is only a
// handle for the runtime, not a device address: it must go through cudaGetSymbolAddress, once
// per device, with that device current.
static int trk_static_scratch(float** out) {
  static std::atomic<float*> cache[kMaxDevices];
  int dev = 0;
  M3R_TRY(check_cuda(cudaGetDevice(&dev)));
  std::atomic<float*>* slot = (dev >= 0 && dev < kMaxDevices) ? &cache[dev] : nullptr;
  float* p = slot ? slot->load() : nullptr;
  if (!p) {
    M3R_TRY(check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_trk_static_part)));
    if (!p) return M3R_ELAUNCH;
    if (slot) slot->store(p);
  }
  *out = p;
  return M3R_OK;
}
}  // namespace m3r

extern "C" int m3r_track_ray_dist_noalloc(const float* Xf, const float* Xk, const float* Qk,
                                          const uint8_t* valid, int n, float sigma_ray,
                                          float sigma_dist, float huber_k, int max_iters,
                                          float rel_error, float delta_norm, float* T_CkCf,
                                          int* status, float* cost, m3r_stream_t stream) {
  M3R_NVTX("m3r_track_ray_dist_noalloc");
  if (n <= 0 || max_iters < 0 || !Xf || !Xk || !Qk || !valid || !T_CkCf || !status) return M3R_EINVAL;
  if (sigma_ray == 0.f || sigma_dist == 0.f) return M3R_EINVAL;
  if (n > M3R_TRACK_NOALLOC_MAX_N) return M3R_ETOOLARGE;
  if (track_grid_size(kTrkRays) > kTrkStaticCtas) return M3R_ETOOLARGE;  // a GPU with > 296 SMs
  float* part = nullptr;
  M3R_TRY(trk_static_scratch(&part));
  TrackArgs a;
  a.Xf = Xf; a.Xk = Xk; a.Qk = Qk; a.valid = valid; a.n = n;
  a.inv_var_ray = (float)(1.0 / ((double)sigma_ray * sigma_ray));
  a.inv_var_dist = (float)(1.0 / ((double)sigma_dist * sigma_dist));
  a.huber_k = huber_k; a.max_iters = max_iters; a.rel_error = rel_error; a.delta_norm = delta_norm;
  a.T = T_CkCf; a.status = status; a.cost = cost; a.part = part;
  a.valid_meas = nullptr; a.K = nullptr;
  return track_launch(kTrkRays, a, sizeof(g_trk_static_part), as_stream(stream));
}
