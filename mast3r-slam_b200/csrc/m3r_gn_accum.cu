// m3r_gn_accum.cu — Sim3 factor-graph Gauss-Newton: index setup, packing and the fused
// residual + Jacobian + normal-equation accumulation kernels (sm_100a).
//
// Replaces ray_align_kernel / calib_proj_kernel / point_align_kernel and the index helpers
// of the reference (gn_kernels.cu:161-170, 455-723, 813-1138, 1231-1543).
//
// What is different from the reference, and why:
//   * The reference pushes every Jacobian row through apply_Sim3_adj_inv(T_i) per point and
//     accumulates a 14x14 upper triangle (105 values, in local memory) plus 2x7 gradients.
//     Since Ji = -Jj and Jj = J_local * Adj(T_i)^-1 with Adj constant per edge, this kernel
//     accumulates only the LOCAL 7x7 (28 unique values) + 7 gradient values in registers,
//     skipping the structural zeros of each residual row at compile time, and the adjoint is
//     applied once per edge afterwards, in fp64, by the last of the edge's CTAs to finish
//     (edge_epilogue).  This moves the pass from ALU/local-memory bound to HBM bound.
//   * Each directed edge is split over S CTAs (the reference runs one 256-thread CTA per edge)
//     and reduced in two fixed-order stages, so the result is deterministic and independent of
//     how edges are distributed over GPUs.
//   * (Xs, Cs) are packed once per solve as float4, so the Xi/Ci gather is one aligned 16-B
//     load and the Xj/Cj stream is a coalesced 128-bit load.
//   * The per-edge streams are read either in the reference's layout (idx i64, valid u8, Q f32:
//     13 B/point) or as 8-byte packed records (m3r_packed_match: idx i32 + sign-folded Q) that
//     FactorGraph stores directly — 40 instead of 45 B per (edge, point), SURVEY 8(f)-3.
#include <type_traits>

#include "m3r_gn.cuh"
#include "m3r_sim3.cuh"

namespace m3r {

// ------------------------------------------------------------------------------------
// Index setup: unique sorted keyframe ids -> pose rows, CSR pose -> incident edge slots.
// ------------------------------------------------------------------------------------
constexpr int kIdxThreads = 1024;
constexpr int kIdxSmemKeys = 4096;

__device__ void bitonic_sort_block(int64_t* buf, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < P; t += blockDim.x) {
        const int x = t ^ j;
        if (x > t) {
          const bool asc = (t & k) == 0;
          const int64_t a = buf[t], b = buf[x];
          if ((a > b) == asc) { buf[t] = b; buf[x] = a; }
        }
      }
      __syncthreads();
    }
  }
}

__device__ __forceinline__ int lower_bound_i64(const int64_t* a, int len, int64_t v) {
  int lo = 0, hi = len;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (a[mid] < v) lo = mid + 1; else hi = mid; }
  return lo;
}

__global__ void __launch_bounds__(kIdxThreads)
gn_index_kernel(const int64_t* __restrict__ ii_in, const int64_t* __restrict__ jj_in, int two_way, int M,
                int N, int P, int64_t* keys, int* rank, int* ie, int* je, int* row_ptr, int* adj,
                int* flags, int nflags, int* tickets, GnHeader* hdr) {
  __shared__ int64_t skeys[kIdxSmemKeys];
  __shared__ int scan[kIdxThreads];
  const int tid = threadIdx.x;
  if (tid == 0) { hdr->done = 0; hdr->chol_fail = 0; hdr->iters = 0; hdr->peer_timeout = 0; }
  for (int t = tid; t < nflags; t += kIdxThreads) flags[t] = 0;
  for (int t = tid; t < M + 2; t += kIdxThreads) tickets[t] = 0;
  // two_way: ii_in / jj_in hold the M/2 UNDIRECTED edges and the directed list is
  // [ii;jj] -> [jj;ii], i.e. FactorGraph.prep_two_way_edges (global_opt.py:104-110) without the cat
  const int half = M >> 1;
  auto ii = [&](int e) { return two_way ? (e < half ? ii_in[e] : jj_in[e - half]) : ii_in[e]; };
  auto jj = [&](int e) { return two_way ? (e < half ? jj_in[e] : ii_in[e - half]) : jj_in[e]; };
  int64_t* buf = (P <= kIdxSmemKeys) ? skeys : keys;
  const int64_t kMax = 0x7fffffffffffffffLL;
  // 1. sorted keyframe ids  (torch::_unique(cat(ii,jj), sorted), gn_kernels.cu:161-164)
  for (int t = tid; t < P; t += kIdxThreads) buf[t] = t < M ? ii(t) : (t < 2 * M ? jj(t - M) : kMax);
  __syncthreads();
  bitonic_sort_block(buf, P);
  // 2. rank[t] = (#distinct ids in buf[0..t]) - 1
  const int per = (P + kIdxThreads - 1) / kIdxThreads;
  const int t0 = tid * per, t1 = min(P, t0 + per);
  int cnt = 0;
  for (int t = t0; t < t1; t++) cnt += (t == 0 || buf[t] != buf[t - 1]) ? 1 : 0;
  scan[tid] = cnt;
  __syncthreads();
  for (int o = 1; o < kIdxThreads; o <<= 1) {
    const int v = tid >= o ? scan[tid - o] : 0;
    __syncthreads();
    scan[tid] += v;
    __syncthreads();
  }
  int run = scan[tid] - cnt;  // exclusive prefix
  for (int t = t0; t < t1; t++) {
    run += (t == 0 || buf[t] != buf[t - 1]) ? 1 : 0;
    rank[t] = run - 1;
  }
  __syncthreads();
  // 3. searchsorted (create_inds, gn_kernels.cu:166-170)
  for (int e = tid; e < M; e += kIdxThreads) {
    ie[e] = rank[lower_bound_i64(buf, 2 * M, ii(e))];
    je[e] = rank[lower_bound_i64(buf, 2 * M, jj(e))];
  }
  __syncthreads();
  // 4. CSR: sort (row, slot) pairs; slot < M: pose is the i end of edge slot, else j end of slot-M
  const int64_t span = 2 * (int64_t)M;
  for (int t = tid; t < P; t += kIdxThreads)
    buf[t] = t < 2 * M ? (int64_t)(t < M ? ie[t] : je[t - M]) * span + t : kMax;
  __syncthreads();
  bitonic_sort_block(buf, P);
  for (int t = tid; t < 2 * M; t += kIdxThreads) adj[t] = (int)(buf[t] % span);
  for (int r = tid; r <= N; r += kIdxThreads) row_ptr[r] = lower_bound_i64(buf, 2 * M, (int64_t)r * span);
}

// (Xs (N,n,3), Cs (N,n,1)) -> float4 (x,y,z,c)
__global__ void __launch_bounds__(256)
gn_pack_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs, size_t count,
               float4* __restrict__ xc) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    xc[i] = make_float4(ld_stream_f(Xs + 3 * i), ld_stream_f(Xs + 3 * i + 1),
                        ld_stream_f(Xs + 3 * i + 2), ld_stream_f(Cs + i));
  }
}

// ------------------------------------------------------------------------------------
// Accumulation
// ------------------------------------------------------------------------------------
struct GnDev {  // kernel-side copy of m3r_gn_params with the sigmas pre-inverted
  float inv_var_a, inv_var_b;  // 1/sigma^2
  float C_thresh, Q_thresh;
  float fx, fy, cx, cy;
  const float* Kdev;  // calib: non-null => fx, fy, cx, cy are read from this device 3x3 (row-major)
  float u_lo, u_hi, v_lo, v_hi;  // pixel_border, width-1-border, ... as floats
  float z_eps;
  int width;
  unsigned width_magic;  // ceil(2^32 / width): idx / width by one multiply-high and a fix-up
};

__device__ __forceinline__ float fast_rcp(float x) {
  float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
}
// rsqrtf() without its denormal pre/post-scaling (FSETP + 2 predicated FMUL around every
// MUFU.RSQ: 24 rsqrt per 4 points were 11 % of the loop).  Same bits for normal inputs; a
// denormal argument (a squared residual below 1.2e-38) flushes to 0 -> +inf -> huber weight 1,
// which is what the scaled form yields as well.
__device__ __forceinline__ float fast_rsqrt(float x) {
#ifdef M3R_GN_PRECISE  // accuracy experiment (profiles/gn_dx_accuracy.py): correctly rounded 1/sqrt
  return __fdiv_rn(1.0f, __fsqrt_rn(x));
#else
  float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#endif
}

// h: upper^T packed 7x7 (index a*(a+1)/2 + c, c <= a); skips Jacobian entries whose bit is
// clear in MASK (structural zeros of the residual row) at compile time.
template <unsigned MASK>
__device__ __forceinline__ void acc_row(float* h, float* g, float w, float e, const float* J) {
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
}

// Huber-robustified weight.  Reference: w = huber(sqrt_w * err) * sqrt_w^2 with
// sqrt_w = sigma^-1 sqrt(q) (gn_kernels.cu:963-978); |sqrt_w err| < k  <=>  wc err^2 < k^2.
__device__ __forceinline__ float robust_weight(float wc, float err) {
  const float s2 = wc * err * err;
  // huber = 1 if s2 < k^2 else k / sqrt(s2)  ==  min(1, k * rsqrt(s2))   (rsqrt(0) = +inf)
  return wc * fminf(1.0f, 1.345f * fast_rsqrt(s2));
}

// The reference reads K on the device (gn_kernels.cu:1274-1279); so does this kernel when the
// caller passes the device matrix: no host read of K, hence no hidden synchronisation.
__device__ __forceinline__ void resolve_intrinsics(GnDev& prm) {
  if (prm.Kdev) {
    prm.fx = __ldg(prm.Kdev + 0); prm.fy = __ldg(prm.Kdev + 4);
    prm.cx = __ldg(prm.Kdev + 2); prm.cy = __ldg(prm.Kdev + 5);
  }
}

template <int MODE>
__device__ __forceinline__ void point_terms(const GnDev& prm, const float* R, const float* t,
                                            const float4 xi, const float4 xj, long long ind,
                                            float qm, float* h, float* g) {
  // qm: the match confidence Q, or -3e38 when valid_match is false / the lane is past the end
  // P = s_ij R_ij Xj + t_ij  (actSim3, gn_kernels.cu:207-219)
  const float Px = fmaf(R[0], xj.x, fmaf(R[1], xj.y, fmaf(R[2], xj.z, t[0])));
  const float Py = fmaf(R[3], xj.x, fmaf(R[4], xj.y, fmaf(R[5], xj.z, t[1])));
  const float Pz = fmaf(R[6], xj.x, fmaf(R[7], xj.y, fmaf(R[8], xj.z, t[2])));
  const bool valid = (qm > prm.Q_thresh) && (fminf(xi.w, xj.w) > prm.C_thresh);
  const float q = valid ? qm : 0.f;  // weight numerator; 0 switches the point off
  if (MODE == M3R_GN_RAYS) {  // gn_kernels.cu:924-1089
    const float n2i = xi.x * xi.x + xi.y * xi.y + xi.z * xi.z;
    const float ni_inv = fast_rsqrt(n2i);
    const float n2j = Px * Px + Py * Py + Pz * Pz;
    const float nj_inv = fast_rsqrt(n2j);
    const float nj = n2j * nj_inv;
    const float rx = Px * nj_inv, ry = Py * nj_inv, rz = Pz * nj_inv;
    const float e0 = rx - xi.x * ni_inv, e1 = ry - xi.y * ni_inv, e2 = rz - xi.z * ni_inv;
    const float e3 = nj - n2i * ni_inv;
    const float wc_ray = q * prm.inv_var_a;
    const float wc_dist = q * prm.inv_var_b;
    const float w0 = robust_weight(wc_ray, e0), w1 = robust_weight(wc_ray, e1);
    const float w2 = robust_weight(wc_ray, e2), w3 = robust_weight(wc_dist, e3);
    // d r / d P = (I - r r^T) / |P|
    const float dxx = nj_inv - rx * rx * nj_inv, dyy = nj_inv - ry * ry * nj_inv;
    const float dzz = nj_inv - rz * rz * nj_inv;
    const float dxy = -rx * ry * nj_inv, dxz = -rx * rz * nj_inv, dyz = -ry * rz * nj_inv;
    { const float J[7] = {dxx, dxy, dxz, 0.f, rz, -ry, 0.f}; acc_row<0x37u>(h, g, w0, e0, J); }
    { const float J[7] = {dxy, dyy, dyz, -rz, 0.f, rx, 0.f}; acc_row<0x2Fu>(h, g, w1, e1, J); }
    { const float J[7] = {dxz, dyz, dzz, ry, -rx, 0.f, 0.f}; acc_row<0x1Fu>(h, g, w2, e2, J); }
    { const float J[7] = {rx, ry, rz, 0.f, 0.f, 0.f, nj}; acc_row<0x47u>(h, g, w3, e3, J); }
  } else if (MODE == M3R_GN_CALIB) {  // gn_kernels.cu:1360-1495
    // (u_t, v_t) = (idx % width, idx / width) for 0 <= idx < 2^31 without a 64-bit division:
    // q = hi32(idx * ceil(2^32 / width)) is the quotient or one above it
    const int i32 = (int)ind;
    int v_t = (int)__umulhi((unsigned)i32, prm.width_magic);
    int u_t = i32 - v_t * prm.width;
    if (u_t < 0) { v_t -= 1; u_t += prm.width; }
    if (u_t >= prm.width) { v_t += 1; u_t -= prm.width; }  // width == 1 (magic saturates)
    const bool valid_z = (Pz > prm.z_eps) && (xi.z > prm.z_eps);
    const float zinv = valid_z ? 1.0f / Pz : 0.f;
    // log z through MUFU.LG2 (__logf, abs. error < 5e-7 for depths of order 1) instead of the
    // ~17-instruction logf, twice per point: the log-depth residual is weighted by 1/sigma_depth^2
    // = 1e-2 against 1 for the pixel residuals, so its 5e-7 moves no block or pose at 1e-8
    const float zj_log = valid_z ? __logf(Pz) : 0.f;
    const float zi_log = valid_z ? __logf(xi.z) : 0.f;
    const float xz = Px * zinv, yz = Py * zinv;
    const float u = fmaf(prm.fx, xz, prm.cx), v = fmaf(prm.fy, yz, prm.cy);
    const bool valid_u = (u > prm.u_lo) && (u < prm.u_hi);
    const bool valid_v = (v > prm.v_lo) && (v < prm.v_hi);
    const float e0 = u - (float)u_t, e1 = v - (float)v_t, e2 = zj_log - zi_log;
    const bool ok = valid_u & valid_v & valid_z;
    const float wc_pix = ok ? q * prm.inv_var_a : 0.f;
    const float wc_dep = ok ? q * prm.inv_var_b : 0.f;
    const float w0 = robust_weight(wc_pix, e0), w1 = robust_weight(wc_pix, e1);
    const float w2 = robust_weight(wc_dep, e2);
    { const float J[7] = {prm.fx * zinv, 0.f, -prm.fx * xz * zinv, -prm.fx * xz * yz,
                          prm.fx * (1.f + xz * xz), -prm.fx * yz, 0.f};
      acc_row<0x3Du>(h, g, w0, e0, J); }
    { const float J[7] = {0.f, prm.fy * zinv, -prm.fy * yz * zinv, -prm.fy * (1.f + yz * yz),
                          prm.fy * xz * yz, prm.fy * xz, 0.f};
      acc_row<0x3Eu>(h, g, w1, e1, J); }
    { const float J[7] = {0.f, 0.f, zinv, yz, -xz, 0.f, 1.f}; acc_row<0x5Cu>(h, g, w2, e2, J); }
  } else {  // points, gn_kernels.cu:567-674
    const float e0 = Px - xi.x, e1 = Py - xi.y, e2 = Pz - xi.z;
    const float wc = q * prm.inv_var_a;
    const float w0 = robust_weight(wc, e0), w1 = robust_weight(wc, e1), w2 = robust_weight(wc, e2);
    { const float J[7] = {1.f, 0.f, 0.f, 0.f, Pz, -Py, Px}; acc_row<0x71u>(h, g, w0, e0, J); }
    { const float J[7] = {0.f, 1.f, 0.f, -Pz, 0.f, Px, Py}; acc_row<0x6Au>(h, g, w1, e1, J); }
    { const float J[7] = {0.f, 0.f, 1.f, Py, -Px, 0.f, Pz}; acc_row<0x5Cu>(h, g, w2, e2, J); }
  }
}

constexpr int kAccThreads = 256;
constexpr int kAccUnroll = 4;

// T_ij = T_i^-1 T_j in fp32 like the reference (relSim3), then as a 3x4 affine map.
__device__ __forceinline__ void edge_transform(const float* __restrict__ Twc, int ix, int jx,
                                               float* R, float* t) {
  float Ti[8], Tj[8], qij[4], sij;
#pragma unroll
  for (int k = 0; k < 8; k++) { Ti[k] = __ldg(Twc + ix * 8 + k); Tj[k] = __ldg(Twc + jx * 8 + k); }
  rel_sim3<float>(Ti, Ti + 3, Ti[7], Tj, Tj + 3, Tj[7], t, qij, sij);
#pragma unroll
  for (int c = 0; c < 3; c++) {
    float ev[3] = {0.f, 0.f, 0.f}, col[3];
    ev[c] = 1.f;
    act_so3<float>(qij, ev, col);
    R[0 + c] = sij * col[0]; R[3 + c] = sij * col[1]; R[6 + c] = sij * col[2];
  }
}

// Peer-memory exchange of the sharded solve (m3r_gn_peers): where the finished blocks go and
// whose flags are raised when this process's pass is complete.
struct GnPeerDst { double* blocks[M3R_MAX_PEERS]; int world; };
struct GnPeerFlags { int* flags[M3R_MAX_PEERS]; int world, rank, epoch; };

__device__ __forceinline__ void peer_signal_all(const GnPeerFlags& f) {
  __threadfence_system();
  for (int r = 0; r < f.world; r++)
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f.flags[r] + f.rank), "r"(f.epoch) : "memory");
}

// Everything an accumulation launch needs.  The per-edge streams address the LOCAL slice in up
// to two segments: local edge el < split is row el of segment 0, the others row el - split of
// segment 1 (FactorGraph keeps the forward and the reverse direction in two blocks; passing both
// replaces prep_two_way_edges' concatenation, global_opt.py:104-110).
struct AccArgs {
  GnDev prm;
  const float* Twc; const float4* xc; const int* ie; const int* je;
  const void* idx[2];        // int64 match index (reference layout) or m3r_packed_match records
  const uint8_t* vmatch[2];  // reference layout only
  const float* Q[2];         // reference layout only
  int split;
  int n, edge_begin, chunk, S, M;
  int pass_edges;            // edges this process finishes per iteration (peer signal), else 0
  float* part; const GnHeader* hdr; int* tickets;
  double* blocks; GnPeerDst dst; GnPeerFlags sig;
};

// Per edge, by the LAST of its S point-split CTAs to arrive (a ticket per edge): sum the S
// partials (fp64, fixed order => deterministic whichever CTA is last), apply the adjoint of T_i once:
//   Jj = J_local * A with A[k][:] = apply_Sim3_adj_inv(T_i, e_k)   (gn_kernels.cu:277-297)
//   S = A^T H_local A,  v = A^T g_local.
// With dst.world > 1 the block is stored into EVERY process's buffer through the NVLink peer
// mappings, and the CTA that finishes this process's last edge raises its flag everywhere: the
// exchange step of the sharded solve costs no launch of its own.  (Until round 2 this was a
// separate 64-thread kernel per pass plus a one-warp signal kernel: 2 of 5 launches per iteration.)
__device__ __forceinline__ void edge_epilogue(const AccArgs& a, int el, int e) {
  __shared__ double Hl[7][7], gl[7], A[7][7], HA[7][7];
  const int t = threadIdx.x;
  if (t < 35) {
    double s = 0.0;
    for (int sp = 0; sp < a.S; sp++) s += (double)__ldcg(a.part + ((size_t)el * a.S + sp) * kGnPartStride + t);
    if (t < 28) {
      int r = 0;
      while ((r + 1) * (r + 2) / 2 <= t) r++;
      const int c = t - r * (r + 1) / 2;
      Hl[r][c] = s; Hl[c][r] = s;
    } else {
      gl[t - 28] = s;
    }
  } else if (t >= 64 && t < 71) {
    const int k = t - 64;
    const int ix = a.ie[e];
    double Ti[8], X[7], Y[7];
    for (int m = 0; m < 8; m++) Ti[m] = (double)a.Twc[ix * 8 + m];
    for (int m = 0; m < 7; m++) X[m] = (m == k) ? 1.0 : 0.0;
    adj_inv_row<double>(Ti, Ti + 3, Ti[7], X, Y);
    for (int m = 0; m < 7; m++) A[k][m] = Y[m];
  }
  __syncthreads();
  if (t < 49) {
    const int r = t / 7, b = t % 7;
    double s = 0.0;
    for (int l = 0; l < 7; l++) s += Hl[r][l] * A[l][b];
    HA[r][b] = s;
  }
  __syncthreads();
  if (t < 56) {
    double s = 0.0;
    if (t < 49) {
      const int r = t / 7, b = t % 7;
      for (int k = 0; k < 7; k++) s += A[k][r] * HA[k][b];
    } else {
      const int m = t - 49;
      for (int k = 0; k < 7; k++) s += gl[k] * A[k][m];
    }
    if (a.dst.world <= 1) a.blocks[(size_t)e * kGnBlock + t] = s;
    else for (int r = 0; r < a.dst.world; r++) a.dst.blocks[r][(size_t)e * kGnBlock + t] = s;
    if (a.dst.world > 1) __threadfence_system();  // peer stores ordered before the flag
  }
  if (a.pass_edges > 0) {
    __syncthreads();
    if (t == 0) {
      const int done_edges = atomicAdd(a.tickets + a.M, 1);
      if (done_edges == a.pass_edges - 1) {
        a.tickets[a.M] = 0;       // ready for the next pass (stream order separates the passes)
        peer_signal_all(a.sig);   // fence.sys, then flags[rank] = epoch in every process
      }
    }
  }
}

// two-stage fixed-order reduction (warp shuffle tree, then 8 warps summed in order) into this
// CTA's partial, then the ticket: the last split of the edge runs the epilogue.
__device__ __forceinline__ void block_reduce_finish(const AccArgs& a, const float* h, const float* g,
                                                    int el, int e, int split) {
  __shared__ float red[kAccThreads / 32][kGnPartStride];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 28; k++) { const float v = warp_sum(h[k]); if (lane == 0) red[warp][k] = v; }
#pragma unroll
  for (int k = 0; k < 7; k++) { const float v = warp_sum(g[k]); if (lane == 0) red[warp][28 + k] = v; }
  __syncthreads();
  if (threadIdx.x < 35) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kAccThreads / 32; w++) s += red[w][threadIdx.x];
    a.part[((size_t)el * a.S + split) * kGnPartStride + threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // one acq_rel atomic is the whole hand-off: its release half publishes this split's partial (the CTA barrier
    // above orders the 35 stores before it, and a release is cumulative), its acquire half lets the last split
    // to arrive read the others' partials.  (It replaced a __threadfence by each of the 35 storing threads, the
    // atomic and a second fence: ~1 us at the tail of every CTA.)
    int arrived;
    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(arrived) : "l"(a.tickets + e) : "memory");
    const int last = arrived == a.S - 1;
    if (last) a.tickets[e] = 0;
    s_last = last;
  }
  __syncthreads();
  if (s_last) edge_epilogue(a, el, e);
}

// A stopped loop (converged or aborted) turns the pass into a no-op; in the peer-memory exchange
// a converged process still raises its flag so that epochs stay consecutive on every rank (the
// two-buffer parity rule of the exchange relies on it).
__device__ __forceinline__ bool acc_stopped(const AccArgs& a) {
  if (!gn_stopped(a.hdr)) return false;
  if (a.pass_edges > 0 && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 &&
      *reinterpret_cast<const volatile int*>(&a.hdr->peer_timeout) == 0)
    peer_signal_all(a.sig);
  return true;
}

template <bool PK>
__device__ __forceinline__ void edge_rows(const AccArgs& a, int el, const unsigned char*& idx_e,
                                          const uint8_t*& vm_e, const float*& q_e) {
  const int seg = el < a.split ? 0 : 1;
  const size_t row = (size_t)(seg ? el - a.split : el) * a.n;
  idx_e = static_cast<const unsigned char*>(a.idx[seg]) + row * 8;  // 8 B per point either way
  vm_e = PK ? nullptr : a.vmatch[seg] + row;
  q_e = PK ? nullptr : a.Q[seg] + row;
}

// A packed record (m3r_packed_match): .x = match index (int32), .y = bits of the sign-folded
// confidence (Q if valid_match else -Q).  valid <=> sign bit clear; an invalid match gathers
// pixel 0 like the reference (gn_kernels.cu:913-914) and fails `q > Q_thresh` (Q_thresh >= 0 on
// this path, see make_dev_params).
__device__ __forceinline__ void unpack_match(const int2 rec, int& ind, float& qm) {
  ind = rec.y >= 0 ? rec.x : 0;
  qm = __int_as_float(rec.y);
}

// Generic accumulation kernel (any n, any alignment): register-staged loads.
template <int MODE, bool PK>
__global__ void __launch_bounds__(kAccThreads, 2)
gn_accum_kernel(const AccArgs a) {
  if (acc_stopped(a)) return;
  GnDev prm = a.prm;
  if (MODE == M3R_GN_CALIB) resolve_intrinsics(prm);
  const int el = blockIdx.y, split = blockIdx.x, n = a.n;
  const int e = a.edge_begin + el;
  const int ix = a.ie[e], jx = a.je[e];
  float R[9], t[3];
  edge_transform(a.Twc, ix, jx, R, t);
  float h[28], g[7];
#pragma unroll
  for (int k = 0; k < 28; k++) h[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 7; k++) g[k] = 0.f;

  const float4* xi_base = a.xc + (size_t)ix * n;
  const float4* xj_base = a.xc + (size_t)jx * n;
  const unsigned char* idx_raw; const uint8_t* vm_e; const float* q_e;
  edge_rows<PK>(a, el, idx_raw, vm_e, q_e);
  const int p_begin = split * a.chunk;
  const int p_end = min(n, p_begin + a.chunk);
  for (int base = p_begin + threadIdx.x; base < p_end; base += kAccThreads * kAccUnroll) {
    int ind[kAccUnroll];
    float qv[kAccUnroll];
    float4 xj[kAccUnroll], xi[kAccUnroll];
#pragma unroll
    for (int u = 0; u < kAccUnroll; u++) {
      const int k = base + u * kAccThreads;
      const bool in = k < p_end;
      const int kk = in ? k : p_begin;
      if (PK) {
        unpack_match(reinterpret_cast<const int2*>(idx_raw)[kk], ind[u], qv[u]);
        if (!in) { ind[u] = 0; qv[u] = -3.0e38f; }
      } else {
        const bool vm = in && (vm_e[kk] != 0);
        ind[u] = vm ? (int)reinterpret_cast<const long long*>(idx_raw)[kk] : 0;  // invalid: pixel 0
        const float qq = ld_stream_f(q_e + kk);
        qv[u] = vm ? qq : -3.0e38f;   // fails q > Q_thresh for any finite threshold
      }
      xj[u] = ld_stream_f4(xj_base + kk);
    }
#pragma unroll
    for (int u = 0; u < kAccUnroll; u++) xi[u] = __ldg(xi_base + ind[u]);
#pragma unroll
    for (int u = 0; u < kAccUnroll; u++)
      point_terms<MODE>(prm, R, t, xi[u], xj[u], (long long)ind[u], qv[u], h, g);
  }
  block_reduce_finish(a, h, g, el, e, split);
}

// ------------------------------------------------------------------------------------
// Streaming form of the same pass (used whenever n % 16 == 0 and the streams are 16-B
// aligned): the per-edge streams of a 1024-point stage are fetched by the bulk-copy engine
// (cp.async.bulk global -> shared, completion on an mbarrier) several stages ahead, so HBM
// stays busy independently of the register file; only the irregular Xi/Ci gather is a
// register load, issued one stage ahead of its use.
//   reference layout: Xj/Cj 16 + idx 8 + Q 4 + valid 1 = 29 B/point streamed, 3 stages (87 KB)
//   packed layout   : Xj/Cj 16 + record 8            = 24 B/point streamed, 4 stages (96 KB)
// (+ 16 B gathered either way: 45 / 40 B per point).  Point -> thread mapping, arithmetic and
// reduction order are identical to gn_accum_kernel and between the two layouts, so all four
// produce the same bits.
// ------------------------------------------------------------------------------------
constexpr int kStagePts = kAccThreads * kAccUnroll;  // 1024
template <bool PK> struct StageCfg;
template <> struct StageCfg<false> {
  static constexpr int kStages = 3;
  static constexpr int kOffIdx = kStagePts * 16, kOffQ = kOffIdx + kStagePts * 8, kOffVm = kOffQ + kStagePts * 4;
  static constexpr int kStageBytes = kOffVm + kStagePts;  // 29 KB
};
template <> struct StageCfg<true> {
  static constexpr int kStages = 4;
  static constexpr int kOffIdx = kStagePts * 16, kOffQ = 0, kOffVm = 0;
  static constexpr int kStageBytes = kOffIdx + kStagePts * 8;  // 24 KB
};
template <bool PK> constexpr int acc_smem_bytes() { return StageCfg<PK>::kStages * StageCfg<PK>::kStageBytes; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                 "selp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int MODE, bool PK>
__global__ void __launch_bounds__(kAccThreads, 2)
gn_accum_stream_kernel(const AccArgs a) {
  using SC = StageCfg<PK>;
  constexpr int kStages = SC::kStages, kStageBytes = SC::kStageBytes;
  constexpr int kOffIdx = SC::kOffIdx, kOffQ = SC::kOffQ, kOffVm = SC::kOffVm;
  extern __shared__ __align__(128) unsigned char stage_mem[];
  __shared__ uint64_t full[kStages];
  if (acc_stopped(a)) return;
  GnDev prm = a.prm;
  if (MODE == M3R_GN_CALIB) resolve_intrinsics(prm);
  const int tid = threadIdx.x, n = a.n;
  const int el = blockIdx.y, split = blockIdx.x;
  const int e = a.edge_begin + el;
  const int ix = a.ie[e], jx = a.je[e];
  const float4* xi_base = a.xc + (size_t)ix * n;
  const float4* xj_base = a.xc + (size_t)jx * n;
  const unsigned char* idx_raw; const uint8_t* vm_e; const float* q_e;
  edge_rows<PK>(a, el, idx_raw, vm_e, q_e);
  const int p_begin = split * a.chunk;
  const int p_end = min(n, p_begin + a.chunk);
  const int T = (p_end - p_begin + kStagePts - 1) / kStagePts;

  auto issue = [&](int st) {  // one thread: arm the barrier, start the copies of stage st
    const int p0 = p_begin + st * kStagePts;
    const uint32_t cnt = (uint32_t)min(kStagePts, p_end - p0);
    unsigned char* buf = stage_mem + (st % kStages) * kStageBytes;
    uint64_t* bar = &full[st % kStages];
    mbar_expect_tx(bar, cnt * (PK ? 24u : 29u));
    bulk_g2s(buf, xj_base + p0, cnt * 16u, bar);
    bulk_g2s(buf + kOffIdx, idx_raw + (size_t)p0 * 8, cnt * 8u, bar);
    if (!PK) {
      bulk_g2s(buf + kOffQ, q_e + p0, cnt * 4u, bar);
      bulk_g2s(buf + kOffVm, vm_e + p0, cnt, bar);
    }
  };
  if (tid == 0) {
    for (int s = 0; s < kStages; s++) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int st = 0; st < kStages && st < T; st++) issue(st);
  }
  float R[9], t[3];
  edge_transform(a.Twc, ix, jx, R, t);
  float h[28], g[7];
#pragma unroll
  for (int k = 0; k < 28; k++) h[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 7; k++) g[k] = 0.f;
  __syncthreads();  // barriers initialised

  // Every stage but the last of the image is full (chunks are multiples of the stage size), and a
  // full stage needs none of the range tests: `stage_full` is CTA-uniform.
  auto stage_full = [&](int st) { return p_begin + (st + 1) * kStagePts <= p_end; };
  // (gathered pixel, confidence or -3e38) of this thread's u-th point of a stage
  auto match_of = [&](const unsigned char* buf, int sl, bool in, int& ind, float& qv) {
    if (PK) {
      unpack_match(reinterpret_cast<const int2*>(buf + kOffIdx)[sl], ind, qv);
      if (!in) { ind = 0; qv = -3.0e38f; }
    } else {
      const bool vm = in && buf[kOffVm + sl] != 0;
      ind = vm ? (int)reinterpret_cast<const long long*>(buf + kOffIdx)[sl] : 0;
      qv = vm ? reinterpret_cast<const float*>(buf + kOffQ)[sl] : -3.0e38f;
    }
  };
  float4 xi_nxt[kAccUnroll];
  auto gather_impl = [&](int st, auto FULL) {
    constexpr bool kFull = decltype(FULL)::value;
    const unsigned char* buf = stage_mem + (st % kStages) * kStageBytes;
#pragma unroll
    for (int u = 0; u < kAccUnroll; u++) {
      const int sl = u * kAccThreads + tid;
      const bool in = kFull || (p_begin + st * kStagePts + sl < p_end);
      int ind; float qv;
      match_of(buf, in ? sl : 0, in, ind, qv);
      xi_nxt[u] = __ldg(xi_base + ind);
    }
  };
  auto gather = [&](int st) {
    if (stage_full(st)) gather_impl(st, std::true_type{}); else gather_impl(st, std::false_type{});
  };
  mbar_wait(&full[0], 0);
  gather(0);
  for (int it = 0; it < T; it++) {
    float4 xi[kAccUnroll];
#pragma unroll
    for (int u = 0; u < kAccUnroll; u++) xi[u] = xi_nxt[u];
    if (it + 1 < T) {
      mbar_wait(&full[(it + 1) % kStages], ((it + 1) / kStages) & 1);
      gather(it + 1);
    }
    const unsigned char* buf = stage_mem + (it % kStages) * kStageBytes;
    auto compute = [&](auto FULL) {
      constexpr bool kFull = decltype(FULL)::value;
#pragma unroll
      for (int u = 0; u < kAccUnroll; u++) {
        const int sl = u * kAccThreads + tid;
        const bool in = kFull || (p_begin + it * kStagePts + sl < p_end);
        const int s0 = in ? sl : 0;  // past-the-end lanes re-read slot 0 with weight 0
        int ind; float qv;           // the stage is still in shared memory: re-read, do not carry
        match_of(buf, s0, in, ind, qv);
        // 32 consecutive points without a usable match (occluded / low-confidence regions are
        // spatially coherent in real maps) contribute exact zeros: skip their arithmetic.  Measured
        // on config 3: +1.3 % with 96 % usable matches, +7 % with half of them dropped in ~30-px
        // blobs (profiles/time_gn_sparse.py); the calib kernel lost 2.6 % to it and goes without.
        if (kFull && MODE != M3R_GN_CALIB && __ballot_sync(0xffffffffu, qv > prm.Q_thresh) == 0u) continue;
        const float4 xj = reinterpret_cast<const float4*>(buf)[s0];
        point_terms<MODE>(prm, R, t, xi[u], xj, (long long)ind, qv, h, g);
      }
    };
    if (stage_full(it)) compute(std::true_type{}); else compute(std::false_type{});
    __syncthreads();  // every thread is done with stage `it`: its buffer may be refilled
    if (tid == 0 && it + kStages < T) issue(it + kStages);
  }
  block_reduce_finish(a, h, g, el, e, split);
}

// Raises flags[rank] = epoch in every process: a pass over zero edges still has to announce itself.
__global__ void gn_peer_signal_kernel(const GnPeerFlags f, const GnHeader* __restrict__ hdr) {
  if (*reinterpret_cast<const volatile int*>(&hdr->peer_timeout)) return;
  if (threadIdx.x == 0) peer_signal_all(f);
}

static GnDev make_dev_params(const m3r_gn_params* p, bool packed) {
  GnDev d;
  d.inv_var_a = (float)(1.0 / ((double)p->sigma_a * (double)p->sigma_a));
  d.inv_var_b = p->sigma_b != 0.f ? (float)(1.0 / ((double)p->sigma_b * (double)p->sigma_b)) : 0.f;
  d.C_thresh = p->C_thresh;
  // packed records carry valid_match in the sign of Q (Q >= 0: it is a square root), so an invalid
  // match must fail `q > Q_thresh` through its sign: clamp the threshold at 0.  A valid match with
  // Q == 0 has weight 0 on either side of the test, so the results do not change.
  d.Q_thresh = (packed && p->Q_thresh < 0.f) ? 0.f : p->Q_thresh;
  d.fx = p->fx; d.fy = p->fy; d.cx = p->cx; d.cy = p->cy;
  d.Kdev = p->mode == M3R_GN_CALIB ? p->K_dev : nullptr;
  d.u_lo = (float)p->pixel_border; d.u_hi = (float)(p->width - 1 - p->pixel_border);
  d.v_lo = (float)p->pixel_border; d.v_hi = (float)(p->height - 1 - p->pixel_border);
  d.z_eps = p->z_eps;
  d.width = p->width > 0 ? p->width : 1;
  d.width_magic = d.width > 1 ? (unsigned)((0x100000000ULL + (unsigned)d.width - 1) / (unsigned)d.width) : 0xFFFFFFFFu;
  return d;
}

template <int MODE, bool PK>
static int launch_accum(const AccArgs& a, dim3 grid, bool stream_ok, cudaStream_t st) {
  if (stream_ok) {  // > 48 KB of dynamic shared memory is an opt-in, per device: set it every time
    M3R_TRY(check_cuda(cudaFuncSetAttribute(gn_accum_stream_kernel<MODE, PK>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            acc_smem_bytes<PK>())));
    gn_accum_stream_kernel<MODE, PK><<<grid, kAccThreads, acc_smem_bytes<PK>(), st>>>(a);
  } else {
    gn_accum_kernel<MODE, PK><<<grid, kAccThreads, 0, st>>>(a);
  }
  return after_launch();
}

}  // namespace m3r

using namespace m3r;

extern "C" size_t m3r_gn_workspace_bytes(int num_poses, int num_edges, int local_edges, int n) {
  if (num_poses < 1 || num_edges < 0 || local_edges < 0 || n < 0) return 0;
  return gn_layout(num_poses, num_edges, local_edges, n).total;
}

extern "C" int* m3r_gn_done_flag(void* workspace) {
  return workspace ? &reinterpret_cast<GnHeader*>(workspace)->done : nullptr;
}

extern "C" int* m3r_gn_status(void* workspace) {
  return workspace ? &reinterpret_cast<GnHeader*>(workspace)->done : nullptr;  // int32[4]: the GnHeader
}

extern "C" double* m3r_gn_blocks(void* workspace, int num_poses, int num_edges, int local_edges,
                                 int n) {
  if (!workspace) return nullptr;
  return ws_ptr<double>(workspace, gn_layout(num_poses, num_edges, local_edges, n).blocks);
}

static int gn_prepare_impl(const float* Xs, const float* Cs, const int64_t* ii, const int64_t* jj,
                           int two_way, int num_poses, int n, int num_edges, int local_edges,
                           void* workspace, size_t workspace_bytes, m3r_stream_t stream) {
  if (num_poses < 1 || n <= 0 || num_edges <= 0 || local_edges < 0 || local_edges > num_edges)
    return M3R_EINVAL;
  if (two_way && (num_edges & 1)) return M3R_EINVAL;
  if (!Xs || !Cs || !ii || !jj || !workspace) return M3R_EINVAL;
  const GnLayout L = gn_layout(num_poses, num_edges, local_edges, n);
  if (workspace_bytes < L.total) return M3R_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  gn_index_kernel<<<1, kIdxThreads, 0, st>>>(
      ii, jj, two_way, num_edges, num_poses, L.P, ws_ptr<int64_t>(workspace, L.keys),
      ws_ptr<int>(workspace, L.rank), ws_ptr<int>(workspace, L.ie), ws_ptr<int>(workspace, L.je),
      ws_ptr<int>(workspace, L.row_ptr), ws_ptr<int>(workspace, L.adj),
      ws_ptr<int>(workspace, L.flags), L.nflags, ws_ptr<int>(workspace, L.tickets),
      ws_ptr<GnHeader>(workspace, L.hdr));
  M3R_TRY(after_launch());
  const size_t count = (size_t)num_poses * n;
  const int grid = (int)((count + 255) / 256 < (size_t)(kNumSMs * 16) ? (count + 255) / 256
                                                                       : (size_t)(kNumSMs * 16));
  gn_pack_kernel<<<grid, 256, 0, st>>>(Xs, Cs, count, ws_ptr<float4>(workspace, L.xc));
  M3R_TRY(after_launch());
  if (L.nbm) return gn_chol64_prepare(workspace, L, st);  // task list of the macro-tile Cholesky
  return M3R_OK;
}

extern "C" int m3r_gn_prepare(const float* Xs, const float* Cs, const int64_t* ii,
                              const int64_t* jj, int num_poses, int n, int num_edges,
                              int local_edges, void* workspace, size_t workspace_bytes,
                              m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_prepare");
  return gn_prepare_impl(Xs, Cs, ii, jj, 0, num_poses, n, num_edges, local_edges, workspace,
                         workspace_bytes, stream);
}

extern "C" int m3r_gn_prepare_two_way(const float* Xs, const float* Cs, const int64_t* ii,
                                      const int64_t* jj, int num_poses, int n, int num_edges,
                                      int local_edges, void* workspace, size_t workspace_bytes,
                                      m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_prepare_two_way");
  return gn_prepare_impl(Xs, Cs, ii, jj, 1, num_poses, n, num_edges, local_edges, workspace,
                         workspace_bytes, stream);
}

// One accumulation pass over local edges [0, edge_count) = global [edge_begin, +edge_count).
// seg: the per-edge streams in up to two segments (AccArgs); packed selects the record layout.
struct EdgeStreams {
  const void* idx[2]; const uint8_t* vm[2]; const float* Q[2];
  int split; bool packed;
};

static int gn_accumulate_impl(const m3r_gn_params* prm, const float* Twc, const EdgeStreams& es,
                              int num_poses, int n, int num_edges, int edge_begin, int edge_count,
                              double* blocks, const m3r_gn_peers* peers, void* workspace,
                              size_t workspace_bytes, m3r_stream_t stream) {
  if (!prm || num_poses < 1 || n <= 0 || num_edges <= 0 || edge_begin < 0 || edge_count < 0 ||
      edge_begin + edge_count > num_edges || edge_count > 65535)
    return M3R_EINVAL;
  AccArgs a;
  a.dst.world = 0; a.sig.world = 0; a.sig.rank = 0; a.sig.epoch = 0; a.pass_edges = 0;
  if (peers) {
    if (peers->world < 1 || peers->world > M3R_MAX_PEERS || peers->rank < 0 || peers->rank >= peers->world)
      return M3R_EINVAL;
    a.dst.world = peers->world;
    a.sig.world = peers->world; a.sig.rank = peers->rank; a.sig.epoch = peers->epoch;
    for (int r = 0; r < peers->world; r++) {
      if (!peers->blocks[r] || !peers->flags[r]) return M3R_EINVAL;
      a.dst.blocks[r] = peers->blocks[r];
      a.sig.flags[r] = peers->flags[r];
    }
    blocks = peers->blocks[peers->rank];
    if (peers->world > 1) a.pass_edges = edge_count;
  }
  cudaStream_t st = as_stream(stream);
  if (edge_count == 0) {
    if (peers && peers->world > 1 && workspace) {  // nothing to add, but the peers still wait for this rank
      const GnLayout L0 = gn_layout(num_poses, num_edges, 0, n);
      gn_peer_signal_kernel<<<1, 32, 0, st>>>(a.sig, ws_ptr<GnHeader>(workspace, L0.hdr));
      return after_launch();
    }
    return M3R_OK;
  }
  if (!Twc || !es.idx[0] || !blocks || !workspace) return M3R_EINVAL;
  if (!es.packed && (!es.vm[0] || !es.Q[0])) return M3R_EINVAL;
  if (es.split < 0 || (es.split < edge_count && !es.idx[1])) return M3R_EINVAL;
  if (prm->sigma_a == 0.f) return M3R_EINVAL;
  // the partial buffer is sized for `local_edges`; accept any slice up to that size
  const GnLayout L = gn_layout(num_poses, num_edges, edge_count, n);  // offsets independent of edge_count
  if (workspace_bytes < L.part + sizeof(float) * (size_t)edge_count * L.S * kGnPartStride)
    return M3R_EWORKSPACE;
  a.prm = make_dev_params(prm, es.packed);
  a.Twc = Twc; a.xc = ws_ptr<float4>(workspace, L.xc);
  a.ie = ws_ptr<int>(workspace, L.ie); a.je = ws_ptr<int>(workspace, L.je);
  for (int s = 0; s < 2; s++) { a.idx[s] = es.idx[s]; a.vmatch[s] = es.vm[s]; a.Q[s] = es.Q[s]; }
  a.split = es.split < edge_count ? es.split : edge_count;
  a.n = n; a.edge_begin = edge_begin; a.chunk = L.chunk; a.S = L.S; a.M = num_edges;
  a.part = ws_ptr<float>(workspace, L.part);
  a.hdr = ws_ptr<GnHeader>(workspace, L.hdr);
  a.tickets = ws_ptr<int>(workspace, L.tickets);
  a.blocks = blocks;
  dim3 grid(L.S, edge_count);
  // streaming (bulk-copy) form when every stage of every stream is 16-B aligned
  bool stream_ok = (n % 16 == 0);
  for (int s = 0; s < 2; s++) {
    stream_ok = stream_ok && ((uintptr_t)es.idx[s] % 16 == 0);
    if (!es.packed) stream_ok = stream_ok && ((uintptr_t)es.vm[s] % 16 == 0) && ((uintptr_t)es.Q[s] % 16 == 0);
  }
#define M3R_ACC(MODE) \
  (es.packed ? launch_accum<MODE, true>(a, grid, stream_ok, st) : launch_accum<MODE, false>(a, grid, stream_ok, st))
  switch (prm->mode) {
    case M3R_GN_RAYS: return M3R_ACC(M3R_GN_RAYS);
    case M3R_GN_CALIB: return M3R_ACC(M3R_GN_CALIB);
    case M3R_GN_POINTS: return M3R_ACC(M3R_GN_POINTS);
    default: return M3R_EINVAL;
  }
#undef M3R_ACC
}

static EdgeStreams reference_streams(const int64_t* idx, const uint8_t* vm, const float* Q) {
  EdgeStreams es;
  es.idx[0] = idx; es.idx[1] = nullptr; es.vm[0] = vm; es.vm[1] = nullptr; es.Q[0] = Q; es.Q[1] = nullptr;
  es.split = 0x7fffffff; es.packed = false;
  return es;
}
static EdgeStreams packed_streams(const m3r_packed_match* pk0, const m3r_packed_match* pk1, int split) {
  EdgeStreams es;
  es.idx[0] = pk0; es.idx[1] = pk1; es.vm[0] = es.vm[1] = nullptr; es.Q[0] = es.Q[1] = nullptr;
  es.split = split; es.packed = true;
  return es;
}

extern "C" int m3r_gn_accumulate(const m3r_gn_params* prm, const float* Twc,
                                 const int64_t* idx_ii2jj, const uint8_t* valid_match,
                                 const float* Q, int num_poses, int n, int num_edges,
                                 int edge_begin, int edge_count, double* blocks, void* workspace,
                                 size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_accumulate");
  return gn_accumulate_impl(prm, Twc, reference_streams(idx_ii2jj, valid_match, Q), num_poses, n,
                            num_edges, edge_begin, edge_count, blocks, nullptr, workspace,
                            workspace_bytes, stream);
}

extern "C" int m3r_gn_accumulate_packed(const m3r_gn_params* prm, const float* Twc,
                                        const m3r_packed_match* pk0, const m3r_packed_match* pk1,
                                        int split, int num_poses, int n, int num_edges,
                                        int edge_begin, int edge_count, double* blocks,
                                        const m3r_gn_peers* peers, void* workspace,
                                        size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_accumulate_packed");
  return gn_accumulate_impl(prm, Twc, packed_streams(pk0, pk1, split), num_poses, n, num_edges,
                            edge_begin, edge_count, blocks, peers, workspace, workspace_bytes, stream);
}

extern "C" int m3r_gn_accumulate_p2p(const m3r_gn_params* prm, const float* Twc,
                                     const int64_t* idx_ii2jj, const uint8_t* valid_match,
                                     const float* Q, int num_poses, int n, int num_edges,
                                     int edge_begin, int edge_count, const m3r_gn_peers* peers,
                                     void* workspace, size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_accumulate_p2p");
  if (!peers) return M3R_EINVAL;
  return gn_accumulate_impl(prm, Twc, reference_streams(idx_ii2jj, valid_match, Q), num_poses, n,
                            num_edges, edge_begin, edge_count, nullptr, peers, workspace,
                            workspace_bytes, stream);
}

extern "C" int m3r_gn_solve_step_p2p(float* Twc, const m3r_gn_peers* peers, int num_poses,
                                     int num_edges, float delta_thresh, float* dx, void* workspace,
                                     size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_solve_step_p2p");
  if (!peers || peers->world < 1 || peers->world > M3R_MAX_PEERS || peers->rank < 0 ||
      peers->rank >= peers->world)
    return M3R_EINVAL;
  if (num_poses < 2 || num_edges <= 0 || !Twc || !dx || !workspace) return M3R_EINVAL;
  const GnLayout L = gn_layout(num_poses, num_edges, 0, 0);
  if (workspace_bytes < L.xc) return M3R_EWORKSPACE;
  return gn_solve_launch(Twc, peers->blocks[peers->rank], num_poses, num_edges, delta_thresh, dx,
                         workspace, L, as_stream(stream), peers->flags[peers->rank], peers->world,
                         peers->epoch);
}

extern "C" int m3r_gn_solve_step(float* Twc, const double* blocks, int num_poses, int num_edges,
                                 float delta_thresh, float* dx, void* workspace,
                                 size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gn_solve_step");
  if (num_poses < 2 || num_edges <= 0 || !Twc || !blocks || !dx || !workspace) return M3R_EINVAL;
  const GnLayout L = gn_layout(num_poses, num_edges, 0, 0);  // offsets up to `x` need (N, M) only
  if (workspace_bytes < L.xc) return M3R_EWORKSPACE;
  return gn_solve_launch(Twc, blocks, num_poses, num_edges, delta_thresh, dx, workspace, L,
                         as_stream(stream));
}

// the whole loop: prepare, then max_iter x (accumulate in grid.y-sized slices, solve step)
static int gauss_newton_impl(const m3r_gn_params* prm, float* Twc, const float* Xs, const float* Cs,
                             const int64_t* ii, const int64_t* jj, int two_way, const EdgeStreams& es,
                             int num_poses, int n, int num_edges, int max_iter, float delta_thresh,
                             float* dx, void* workspace, size_t workspace_bytes, m3r_stream_t stream) {
  if (!prm || num_poses < 2 || n <= 0 || num_edges <= 0 || max_iter < 0 || !dx) return M3R_EINVAL;
  if (workspace_bytes < m3r_gn_workspace_bytes(num_poses, num_edges, num_edges, n))
    return M3R_EWORKSPACE;
  M3R_TRY(check_cuda(cudaMemsetAsync(dx, 0, sizeof(float) * 7 * (size_t)(num_poses - 1),
                                     as_stream(stream))));
  if (max_iter == 0) return M3R_OK;
  M3R_TRY(gn_prepare_impl(Xs, Cs, ii, jj, two_way, num_poses, n, num_edges, num_edges, workspace,
                          workspace_bytes, stream));
  double* blocks = m3r_gn_blocks(workspace, num_poses, num_edges, num_edges, n);
  const int kSlice = 32768;  // grid.y limit is 65535
  for (int it = 0; it < max_iter; it++) {
    for (int e0 = 0; e0 < num_edges; e0 += kSlice) {
      const int cnt = num_edges - e0 < kSlice ? num_edges - e0 : kSlice;
      EdgeStreams sl = es;  // the slice [e0, e0+cnt) of the local (= whole) edge list
      const size_t esz = (size_t)n;
      auto adv = [&](int seg, size_t rows) {
        sl.idx[seg] = static_cast<const unsigned char*>(es.idx[seg]) + rows * esz * 8;
        if (!es.packed) { sl.vm[seg] = es.vm[seg] + rows * esz; sl.Q[seg] = es.Q[seg] + rows * esz; }
      };
      if (e0 >= es.split) {  // wholly inside segment 1
        adv(1, (size_t)(e0 - es.split));
        sl.idx[0] = sl.idx[1]; sl.vm[0] = sl.vm[1]; sl.Q[0] = sl.Q[1];
        sl.split = 0x7fffffff;
      } else {
        adv(0, (size_t)e0);
        sl.split = es.split - e0;
      }
      M3R_TRY(gn_accumulate_impl(prm, Twc, sl, num_poses, n, num_edges, e0, cnt, blocks, nullptr,
                                 workspace, workspace_bytes, stream));
    }
    M3R_TRY(m3r_gn_solve_step(Twc, blocks, num_poses, num_edges, delta_thresh, dx, workspace,
                              workspace_bytes, stream));
  }
  return M3R_OK;
}

extern "C" int m3r_gauss_newton(const m3r_gn_params* prm, float* Twc, const float* Xs,
                                const float* Cs, const int64_t* ii, const int64_t* jj,
                                const int64_t* idx_ii2jj, const uint8_t* valid_match,
                                const float* Q, int num_poses, int n, int num_edges, int max_iter,
                                float delta_thresh, float* dx, void* workspace,
                                size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gauss_newton");
  if (max_iter > 0 && (!idx_ii2jj || !valid_match || !Q)) return M3R_EINVAL;
  return gauss_newton_impl(prm, Twc, Xs, Cs, ii, jj, 0, reference_streams(idx_ii2jj, valid_match, Q),
                           num_poses, n, num_edges, max_iter, delta_thresh, dx, workspace,
                           workspace_bytes, stream);
}

extern "C" int m3r_gauss_newton_packed(const m3r_gn_params* prm, float* Twc, const float* Xs,
                                       const float* Cs, const int64_t* ii, const int64_t* jj,
                                       int two_way, const m3r_packed_match* pk0,
                                       const m3r_packed_match* pk1, int split, int num_poses, int n,
                                       int num_edges, int max_iter, float delta_thresh, float* dx,
                                       void* workspace, size_t workspace_bytes, m3r_stream_t stream) {
  M3R_NVTX("m3r_gauss_newton_packed");
  if (max_iter > 0 && (!pk0 || split < 0 || (split < num_edges && !pk1))) return M3R_EINVAL;
  return gauss_newton_impl(prm, Twc, Xs, Cs, ii, jj, two_way, packed_streams(pk0, pk1, split),
                           num_poses, n, num_edges, max_iter, delta_thresh, dx, workspace,
                           workspace_bytes, stream);
}
