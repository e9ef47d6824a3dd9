// m3r_frontend.cu — the two streaming steps either side of the solvers, sm_100a.
//
//  * m3r_edge_build: the per-edge tensor construction of FactorGraph.add_factors (reference
//    global_opt.py:55-67): Qj = sqrt(Qii[idx_i2j] * Qji), valid_j = valid_match_j & (Qj > Q_conf),
//    and the per-edge count of valid_j that match_frac is made of.  The reference runs 6 PyTorch
//    kernels (arange/repeat index tensor, advanced-index gather, mul, sqrt, compare, and, sum)
//    and materialises a (B,n) int64 batch-index tensor; here it is one pass, 21 B per point.
//  * m3r_pointmap_update: Frame.update_pointmap (frame.py:41-105) for the per-point filtering
//    modes, optionally preceded by T.act(X) (tracker.py:97-98: `Xkk = T_CkCf.act(Xkf)` then
//    `keyframe.update_pointmap(Xkk, Ckf)`), in place: 48 B per point instead of ~8 elementwise
//    kernels with their temporaries.
//    The weighted mean is evaluated with the reference's operation order and without FMA
//    contraction, so the weighted_pointmap / indep_conf modes are bit-identical to torch.
#include "m3r_common.cuh"
#include "m3r_sim3.cuh"

namespace m3r {

constexpr int kFeThreads = 256;
constexpr int kFeUnroll = 4;

// packed record of one correspondence (include/m3r_b200.h, m3r_packed_match)
__device__ __forceinline__ m3r_packed_match pack_match(long long idx, bool valid, float q) {
  m3r_packed_match r;
  r.idx = (int32_t)idx;
  r.q = valid ? q : -q;  // -0.0f when q == 0: the sign bit alone carries valid_match
  return r;
}

__global__ void __launch_bounds__(kFeThreads)
edge_build_kernel(const int64_t* __restrict__ idx, const uint8_t* __restrict__ vmatch,
                  const float* __restrict__ Qsrc, const float* __restrict__ Qdst, int n,
                  float Q_conf, float* __restrict__ Qout, m3r_packed_match* __restrict__ pk,
                  int* __restrict__ count) {
  const int b = blockIdx.y;
  const size_t row = (size_t)b * n;
  int local = 0;
  const int base = blockIdx.x * (kFeThreads * kFeUnroll) + threadIdx.x;
  int64_t id[kFeUnroll];
  float qd[kFeUnroll];
  bool vm[kFeUnroll];
#pragma unroll
  for (int u = 0; u < kFeUnroll; u++) {
    const int k = base + u * kFeThreads;
    const bool in = k < n;
    id[u] = in ? idx[row + k] : 0;
    qd[u] = in ? ld_stream_f(Qdst + row + k) : 0.f;
    vm[u] = in && vmatch[row + k] != 0;
  }
#pragma unroll
  for (int u = 0; u < kFeUnroll; u++) {
    const int k = base + u * kFeThreads;
    if (k < n) {
      const float q = sqrtf(__fmul_rn(__ldg(Qsrc + row + id[u]), qd[u]));  // torch.sqrt(Qii[.., idx] * Qji)
      if (Qout) Qout[row + k] = q;
      if (pk) pk[row + k] = pack_match(id[u], vm[u], q);
      local += (vm[u] && q > Q_conf) ? 1 : 0;
    }
  }
  // block count -> one integer atomic per CTA (order-independent, hence deterministic)
  __shared__ int wsum[kFeThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < kFeThreads / 32; w++) s += wsum[w];
    if (s) atomicAdd(count + b, s);
  }
}

// Same pass with four CONSECUTIVE points per thread (n % 4 == 0, 16-byte aligned rows): idx as
// two 128-bit loads, valid as one 32-bit load, Q_dst / Q_out as one 128-bit access each; only the
// Q_src gather stays scalar.  A 256-thread CTA covers 4096 points.
constexpr int kEbVec = 2;  // groups of 4 points per thread
__global__ void __launch_bounds__(kFeThreads)
edge_build_vec4_kernel(const int64_t* __restrict__ idx, const uint8_t* __restrict__ vmatch,
                       const float* __restrict__ Qsrc, const float* __restrict__ Qdst, int n,
                       float Q_conf, float* __restrict__ Qout, m3r_packed_match* __restrict__ pk,
                       int* __restrict__ count) {
  const int b = blockIdx.y;
  const size_t row = (size_t)b * n;
  int local = 0;
  const int n4 = n >> 2;
  int g[kEbVec];
  longlong2 ia[kEbVec], ib[kEbVec];
  float4 qd[kEbVec];
  uint32_t vm[kEbVec];
#pragma unroll
  for (int u = 0; u < kEbVec; u++) {
    g[u] = (blockIdx.x * kEbVec + u) * kFeThreads + threadIdx.x;
    const bool in = g[u] < n4;
    const size_t k = row + 4 * (size_t)(in ? g[u] : 0);
    const uint4 a = ld_stream_u4(idx + k), c = ld_stream_u4(idx + k + 2);
    ia[u].x = (long long)(((unsigned long long)a.y << 32) | a.x); ia[u].y = (long long)(((unsigned long long)a.w << 32) | a.z);
    ib[u].x = (long long)(((unsigned long long)c.y << 32) | c.x); ib[u].y = (long long)(((unsigned long long)c.w << 32) | c.z);
    qd[u] = ld_stream_f4(Qdst + k);
    vm[u] = in ? *reinterpret_cast<const uint32_t*>(vmatch + k) : 0u;
  }
#pragma unroll
  for (int u = 0; u < kEbVec; u++) {
    if (g[u] < n4) {
      const size_t k = row + 4 * (size_t)g[u];
      float4 q;
      q.x = sqrtf(__fmul_rn(__ldg(Qsrc + row + ia[u].x), qd[u].x));  // torch.sqrt(Qii[.., idx] * Qji)
      q.y = sqrtf(__fmul_rn(__ldg(Qsrc + row + ia[u].y), qd[u].y));
      q.z = sqrtf(__fmul_rn(__ldg(Qsrc + row + ib[u].x), qd[u].z));
      q.w = sqrtf(__fmul_rn(__ldg(Qsrc + row + ib[u].y), qd[u].w));
      if (Qout) *reinterpret_cast<float4*>(Qout + k) = q;
      if (pk) {  // four 8-byte records = two 128-bit stores
        const m3r_packed_match r0 = pack_match(ia[u].x, vm[u] & 0xFFu, q.x), r1 = pack_match(ia[u].y, vm[u] & 0xFF00u, q.y);
        const m3r_packed_match r2 = pack_match(ib[u].x, vm[u] & 0xFF0000u, q.z), r3 = pack_match(ib[u].y, vm[u] & 0xFF000000u, q.w);
        float4* o = reinterpret_cast<float4*>(pk + k);
        o[0] = make_float4(__int_as_float(r0.idx), r0.q, __int_as_float(r1.idx), r1.q);
        o[1] = make_float4(__int_as_float(r2.idx), r2.q, __int_as_float(r3.idx), r3.q);
      }
      local += ((vm[u] & 0xFFu) && q.x > Q_conf) + ((vm[u] & 0xFF00u) && q.y > Q_conf) +
               ((vm[u] & 0xFF0000u) && q.z > Q_conf) + ((vm[u] & 0xFF000000u) && q.w > Q_conf);
    }
  }
  __shared__ int wsum[kFeThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < kFeThreads / 32; w++) s += wsum[w];
    if (s) atomicAdd(count + b, s);
  }
}

// reference layout (idx i64, valid u8, Q f32) <-> packed records; grid-stride, one point per thread
__global__ void __launch_bounds__(kFeThreads)
edge_pack_kernel(const int64_t* __restrict__ idx, const uint8_t* __restrict__ vmatch,
                 const float* __restrict__ Q, size_t count, m3r_packed_match* __restrict__ pk) {
  const size_t stride = (size_t)gridDim.x * kFeThreads;
  for (size_t k = (size_t)blockIdx.x * kFeThreads + threadIdx.x; k < count; k += stride)
    pk[k] = pack_match(idx[k], vmatch[k] != 0, Q[k]);
}
__global__ void __launch_bounds__(kFeThreads)
edge_unpack_kernel(const m3r_packed_match* __restrict__ pk, size_t count, int64_t* __restrict__ idx,
                   uint8_t* __restrict__ vmatch, float* __restrict__ Q) {
  const size_t stride = (size_t)gridDim.x * kFeThreads;
  for (size_t k = (size_t)blockIdx.x * kFeThreads + threadIdx.x; k < count; k += stride) {
    const m3r_packed_match r = pk[k];
    if (idx) idx[k] = (int64_t)r.idx;
    if (vmatch) vmatch[k] = (__float_as_int(r.q) >= 0) ? 1 : 0;
    if (Q) Q[k] = fabsf(r.q);
  }
}

enum { kPmWeighted = 0, kPmSpherical = 1, kPmIndepConf = 2 };

__device__ __forceinline__ void to_spherical(float x, float y, float z, float& r, float& phi, float& th) {
  r = sqrtf(x * x + y * y + z * z);  // torch.linalg.norm
  phi = atan2f(y, x);
  th = acosf(z / r);
}

template <int MODE, bool ACT>
__global__ void __launch_bounds__(kFeThreads)
pointmap_update_kernel(float* __restrict__ Xc, float* __restrict__ Cc, const float* __restrict__ Xn,
                       const float* __restrict__ Cn, const float* __restrict__ T, int n) {
  float R[9], t[3];
  if (ACT) {  // s R X + t  (Sim3.act; actSim3 gn_kernels.cu:207-219)
    float q[4] = {__ldg(T + 3), __ldg(T + 4), __ldg(T + 5), __ldg(T + 6)};
    const float s = __ldg(T + 7);
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float ev[3] = {0.f, 0.f, 0.f}, col[3];
      ev[c] = 1.f;
      act_so3<float>(q, ev, col);
      R[0 + c] = s * col[0]; R[3 + c] = s * col[1]; R[6 + c] = s * col[2];
    }
    t[0] = __ldg(T); t[1] = __ldg(T + 1); t[2] = __ldg(T + 2);
  }
  const int stride = gridDim.x * kFeThreads;
  for (int p = blockIdx.x * kFeThreads + threadIdx.x; p < n; p += stride) {
    float x = ld_stream_f(Xn + 3 * (size_t)p), y = ld_stream_f(Xn + 3 * (size_t)p + 1),
          z = ld_stream_f(Xn + 3 * (size_t)p + 2);
    const float cn = ld_stream_f(Cn + p);
    if (ACT) {
      const float px = fmaf(R[0], x, fmaf(R[1], y, fmaf(R[2], z, t[0])));
      const float py = fmaf(R[3], x, fmaf(R[4], y, fmaf(R[5], z, t[1])));
      const float pz = fmaf(R[6], x, fmaf(R[7], y, fmaf(R[8], z, t[2])));
      x = px; y = py; z = pz;
    }
    float* xo = Xc + 3 * (size_t)p;
    const float c0 = Cc[p];
    const float ox = xo[0], oy = xo[1], oz = xo[2];
    if (MODE == kPmWeighted) {  // ((C * X_canon) + (C_new * X)) / (C + C_new), frame.py:74-77
      const float den = __fadd_rn(c0, cn);
      xo[0] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, ox), __fmul_rn(cn, x)), den);
      xo[1] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, oy), __fmul_rn(cn, y)), den);
      xo[2] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, oz), __fmul_rn(cn, z)), den);
      Cc[p] = den;
    } else if (MODE == kPmSpherical) {  // frame.py:78-101
      float r1, p1, t1, r2, p2, t2;
      to_spherical(ox, oy, oz, r1, p1, t1);
      to_spherical(x, y, z, r2, p2, t2);
      const float den = c0 + cn;
      const float r = (c0 * r1 + cn * r2) / den, ph = (c0 * p1 + cn * p2) / den, th = (c0 * t1 + cn * t2) / den;
      float st, ct, sp, cp;
      sincosf(th, &st, &ct);
      sincosf(ph, &sp, &cp);
      xo[0] = r * st * cp; xo[1] = r * st * sp; xo[2] = r * ct;
      Cc[p] = den;
    } else {  // indep_conf: keep the more confident observation per point, frame.py:69-72
      if (cn > c0) { xo[0] = x; xo[1] = y; xo[2] = z; Cc[p] = cn; }
    }
  }
}

// Four points per thread through 128-bit accesses (n % 4 == 0, 16-byte aligned arrays): the
// (n,3) arrays are read as 3 float4 per 4 points, so each thread keeps 8 LDG.128 in flight instead
// of 8 scalar loads and issues 4 STG.128 instead of 16 scalar stores.  Same per-point arithmetic.
template <int MODE, bool ACT>
__global__ void __launch_bounds__(kFeThreads)
pointmap_update_vec4_kernel(float* __restrict__ Xc, float* __restrict__ Cc, const float* __restrict__ Xn,
                            const float* __restrict__ Cn, const float* __restrict__ T, int n4) {
  float R[9], t[3];
  if (ACT) {
    float q[4] = {__ldg(T + 3), __ldg(T + 4), __ldg(T + 5), __ldg(T + 6)};
    const float s = __ldg(T + 7);
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float ev[3] = {0.f, 0.f, 0.f}, col[3];
      ev[c] = 1.f;
      act_so3<float>(q, ev, col);
      R[0 + c] = s * col[0]; R[3 + c] = s * col[1]; R[6 + c] = s * col[2];
    }
    t[0] = __ldg(T); t[1] = __ldg(T + 1); t[2] = __ldg(T + 2);
  }
  const int stride = gridDim.x * kFeThreads;
  for (int g = blockIdx.x * kFeThreads + threadIdx.x; g < n4; g += stride) {
    float xn[12], xo[12], cn[4], co[4];
    {
      const float4* pn = reinterpret_cast<const float4*>(Xn) + 3 * (size_t)g;
      const float4* po = reinterpret_cast<const float4*>(Xc) + 3 * (size_t)g;
      const float4 a0 = ld_stream_f4(pn), a1 = ld_stream_f4(pn + 1), a2 = ld_stream_f4(pn + 2);
      const float4 b0 = po[0], b1 = po[1], b2 = po[2];
      const float4 c4 = ld_stream_f4(reinterpret_cast<const float4*>(Cn) + g);
      const float4 d4 = reinterpret_cast<const float4*>(Cc)[g];
      xn[0] = a0.x; xn[1] = a0.y; xn[2] = a0.z; xn[3] = a0.w; xn[4] = a1.x; xn[5] = a1.y; xn[6] = a1.z; xn[7] = a1.w;
      xn[8] = a2.x; xn[9] = a2.y; xn[10] = a2.z; xn[11] = a2.w;
      xo[0] = b0.x; xo[1] = b0.y; xo[2] = b0.z; xo[3] = b0.w; xo[4] = b1.x; xo[5] = b1.y; xo[6] = b1.z; xo[7] = b1.w;
      xo[8] = b2.x; xo[9] = b2.y; xo[10] = b2.z; xo[11] = b2.w;
      cn[0] = c4.x; cn[1] = c4.y; cn[2] = c4.z; cn[3] = c4.w;
      co[0] = d4.x; co[1] = d4.y; co[2] = d4.z; co[3] = d4.w;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      float x = xn[3 * k], y = xn[3 * k + 1], z = xn[3 * k + 2];
      if (ACT) {
        const float px = fmaf(R[0], x, fmaf(R[1], y, fmaf(R[2], z, t[0])));
        const float py = fmaf(R[3], x, fmaf(R[4], y, fmaf(R[5], z, t[1])));
        const float pz = fmaf(R[6], x, fmaf(R[7], y, fmaf(R[8], z, t[2])));
        x = px; y = py; z = pz;
      }
      const float c0 = co[k], c1 = cn[k];
      const float ox = xo[3 * k], oy = xo[3 * k + 1], oz = xo[3 * k + 2];
      if (MODE == kPmWeighted) {
        const float den = __fadd_rn(c0, c1);
        xo[3 * k] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, ox), __fmul_rn(c1, x)), den);
        xo[3 * k + 1] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, oy), __fmul_rn(c1, y)), den);
        xo[3 * k + 2] = __fdiv_rn(__fadd_rn(__fmul_rn(c0, oz), __fmul_rn(c1, z)), den);
        co[k] = den;
      } else {  // indep_conf
        if (c1 > c0) { xo[3 * k] = x; xo[3 * k + 1] = y; xo[3 * k + 2] = z; co[k] = c1; }
      }
    }
    float4* po = reinterpret_cast<float4*>(Xc) + 3 * (size_t)g;
    po[0] = make_float4(xo[0], xo[1], xo[2], xo[3]);
    po[1] = make_float4(xo[4], xo[5], xo[6], xo[7]);
    po[2] = make_float4(xo[8], xo[9], xo[10], xo[11]);
    reinterpret_cast<float4*>(Cc)[g] = make_float4(co[0], co[1], co[2], co[3]);
  }
}

}  // namespace m3r

using namespace m3r;

static int edge_build_impl(const int64_t* idx_src2dst, const uint8_t* valid_match, const float* Q_src,
                           const float* Q_dst, int num_edges, int n, float Q_conf, float* Q_out,
                           m3r_packed_match* pk, int* valid_count, m3r_stream_t stream) {
  if (num_edges < 0 || n < 0 || num_edges > 65535) return M3R_EINVAL;
  if (num_edges == 0) return M3R_OK;
  if (!valid_count) return M3R_EINVAL;
  cudaStream_t st = as_stream(stream);
  M3R_TRY(check_cuda(cudaMemsetAsync(valid_count, 0, sizeof(int) * (size_t)num_edges, st)));
  if (n == 0) return M3R_OK;
  if (!idx_src2dst || !valid_match || !Q_src || !Q_dst || (!Q_out && !pk)) return M3R_EINVAL;
  const bool vec4 = (n % 4) == 0 &&
                    ((reinterpret_cast<uintptr_t>(idx_src2dst) | reinterpret_cast<uintptr_t>(valid_match) |
                      reinterpret_cast<uintptr_t>(Q_dst) | reinterpret_cast<uintptr_t>(Q_out) |
                      reinterpret_cast<uintptr_t>(pk)) & 15) == 0;
  if (vec4) {
    const int per = kFeThreads * kEbVec * 4;
    dim3 grid4((n + per - 1) / per, num_edges);
    edge_build_vec4_kernel<<<grid4, kFeThreads, 0, st>>>(idx_src2dst, valid_match, Q_src, Q_dst, n, Q_conf,
                                                         Q_out, pk, valid_count);
    return after_launch();
  }
  dim3 grid((n + kFeThreads * kFeUnroll - 1) / (kFeThreads * kFeUnroll), num_edges);
  edge_build_kernel<<<grid, kFeThreads, 0, st>>>(idx_src2dst, valid_match, Q_src, Q_dst, n, Q_conf,
                                                 Q_out, pk, valid_count);
  return after_launch();
}

extern "C" int m3r_edge_build(const int64_t* idx_src2dst, const uint8_t* valid_match,
                              const float* Q_src, const float* Q_dst, int num_edges, int n,
                              float Q_conf, float* Q_out, int* valid_count, m3r_stream_t stream) {
  M3R_NVTX("m3r_edge_build");
  if (num_edges > 0 && n > 0 && !Q_out) return M3R_EINVAL;
  return edge_build_impl(idx_src2dst, valid_match, Q_src, Q_dst, num_edges, n, Q_conf, Q_out, nullptr,
                         valid_count, stream);
}

extern "C" int m3r_edge_build_packed(const int64_t* idx_src2dst, const uint8_t* valid_match,
                                     const float* Q_src, const float* Q_dst, int num_edges, int n,
                                     float Q_conf, m3r_packed_match* packed_out, float* Q_out,
                                     int* valid_count, m3r_stream_t stream) {
  M3R_NVTX("m3r_edge_build_packed");
  if (num_edges > 0 && n > 0 && !packed_out) return M3R_EINVAL;
  return edge_build_impl(idx_src2dst, valid_match, Q_src, Q_dst, num_edges, n, Q_conf, Q_out, packed_out,
                         valid_count, stream);
}

extern "C" int m3r_edge_pack(const int64_t* idx, const uint8_t* valid_match, const float* Q, size_t count,
                             m3r_packed_match* packed_out, m3r_stream_t stream) {
  M3R_NVTX("m3r_edge_pack");
  if (count == 0) return M3R_OK;
  if (!idx || !valid_match || !Q || !packed_out) return M3R_EINVAL;
  const size_t blocks = (count + kFeThreads - 1) / kFeThreads;
  const int grid = (int)(blocks < (size_t)kNumSMs * 16 ? blocks : (size_t)kNumSMs * 16);
  edge_pack_kernel<<<grid, kFeThreads, 0, as_stream(stream)>>>(idx, valid_match, Q, count, packed_out);
  return after_launch();
}

extern "C" int m3r_edge_unpack(const m3r_packed_match* packed, size_t count, int64_t* idx,
                               uint8_t* valid_match, float* Q, m3r_stream_t stream) {
  M3R_NVTX("m3r_edge_unpack");
  if (count == 0) return M3R_OK;
  if (!packed) return M3R_EINVAL;
  const size_t blocks = (count + kFeThreads - 1) / kFeThreads;
  const int grid = (int)(blocks < (size_t)kNumSMs * 16 ? blocks : (size_t)kNumSMs * 16);
  edge_unpack_kernel<<<grid, kFeThreads, 0, as_stream(stream)>>>(packed, count, idx, valid_match, Q);
  return after_launch();
}

extern "C" int m3r_pointmap_update(float* X_canon, float* C, const float* X_new, const float* C_new,
                                   const float* T_act, int n, int mode, m3r_stream_t stream) {
  M3R_NVTX("m3r_pointmap_update");
  if (n < 0 || mode < 0 || mode > 2) return M3R_EINVAL;
  if (n == 0) return M3R_OK;
  if (!X_canon || !C || !X_new || !C_new) return M3R_EINVAL;
  cudaStream_t st = as_stream(stream);
  const bool vec4 = mode != kPmSpherical && (n % 4) == 0 &&
                    ((reinterpret_cast<uintptr_t>(X_canon) | reinterpret_cast<uintptr_t>(C) |
                      reinterpret_cast<uintptr_t>(X_new) | reinterpret_cast<uintptr_t>(C_new)) & 15) == 0;
  if (vec4) {
    const int n4 = n / 4;
    int g4 = (n4 + kFeThreads - 1) / kFeThreads;
    if (g4 > kNumSMs * 8) g4 = kNumSMs * 8;
#define M3R_PMV(MODE)                                                                                  \
  do {                                                                                                 \
    if (T_act) pointmap_update_vec4_kernel<MODE, true><<<g4, kFeThreads, 0, st>>>(X_canon, C, X_new,   \
                                                                                  C_new, T_act, n4);   \
    else pointmap_update_vec4_kernel<MODE, false><<<g4, kFeThreads, 0, st>>>(X_canon, C, X_new,        \
                                                                             C_new, T_act, n4);        \
  } while (0)
    if (mode == kPmWeighted) M3R_PMV(kPmWeighted); else M3R_PMV(kPmIndepConf);
#undef M3R_PMV
    return after_launch();
  }
  int grid = (n + kFeThreads - 1) / kFeThreads;
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
#define M3R_PM(MODE)                                                                              \
  do {                                                                                            \
    if (T_act) pointmap_update_kernel<MODE, true><<<grid, kFeThreads, 0, st>>>(X_canon, C, X_new, \
                                                                               C_new, T_act, n);  \
    else pointmap_update_kernel<MODE, false><<<grid, kFeThreads, 0, st>>>(X_canon, C, X_new,      \
                                                                          C_new, T_act, n);       \
  } while (0)
  switch (mode) {
    case kPmWeighted: M3R_PM(kPmWeighted); break;
    case kPmSpherical: M3R_PM(kPmSpherical); break;
    default: M3R_PM(kPmIndepConf); break;
  }
#undef M3R_PM
  return after_launch();
}
