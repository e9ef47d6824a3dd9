// m3r_refine.cu — dilated-window descriptor argmax (sm_100a).
//
// Replaces refine_matches_kernel / refine_matches_cuda (reference
// matching_kernels.cu:25-116).  The output is an integer pixel, so the reference's
// score arithmetic is reproduced bit-for-bit:
//   score_k = rn16( float(score_{k-1}) + float( rn16( float(q_k) * float(c_k) ) ) ),  k = 0..F-1
// (c10::Half operators: every product and every partial sum is formed in fp32 and rounded
// to fp16 — SURVEY H1).  fp32 has >= 2*11+2 significand bits, so the double rounding
// fp32->fp16 is innocuous for one + or * of fp16 operands: rn16(float(a) op float(b)) ==
// the correctly rounded fp16 result == HADD2.RN / HMUL2.RN per lane.  Two candidates are
// therefore scored at once in the two lanes of half2 registers (mul.rn/add.rn, which
// ptxas may not fuse into an fma).
//
// What bounds this kernel (ncu, profiles/r1q_*): not HBM.  Per (query, candidate) pair the exact
// rounding chain needs F HMUL2-lanes + F HADD2-lanes; fp16x2 instructions issue on the
// fma-heavy pipe only (one warp instruction per 2 cycles per SM sub-partition), and every
// candidate is a 2F-byte gather through the L1 data pipe (128 B/clk/SM when the lanes' 16-B
// words fall in distinct banks: profiles/micro/l1_align.cu; ~4.9 wavefronts per LDG.128 on real
// match fields, whose lanes are not perfectly consecutive).  At F = 24, 241 candidates per
// query the two floors are 0.49 ms (fp16 pipe) and 0.67 ms (L1) for 8 x 512x384 queries;
// everything else (index arithmetic, bounds tests, tie-breaking) is kept off both pipes:
//   * the argmax keeps ONE signed 32-bit key per query, (score_fp16_bits << 16) |
//     (0xFFFE - scan_index): a signed integer max then realises "strict > in (u outer,
//     v inner) scan order"  in any evaluation order — negative scores have the
//     sign bit set and lose against the initial +0, equal scores lose against an
//     earlier scan index, and a maximum carried over from a previous dilation carries the
//     tie field 0xFFFF so that no later tie can displace it;
//   * the evaluation order is therefore free: interior windows at dilation 1..5 take their
//     candidates in point-symmetric pairs with compile-time column offsets (no per-candidate
//     address arithmetic) and skip the centre after the first dilation (it is the previous
//     winner or an earlier loser: re-scoring it cannot change a strict-> argmax);
//   * a CTA is a 32x24 tile of queries, one CTA per SM, so that the union of its windows
//     (<= 62x54 pixels x 48 B at dilation 5) stays resident in that SM's L1 (hit rate 97.8 %);
//   * windows that touch the image border take the bounds-tested row-major form.
// Scores must not be NaN (a NaN never wins in the reference; its bit pattern would here).
#include "m3r_common.cuh"

namespace m3r {

#ifndef M3R_REFINE_TY
#define M3R_REFINE_TY 24  // one 32x24-query CTA per SM: its windows (<= 62x54 pixels) stay L1-resident
#endif
constexpr int kRefTX = 32, kRefTY = M3R_REFINE_TY;
constexpr int kRefineThreads = kRefTX * kRefTY;
constexpr int kKeyInvalid = (int)0x80000000;
// 0xFFFE - 7*i: tie field of window column i (row part subtracted per row).  Read from constant
// memory so that ptxas keeps the values in registers and PRMT's immediate is its selector.
__constant__ uint32_t kColumnTie[7] = {0xFFFE, 0xFFF7, 0xFFF0, 0xFFE9, 0xFFE2, 0xFFDB, 0xFFD4};

// Scores of two candidates (A -> low lane, B -> high lane) against the packed query
// q[t] = (q_2t, q_2t+1).  F8 = F/8 uint4 words per descriptor.
template <int F8>
__device__ __forceinline__ uint32_t score_pair(const uint4* __restrict__ ca,
                                               const uint4* __restrict__ cb,
                                               const uint32_t* q) {
  uint4 a[F8], b[F8];
#pragma unroll
  for (int w = 0; w < F8; w++) { a[w] = __ldg(ca + w); b[w] = __ldg(cb + w); }
  __half2 acc = __float2half2_rn(0.f);
#pragma unroll
  for (int w = 0; w < F8; w++) {
    const uint32_t aw[4] = {a[w].x, a[w].y, a[w].z, a[w].w};
    const uint32_t bw[4] = {b[w].x, b[w].y, b[w].z, b[w].w};
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const uint32_t lo = __byte_perm(aw[t], bw[t], 0x5410);  // (A_2t  , B_2t  )
      const uint32_t hi = __byte_perm(aw[t], bw[t], 0x7632);  // (A_2t+1, B_2t+1)
      const __half2 qq = *reinterpret_cast<const __half2*>(&q[w * 4 + t]);
      acc = __hadd2_rn(acc, __hmul2_rn(*reinterpret_cast<const __half2*>(&lo), __low2half2(qq)));
      acc = __hadd2_rn(acc, __hmul2_rn(*reinterpret_cast<const __half2*>(&hi), __high2half2(qq)));
    }
  }
  return *reinterpret_cast<uint32_t*>(&acc);
}

// key = (score bits << 16) | tie, as a signed int (see the header comment)
__device__ __forceinline__ int key_lo(uint32_t s2, uint32_t tie) { return (int)__byte_perm(s2, tie, 0x1054); }
__device__ __forceinline__ int key_hi(uint32_t s2, uint32_t tie) { return (int)__byte_perm(s2, tie, 0x3254); }

// One dilation of the search for one query, window side S = 2R+1 known at compile time.
// Candidates are visited row-major (c' = j*S + i), two per step; `best` is the running key.
template <int F8, int R, bool CHECK>
__device__ __forceinline__ void dilation_rows(const uint16_t* __restrict__ D11b, const uint32_t* q,
                                              int H, int W, int d, int us, int vs, int& best) {
  constexpr int S = 2 * R + 1, FB = F8 * 16;  // descriptor bytes
  const char* base = reinterpret_cast<const char*>(D11b);
  // byte offsets in wrap-around unsigned arithmetic: every offset that is dereferenced
  // lies inside the image (< 2^31, checked at the entry point)
  const unsigned cs = (unsigned)d * FB;               // step between window columns
  const unsigned rs = (unsigned)d * (unsigned)W * FB; // step between window rows
  const unsigned org = ((unsigned)vs * (unsigned)W + (unsigned)us) * FB;
  unsigned um = 0, vm = 0;        // CHECK: validity bit per window column / row
  if (CHECK) {
#pragma unroll
    for (int i = 0; i < S; i++) {
      um |= (unsigned)((us + i * d) >= 0 && (us + i * d) < W) << i;
      vm |= (unsigned)((vs + i * d) >= 0 && (vs + i * d) < H) << i;
    }
  }
  auto eval2 = [&](int ja, int ia, int jb, int ib, bool have_b) {
    // scan index of (i, j) is i*S + j (u outer, v inner: matching_kernels.cu:52-53)
    const uint32_t tie_a = 0xFFFEu - (uint32_t)(ia * S + ja);
    const uint32_t tie_b = 0xFFFEu - (uint32_t)(ib * S + jb);
    unsigned off_a = org + (unsigned)ja * rs + (unsigned)ia * cs;
    unsigned off_b = org + (unsigned)jb * rs + (unsigned)ib * cs;
    bool ok_a = true, ok_b = have_b;
    if (CHECK) {
      ok_a = ((um >> ia) & (vm >> ja) & 1u) != 0;
      ok_b = have_b && ((um >> ib) & (vm >> jb) & 1u) != 0;
      if (!ok_a && !ok_b) return;
      if (!ok_a) off_a = ok_b ? off_b : 0;
      if (!ok_b) off_b = off_a;
    } else if (!have_b) {
      off_b = off_a;
    }
    const uint32_t s2 = score_pair<F8>(reinterpret_cast<const uint4*>(base + off_a),
                                       reinterpret_cast<const uint4*>(base + off_b), q);
    const int ka = ok_a ? key_lo(s2, tie_a) : kKeyInvalid;
    const int kb = ok_b ? key_hi(s2, tie_b) : kKeyInvalid;
    best = max(best, max(ka, kb));
  };
  // row pairs (2 rows = 2S candidates = S pairs), then the last row (S odd)
#pragma unroll 1
  for (int jp = 0; jp < S / 2; jp++) {
    const int j0 = 2 * jp;
#pragma unroll
    for (int c = 0; c < 2 * S; c += 2) {
      const int ja = c / S, ia = c % S, jb = (c + 1) / S, ib = (c + 1) % S;  // compile time
      eval2(j0 + ja, ia, j0 + jb, ib, true);
    }
  }
#pragma unroll
  for (int i = 0; i < S; i += 2) eval2(S - 1, i, S - 1, (i + 1 < S) ? i + 1 : i, i + 1 < S);
}

// Interior window of side 7 at a compile-time dilation D (the hot case).  Candidates are taken in
// point-symmetric pairs about the centre: A = (row j, col i), B = (row 6-j, col 6-i), j = 0..2,
// then the three pairs of the centre row; the centre itself is the previous winner (or a
// candidate that already lost), so it is only scored at the first dilation.  Column offsets
// i*D*FB are instruction immediates; the two row pointers move by one 64-bit add per row.
// Keys are formed with column-only ties CT[i] = 0xFFFE - 7i held in registers; the row part of
// the scan index (i*7 + j) is subtracted once per row from the row maximum.
__device__ __forceinline__ uint4 ld_cand(const char* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

template <int F8>
__device__ __forceinline__ uint32_t score_pair_at(const char* pa, const char* pb, int offa, int offb,
                                                  const uint32_t* q) {
  uint4 a[F8], b[F8];
#pragma unroll
  for (int w = 0; w < F8; w++) {
    a[w] = ld_cand(pa + offa + 16 * w);
    b[w] = ld_cand(pb + offb + 16 * w);
  }
  __half2 acc;
#pragma unroll
  for (int w = 0; w < F8; w++) {
    const uint32_t aw[4] = {a[w].x, a[w].y, a[w].z, a[w].w};
    const uint32_t bw[4] = {b[w].x, b[w].y, b[w].z, b[w].w};
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const uint32_t lo = __byte_perm(aw[t], bw[t], 0x5410);
      const uint32_t hi = __byte_perm(aw[t], bw[t], 0x7632);
      const __half2 qq = *reinterpret_cast<const __half2*>(&q[w * 4 + t]);
      const __half2 plo = __hmul2_rn(*reinterpret_cast<const __half2*>(&lo), __low2half2(qq));
      // 0 + p == p up to the sign of a zero, which cannot change a non-zero sum or any comparison
      acc = (w == 0 && t == 0) ? plo : __hadd2_rn(acc, plo);
      acc = __hadd2_rn(acc, __hmul2_rn(*reinterpret_cast<const __half2*>(&hi), __high2half2(qq)));
    }
  }
  return *reinterpret_cast<uint32_t*>(&acc);
}

template <int F8, int D>
__device__ __forceinline__ void dilation_sym(const char* centre, const uint32_t* q, int W, bool first,
                                             const uint32_t (&CT)[7], int& best) {
  constexpr int FB = F8 * 16, CS = D * FB;
  const long long rs = (long long)D * W * FB;
  const char* pa = centre - 3 * rs - 3 * CS;  // (row j, col 0)
  const char* pb = centre + 3 * rs + 3 * CS;  // (row 6-j, col 6)
#pragma unroll 1
  for (int j = 0; j < 3; j++) {
    int ra = kKeyInvalid + 64, rb = kKeyInvalid + 64;
#pragma unroll
    for (int i = 0; i < 7; i++) {
      const uint32_t s2 = score_pair_at<F8>(pa, pb, i * CS, -i * CS, q);
      ra = max(ra, key_lo(s2, CT[i]));
      rb = max(rb, key_hi(s2, CT[6 - i]));
    }
    best = max(best, max(ra - j, rb - 6 + j));  // scan index = 7*col + row
    pa += rs;
    pb -= rs;
  }
  {  // centre row: pa is (row 3, col 0)
    int rc = kKeyInvalid + 64;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const uint32_t s2 = score_pair_at<F8>(pa, pa, i * CS, (6 - i) * CS, q);
      rc = max(rc, max(key_lo(s2, CT[i]), key_hi(s2, CT[6 - i])));
    }
    best = max(best, rc - 3);
  }
  if (first) {
    const uint32_t s2 = score_pair_at<F8>(pa, pa, 3 * CS, 3 * CS, q);
    best = max(best, key_lo(s2, CT[3]) - 3);
  }
}

// Generic window side (radius != 3): same key logic, runtime loops, scan order.
template <int F8>
__device__ __forceinline__ void dilation_generic(const uint16_t* __restrict__ D11b, const uint32_t* q,
                                                 int H, int W, int radius, int d, int us, int vs,
                                                 int& best) {
  constexpr int FB = F8 * 16;
  const char* base = reinterpret_cast<const char*>(D11b);
  const int side = 2 * radius + 1;
  for (int i = 0; i < side; i++) {
    const int u = us + i * d;
    if (u < 0 || u >= W) continue;
    for (int j = 0; j < side; j += 2) {
      const int va = vs + j * d, vb = va + d;
      const bool ok_a = va >= 0 && va < H;
      const bool ok_b = (j + 1 < side) && vb >= 0 && vb < H;
      if (!ok_a && !ok_b) continue;
      unsigned off_a = ((unsigned)va * W + u) * FB, off_b = ((unsigned)vb * W + u) * FB;
      if (!ok_a) off_a = off_b;
      if (!ok_b) off_b = off_a;
      const uint32_t s2 = score_pair<F8>(reinterpret_cast<const uint4*>(base + off_a),
                                         reinterpret_cast<const uint4*>(base + off_b), q);
      const uint32_t tie = 0xFFFEu - (uint32_t)(i * side + j);
      const int ka = ok_a ? key_lo(s2, tie) : kKeyInvalid;
      const int kb = ok_b ? key_hi(s2, tie - 1u) : kKeyInvalid;
      best = max(best, max(ka, kb));
    }
  }
}

// Whole search for one query: dilation_max..1, each re-centred on the previous winner
// (matching_kernels.cu:47-77).  q: packed fp16 query; (u0,v0) in/out.
template <int F8>
__device__ __forceinline__ void refine_query(const uint16_t* __restrict__ D11b, const uint32_t* q,
                                             int H, int W, int radius, int dilation_max, int& u0,
                                             int& v0) {
  int best = 0x0000FFFF;  // score +0 (c10::Half() == 0), carried
  const int side = 2 * radius + 1;
  uint32_t CT[7];
#pragma unroll
  for (int i = 0; i < 7; i++) CT[i] = kColumnTie[i];  // from constant memory: stays in registers
  for (int d = dilation_max; d > 0; d--) {
    const int us = u0 - radius * d, vs = v0 - radius * d;
    best |= 0xFFFF;  // the running maximum now predates every candidate of this dilation
    if (radius == 3) {
      const bool interior = us >= 0 && us + 6 * d < W && vs >= 0 && vs + 6 * d < H;
#ifdef M3R_REFINE_OLD
      if (false) {
#else
      if (interior && d <= 5) {
#endif
        const char* centre = reinterpret_cast<const char*>(D11b) +
                             (size_t)((unsigned)v0 * (unsigned)W + (unsigned)u0) * (F8 * 16);
        const bool first = d == dilation_max;
        switch (d) {
          case 5: dilation_sym<F8, 5>(centre, q, W, first, CT, best); break;
          case 4: dilation_sym<F8, 4>(centre, q, W, first, CT, best); break;
          case 3: dilation_sym<F8, 3>(centre, q, W, first, CT, best); break;
          case 2: dilation_sym<F8, 2>(centre, q, W, first, CT, best); break;
          default: dilation_sym<F8, 1>(centre, q, W, first, CT, best); break;
        }
      } else if (interior) {
        dilation_rows<F8, 3, false>(D11b, q, H, W, d, us, vs, best);
      } else {
        dilation_rows<F8, 3, true>(D11b, q, H, W, d, us, vs, best);
      }
    } else {
      dilation_generic<F8>(D11b, q, H, W, radius, d, us, vs, best);
    }
    const int tie = best & 0xFFFF;
    if (tie != 0xFFFF) {
      const int sidx = 0xFFFE - tie;
      const int i = sidx / side, j = sidx - i * side;
      u0 = us + i * d; v0 = vs + j * d;
    }
  }
}

// Query index of this thread.  When the queries form the H x W grid (n == H*W) CTAs are
// 32x8 pixel tiles so that neighbouring queries' windows share L1 lines; otherwise 1-D.
__device__ __forceinline__ int query_index(int n, int W, int tiles_x, bool tiled) {
  if (!tiled) {
    const int k = blockIdx.x * kRefineThreads + threadIdx.x;
    return k < n ? k : -1;
  }
  const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
  const int x = tx * kRefTX + (threadIdx.x & (kRefTX - 1));
  const int y = ty * kRefTY + (threadIdx.x / kRefTX);
  return (x < W && y * W + x < n) ? y * W + x : -1;
}

// p1 coordinates outside int32 cannot be inside the image; clamp them far outside so that
// 32-bit window arithmetic stays exact (the reference truncates long -> int, H3).
__device__ __forceinline__ int clamp_coord(long long c) {
  return c < -(1 << 20) ? -(1 << 20) : (c > (1 << 20) ? (1 << 20) : (int)c);
}

// Plain form: D21 already fp16, output (u,v) pairs.
template <int F8>
__global__ void __launch_bounds__(kRefineThreads, 1)
refine_f16_kernel(const uint16_t* __restrict__ D11, const uint16_t* __restrict__ D21,
                  const int64_t* __restrict__ p1, int H, int W, int n, int radius,
                  int dilation_max, int tiles_x, int tiled, int64_t* __restrict__ p1_new) {
  constexpr int F = F8 * 8;
  const int k = query_index(n, W, tiles_x, tiled != 0);
  if (k < 0) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  uint32_t q[F / 2];
  const uint4* qp = reinterpret_cast<const uint4*>(D21 + bk * F);
#pragma unroll
  for (int w = 0; w < F8; w++) {
    const uint4 qv = ld_stream_u4(qp + w);
    q[w * 4 + 0] = qv.x; q[w * 4 + 1] = qv.y; q[w * 4 + 2] = qv.z; q[w * 4 + 3] = qv.w;
  }
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  int u0 = clamp_coord(p.x), v0 = clamp_coord(p.y);
  const int u_in = u0, v_in = v0;
  refine_query<F8>(D11 + (size_t)blockIdx.y * H * W * F, q, H, W, radius, dilation_max, u0, v0);
  longlong2 o;  // unchanged coordinates are returned as given (even outside int32)
  o.x = (u0 == u_in) ? p.x : (long long)u0;
  o.y = (v0 == v_in) ? p.y : (long long)v0;
  reinterpret_cast<longlong2*>(p1_new)[bk] = o;
}

// Fused form: D21 fp32 rounded to fp16 on the fly (== .half()), output linear index.
template <int F8>
__global__ void __launch_bounds__(kRefineThreads, 1)
refine_fused_kernel(const uint16_t* __restrict__ D11h, const float* __restrict__ D21,
                    const int64_t* __restrict__ p1, int H, int W, int radius,
                    int dilation_max, int tiles_x, int64_t* __restrict__ idx_out) {
  constexpr int F = F8 * 8;
  const int n = H * W;
  const int k = query_index(n, W, tiles_x, true);
  if (k < 0) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  uint32_t q[F / 2];
  const float4* qp = reinterpret_cast<const float4*>(D21 + bk * F);
#pragma unroll
  for (int w = 0; w < F / 4; w++) {
    const float4 qv = ld_stream_f4(qp + w);
    const __half2 h0 = __floats2half2_rn(qv.x, qv.y), h1 = __floats2half2_rn(qv.z, qv.w);
    q[w * 2 + 0] = *reinterpret_cast<const uint32_t*>(&h0);
    q[w * 2 + 1] = *reinterpret_cast<const uint32_t*>(&h1);
  }
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  int u0 = clamp_coord(p.x), v0 = clamp_coord(p.y);
  const int u_in = u0, v_in = v0;
  if (radius > 0)
    refine_query<F8>(D11h + (size_t)blockIdx.y * n * F, q, H, W, radius, dilation_max, u0, v0);
  const long long uo = (u0 == u_in) ? p.x : (long long)u0;
  const long long vo = (v0 == v_in) ? p.y : (long long)v0;
  idx_out[bk] = uo + (long long)W * vo;  // pixel_to_lin, matching.py:13-15
}

// <float> instantiation of the reference: sequential fp32 accumulate (FFMA), max from FLT_MIN.
__global__ void __launch_bounds__(128)
refine_f32_kernel(const float* __restrict__ D11, const float* __restrict__ D21,
                  const int64_t* __restrict__ p1, int H, int W, int n, int F, int radius,
                  int dilation_max, int64_t* __restrict__ p1_new) {
  const int k = blockIdx.x * 128 + threadIdx.x;
  if (k >= n) return;
  const size_t bk = (size_t)blockIdx.y * n + k;
  const float* q = D21 + bk * F;
  const float* D11b = D11 + (size_t)blockIdx.y * H * W * F;
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + bk);
  long long u0 = p.x, v0 = p.y, u_new = p.x, v_new = p.y;
  float max_score = 1.17549435e-38f;  // numeric_limits<float>::min()
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i, v = v0 - rd + j;
        if ((int)v >= 0 && (int)v < H && (int)u >= 0 && (int)u < W) {
          const float* c = D11b + ((size_t)(int)v * W + (int)u) * F;
          float score = 0.f;
          for (int t = 0; t < F; t++) score = fmaf(__ldg(q + t), __ldg(c + t), score);
          if (score > max_score) { max_score = score; u_new = u; v_new = v; }
        }
      }
    }
    u0 = u_new; v0 = v_new;
  }
  longlong2 o; o.x = u_new; o.y = v_new;
  reinterpret_cast<longlong2*>(p1_new)[bk] = o;
}

}  // namespace m3r

using namespace m3r;

#define M3R_DISPATCH_F8(F8v, CALL)                   \
  switch (F8v) {                                     \
    case 1: { constexpr int F8 = 1; CALL; } break;   \
    case 2: { constexpr int F8 = 2; CALL; } break;   \
    case 3: { constexpr int F8 = 3; CALL; } break;   \
    case 4: { constexpr int F8 = 4; CALL; } break;   \
    default: return M3R_EINVAL;                      \
  }

static bool refine_args_ok(int b, int H, int W, int F, int radius) {
  if (b < 0 || H <= 0 || W <= 0 || F <= 0 || (F % 8) != 0 || F > 32 || radius < 0 || radius > 100 ||
      b > 65535)
    return false;
  return (long long)H * W * F * 2 < (1LL << 31);  // 32-bit byte offsets inside one image
}

extern "C" int m3r_refine_matches_f16(const uint16_t* D11, const uint16_t* D21,
                                      const int64_t* p1, int b, int H, int W, int n, int F,
                                      int radius, int dilation_max, int64_t* p1_new,
                                      m3r_stream_t stream) {
  M3R_NVTX("m3r_refine_matches_f16");
  if (!refine_args_ok(b, H, W, F, radius) || n < 0) return M3R_EINVAL;
  if (b == 0 || n == 0) return M3R_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_EINVAL;
  const int tiled = (n == H * W) ? 1 : 0;
  const int tiles_x = (W + kRefTX - 1) / kRefTX;
  const int ctas = tiled ? tiles_x * ((H + kRefTY - 1) / kRefTY)
                         : (n + kRefineThreads - 1) / kRefineThreads;
  dim3 grid(ctas, b);
  M3R_DISPATCH_F8(F / 8, (refine_f16_kernel<F8><<<grid, kRefineThreads, 0, as_stream(stream)>>>(
                             D11, D21, p1, H, W, n, radius, dilation_max, tiles_x, tiled, p1_new)));
  return after_launch();
}

extern "C" int m3r_match_refine_fused(const uint16_t* D11h, const float* D21,
                                      const int64_t* p1, int b, int H, int W, int F, int radius,
                                      int dilation_max, int64_t* idx_1_to_2,
                                      m3r_stream_t stream) {
  M3R_NVTX("m3r_match_refine_fused");
  if (!refine_args_ok(b, H, W, F, radius)) return M3R_EINVAL;
  if (b == 0) return M3R_OK;
  if (!D11h || !D21 || !p1 || !idx_1_to_2) return M3R_EINVAL;
  const int tiles_x = (W + kRefTX - 1) / kRefTX;
  dim3 grid(tiles_x * ((H + kRefTY - 1) / kRefTY), b);
  M3R_DISPATCH_F8(F / 8, (refine_fused_kernel<F8><<<grid, kRefineThreads, 0, as_stream(stream)>>>(
                             D11h, D21, p1, H, W, radius, dilation_max, tiles_x, idx_1_to_2)));
  return after_launch();
}

extern "C" int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                                      int b, int H, int W, int n, int F, int radius,
                                      int dilation_max, int64_t* p1_new, m3r_stream_t stream) {
  M3R_NVTX("m3r_refine_matches_f32");
  if (b < 0 || n < 0 || H <= 0 || W <= 0 || F <= 0 || radius < 0 || b > 65535) return M3R_EINVAL;
  if (b == 0 || n == 0) return M3R_OK;
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_EINVAL;
  dim3 grid((n + 127) / 128, b);
  refine_f32_kernel<<<grid, 128, 0, as_stream(stream)>>>(D11, D21, p1, H, W, n, F, radius,
                                                         dilation_max, p1_new);
  return after_launch();
}
