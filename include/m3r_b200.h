/*
 * m3r_b200.h — C ABI of libm3r_b200.so: B200 (sm_100a) dense matching and Sim3
 * Gauss-Newton registration, the drop-in for MASt3R-SLAM's `mast3r_slam_backends`
 * extension (reference: mast3r_slam/backend/include/gn.h, src/gn.cpp:116-122).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to a dense, contiguous, row-major array
 *     (the reference asserts contiguity only: gn.h:5, gn.cpp:14-21,92-94,108-110);
 *   - `stream` is a cudaStream_t passed as an opaque pointer; all work is enqueued
 *     on it and the call returns without synchronising (the reference launches on
 *     the legacy default stream and blocks: matching_kernels.cu:104,303,
 *     gn_kernels.cu:1183-1220);
 *   - the library never allocates or frees device memory: outputs and workspace are
 *     caller-owned (the reference allocates with torch::zeros per call:
 *     matching_kernels.cu:99-101,295-301; gn_kernels.cu:1173-1174);
 *   - return value: 0 on success, a negative M3R_E* code for usage/launch errors.
 *     Numerical failure (non-SPD system) is NOT an error: like the reference
 *     (gn_kernels.cu:142-150) the step is zero and the poses are left unchanged;
 *   - bool tensors are one byte per element (torch.bool).
 *
 * No torch types appear here; the Python package `mast3r_slam_backends` binds these
 * with ctypes and owns all allocations.
 */
#ifndef M3R_B200_H
#define M3R_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* m3r_stream_t; /* cudaStream_t */

enum {
  M3R_OK = 0,
  M3R_EINVAL = -1,    /* bad shape / null pointer / unsupported parameter */
  M3R_ELAUNCH = -2,   /* CUDA launch or runtime error (see m3r_last_cuda_error) */
  M3R_EWORKSPACE = -3, /* workspace too small */
  M3R_ETOOLARGE = -4   /* size beyond the fixed maximum of a *_noalloc entry point */
};

/* Residual family of a Gauss-Newton accumulation pass. */
enum {
  M3R_GN_RAYS = 0,  /* ray_align_kernel   gn_kernels.cu:813-1138  (4 residuals / point) */
  M3R_GN_CALIB = 1, /* calib_proj_kernel  gn_kernels.cu:1231-1543 (3 residuals / point) */
  M3R_GN_POINTS = 2 /* point_align_kernel gn_kernels.cu:455-723   (3 residuals / point) */
};

int m3r_version(void);
const char* m3r_error_string(int code);
/* cudaError_t of the last failing runtime call on this thread (0 if none). */
int m3r_last_cuda_error(void);
/* Number of kernels this library has launched in this process (bench `gpu_launches`). */
unsigned long long m3r_launch_count(void);

/* ------------------------------------------------------------------------- *
 * Dense matching
 * ------------------------------------------------------------------------- */

/* iter_proj — replaces mast3r_slam_backends.iter_proj (gn.cpp:84-99,
 * matching_kernels.cu:119-315).  Per query pixel: <= max_iter Levenberg-Marquardt
 * steps on (u,v) minimising |normalize(bilinear(ray_img)(u,v)) - target_ray|^2.
 *   rays_img_with_grad (b,H,W,9) f32   [ray(3), d ray/du (3), d ray/dv (3)]
 *   pts_3d_norm        (b,n,3)   f32   unit target rays
 *   p_init             (b,n,2)   f32   initial (u,v)
 *   p_new              (b,n,2)   f32   out
 *   converged          (b,n)     u8    out (torch.bool)
 * Unlike the reference (no bounds guard, matching_kernels.cu:131) any n is safe. */
int m3r_iter_proj(const float* rays_img_with_grad, const float* pts_3d_norm,
                  const float* p_init, int b, int H, int W, int n, int max_iter,
                  float lambda_init, float cost_thresh, float* p_new,
                  uint8_t* converged, m3r_stream_t stream);

/* refine_matches, fp16 descriptors — replaces mast3r_slam_backends.refine_matches
 * for the dtype its only caller uses (matching.py:79-85; gn.cpp:101-114,
 * matching_kernels.cu:25-116).  Dilated (2*radius+1)^2 window argmax of the
 * descriptor dot product, dilation_max..1, re-centred per dilation; scores are
 * accumulated in the reference's sequential fp16-rounded order and compared with
 * strict `>` starting from fp16 zero, so p1_new is bit-exact.
 *   D11 (b,H,W,F) f16 ; D21 (b,n,F) f16 ; p1 (b,n,2) i64 (u,v) ; p1_new (b,n,2) i64
 * F must be a multiple of 8 and <= 32 (MASt3R: 24). */
int m3r_refine_matches_f16(const uint16_t* D11, const uint16_t* D21,
                           const int64_t* p1, int b, int H, int W, int n, int F,
                           int radius, int dilation_max, int64_t* p1_new,
                           m3r_stream_t stream);

/* refine_matches, fp32 descriptors (the reference's <float> instantiation:
 * sequential `score += a*b` in fp32, initial max = FLT_MIN). */
int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                           int b, int H, int W, int n, int F, int radius,
                           int dilation_max, int64_t* p1_new, m3r_stream_t stream);

/* Fused caller-side glue of matching.match_iterative_proj (matching.py:25-90,
 * image.py:5-38): everything around the two kernels above in three launches.
 *
 * m3r_match_prep: X11 (b,H,W,3) f32 -> rays_img_with_grad (b,H,W,9) f32
 *   (F.normalize eps 1e-12, reflect-padded 3x3 [-3,0,3;-10,0,10;-3,0,3]/32 gradients),
 *   and D11 (b,H,W,F) f32 -> D11h (b,H,W,F) f16 (round-to-nearest-even, == .half()).
 *   Either half may be skipped by passing NULL for its output. */
int m3r_match_prep(const float* X11, const float* D11, int b, int H, int W, int F,
                   float* rays_img_with_grad, uint16_t* D11h, m3r_stream_t stream);

/* m3r_match_iter_proj_fused: iter_proj with the caller's pre/post-processing fused:
 *   target rays = normalize(X21) computed on the fly; p_init = (idx%W, idx/W) from
 *   idx_init (b,n) i64 or the identity map when idx_init == NULL; after the LM loop
 *   p1 = trunc(p_new) and valid = converged & (|X11[p1] - X21| < dist_thresh).
 *   Outputs: p1 (b,n,2) i64 and valid (b,n) u8; p_new (b,n,2) f32 optional (NULL ok). */
int m3r_match_iter_proj_fused(const float* rays_img_with_grad, const float* X11,
                              const float* X21, const int64_t* idx_init, int b, int H,
                              int W, int max_iter, float lambda_init,
                              float cost_thresh, float dist_thresh, int64_t* p1,
                              uint8_t* valid, float* p_new, m3r_stream_t stream);

/* image.img_gradient (mast3r_slam/image.py:5-38) on its own: reflect-padded depthwise 3x3
 * gradients (1/32)[[-3,0,3],[-10,0,10],[-3,0,3]] and its transpose of img (planes,H,W) f32,
 * planes = batch*channels, into gx, gy of the same shape.  H, W >= 2; planes <= 65535.
 * (match_prep computes the same taps fused with the normalisation; this entry serves callers
 * of img_gradient itself.) */
int m3r_img_gradient(const float* img, int planes, int H, int W, float* gx, float* gy,
                     m3r_stream_t stream);

/* The same two steps over the PRIVATE float4-packed ray image of the fused pipeline:
 * rays_packed (b,H,W,12) f32, 16-byte aligned = three 16-B words per pixel
 * [ray.xyz, 0 | d(ray)/du.xyz, 0 | d(ray)/dv.xyz, 0], so that a bilinear sample of one group is
 * four 128-bit gathers instead of twelve scalar ones.  Values and arithmetic are those of the
 * 9-channel form (same results bit for bit); only the layout differs. */
int m3r_match_prep_packed(const float* X11, const float* D11, int b, int H, int W, int F,
                          float* rays_packed, uint16_t* D11h, m3r_stream_t stream);
int m3r_match_iter_proj_packed(const float* rays_packed, const float* X11, const float* X21,
                               const int64_t* idx_init, int b, int H, int W, int max_iter,
                               float lambda_init, float cost_thresh, float dist_thresh,
                               int64_t* p1, uint8_t* valid, float* p_new, m3r_stream_t stream);

/* m3r_match_refine_fused: refine_matches reading D21 as f32 (rounded to f16 on the
 * fly, == .half()) against the f16 copy of D11 made by m3r_match_prep, and writing
 * the linear index idx_1_to_2 = u + W*v (pixel_to_lin, matching.py:13-15).
 * radius == 0 skips the search (idx from p1 directly, matching.py:78). */
int m3r_match_refine_fused(const uint16_t* D11h, const float* D21, const int64_t* p1,
                           int b, int H, int W, int F, int radius, int dilation_max,
                           int64_t* idx_1_to_2, m3r_stream_t stream);

/* ------------------------------------------------------------------------- *
 * Sim3 retraction
 * ------------------------------------------------------------------------- */

/* pose_retr — T_k <- Exp(dx_{k-num_fix}) * T_k for k >= num_fix
 * (pose_retr_kernel / retrSim3 / expSim3 / expSO3, gn_kernels.cu:299-453).
 *   poses (N,8) f32 [t(3), q(xyzw), s] in place ; dx (N-num_fix,7) f32 [tau, phi, sigma] */
int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix,
                  m3r_stream_t stream);

/* ------------------------------------------------------------------------- *
 * Sim3 factor-graph Gauss-Newton
 * ------------------------------------------------------------------------- */

typedef struct m3r_gn_params {
  int mode;           /* M3R_GN_RAYS | M3R_GN_CALIB | M3R_GN_POINTS */
  float sigma_a;      /* rays: sigma_ray   | calib: sigma_pixel | points: sigma_point */
  float sigma_b;      /* rays: sigma_dist  | calib: sigma_depth | points: unused      */
  float C_thresh;
  float Q_thresh;
  /* calib only (gn.h:74-87) */
  float fx, fy, cx, cy; /* K[0][0], K[1][1], K[0][2], K[1][2]  (gn_kernels.cu:1274-1279) */
  int height, width, pixel_border;
  float z_eps;
  /* calib only: when non-NULL, a DEVICE pointer to K (3,3) f32 row-major; the kernels read
   * fx, fy, cx, cy from it (as the reference does, gn_kernels.cu:1274-1279) and the four host
   * fields above are ignored — the caller never copies K to the host. */
  const float* K_dev;
} m3r_gn_params;

/* One correspondence of a directed edge in the PACKED layout FactorGraph stores
 * (SURVEY 8(f)-3): 8 bytes instead of the reference's idx i64 + valid u8 + Q f32 = 13.
 *   idx : match index into the other keyframe's pointmap (kept as given, also when invalid)
 *   q   : valid_match ? Q : -Q   (Q = sqrt(Q_src[idx] * Q_dst) >= 0, so the sign bit IS
 *         valid_match; -0.0f encodes an invalid match with Q == 0)
 * m3r_edge_pack / m3r_edge_unpack convert losslessly to and from the reference's three tensors. */
typedef struct m3r_packed_match {
  int32_t idx;
  float q;
} m3r_packed_match;

/* Bytes of caller-provided scratch needed by the m3r_gn_* calls for a graph with
 * num_poses keyframes, num_edges directed edges (of which at most local_edges are
 * accumulated by this process) and n points per keyframe. */
size_t m3r_gn_workspace_bytes(int num_poses, int num_edges, int local_edges, int n);

/* One call == mast3r_slam_backends.gauss_newton_{rays,calib,points}
 * (gn.cpp:3-82; host loops gn_kernels.cu:725-811,1140-1228,1546-1638).
 *   Twc (N,8) f32 in/out ; Xs (N,n,3) f32 ; Cs (N,n,1) f32 ; ii,jj (M) i64 keyframe ids ;
 *   idx_ii2jj (M,n) i64 ; valid_match (M,n,1) u8 ; Q (M,n,1) f32 ;
 *   dx (N-1,7) f32 out: the last iteration's update (zeros if max_iter == 0).
 * Rows of Twc/Xs/Cs follow the sorted unique keyframe ids of cat(ii,jj); the first
 * row is the gauge (num_fix = 1, gn_kernels.cu:1157).  The whole loop — index setup,
 * accumulation, assembly, fp64 Cholesky, retraction, |dx| < delta_thresh early exit —
 * runs on `stream` with no host synchronisation. */
int m3r_gauss_newton(const m3r_gn_params* prm, float* Twc, const float* Xs,
                     const float* Cs, const int64_t* ii, const int64_t* jj,
                     const int64_t* idx_ii2jj, const uint8_t* valid_match,
                     const float* Q, int num_poses, int n, int num_edges,
                     int max_iter, float delta_thresh, float* dx, void* workspace,
                     size_t workspace_bytes, m3r_stream_t stream);

/* Stepwise form of the same loop, for edge-sharded multi-GPU solves (each process
 * accumulates its slice of the edge list, the per-edge blocks are exchanged with an
 * NCCL all-gather, every process then runs the identical replicated solve) and for
 * block-level parity tests.
 *
 * m3r_gn_prepare: once per solve.  Maps keyframe ids to pose rows
 *   (get_unique_kf_idx/create_inds, gn_kernels.cu:161-170) for ALL num_edges edges,
 *   builds the pose->edge adjacency used by the deterministic assembly, packs
 *   (Xs,Cs) as float4 and resets the convergence flag.
 * m3r_gn_accumulate: one pass over directed edges [edge_begin, edge_begin+edge_count).
 *   The per-edge stream pointers (idx_ii2jj, valid_match, Q) address the LOCAL slice,
 *   i.e. row 0 is edge `edge_begin`.  Writes blocks[e] for those edges, e global:
 *   56 doubles per edge = S (7x7, row-major) followed by v (7), where in the
 *   reference's layout (gn_kernels.cu:1096-1137)
 *     Hs[0][e] = Hs[3][e] = S, Hs[1][e] = Hs[2][e] = -S, gs[1][e] = v, gs[0][e] = -v.
 * m3r_gn_solve_step: assemble all num_edges blocks, factorise, dx = -x, retract Twc,
 *   set the convergence flag if |dx| < delta_thresh.  Once the flag is set,
 *   m3r_gn_accumulate and m3r_gn_solve_step become no-ops (the reference's `break`). */
int m3r_gn_prepare(const float* Xs, const float* Cs, const int64_t* ii,
                   const int64_t* jj, int num_poses, int n, int num_edges,
                   int local_edges, void* workspace, size_t workspace_bytes,
                   m3r_stream_t stream);
int m3r_gn_accumulate(const m3r_gn_params* prm, const float* Twc,
                      const int64_t* idx_ii2jj, const uint8_t* valid_match,
                      const float* Q, int num_poses, int n, int num_edges,
                      int edge_begin, int edge_count, double* blocks, void* workspace,
                      size_t workspace_bytes, m3r_stream_t stream);
int m3r_gn_solve_step(float* Twc, const double* blocks, int num_poses, int num_edges,
                      float delta_thresh, float* dx, void* workspace,
                      size_t workspace_bytes, m3r_stream_t stream);

/* The same loop over PACKED edge records held in up to two row blocks: edge e < split is row e of
 * pk0, edge e >= split row e - split of pk1 (each row n records).  FactorGraph keeps the forward
 * and the reverse direction of its E undirected edges in two blocks, so
 *   m3r_gauss_newton_packed(.., ii, jj, two_way = 1, fwd, rev, split = E, .., num_edges = 2E, ..)
 * is solve_GN_rays / solve_GN_calib (global_opt.py:112-213) WITHOUT prep_two_way_edges'
 * five concatenations (global_opt.py:104-110): with two_way != 0, ii / jj hold the E undirected
 * edges and the directed list [ii;jj] -> [jj;ii] is formed inside the index kernel.
 * 24 B streamed + 16 B gathered = 40 B per (directed edge, point) instead of 45; arithmetic and
 * reduction order equal the reference-layout path, so blocks, dx and poses are bit-identical.
 * m3r_gn_accumulate_packed: as m3r_gn_accumulate with pk0 / pk1 / split addressing the LOCAL slice
 * (row 0 = edge `edge_begin`); peers != NULL selects the peer-memory exchange (then `blocks` is
 * ignored, as in m3r_gn_accumulate_p2p). */
int m3r_gn_prepare_two_way(const float* Xs, const float* Cs, const int64_t* ii, const int64_t* jj,
                           int num_poses, int n, int num_edges, int local_edges, void* workspace,
                           size_t workspace_bytes, m3r_stream_t stream);
struct m3r_gn_peers;
int m3r_gn_accumulate_packed(const m3r_gn_params* prm, const float* Twc,
                             const m3r_packed_match* pk0, const m3r_packed_match* pk1, int split,
                             int num_poses, int n, int num_edges, int edge_begin, int edge_count,
                             double* blocks, const struct m3r_gn_peers* peers, void* workspace,
                             size_t workspace_bytes, m3r_stream_t stream);
int m3r_gauss_newton_packed(const m3r_gn_params* prm, float* Twc, const float* Xs, const float* Cs,
                            const int64_t* ii, const int64_t* jj, int two_way,
                            const m3r_packed_match* pk0, const m3r_packed_match* pk1, int split,
                            int num_poses, int n, int num_edges, int max_iter, float delta_thresh,
                            float* dx, void* workspace, size_t workspace_bytes, m3r_stream_t stream);
/* Edge-sharded solve with the block exchange FUSED into the kernels over NVLink peer memory
 * (no collective call): every process maps every other process's `blocks` buffer and a small
 * int32 flag array (CUDA peer / symmetric memory; torch.distributed._symmetric_memory in the
 * Python binding).  m3r_gn_accumulate_p2p stores each finished per-edge block straight into
 * ALL processes' buffers (plain stores to the peer pointers) and then raises flags[rank] = epoch
 * in every process; m3r_gn_solve_step_p2p waits, on the device, until all `world` flags of its
 * own array have reached `epoch` before it assembles.  `epoch` must increase by one per
 * iteration across the whole life of the buffers (the caller counts).  The wait is bounded
 * (about 10 s): on expiry the step becomes zero like a failed factorisation. */
#define M3R_MAX_PEERS 16
typedef struct m3r_gn_peers {
  double* blocks[M3R_MAX_PEERS]; /* blocks[r]: rank r's (num_edges*56) buffer as mapped HERE  */
  int* flags[M3R_MAX_PEERS];     /* flags[r]: rank r's int32[world] flag array as mapped HERE */
  int world, rank;
  int epoch;
} m3r_gn_peers;
int m3r_gn_accumulate_p2p(const m3r_gn_params* prm, const float* Twc,
                          const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q,
                          int num_poses, int n, int num_edges, int edge_begin, int edge_count,
                          const m3r_gn_peers* peers, void* workspace, size_t workspace_bytes,
                          m3r_stream_t stream);
int m3r_gn_solve_step_p2p(float* Twc, const m3r_gn_peers* peers, int num_poses, int num_edges,
                          float delta_thresh, float* dx, void* workspace,
                          size_t workspace_bytes, m3r_stream_t stream);
/* Device pointer (inside workspace) of the int32 convergence flag: 0 running, 1 done. */
int* m3r_gn_done_flag(void* workspace);
/* Device pointer (inside workspace) of the solve's status words, int32[4]:
 *   [0] done (|dx| < delta_thresh reached), [1] the last factorisation failed (zero step),
 *   [2] iterations executed, [3] peer_timeout: a peer's blocks did not arrive within ~10 s in a
 *   peer-memory exchange — that step was zero, `done` was NOT set and every later launch of the
 *   solve was a no-op; the caller must treat the solve as failed (the Python binding raises). */
int* m3r_gn_status(void* workspace);
/* Device pointer (inside workspace) of the per-edge block buffer (num_edges*56 doubles). */
double* m3r_gn_blocks(void* workspace, int num_poses, int num_edges, int local_edges, int n);

/* Diagnostics: when device_buf != NULL, every later m3r_gn_solve_step records a timeline of
 * its tile tasks (4 x u64 per task: {kind<<32 | i<<16 | j, t_begin, t_inputs_ready, t_end},
 * globaltimer ns) into device_buf (capacity_tasks entries).  NULL switches it off (default). */
void m3r_debug_chol_trace(unsigned long long* device_buf, int capacity_tasks);
/* Diagnostics: every later m3r_gn_solve_step stalls some tile tasks of the dataflow Cholesky for
 * delay_us microseconds before they start — which = 1: owners of diagonal tiles (j > 0),
 * which = 2: owners of off-diagonal tiles; which >= 3: every task, for a pseudo-random
 * 0..delay_us (seeded by `which`) before it starts and again before it stores its tile.  The flag
 * protocol must make the result independent of such delays (tests/test_gpu_gn.py).
 * delay_us <= 0 switches it off (default). */
void m3r_debug_chol_delay(int delay_us, int which);

/* Diagnostics / tests: systems of more than 5 panels of 32 (more than 23 keyframes) are
 * factorised in 64x64 macro tiles by a kernel whose work is a statically ordered task list
 * (csrc/m3r_gn_chol64.cu).  This returns that list for `nbm` 64-wide block columns, built on the
 * host by the same enumeration the device runs: 4 ints per task {I | J << 16, k0 | k1 << 16,
 * kind (0 chunk update, 1 final = last chunk + TRSM), chunk index}.  Returns the task count
 * (-1 if the enumeration and the workspace formula disagree); writes at most `capacity` tasks. */
int m3r_debug_chol64_tasklist(int nbm, int* out, int capacity);

/* ------------------------------------------------------------------------- *
 * Two-frame tracker solve
 * ------------------------------------------------------------------------- */

/* Whole FrameTracker.opt_pose_ray_dist_sim3 loop (tracker.py:156-214 with
 * geometry.py:17-52, nonlinear_optimizer.py:5-33) in one launch sequence:
 *   Xf (n,3), Xk (n,3) f32 ; Qk (n,1) f32 ; valid (n,1) u8 ;
 *   T_CkCf (8) f32 in/out [t, q(xyzw), s]: relative pose, updated in place ;
 *   status (4) i32 out: [iterations run, converged(0/1), cholesky_failed(0/1), reserved] ;
 *   cost (max_iters) f32 out: cost at each iteration (0.5*b^T b) — optional (NULL ok).
 * Stops when |d cost / cost| < rel_error or |tau| < delta_norm, or after max_iters. */
size_t m3r_track_workspace_bytes(int n);
int m3r_track_ray_dist(const float* Xf, const float* Xk, const float* Qk,
                       const uint8_t* valid, int n, float sigma_ray, float sigma_dist,
                       float huber_k, int max_iters, float rel_error, float delta_norm,
                       float* T_CkCf, int* status, float* cost, void* workspace,
                       size_t workspace_bytes, m3r_stream_t stream);

/* m3r_track_ray_dist for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
 * reference.  Same arguments, arithmetic and
 * results as m3r_track_ray_dist, minus the workspace: the per-CTA partial sums live in a static
 * device array inside the library (85 248 bytes, held by every GPU that loads the library until
 * the process exits), and the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limits: n <= M3R_TRACK_NOALLOC_MAX_N points and a device of at most 296 SMs, otherwise
 * M3R_ETOOLARGE and nothing is launched.  All calls on one device share the array: calls that
 * overlap on the GPU (two streams) race — serialise them on one stream or with events.  Nothing
 * in the array needs to start at zero and it keeps the last call's partials. */
#define M3R_TRACK_NOALLOC_MAX_N (1 << 22) /* 4 194 304 points = a 2048 x 2048 pointmap */
int m3r_track_ray_dist_noalloc(const float* Xf, const float* Xk, const float* Qk,
                               const uint8_t* valid, int n, float sigma_ray, float sigma_dist,
                               float huber_k, int max_iters, float rel_error, float delta_norm,
                               float* T_CkCf, int* status, float* cost, m3r_stream_t stream);

/* Calibrated twin: FrameTracker.opt_pose_calib_sim3 (tracker.py:216-266 with
 * geometry.project_calib, geometry.py:63-104): residuals [u - u_k, v - v_k, log z - log z_k]
 * with u = fx x/z + cx, weights sigma_pixel (x2) / sigma_depth, validity
 * valid & valid_meas_k & (pixel_border < u < width-1-pixel_border, same for v) & (z > z_eps).
 *   Xf (n,3) f32: the frame's points, already gathered by the match index and ray-constrained ;
 *   meas_k (n,3) f32 = [u_k, v_k, log z_k] and valid_meas_k (n,1) u8 as built by
 *   FrameTracker.get_points_poses (tracker.py:143-152) ; K_dev (3,3) f32 row-major on the DEVICE
 *   (the reference keeps K on the device; nothing is read back) ; everything else as above. */
int m3r_track_calib(const float* Xf, const float* meas_k, const float* Qk,
                    const uint8_t* valid, const uint8_t* valid_meas_k, int n,
                    const float* K_dev, int height, int width, int pixel_border,
                    float z_eps, float sigma_pixel, float sigma_depth, float huber_k,
                    int max_iters, float rel_error, float delta_norm, float* T_CkCf,
                    int* status, float* cost, void* workspace, size_t workspace_bytes,
                    m3r_stream_t stream);

/* ------------------------------------------------------------------------- *
 * Streaming steps either side of the solvers
 * ------------------------------------------------------------------------- */

/* Edge-tensor construction of FactorGraph.add_factors for ONE direction (global_opt.py:55-67):
 *   Q_out[e][k] = sqrt(Q_src[e][idx_src2dst[e][k]] * Q_dst[e][k])
 *   valid_count[e] = #{k : valid_match[e][k] && Q_out[e][k] > Q_conf}   (match_frac = count / n)
 * idx (E,n) i64 ; valid_match (E,n,1) u8 ; Q_src, Q_dst, Q_out (E,n,1) f32 ; valid_count (E) i32.
 * The reference's call for direction j is (idx_i2j, valid_match_j, Qii, Qji). */
int m3r_edge_build(const int64_t* idx_src2dst, const uint8_t* valid_match, const float* Q_src,
                   const float* Q_dst, int num_edges, int n, float Q_conf, float* Q_out,
                   int* valid_count, m3r_stream_t stream);
/* The same pass writing the PACKED records FactorGraph appends (global_opt.py:88-97 keeps
 * idx, valid_match and Q as three tensors; here they are one): packed_out (E,n) records with
 * q = valid_match ? Q : -Q.  Q_out is optional (NULL ok).  16 instead of 21 B per point. */
int m3r_edge_build_packed(const int64_t* idx_src2dst, const uint8_t* valid_match,
                          const float* Q_src, const float* Q_dst, int num_edges, int n,
                          float Q_conf, m3r_packed_match* packed_out, float* Q_out,
                          int* valid_count, m3r_stream_t stream);
/* Reference layout <-> packed records, lossless both ways for 0 <= idx < 2^31 and Q >= 0
 * (count = num_edges * n points; any of the three outputs of unpack may be NULL). */
int m3r_edge_pack(const int64_t* idx, const uint8_t* valid_match, const float* Q, size_t count,
                  m3r_packed_match* packed_out, m3r_stream_t stream);
int m3r_edge_unpack(const m3r_packed_match* packed, size_t count, int64_t* idx,
                    uint8_t* valid_match, float* Q, m3r_stream_t stream);

/* Frame.update_pointmap for the per-point filtering modes (frame.py:41-105), in place on
 * X_canon (n,3) / C (n,1), optionally applying a Sim3 to the new points first
 * (tracker.py:97-98).  mode: 0 weighted_pointmap (bit-identical to the reference's expression),
 * 1 weighted_spherical, 2 indep_conf.  T_act: (8) f32 [t, q(xyzw), s] device pointer or NULL. */
int m3r_pointmap_update(float* X_canon, float* C, const float* X_new, const float* C_new,
                        const float* T_act, int n, int mode, m3r_stream_t stream);

/* ------------------------------------------------------------------------- *
 * Headless splat of the viewer's pointmap shaders (last row)
 * ------------------------------------------------------------------------- *
 * Replaces, for a device buffer instead of a window, the two GLSL programs the viewer draws
 * keyframe pointmaps with — resources/programs/surfelmap.glsl and trianglemap.glsl — as driven
 * by Window.render_pointmap (visualization.py:334-356) under DEPTH_TEST + CULL_FACE
 * (visualization.py:100-103).  No graphics API is involved.
 *   fb (height*width) u64, caller-owned: word = (float bits of window depth << 32) | RGBA8
 *   (R in the low byte), row 0 = BOTTOM row (GL window coordinates).  m3r_render_clear sets
 *   depth 1.0 + the clear colour; the render calls depth-test with one 64-bit atomicMin per
 *   fragment, so any number of them may be enqueued on the same buffer in any order;
 *   m3r_render_resolve unpacks to a top-down RGBA8 image and/or a float depth image.
 * Matrices are ROW-MAJOR 4x4 (clip = proj * camera * model * [p,1]); the reference passes the
 * transposes to GL (visualization.py:340-344).  models: (K,4,4) f32 on the DEVICE, one per
 * keyframe = T_WC.matrix(); X (K,H,W,3), C (K,H,W), img (K,H,W,3) f32 device.
 * Deviations from a GL rasteriser are listed at the top of csrc/m3r_render.cu. */
typedef struct m3r_render_view {
  float camera[16];  /* world -> view  (m_camera) */
  float proj[16];    /* view -> clip   (m_proj)   */
  int width, height; /* target size in pixels */
} m3r_render_view;

typedef struct m3r_render_opts {
  float conf_threshold;  /* uniform conf_threshold (WindowMsg.C_conf_threshold = 1.5) */
  float radius;          /* surfelmap.glsl:17, 0.003 */
  float slant_threshold; /* trianglemap.glsl:17, 0.1 */
  float depth_bias;      /* trianglemap.glsl:18 */
  float lightpos[3];     /* (0.1, 0.1, 0) */
  float phong[3];        /* (0.3, 0.5, 0.4) */
  float spec;            /* 32 */
  float base_color[3];   /* (1, 1, 1) */
  int use_img, show_normal, cull_face;
} m3r_render_opts;

void m3r_render_default_opts(m3r_render_opts* opts);
int m3r_render_clear(uint64_t* fb, int width, int height, uint32_t clear_rgba,
                     m3r_stream_t stream);
int m3r_render_surfels(const float* X, const float* C, const float* img, const float* models,
                       int num_keyframes, int H, int W, const m3r_render_view* view,
                       const m3r_render_opts* opts, uint64_t* fb, m3r_stream_t stream);
int m3r_render_triangles(const float* X, const float* C, const float* img, const float* models,
                         int num_keyframes, int H, int W, const m3r_render_view* view,
                         const m3r_render_opts* opts, uint64_t* fb, m3r_stream_t stream);
int m3r_render_resolve(const uint64_t* fb, int width, int height, uint8_t* rgba, float* depth,
                       m3r_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* M3R_B200_H */
