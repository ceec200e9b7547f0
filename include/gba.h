/*
 * gba.h -- C ABI of the B200-native dense bundle-adjustment / DSPO library (libgba.so).
 *
 * This is the drop-in boundary for ONE hot path of GlORIE-SLAM: the call
 *     droid_backends.ba(poses, disps, intrinsics, disps_sens, targets, weights, eta,
 *                       ii, jj, t0, t1, iterations, lm, ep, motion_only, depth_only)
 * (reference src/lib/droid.cpp:89-119 -> src/lib/droid_kernels.cu:1314-1437) and its DSPO
 * stage-2 variant src/geom/ba.py:127-216 (BA_with_scale_shift), both dispatched from
 * reference src/depth_video.py:198-288.
 *
 * Conventions
 *   - plain C, no C++ / torch types; every array argument is a raw DEVICE pointer owned by
 *     the caller (e.g. torch.Tensor.data_ptr()), contiguous, fp32 unless stated; ii/jj are
 *     int64 as in the reference.  `stream` is a cudaStream_t passed as void*.
 *   - functions return 0 on success, a negative GBA_E* code otherwise; gba_last_error() gives
 *     the text.  A non-SPD reduced camera system is NOT an error: like the reference
 *     (droid_kernels.cu:1207-1210, chol.py:13-15) the pose update is zero; the count of such
 *     iterations is reported by gba_query().
 *   - a context is bound to one device and is not re-entrant; calls are asynchronous on
 *     `stream` except for one small status read-back after the index build.  Inside an iteration
 *     one kernel (the Gram of the few keyframes with more than 21 out-edges) runs on a non-blocking
 *     stream owned by the context, forked from and joined to `stream` by events: everything the call
 *     does is ordered before whatever the caller enqueues on `stream` afterwards.
 *   - per-pixel arrays (disps, disps_sens, targets, weights, eta, mono_disps, Q, W) are read with
 *     16-byte vector loads: their base pointers must be 16-byte aligned (torch allocations are);
 *     a misaligned pointer is refused with GBA_EINVAL.
 *   - no CPU fallback exists: every entry point fails with GBA_ECUDA when no device is usable.
 */
#ifndef GBA_H_
#define GBA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gba_ctx gba_ctx;

/* semantics modes (the reference's two BA implementations differ) */
#define GBA_MODE_DROID_CUDA 0 /* src/lib/droid_kernels.cu::ba_cuda -- what droid_backends.ba runs */
#define GBA_MODE_GEOM_PY 1    /* src/geom/ba.py::BA + src/geom/chol.py::schur_solve              */
/* Layout flag, OR-ed into `mode` of gba_ba / gba_ba_begin: targets and weights
 * are [E,ht,wd,2] -- (u,v) interleaved, the layout the update network emits and src/geom/ba.py takes --
 * instead of the [E,2,ht,wd] that droid_backends.ba requires.  The caller can then drop the
 * permute + contiguous copy of src/depth_video.py:215-216 (2 x 16 B per edge pixel of extra HBM
 * traffic per call).  Same per-pixel arithmetic as the planar call.  gba_edge_blocks_static is planar only. */
#define GBA_LAYOUT_INTERLEAVED 0x100

#define GBA_OK 0
#define GBA_EINVAL -1   /* bad argument / index out of range / eta rows != K */
#define GBA_ECUDA -2    /* CUDA runtime error (text in gba_last_error)       */
#define GBA_ECAPACITY -3 /* problem exceeds the sizes given to gba_create    */

/* Creates a context on `device` with workspaces for up to max_frames keyframes, max_edges
 * factor-graph edges and ht x wd disparity maps (1/8 resolution, e.g. 48 x 64).  All device
 * memory the library ever uses is allocated here.  Replaces the per-call torch::zeros /
 * Eigen allocations of droid_kernels.cu:1351-1356,1117-1129. */
int gba_create(gba_ctx **out, int device, int max_frames, int max_edges, int ht, int wd);
void gba_destroy(gba_ctx *ctx);
/* ctx may be NULL: returns the text of the last gba_create failure on this thread. */
const char *gba_last_error(const gba_ctx *ctx);

/* Dense bundle adjustment, `iters` Gauss-Newton iterations; replaces
 * droid_backends.ba == ba_cuda (src/lib/droid_kernels.cu:1314-1437).
 *   poses      [n_frames,7]  in/out  (tx,ty,tz,qx,qy,qz,qw), rows [t0,t1) updated in place
 *   disps      [n_frames,ht,wd] in/out, rows kx[] updated in place (no clamp in droid mode;
 *                              clamp(min=0) per iteration in geom mode, ba.py:119)
 *   intrinsics [4] fx,fy,cx,cy;  disps_sens [n_frames,ht,wd] (zeros = no sensor depth)
 *   targets, weights [n_edges,2,ht,wd] planar (u plane, v plane)
 *   eta        [eta_rows,ht,wd], eta_rows must equal K = |unique([t0,t1) U ii)| (droid) or
 *              |unique(ii)| (geom)
 *   ii, jj     [n_edges] int64 source / destination keyframe of each edge
 *   dx_out     [t1-t0,6] or NULL : last iteration's pose update
 *   dz_out     [K,ht*wd] or NULL : last iteration's disparity update (untouched if motion_only)
 */
int gba_ba(gba_ctx *ctx, float *poses, float *disps, const float *intrinsics, const float *disps_sens,
           const float *targets, const float *weights, const float *eta, const int64_t *ii, const int64_t *jj,
           int n_frames, int n_edges, int eta_rows, int t0, int t1, int iters, float lm, float ep, int motion_only,
           int depth_only, int mode, float *dx_out, float *dz_out, void *stream);

/* DSPO stage 2: one iteration of src/geom/ba.py:127-216 (BA_with_scale_shift): optimises
 * disparities and per-keyframe (scale, shift) of the mono-depth prior, poses fixed.
 *   targets, weights: [n_edges,ht,wd,2] when interleaved != 0 (the layout this reference entry
 *                     point receives), else planar [n_edges,2,ht,wd]
 *   eta [eta_rows,ht,wd] with eta_rows == |unique(ii)|;  mono_disps [n_frames,ht,wd]
 *   scales, shifts [n_frames] in/out;  valid_mask [n_frames,ht,wd] one byte per pixel
 *   disps updated in place and clamped at 0 (ba.py:214).
 * The reference returns new tensors; the Python mirror clones before calling. */
int gba_dspo(gba_ctx *ctx, const float *poses, float *disps, const float *intrinsics, const float *targets,
             const float *weights, const float *eta, const int64_t *ii, const int64_t *jj, const float *mono_disps,
             float *scales, float *shifts, const uint8_t *valid_mask, int n_frames, int n_edges, int eta_rows,
             int ignore_frames, float lm, float ep, float alpha, int interleaved, void *stream);

/* ---- split-phase form of gba_ba for edge-sharded multi-GPU runs --------
 * Every rank calls begin with ITS edges (all edges of a source keyframe live on one rank) plus
 *   kx_global [n_kx]      the global sorted list of depth keyframes, so that eta rows and dz rows
 *                         mean the same on every rank (NULL = derive from the local edges),
 *   owned [n_kx]          one byte per kx entry: whose disparities this rank updates (NULL = all),
 *   ii_all, jj_all [n_all] the edge list of the WHOLE graph (indices only, replicated): the pose
 *                         ordering and the tile structure of the reduced camera system are
 *                         derived from it, so that the packed system has the same layout on
 *                         every rank (NULL = the local edges are the whole graph).
 * Per iteration:
 *   gba_ba_build  -> local contribution to the reduced camera system, packed in a buffer
 *   (caller sums the buffer over ranks, e.g. ncclAllReduce / torch.distributed.all_reduce)
 *   gba_ba_solve  -> damping, factorisation, solve, back-substitution, retraction
 * gba_ba() is begin + iters x (build, solve) + end on one rank. */
int gba_ba_begin(gba_ctx *ctx, float *poses, float *disps, const float *intrinsics, const float *disps_sens,
                 const float *targets, const float *weights, const float *eta, const int64_t *ii, const int64_t *jj,
                 int n_frames, int n_edges, int eta_rows, int t0, int t1, float lm, float ep, int motion_only,
                 int depth_only, int mode, const int64_t *kx_global, int n_kx, const uint8_t *owned,
                 const int64_t *ii_all, const int64_t *jj_all, int n_all, void *stream);
int gba_ba_build(gba_ctx *ctx);
/* device pointer + element count (doubles) of the packed reduced system of the last build */
int gba_ba_system(gba_ctx *ctx, double **buf, size_t *count);
int gba_ba_solve(gba_ctx *ctx);
int gba_ba_end(gba_ctx *ctx, float *dx_out, float *dz_out);

/* ---- sharded SOLVE (optional second stage of the split-phase form) ---------------------------------
 * With gba_ba_solve every rank factorises the whole summed system.  When the ordering found independent band
 * segments (gba_query: n_segments > 1) and every source keyframe lives on the rank that owns the band segment its
 * pose blocks fall into, only the closure block of the reduced system is shared between ranks:
 *   gba_plan            index build + ordering + tile structure from the edge indices alone; read the pose ->
 *                       position map with gba_get(GBA_GET_PERM) and n_segments / segment_tiles with gba_query to
 *                       decide the ownership (glorie-slam_b200/sharded.py::plan does it)
 *   gba_ba_begin        as above (owned = the keyframes of this rank's segments)
 *   gba_ba_set_shard    this rank eliminates band segments [seg_lo, seg_lo + seg_n); first_rank != 0 on exactly one rank
 *   per iteration:
 *     gba_ba_build      local contributions
 *     gba_ba_eliminate  damping + factorisation of this rank's band segments (their Schur complement lands in
 *                       the closure tiles)
 *     (caller sums gba_ba_system_part(ctx, 1, ..) over ranks: closure x closure tiles + rhs, one contiguous range)
 *     gba_ba_finish     closure factorisation (replicated), backward substitution of the closure and of this rank's
 *                       segments, disparity update of the owned keyframes, retraction of the closure poses and of
 *                       this rank's poses (the other poses get a zero update here: merge them after the last iteration)
 * No counterpart in the reference (it has no multi-GPU path). */
int gba_plan(gba_ctx *ctx, const int64_t *ii, const int64_t *jj, int n_edges, int n_frames, int t0, int t1, void *stream);
int gba_ba_set_shard(gba_ctx *ctx, int seg_lo, int seg_n, int first_rank);
int gba_ba_eliminate(gba_ctx *ctx);
int gba_ba_system_part(gba_ctx *ctx, int which /* 0 whole system, 1 closure block + rhs */, double **buf, size_t *count);
int gba_ba_finish(gba_ctx *ctx);

/* ---- introspection (used by the parity tests; synchronises the stream) ----------------- */
typedef struct gba_info {
  int n_depth_frames;   /* K                                                  */
  int n_poses;          /* P = t1 - t0                                        */
  int max_out_degree;   /* largest number of edges sharing a source keyframe  */
  int chol_failures;    /* iterations whose factorisation hit a pivot <= 0    */
  int n_tiles;          /* tile rows of the reduced camera matrix             */
  int n_active_tiles;   /* lower-triangular tiles touched after symbolic fill */
  int kernel_launches;  /* kernels launched by the last gba_ba / gba_dspo call */
  int chain_columns;    /* tile columns on the factorisation's critical path: one band segment + the closure block */
  int last_fail_code;   /* why the most recent failed factorisation failed: 1 = pivot <= 0 (the reference's
                           NumericalIssue, droid_kernels.cu:1202-1210), 3 / 4 = an in-kernel wait of the tile
                           Cholesky / the tensor-core Gram pipeline exceeded its watchdog; 0 = none since begin */
  int watchdog_trips;   /* GBA_GET_WATCHDOG[0]: watchdog records since gba_create (0 in a healthy process)       */
  int n_segments;       /* independent band segments of the ordered reduced system (1: none)                     */
  int segment_tiles;    /* tile columns per segment                                                              */
} gba_info;
int gba_query(gba_ctx *ctx, gba_info *info);
/* K = number of depth keyframes of the last gba_ba / gba_ba_begin / gba_dspo call on this context, from the host copy
 * of the status the index build already read back: no device access, no synchronisation.  Sizes dz_out. */
int gba_last_depth_frames(const gba_ctx *ctx);

/* Copies an internal device array of the last call to HOST memory `dst` (capacity `bytes`).
 * what: */
#define GBA_GET_KX 1       /* int32 [K]   sorted depth keyframes (droid_kernels.cu:1341-1345) */
#define GBA_GET_EPTR 2     /* int32 [K+1] CSR row pointers by source keyframe (:948-984)      */
#define GBA_GET_EIDX 3     /* int32 [E]   edge ids grouped by source keyframe, edge order     */
#define GBA_GET_EDGE_HV 4  /* float [E,32] per edge: 21 upper-tri of H_jj then 6 of v_j       */
#define GBA_GET_Q 5        /* float [K,HW] 1/C of the last iteration                          */
#define GBA_GET_W 6        /* float [K,HW] w of the last iteration                            */
#define GBA_GET_SYSTEM 7   /* double [6P*6P + 6P] row-major lower triangle of the reduced camera matrix A - S as
                              left by the last gba_ba_build (geom mode: damping already on A; droid mode: not
                              yet damped), then the rhs b - vS.  Valid between gba_ba_build and gba_ba_solve. */
#define GBA_GET_TILEMASK 8 /* uint8 [nT,nT] lower-triangular tile structure after fill        */
#define GBA_GET_GRAM 9     /* float: per depth keyframe k (in GBA_GET_KX order) one Rp x Rp block, Rp = 8*ceil((6 deg_k
                              + 1)/8), concatenated; entry (r, c >= r) = sum_pix E_r Q E_c, column 6 deg_k = sum_pix
                              E_r Q w; entries below the diagonal are unspecified (last iteration)                  */
#define GBA_GET_WATCHDOG 10 /* int32 [256]: [0] = number of pipeline waits in the tensor-core Gram kernel that exceeded
                               the watchdog limit since gba_create (0 in a healthy run), then up to 42 records
                               {keyframe, split, warp, wait site, index, group}                                  */
#define GBA_GET_PERM 12     /* int32 [P]: pose (frame - t0) -> position in the ordered reduced system; positions below
                               n_segments * segment_tiles * 64 / 6 belong to band segment position / (segment_tiles * 64 / 6),
                               the others to the closure block (gba_query gives n_segments, segment_tiles)          */
#define GBA_GET_CHAIN_TRACE 11 /* uint64: development aid, only when the context was created with the environment
                                  variable GBA_CHAIN_TRACE set: [0] = tasks of the last factorisation kernel, then per task
                                  {role<<56 | column<<40 | a<<28 | b<<16 | slab<<12 | SM, t_taken, t_inputs_ready, t_done}
                                  in %globaltimer nanoseconds                                                           */
int gba_get(gba_ctx *ctx, int what, void *dst, size_t bytes);

/* ---- optional per-stage device timing ---------------------------------------------------
 * When enabled, the library brackets every stage of gba_ba / gba_dspo with CUDA events on the
 * call's own stream (the stream the kernels are launched on).  gba_get_timing synchronises the
 * stream, returns the accumulated milliseconds and kernel-launch counts per stage since the last
 * read and clears them.  The reference has no timers around BA. */
#define GBA_STAGE_INDEX 0     /* index build + tile structure (once per call)         */
#define GBA_STAGE_LINEARIZE 1 /* edge geometry + k_linearize (per-edge HBM pass)      */
#define GBA_STAGE_POSE 2      /* per-edge pose blocks into the reduced system         */
#define GBA_STAGE_GRAM 3      /* Schur Gram accumulation                              */
#define GBA_STAGE_SCHUR 4     /* Schur assembly (+ damping)                           */
#define GBA_STAGE_FACTOR 5    /* tile Cholesky                                        */
#define GBA_STAGE_SOLVE 6     /* triangular solves                                    */
#define GBA_STAGE_BACKSUB 7   /* back-substitution + disparity retraction             */
#define GBA_STAGE_RETR 8      /* pose retraction                                      */
#define GBA_STAGE_DSPO 9      /* DSPO stage-2 kernels                                 */
#define GBA_N_STAGES 10
int gba_set_timing(gba_ctx *ctx, int enable);
int gba_get_timing(gba_ctx *ctx, float *ms /* [GBA_N_STAGES] */, int *launches /* [GBA_N_STAGES] */);

/* ---- frame_distance -------------------------------------------------
 * Replaces droid_backends.frame_distance (src/lib/droid.cpp:122-138 ->
 * src/lib/droid_kernels.cu:518-657,1441-1463): for every pair (ii[n] -> jj[n]) the mean
 * induced-flow magnitude, beta * full relative motion + (1 - beta) * translation only, over
 * the pixels that land in front of the target camera; 1000 when fewer than 75 % do.
 * poses [*,7], disps [*,ht,wd], intrinsics [4], ii/jj [n_pairs] int64, dist [n_pairs] out;
 * device pointers, asynchronous on `stream`, needs no gba_ctx.  wd must be a multiple of 4. */
int gba_frame_distance(const float *poses, const float *disps, const float *intrinsics, const int64_t *ii,
                       const int64_t *jj, int n_pairs, int ht, int wd, float beta, float *dist, void *stream);

/* ---- depth_filter ---------------------------------------------------
 * Replaces droid_backends.depth_filter (src/lib/droid.cpp:222-236 ->
 * src/lib/droid_kernels.cu:661-775,1494-1518; caller src/depth_video.py:326-361): for every
 * listed keyframe inds[b] and pixel, how many of the 6 neighbouring keyframes (inds[b]-1,-2,-3,
 * +3,+4,+5, those inside [0, n_frames)) hold, at one of the 4 pixels around the reprojection,
 * a depth within thresh[b] of the reprojected one.  counter [n_inds,ht,wd] is fully written
 * (values 0..6 as float, like the reference's float counter); it need not be zeroed.
 * poses [n_frames,7], disps [n_frames,ht,wd], intrinsics [4], inds [n_inds] int64,
 * thresh [n_inds]; device pointers, asynchronous on `stream`, needs no gba_ctx. */
int gba_depth_filter(const float *poses, const float *disps, const float *intrinsics, const int64_t *inds,
                     const float *thresh, int n_frames, int n_inds, int ht, int wd, float *counter, void *stream);

/* ---- align_scale_shift ----------------------------------------------
 * Replaces src/utils/common.py:402-438 align_scale_and_shift (callers src/depth_video.py:233,
 * 308): per keyframe b the weighted least-squares fit  min sum w (scale*prediction + shift -
 * target)^2  and the weighted mean absolute residual of that fit.  prediction/target/weights
 * [n,ht,wd] float32; scale/shift/avg_error [n] out.  A keyframe whose weights make the 2x2
 * normal matrix singular yields inf/nan exactly like the reference's division by det = 0.
 * Device pointers, asynchronous on `stream`, needs no gba_ctx. */
int gba_align_scale_shift(const float *prediction, const float *target, const float *weights, int n, int ht, int wd,
                          float *scale, float *shift, float *avg_error, void *stream);

/* ---- iproj / projmap ------------------------------------------------
 * Replace droid_backends.iproj (src/lib/droid.cpp:159-168 -> src/lib/droid_kernels.cu:779-850,
 * 1520-1545; caller src/neural_point.py:156) and droid_backends.projmap (src/lib/droid.cpp:
 * 141-156 -> src/lib/droid_kernels.cu:427-516,1466-1492).
 * gba_iproj: points [n_frames,ht,wd,3] = (R_b (x, y, 1) + t_b d) / d for every pixel of every
 * keyframe b, pose b as given (the caller passes inverse poses to get world points); d = 0 gives
 * inf / nan like the reference.
 * gba_projmap: for pair n the pixels of keyframe ii[n] reprojected into keyframe jj[n]:
 * coords [n_pairs,ht,wd,3] = (u', v', 0), the pixel's own (u, v) where Z <= 0.01;
 * valid [n_pairs,ht,wd,1] = 1 where Z > 0.25, else 0.  Both outputs are fully written.
 * Device pointers, asynchronous on `stream`, no gba_ctx; n_frames, n_pairs <= 65535. */
int gba_iproj(const float *poses, const float *disps, const float *intrinsics, int n_frames, int ht, int wd,
              float *points, void *stream);
int gba_projmap(const float *poses, const float *disps, const float *intrinsics, const int64_t *ii, const int64_t *jj,
                int n_pairs, int ht, int wd, float *coords, float *valid, void *stream);

/* ---- proximity edge selection ------------------------------------------------
 * Replaces the host loops of FactorGraph.add_backend_proximity_factors (src/factor_graph.py:386-462): thresholds on
 * the bidirectional frame distance, the local-window edges, sort, greedy selection with suppression of the
 * (2 nms + 1)^2 neighbourhood, the max_factors cap, and (loop != 0) the loop-closure variant.
 *   d [ilen*jlen]  0.5 (frame_distance(i->j) + frame_distance(j->i)) for i in [t_start_loop, t_end) (rows) and
 *                  j in [t_start, t_end) (columns), as DepthVideo.distance returns it (src/depth_video.py:166-196);
 *                  loop == 0 forces t_start_loop = t_start like the reference.  Not modified.
 *   scratch        gba_proximity_scratch_bytes(ilen, jlen) bytes of device memory
 *   out_ii/out_jj  [cap] int32 edge list in the reference's order; out_n[0] = edges written, out_n[1] = 0, or 1 when
 *                  the reference returns without adding edges (fewer than 3, or no loop edge), or 2 when cap was too small
 * One CTA; equal distances are taken in ascending pair index (a stable sort).  Asynchronous on `stream`. */
size_t gba_proximity_scratch_bytes(int ilen, int jlen);
int gba_proximity_select(const float *d, int t_start, int t_start_loop, int t_end, int nms, int radius, float thresh,
                         int max_factors, int loop, void *scratch, size_t scratch_bytes, int *out_ii, int *out_jj,
                         int cap, int *out_n, void *stream);

/* ---- valid-depth mask and mono_thres edge filter -----------------------------
 * gba_valid_depth_mask replaces src/depth_video.py:340-361 after depth_filter: per listed keyframe, depth = 1 / disp,
 * pixels with count < visible_num dropped, median = torch.nanmedian (lower middle value), mask = depth < 3 median.
 *   disps, count [n,ht,wd] (the selected keyframes' rows); mask [n,ht,wd] one byte per pixel;
 *   valid_cnt [n] = mask.sum(), mean_disp [n] = disps.mean() per keyframe (inputs of the filter below).
 * gba_mono_filter replaces the loops of src/depth_video.py:245-261: invalid[f] = error/avg_disp > mono_thres | isnan(error)
 * | scale < 0 | valid_cnt < ht*wd/2 for the first n keyframes; keep[e] = neither end of edge e is invalid;
 * src_kept[f] (f < n_frames) = f is still the source of a kept edge (eta rows follow unique(ii): mask them with it).
 * Device pointers, asynchronous on `stream`, no gba_ctx. */
int gba_valid_depth_mask(const float *disps, const float *count, int n, int ht, int wd, int visible_num, uint8_t *mask,
                         int *valid_cnt, float *mean_disp, void *stream);
int gba_mono_filter(const float *error, const float *avg_disp, const float *scale, const int *valid_cnt, int n, int ht,
                    int wd, float mono_thres, const int64_t *ii, const int64_t *jj, int n_edges, int n_frames,
                    uint8_t *invalid, uint8_t *keep, uint8_t *src_kept, void *stream);

/* ---- per-edge pose blocks for callers that cannot allocate --------------------------------
 * No counterpart in the reference: there, these numbers only exist inside ba_cuda.
 *
 * Computes, for every edge i -> j, H_jj = sum_pix w J_j^T J_j (21 upper-triangular entries, row
 * major) followed by v_j = sum_pix w J_j^T r (6 entries) into out[e*32 .. e*32+26] (entries
 * 27..31 are written as 0) -- the layout of gba_get(GBA_GET_EDGE_HV).  Arguments are device
 * pointers shaped as in gba_ba: poses [n_frames,7], disps [n_frames,ht,wd], intrinsics [4],
 * targets / weights [n_edges,2,ht,wd], ii / jj [n_edges] int64; out [n_edges,32] fp32.
 *
 * Unlike every other entry point it needs no gba_ctx and performs no allocation of any kind
 * (no cudaMalloc / cudaFree, no allocating library call): it is meant for callers that must not
 * or cannot allocate when they need these blocks -- a latency-critical tracking thread that may
 * not block on the driver's allocator lock, or a process that has already given all device memory
 * to another allocator.  Its scratch memory is a fixed array inside libgba.so (about 2.6 MB on
 * every GPU that loads the library), which sets the limits below; larger problems are refused
 * with GBA_ECAPACITY (text via gba_last_error(NULL)) -- split the edge list or use gba_ba.
 *   - Works on the CURRENT device.  The first call on each device also looks up the device
 *     address of the built-in array; make that first call outside any stream capture.
 *   - All calls share the one array: calls whose kernels may overlap on the GPU (different
 *     streams, different threads) race with each other.  Serialise them on one stream.
 *   - ii / jj must lie in [0, n_frames); an edge that does not gets an all-zero block.
 *   - disps / targets / weights must be 16-byte aligned (vector loads); otherwise GBA_EINVAL.
 *   - Asynchronous on `stream`, like gba_ba. */
#define GBA_STATIC_MAX_EDGES 4096
#define GBA_STATIC_MAX_PIXELS 4800 /* ht*wd, e.g. 60 x 80 */
int gba_edge_blocks_static(const float *poses, const float *disps, const float *intrinsics, const float *targets,
                           const float *weights, const int64_t *ii, const int64_t *jj, int n_frames, int n_edges,
                           int ht, int wd, int mode, float *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GBA_H_ */
