/*
 * deepclr_b200.h -- C ABI of libdeepclr_b200.so (hand-written CUDA for sm_100a).
 *
 * Drop-in boundary for the registration forward path of Leedk3/deepclr.  Each entry point names the reference
 * interface it replaces (paths relative to the reference tree).  Conventions:
 *   - every pointer is a DEVICE pointer owned by the caller (the reference's wrappers also only borrow data_ptrs:
 *     extern/Pointnet2.PyTorch/pointnet2/pointnet2_utils.py:25-28,55-57,172-174,218-220);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is enqueued
 *     asynchronously on it, nothing synchronises, so every call is CUDA-graph capturable;
 *   - return value: 0 on success, negative DCLR_E_* otherwise; dclr_last_error() gives a thread-local message.
 *     (The reference prints and exit(-1)s on launch failure: sampling_gpu.cu:39-43.)
 *   - index dtype int32 for the PointNet++ ops, int64 for kNN (torch-cluster layout);
 *   - the current CUDA device of the calling thread is used (same as the reference: no device guard).
 */
#ifndef DEEPCLR_B200_H
#define DEEPCLR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCLR_OK 0
#define DCLR_E_INVALID (-1)   /* bad argument (shape, null pointer, unsupported size) */
#define DCLR_E_CUDA (-2)      /* CUDA runtime error at launch */
#define DCLR_E_UNSUPPORTED (-3)

/* Library identification; also used by the loader to prove the native code is the one running. */
const char *dclr_version(void);
const char *dclr_last_error(void);

/*
 * Farthest point sampling.  Replaces pointnet2_cuda.furthest_point_sampling_wrapper
 * (src/sampling.cpp:36-46 -> src/sampling_gpu.cu:94-253) and, when new_xyz_out != NULL, the gather_operation that
 * always follows it (pointnet2_modules.py:32-35 -> sampling_gpu.cu:8-44).
 *   pts        [B, N, pt_stride] fp32, xyz in the first three floats of each point (pt_stride >= 3)
 *   tie_rule   how two points holding the running maximum with bit-equal fp32 distances are ordered (it happens about
 *              once per 10^4 rounds on 16k-point scans):
 *                DCLR_TIE_LOWEST_INDEX  lowest point index (the rule BASELINE.json states)
 *                DCLR_TIE_REFERENCE     the reference kernel's own order, furthest_point_sampling_kernel
 *                                       (sampling_gpu.cu:113-203): thread k mod T keeps its first strict maximum and
 *                                       the shared-memory tree keeps the lower slot, i.e. the smallest bit-reversed
 *                                       slot, then the smallest k; T = opt_n_threads(N) (cuda_utils.h:10-14).  With
 *                                       it the indices are bit-identical to the reference kernel on every input.
 *              There is no process-wide setting: the rule is an argument here and a property of a model handle
 *              (dclr_model_set_fps_tie_rule, default DCLR_TIE_REFERENCE).
 *   idx_out    [B, npoint] int32; idx[.,0] = 0
 *   new_xyz_out[B, npoint, 3] fp32 or NULL
 * The min-distance scratch the reference makes the caller allocate and fill with 1e10 (pointnet2_utils.py:26)
 * lives in registers here.  npoint > N is allowed (reference test tests/model/test_deepclr.py:10-25).
 */
#define DCLR_TIE_LOWEST_INDEX 0
#define DCLR_TIE_REFERENCE 1
int dclr_fps(const float *pts, int pt_stride, int B, int N, int npoint, int tie_rule, int32_t *idx_out,
             float *new_xyz_out, void *stream);

/*
 * dclr_fps without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Arguments, kernels and results are those of dclr_fps; the scratch dclr_fps takes from the stream-ordered pool (the
 * Z-ordered records of the bucketed kernels, 16 bytes per point) is a file-scope device array of the library
 * (csrc/fps_static.cu, 2 MiB, held by every GPU that loads the library until the process exits).  Limits, reported by
 * dclr_fps_static_limits: N <= 16384 points per cloud and B * N <= 131072 points per call; anything larger returns
 * DCLR_E_INVALID with a message naming the limit.  The call makes no cudaMalloc / cudaFree / pool allocation.
 * Every call shares the one array: calls that overlap on the GPU (different streams, or several host threads without
 * ordering) race on it; calls on one stream are ordered and safe.  Nothing in it has to be cleared between calls.
 */
int dclr_fps_static(const float *pts, int pt_stride, int B, int N, int npoint, int tie_rule, int32_t *idx_out,
                    float *new_xyz_out, void *stream);
int dclr_fps_static_limits(int *max_points_per_cloud, int *max_points_per_call);

/* out[b,c,j] = feat[b,c,idx[b,j]].  Replaces pointnet2_cuda.gather_points_wrapper (sampling.cpp:11-20). */
int dclr_gather(const float *feat, const int32_t *idx, int B, int C, int N, int M, float *out, void *stream);

/*
 * Ball query.  Replaces pointnet2_cuda.ball_query_wrapper (src/ball_query.cpp:14-25 -> ball_query_gpu.cu:9-67).
 *   new_xyz [B,M,3], pts [B,N,pt_stride] (xyz first) -> idx_out [B,M,nsample] int32: first nsample points with
 *   d^2 < radius^2 in ascending index order, remaining slots = first hit, all zeros when the ball is empty
 *   (the caller-side zero fill of pointnet2_utils.py:218 is done here).  cnt_out [B,M] (nullable) = genuine hits.
 *   pad = 0 skips the padding (slots >= cnt are left untouched): the fused path's compact form.
 */
int dclr_ball_query(const float *new_xyz, const float *pts, int pt_stride, int B, int N, int M, float radius,
                    int nsample, int32_t *idx_out, int32_t *cnt_out, int pad, void *stream);

/* out[b,c,i,j] = feat[b,c,idx[b,i,j]].  Replaces pointnet2_cuda.group_points_wrapper (group_points.cpp:25-35).
 * Compatibility/parity only: the fused path never materialises this tensor. */
int dclr_group(const float *feat, const int32_t *idx, int B, int C, int N, int M, int nsample, float *out,
               void *stream);

/*
 * Backward of gather / group: grad_points[b,c,idx[...]] += grad_out[...]  (fp32 atomicAdd; grad_points [B,C,N] must be
 * zero-filled by the caller, as the reference's autograd Functions do: pointnet2_utils.py:65-70, 189-194).
 * Replace pointnet2_cuda.gather_points_grad_wrapper (sampling.cpp:23-33 -> sampling_gpu.cu:46-72) and
 * group_points_grad_wrapper (group_points.cpp:11-22 -> group_points_gpu.cu:8-44).  Equal to the reference up to the
 * order of the floating-point additions (both sides use atomics).
 */
int dclr_gather_grad(const float *grad_out, const int32_t *idx, int B, int C, int N, int M, float *grad_points,
                     void *stream);
int dclr_group_grad(const float *grad_out, const int32_t *idx, int B, int C, int N, int M, int nsample,
                    float *grad_points, void *stream);

/*
 * Batched kNN.  Replaces torch_cluster.knn(x, y, k, batch_x, batch_y) as called at deepclr/models/deepclr.py:164-166
 * (equal-sized, sorted batches).  x [B*Nx,3] candidates, y [B*Ny,3] queries.  row_out/col_out [B*Ny*k] int64:
 * row = global y row id, col = global x row id, ascending distance, ties -> lower x index.  Needs 0 < k <= min(Nx,128).
 */
int dclr_knn(const float *x, const float *y, int B, int Nx, int Ny, int k, int64_t *row_out, int64_t *col_out,
             void *stream);

/* ---- fused model path --------------------------------------------------------------------------------- */

typedef struct dclr_model dclr_model;   /* opaque; device-resident weights + workspace */

/* Architecture = the `params` block of a reference model_config.yaml (models/kitti_00-10/model_config.yaml). */
typedef struct {
    int input_dim;            /* 3 or 4 (xyz [+ 1 feature])                            model_config.yaml:2 */
    int npoint;               /* set-abstraction samples                               :12-13 */
    int num_scales;           /* 1 or 2 */
    float radii[2];           /* :14-16 */
    int nsamples[2];          /* :17-19 */
    int sa_mlp[2][3];         /* hidden sizes per scale, e.g. {16,16,32}               :20-26 */
    int knn_k;                /* :30 */
    float merge_radius;       /* :31 (<= 0 disables the mask, deepclr.py:220) */
    int merge_mlp[3];         /* :32-35 */
    int out_mlp[5];           /* :39-44 */
    int out_linear[3];        /* :45-48 (first entry == out_mlp[4]) */
    int label_type;           /* 0 EULER(6) 1 QUAT(7) 2 DUAL_QUAT(8)                   deepclr/data/labels.py:12-24 */
} dclr_arch;

/*
 * weights: HOST pointer to all parameters as fp32, concatenated in the reference's state_dict order
 * (SURVEY.md section 3.3): SA scale0 (W,b)x3, SA scale1 (W,b)x3, merge (W,b)x3, output conv (W,b)x5,
 * linear (W,b)x2, final (W,b); each W row-major [Cout,Cin].  Copied to the device (synchronously).
 */
int dclr_model_create(const dclr_arch *arch, const float *weights, size_t num_floats, dclr_model **out);
void dclr_model_destroy(dclr_model *m);
/* numerics: 0 = fp32 SIMT everywhere (fp32-faithful); 1 = bf16, 2 = fp16 operands on the tcgen05 tensor cores
 * (fp32 accumulation in TMEM) for the flow-embedding and pose-head MLP stacks */
int dclr_model_set_numerics(dclr_model *m, int mode);
/* FPS tie order of this handle (see dclr_fps); default DCLR_TIE_REFERENCE = the reference kernel's own order. */
int dclr_model_set_fps_tie_rule(dclr_model *m, int rule);
int dclr_model_get_fps_tie_rule(const dclr_model *m);

/*
 * Set abstraction (deepclr.py:510-521 -> :88-94 -> pointnet2_modules.py:19-55).
 *   x [num_clouds, N, input_dim] fp32 -> feat_out [num_clouds, 3+F, npoint] fp32 (reference layout, xyz rows first).
 */
int dclr_cloud_features(dclr_model *m, const float *x, int num_clouds, int N, float *feat_out, void *stream);

/*
 * Merge + output on precomputed features (DeepCLR.forward(is_feat=True), deepclr.py:492-496; used by the sequential
 * ModelInferenceHelper, deepclr/models/base.py:97-112).  feat [2B, 3+F, npoint] templates first -> y_out [B, label_dim].
 * motion_out (nullable) [B, 3+merge_mlp[2], npoint] = MotionEmbedding output.
 */
int dclr_forward_features(dclr_model *m, const float *feat, int B, float *y_out, float *motion_out, void *stream);

/*
 * Whole forward (DeepCLR.forward, deepclr.py:488-508): x [2B, N, input_dim] templates first -> y_out [B, label_dim].
 * feat_out (nullable) as in dclr_cloud_features; motion_out (nullable) as above.
 */
int dclr_forward_pairs(dclr_model *m, const float *x, int B, int N, float *y_out, float *feat_out,
                       float *motion_out, void *stream);

/*
 * Stage-level entries: the three fused stacks of the path one by one (op-level parity and sweeps).
 *
 * dclr_sa_msg_fused: ball query of both radii + grouping + shared MLP + max-pool around GIVEN centres = what
 *   PointnetSAModuleMSG.forward does after its sampling (pointnet2_modules.py:37-55: QueryAndGroup -> SharedMLP ->
 *   max_pool2d per scale, concatenated).  x [num_clouds, N, input_dim], new_xyz [num_clouds, npoint, 3] ->
 *   feat_out [num_clouds, F, npoint] (features only, the reference's new_features layout).
 * dclr_flow_embed_fused: kNN grouping + flow-embedding MLP + radius mask + max over the neighbours =
 *   MotionEmbedding.forward (deepclr/models/deepclr.py:201-246).  feat [2B, 3+F, npoint] templates first ->
 *   motion_out [B, 3+merge_mlp[2], npoint].
 * dclr_pose_head_fused: OutputSimple.forward (deepclr.py:284-294).  motion [B, 3+merge_mlp[2], npoint] -> y_out
 *   [B, label_dim].
 * dclr_workspace_bytes: device bytes a call with num_clouds clouds of N points makes the handle hold (0 on bad arguments);
 *   the workspace is allocated lazily and only ever grows (see dclr_model_generation).
 *
 * Threading contract of a handle: calls on ONE dclr_model must not overlap (one workspace, one side stream; they are
 * ordered by the caller's stream); different handles are independent and may be driven from different threads and
 * streams concurrently.  The library keeps no process-wide mutable state besides a per-device scratch pool behind a lock
 * (op-level entries) and per-device one-time kernel attributes.
 */
size_t dclr_workspace_bytes(const dclr_model *m, int num_clouds, int N);
int dclr_sa_msg_fused(dclr_model *m, const float *x, const float *new_xyz, int num_clouds, int N, float *feat_out,
                      void *stream);
int dclr_flow_embed_fused(dclr_model *m, const float *feat, int B, float *motion_out, void *stream);
int dclr_pose_head_fused(dclr_model *m, const float *motion, int B, float *y_out, void *stream);

/* Number of kernels the last dclr_forward_* / dclr_cloud_features call on this model enqueued. */
int dclr_model_last_launches(const dclr_model *m);
/* Workspace generation: grows by one each time a call had to re-allocate the handle's workspace (a larger shape than
 * any before).  A CUDA graph captured from calls on this handle holds workspace pointers: it is valid only while the
 * generation it was captured at is still current (deepclr_b200/models/stream.py re-captures on a mismatch). */
int dclr_model_generation(const dclr_model *m);
/*
 * Per-stage device timing with CUDA events recorded on the caller's stream (the reference times with an event pair
 * around predict: scripts/timing.py:32-47; this splits that interval by kernel stage).  When on, every
 * dclr_cloud_features / dclr_forward_* call records begin/end events per stage into a ring of the last 64 calls.
 * dclr_model_collect_timing synchronises on them, writes the per-stage sum of milliseconds over the recorded calls
 * into stage_ms_sum[dclr_model_stage_count()] and the number of calls summed into *calls_out, then resets the ring.
 */
int dclr_model_set_timing(dclr_model *m, int on);
int dclr_model_stage_count(void);
const char *dclr_model_stage_name(int stage);
int dclr_model_collect_timing(dclr_model *m, float *stage_ms_sum, int *calls_out);
/* Debug/parity taps of the last forward: copies `nbytes` of a workspace array into dst (device) on `stream`.
 * which: 0 fps idx [2B,npoint] i32; 1 ball cnt scale0 [2B,npoint] i32; 2 ball cnt scale1; 3 knn col [B*npoint*k] i32
 * (source index local to the pair); 4 ball idx scale0 (compact, only the first cnt slots of a row are defined)
 * [2B,npoint,ns0] i32; 5 ball idx scale1. */
int dclr_model_debug_copy(const dclr_model *m, int which, void *dst, size_t nbytes, void *stream);

/* ------------------------------------------------------------------ post-processing on the device
 * Pose labels -> 4x4 matrices, batched, float64: LabelType.to_matrix (deepclr/data/labels.py:80-101; dual quaternions
 * are normalised by |real| + 1e-8 first, :46-51).  labels [B, 6|7|8] fp32 as the network emits them, label_type as in
 * dclr_arch, scale = the label scale the reference divides by (<= 0: none).  T_out [B,16] row-major.
 * Replaces the per-pair host loop of scripts/inference.py:113-119. */
int dclr_pose_to_matrix(const float *labels, int B, int label_type, float scale, double *T_out, void *stream);
/* KITTI odometry errors between two batches of 4x4 matrices (deepclr/evaluation/metrics.py:8-49): for each pair the
 * smaller of err(m1 inv(m2)) and err(m2 inv(m1)); rotation in RADIANS (acos((trace-1)/2) clamped), translation = norm
 * of the translation part. */
int dclr_pose_errors(const double *T1, const double *T2, int B, double *rot_rad_out, double *trans_out, void *stream);

/* ------------------------------------------------------------------ pre-processing on the device
 * A ragged batch of raw scans: pts [total, C_in] fp32 = the clouds back to back, offsets [B+1] i32 (DEVICE) = first
 * row of each cloud, offsets[B] = total.
 *
 * dclr_filter_clouds: SystematicErasing(nth, start) (cloud[start::nth], deepclr/data/transforms/transforms.py:244-268)
 * -> RangeSelection(min_range, max_range) (keep rows with min <= max(|x|,|y|) <= max compared in fp32, rows with a NaN
 * in x or y drop; (0, +inf) = no test at all, as the reference returns the cloud untouched; :90-113) ->
 * TruncateDimension(C_out) (:271-283) -- the chain of deepclr/data/transforms/build.py:23-24,37-38 -- as one
 * order-preserving compaction per cloud.  Cloud b's surviving rows are written to out [total, C_out] from row
 * offsets[b] on (same offsets as the input; out may not alias pts), their number to counts_out[b].  start = -1 (random
 * start) is drawn by the caller.  max_points = an upper bound of the largest cloud (it sizes the grid and the per-tile
 * scratch); a cloud that is longer is not processed and gets counts_out[b] = -1. */
int dclr_filter_clouds(const float *pts, const int32_t *offsets, int B, int max_points, int C_in, int C_out, int nth,
                       int start, float min_range, float max_range, float *out, int32_t *counts_out, void *stream);
/* dclr_subsample_clouds: N_out distinct random rows of every cloud into the dense batch the forward reads,
 * x_out [B, N_out, C_out] -- ModelInferenceHelper.stack (deepclr/models/base.py:130-135: cloud[randperm(n)[:N_out]])
 * and RandomErasing(max_points) (transforms.py:128-131), batched.  Cloud b = rows offsets[b] .. + n_b with
 * n_b = counts[b] (counts = NULL: offsets[b+1] - offsets[b]).  Row j of cloud b is source row P_b(j), P_b a keyed
 * bijection of [0, n_b) (4-round Feistel network on ceil(log2 n_b) bits + cycle walking; key from seed and b), so the rows are
 * distinct; n_b == N_out with ordered_if_equal != 0 copies the cloud in order (the reference does not permute equal
 * sizes).  n_b < N_out: rows j >= n_b repeat P_b(j mod n_b) and *short_out (optional, DEVICE i32) counts such clouds;
 * n_b == 0 gives zeros and perm -1.  perm_out (optional) [B, N_out] i32 = the chosen source rows. */
int dclr_subsample_clouds(const float *pts, const int32_t *offsets, const int32_t *counts, int B, int C_in, int C_out,
                          int N_out, uint64_t seed, int ordered_if_equal, float *x_out, int32_t *perm_out,
                          int32_t *short_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPCLR_B200_H */
