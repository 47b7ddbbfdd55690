/*
 * p2c_b200.h -- C ABI of the B200-native FK + projection + loc_2d_3d hot path.
 *
 * The reference (wielgosz-info/pedestrians-video-2-carla @ 410bba4) is pure Python and has
 * no FFI; each entry point below names the reference Python interface it replaces
 * (paths relative to src/pedestrians_video_2_carla/).  A maintainer binds these with
 * ctypes; the in-tree binding is pedestrians_video_2_carla_b200/_lib.py.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous row-major fp32 unless stated;
 *     the caller (PyTorch's caching allocator) owns all buffers incl. workspaces;
 *     the library never allocates, frees, or synchronises (one exception, at set-up time:
 *     the 8 KB IPC-shareable exchange block of p2c_exchange_create / _destroy).
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*); pass PyTorch's
 *     current stream so autograd ordering holds.
 *   - rotation matrices act on ROW vectors from the right (p' = p R + t); the bone axis is
 *     the DFS order of structure.yaml (walker_control/pose.py:41-47), parent[b] < b.
 *   - return value: 0 = ok, < 0 = bad argument (P2C_E_*), > 0 = cudaError_t of the launch.
 *     Never throws, never exits.  Stateless and re-entrant.
 */
#ifndef P2C_B200_H
#define P2C_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define P2C_MAX_BONES 32
#define P2C_E_BADARG (-1)
#define P2C_E_BONES (-2)   /* n_bones outside 1..32 or parent[] not in DFS order */
#define P2C_E_NULL (-3)    /* a required pointer is NULL */
#define P2C_E_UNSUPPORTED (-4)
#define P2C_E_TOO_LARGE (-5)   /* beyond the stated maximum of an entry point with built-in scratch (p2c_loc2d_step_static) */

/* Bone tree: walker_control/pose.py:29-47 (structure.yaml walked depth-first). */
typedef struct {
    int32_t n_bones;
    int32_t parent[P2C_MAX_BONES]; /* -1 for the root */
} p2c_tree_t;

/* Pinhole camera of walker_control/p3d_pose_projection.py:37-69 (pytorch3d PerspectiveCameras,
 * in_ndc=False): c = w R + T ; u = px - f c.x / c.z ; v = py - f c.y / c.z ; out.z = 1 / c.z */
typedef struct {
    float R[9];
    float T[3];
    float f, px, py;
} p2c_camera_t;

enum { P2C_MODE_POSE_CHANGES = 0, /* modules/layers/projection.py:170-195: rel_rot[t] = in[t] . rel_rot[t-1] */
       P2C_MODE_RELATIVE_ROT = 1  /* modules/layers/projection.py:144-168: rel_rot[t] = in[t]               */ };

/* ---- P3dPose.forward / relative_to_absolute over clips (walker_control/p3d_pose.py:98-213,
 *      looped by modules/layers/projection.py:144-195) with the optional fused
 *      P3dPoseProjection.forward (walker_control/p3d_pose_projection.py:115-152).           */
typedef struct {
    p2c_tree_t tree;
    int32_t n_clips, clip_len;  /* N, L ; P3dPose.forward itself is L = 1                          */
    int32_t mode;               /* P2C_MODE_*                                                       */
    const float* pose_in;       /* (N,L,B,3,3) changes (mode 0) or relative rotations (mode 1)      */
    const float* ref_loc;       /* (S,B,3)   relative bone locations, p3d space                     */
    const float* ref_rot;       /* (S,B,3,3) relative bone rotations before frame 0 (mode 0 only)   */
    const int32_t* ref_index;   /* (N) row of ref_* per clip, or NULL: row = clip index (S = N)     */
    const float* world_loc;     /* (N,L,3)   or NULL = zeros      } absolute world transform per    */
    const float* world_rot;     /* (N,L,3,3) or NULL = identity   } frame (utils/world.py output)   */
    p2c_camera_t camera;
    float* rel_rot;             /* out (N,L,B,3,3) or NULL                                          */
    float* abs_rot;             /* out (N,L,B,3,3) or NULL                                          */
    float* abs_loc;             /* out (N,L,B,3)   or NULL                                          */
    float* proj;                /* out (N,L,B,3)   or NULL (then world_* and camera are unused)     */
} p2c_fk_fwd_t;
int p2c_fk_fwd(const p2c_fk_fwd_t* p, void* stream);

/* Backward of the above (replaces torch autograd over p3d_pose.py:98-184 and
 * p3d_pose_projection.py:115-152).  Upstream gradients may each be NULL (= zero).           */
typedef struct {
    p2c_tree_t tree;
    int32_t n_clips, clip_len, mode;
    const float* pose_in;       /* saved input                                                      */
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;
    const float* world_rot;
    p2c_camera_t camera;
    const float* rel_rot;       /* saved output (N,L,B,3,3); NULL allowed in mode 1 (= pose_in)     */
    const float* abs_rot;       /* saved output (N,L,B,3,3)                                         */
    const float* abs_loc;       /* saved output (N,L,B,3); only read when g_proj != NULL            */
    const float* g_rel_rot;     /* upstream (N,L,B,3,3) or NULL                                     */
    const float* g_abs_rot;     /* upstream (N,L,B,3,3) or NULL                                     */
    const float* g_abs_loc;     /* upstream (N,L,B,3)   or NULL                                     */
    const float* g_proj;        /* upstream (N,L,B,3)   or NULL                                     */
    float* g_pose_in;           /* out (N,L,B,3,3)                                                  */
    float* g_ref_rot;           /* out (N,B,3,3) or NULL; needs ref_index == NULL, mode 0           */
    float* g_ref_loc;           /* out (N,B,3)   or NULL; needs ref_index == NULL                   */
    float* g_world_loc;         /* out (N,L,3)   or NULL                                            */
    float* g_world_rot;         /* out (N,L,3,3) or NULL                                            */
} p2c_fk_bwd_t;
int p2c_fk_bwd(const p2c_fk_bwd_t* p, void* stream);

/* ---- P3dPoseProjection.forward alone (walker_control/p3d_pose_projection.py:115-152):
 *      x (M,J,3), loc (M,3), rot (M,3,3) -> out (M,J,3).                                    */
int p2c_project_fwd(const float* x, const float* loc, const float* rot, const p2c_camera_t* cam,
                    float* out, int64_t m, int32_t j, void* stream);
int p2c_project_bwd(const float* x, const float* loc, const float* rot, const p2c_camera_t* cam,
                    const float* g_out, float* g_x, float* g_loc, float* g_rot,
                    int64_t m, int32_t j, void* stream);

/* ---- utils/world.py:16-63 calculate_world_from_changes: cumulative world transform per clip.
 *      init_* (N,3)/(N,3,3) may be NULL (zeros / identity).                                 */
int p2c_world_scan_fwd(const float* loc_change, const float* rot_change, const float* init_loc,
                       const float* init_rot, float* world_loc, float* world_rot,
                       int32_t n_clips, int32_t clip_len, void* stream);
int p2c_world_scan_bwd(const float* rot_change, const float* world_rot, const float* init_rot,
                       const float* g_world_loc, const float* g_world_rot,
                       float* g_loc_change, float* g_rot_change,
                       int32_t n_clips, int32_t clip_len, void* stream);

/* ---- pytorch3d rotation conversions at the reference's call sites
 *      (modules/movements/movements.py:113-116; walker_control/p3d_pose.py:54,78,259).      */
int p2c_rot6d_to_matrix_fwd(const float* d6, float* out, int64_t n, void* stream);
int p2c_rot6d_to_matrix_bwd(const float* d6, const float* g_out, float* g_d6, int64_t n, void* stream);
int p2c_euler_xyz_to_matrix(const float* euler, float* out, int64_t n, void* stream);
int p2c_matrix_to_euler_xyz(const float* mat, float* euler, int64_t n, void* stream);

/* ---- Normalizer(HipsNeckExtractor) (transforms/pose/normalization/normalizer.py:20-41,
 *      extractor.py:23-36, hips_neck_extractor.py:6-13): sample (M,J,C), C in {2,3}.        */
int p2c_hips_neck_normalize_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c,
                                int32_t hips, int32_t neck, float near_zero, void* stream);
int p2c_hips_neck_normalize_bwd(const float* sample, const float* g_out, float* g_sample, int64_t m,
                                int32_t j, int32_t c, int32_t hips, int32_t neck, float near_zero,
                                void* stream);

/* ---- nn.MSELoss(reduction='mean') over the first `c_used` channels with the missing-joints
 *      mask of utils/tensors.py:29-40 (loss/base_pose_loss.py:36-66, loss/loc_2d.py:69-89,
 *      loss/loc_3d.py:12-40).  pred (M,J,c_pred), gt (M,J,c_gt).  mask_hips < 0: no mask.
 *      Deterministic two-stage reduction: `partials` is a caller-owned workspace of
 *      p2c_mse_workspace_floats(m) floats; `out` receives [loss, sum_sq, count].            */
int64_t p2c_mse_workspace_floats(int64_t m);
int p2c_masked_mse_fwd(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_pred,
                       int32_t c_gt, int32_t c_used, int32_t mask_hips, float* partials, float* out,
                       void* stream);
int p2c_masked_mse_bwd(const float* pred, const float* gt, const float* out, const float* g_loss,
                       float* g_pred, int64_t m, int32_t j, int32_t c_pred, int32_t c_gt,
                       int32_t c_used, int32_t mask_hips, void* stream);

/* ---- The autoencoder flow's step on this path (modules/flow/autoencoder.py:102-112: the model emits 2-D poses, the
 *      loss is Loc2DPoseLoss, loss/loc_2d.py:69-89): the masked MSE of p2c_masked_mse_fwd over two channels AND
 *      d(upstream * loss)/d pred in ONE cooperative launch -- predictions and targets read once, gradient written once.
 *      pred, g_pred (M,J,2), 16-byte aligned; gt (M,J,c_gt), c_gt 2 (16-byte aligned) or 3 (8-byte aligned).
 *      workspace: p2c_loc2d_step_workspace_floats() floats whose first 16 are self-resetting counters (zero them once).
 *      out[4] = {loss, sum_sq, count, 0}.  exchange: as in p2c_fused_step_t (publishes {sum_sq, count, 0, 0}) or NULL.
 *      p2c_loc2d_step_supported == 0 (the errors of the batch do not fit on chip): use p2c_masked_mse_fwd / _bwd.      */
int64_t p2c_loc2d_step_workspace_floats(void);
int p2c_loc2d_step_supported(int64_t m, int32_t j, int32_t c_gt);
int p2c_loc2d_step(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_gt, int32_t mask_hips,
                   float upstream, float* g_pred, float* workspace, float* out, float* exchange,
                   int32_t exchange_rank, int32_t exchange_world, void* stream);

/* ---- p2c_loc2d_step without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 *      that forbid it).  Same kernel, arguments and results as p2c_loc2d_step (no workspace, no exchange); the scratch
 *      (barrier words + per-CTA partial sums, 8704 bytes) is a file-scope device array that every GPU loading the library
 *      holds until exit.  Calls no cudaMalloc / cudaFree / allocating routine.
 *      Limit: m*j <= p2c_loc2d_step_static_max_joints() = 3 788 800 joints (145 723 poses of 26 joints: what one B200
 *      keeps on chip); beyond it, or on a GPU too small for the batch, returns P2C_E_TOO_LARGE and launches nothing.
 *      All calls on a device SHARE the scratch: calls that overlap on the GPU (different streams, no ordering) race;
 *      calls on one stream, or ordered by events, are safe (the barrier words are cleared on the stream before every
 *      launch, the partial sums are overwritten before they are read).                                              */
int64_t p2c_loc2d_step_static_max_joints(void);
int p2c_loc2d_step_static(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c_gt, int32_t mask_hips,
                          float upstream, float* g_pred, float* out, void* stream);

/* ---- One pose_lifting step of the hot path, fused: ProjectionModule.forward
 *      (modules/layers/projection.py:73-123) -> Normalizer (modules/flow/pose_lifting.py:167-170) ->
 *      loc_2d, loc_3d, loc_2d_3d (modules/flow/base.py:440-469; loss/loc_2d.py, loc_3d.py, loc_2d_3d.py)
 *      AND the gradient of  weight_2d*loc_2d + weight_3d*loc_3d  w.r.t. pose_in, so that
 *      loss.backward() only has to scale g_pose_in by the upstream scalar.
 *      losses[8] = { loc_2d, loc_3d, weight_2d*loc_2d + weight_3d*loc_3d, sum_sq_2d, count_2d,
 *                    sum_sq_3d, count_3d, 0 }  (means are LOCAL to this call: DDP semantics).     */
typedef struct {
    p2c_tree_t tree;
    int32_t hips, neck;          /* joints used by HipsNeckExtractor and never masked               */
    int32_t n_clips, clip_len;
    int32_t mode;                /* P2C_MODE_*                                                       */
    int32_t input_6d;            /* 1: pose_in / g_pose_in are (N,L,B,6) rotation_6d (movements.py:113-116) */
    int32_t normalize;           /* 1: hips/neck normalise the projection before loc_2d              */
    int32_t mask_missing_joints; /* utils/tensors.py:29-40                                           */
    const float* pose_in;
    const float* ref_loc;
    const float* ref_rot;
    const int32_t* ref_index;
    const float* world_loc;      /* (N,L,3) or NULL                                                  */
    const float* world_rot;      /* (N,L,3,3) or NULL                                                */
    p2c_camera_t camera;
    const float* gt_2d;          /* targets['projection_2d(_transformed)'] (N,L,B,gt_2d_channels)    */
    int32_t gt_2d_channels;      /* >= 2                                                             */
    const float* gt_3d;          /* targets['absolute_pose_loc'] (N,L,B,3)                           */
    float weight_2d, weight_3d;
    float* rel_rot;              /* out (N,L,B,3,3)                                                  */
    float* abs_rot;              /* out (N,L,B,3,3)                                                  */
    float* abs_loc;              /* out (N,L,B,3)                                                    */
    float* proj;                 /* out (N,L,B,3)                                                    */
    float* proj_norm;            /* out (N,L,B,3) normalised projection or NULL                      */
    float* g_pose_in;            /* out, same shape as pose_in                                       */
    float* workspace;            /* p2c_fused_step_workspace_floats(N*L, B, input_6d) floats; its first 16 floats
                                    are counters the kernels maintain themselves: zero them ONCE before the first
                                    call and leave them alone afterwards (every call leaves its counters zero again;
                                    word 5 is a launch sequence that keeps counting); do not share one workspace
                                    between concurrent streams                                                    */
    float* losses;               /* out, 8 floats (12 with gt_abs_rot, see below)                    */
    /* data-parallel ranks (one process per GPU, clips sharded by batch): when `exchange` is a block made by
       p2c_exchange_create + p2c_exchange_set_peers, the kernel that finishes the loss sums also stores this rank's
       {sum_sq_2d, count_2d, sum_sq_3d, count_3d} into every rank's block (peer stores over NVLink; no collective
       kernel, nothing waits for a peer).  NULL: single rank.  One block per stream, like the workspace.            */
    float* exchange;
    int32_t exchange_rank, exchange_world;
    /* rot_3d fused (loss/rot_3d.py:9-37; loc_rot_3d, loc_2d_loc_rot_3d, weighted_loc_2d_loc_rot_3d of loss/__init__.py:
       28-48): when gt_abs_rot = targets['absolute_pose_rot'] (N,L,B,3,3) is given, the step also forms
       rot_3d = MSE(abs_rot, gt_abs_rot), adds weight_rot * rot_3d to losses[2] and its gradient to g_pose_in, and
       `losses` must hold 12 floats: [8] rot_3d, [9] sum_sq_rot, [10] count_rot, [11] 0.  NULL: as before (8 floats). */
    const float* gt_abs_rot;
    float weight_rot;
} p2c_fused_step_t;
int64_t p2c_fused_step_workspace_floats(int64_t n_poses, int32_t n_bones, int32_t input_6d);
int p2c_fused_step(const p2c_fused_step_t* p, void* stream);
/* number of kernels p2c_fused_step(p) enqueues (for launch accounting; < 0 = bad argument) */
int p2c_fused_step_launches(const p2c_fused_step_t* p);

/* ---- The path's only cross-GPU exchange (replaces Lightning's DDP-side reduction of the logged losses,
 *      modules/flow/base.py:467-468 with sync_dist; README.md:74-75 run shape): every rank owns one small block in its
 *      own HBM, maps the blocks of its peers (CUDA IPC), and the step kernels write their four loss sums into all of
 *      them.  p2c_exchange_create allocates the block (the ONE allocation this library makes: IPC-shareable memory
 *      cannot come from a sub-allocating caller) and returns its 64-byte IPC handle for the peers;
 *      p2c_exchange_open maps a peer's block; p2c_exchange_set_peers stores the table of mapped pointers (entry r =
 *      rank r's block as this process addresses it; entry `rank` = the own block); p2c_exchange_read sums the entries
 *      of step `seq` (1-based count of steps published through the block; 0 = the latest) over the ranks IN RANK ORDER:
 *      out[8] = {sum_sq_2d, count_2d, sum_sq_3d, count_3d, ranks whose entry carries `seq`, global loc_2d,
 *      global loc_3d, seq}.  The reader never waits: a rank that has not published yet is simply not counted (the
 *      caller orders its read after the peers' steps, e.g. behind the barrier it already has).                       */
int64_t p2c_exchange_block_bytes(void);
int32_t p2c_exchange_max_ranks(void);
int p2c_exchange_create(void** block, unsigned char* ipc_handle /* 64 bytes out, or NULL */);
int p2c_exchange_open(const unsigned char* ipc_handle, void** peer_block);
int p2c_exchange_close(void* peer_block);
int p2c_exchange_destroy(void* block);
int p2c_exchange_set_peers(void* block, void* const* peer_blocks, int32_t world);
int p2c_exchange_read(const void* block, int32_t world, uint32_t seq, float* out, void* stream);
/* The step kernels on the latency-critical path (p2c_fused_step's tile kernel, p2c_loc2d_step) leave their sums in the
 * rank's own block and SEND them at the start of the next step's kernel (remote stores on a kernel's tail delay its
 * completion by an NVLink round trip).  p2c_exchange_flush sends what is still waiting: enqueue it behind the last
 * step, before the synchronisation point that precedes p2c_exchange_read.                                           */
int p2c_exchange_flush(void* block, int32_t rank, int32_t world, void* stream);

/* ---- pytorch3d 0.6.0 quaternion_to_matrix (real part first, not normalised first): the quaternion entry the
 *      north star names next to euler_angles_to_matrix (walker_control/p3d_pose.py:54) and rotation_6d_to_matrix
 *      (modules/movements/movements.py:113-116); the reference itself has no quaternion call site.
 *      quat (n,4) -> out (n,3,3); both quaternion pointers must be 16-byte aligned.                             */
int p2c_quaternion_to_matrix_fwd(const float* quat, float* out, int64_t n, void* stream);
int p2c_quaternion_to_matrix_bwd(const float* quat, const float* g_out, float* g_quat, int64_t n, void* stream);

/* ---- loss/cum_pose_changes.py:9-56: per-bone cumulative product of the pose changes over the frames of a clip,
 *      cumulative[n,t,b] = cumulative[n,t-1,b] . changes[n,t,b] (identity before frame 0); (N,L,B,3,3) both.
 *      The loss itself is p2c_masked_mse_fwd over 9 channels of the two scans (prediction and target).          */
int p2c_cum_changes_fwd(const float* changes, float* cumulative, int32_t n_clips, int32_t clip_len,
                        int32_t n_bones, void* stream);
int p2c_cum_changes_bwd(const float* changes, const float* cumulative, const float* g_cumulative,
                        float* g_changes, int32_t n_clips, int32_t clip_len, int32_t n_bones, void* stream);

/* ---- Normalizer.__call__(sample, dim) (transforms/pose/normalization/normalizer.py:20-41) for dim 2 or 3, optionally
 *      followed by DeNormalizer.from_reference (denormalizer.py:13-37) with a per-clip reference shift / scale, i.e.
 *      ReferenceSkeletonsDeNormalizer.from_abs / from_projection (reference_skeletons_denormalizer.py:41-91), the
 *      entry of the absolute_loc(_rot) movements output (modules/layers/projection.py:125-142).
 *      sample, out (M,J,c), c == dim or (dim 2, c 3: channel 2 carried).  autonormalize 0: denormalise only.
 *      ref_shift (rows,dim) / ref_scale (rows): both NULL = normalise only.  Pose m belongs to clip m / clip_len;
 *      its row is ref_index[clip] (int32) or, with ref_index NULL, the clip itself.                              */
int p2c_renormalize_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c, int32_t dim,
                        int32_t hips, int32_t neck, int32_t autonormalize, float near_zero,
                        const float* ref_shift, const float* ref_scale, const int32_t* ref_index,
                        int32_t clip_len, void* stream);
int p2c_renormalize_bwd(const float* sample, const float* g_out, float* g_sample, int64_t m, int32_t j,
                        int32_t c, int32_t dim, int32_t hips, int32_t neck, int32_t autonormalize,
                        float near_zero, const float* ref_scale, const int32_t* ref_index, int32_t clip_len,
                        void* stream);

/* ---- Normalizer.__call__(sample, dim=2) with BBoxExtractor (extractor 1; bbox_extractor.py:6-18) or
 *      HipsNeckBBoxFallbackExtractor (extractor 2; hips_neck_bbox_fallback_extractor.py:10-46, the reference's default
 *      `hips_neck_bbox` transform, base_datamodule.py:61,276).  Forward only: these normalise dataset inputs.
 *      sample, out (M,J,c), c 2 or 3 (channel 2 = confidence, carried; xy zeroed where it is < near_zero).      */
int p2c_normalize_extractor_fwd(const float* sample, float* out, int64_t m, int32_t j, int32_t c,
                                int32_t hips, int32_t neck, int32_t extractor, float near_zero, void* stream);

/* ---- Extractor.get_shift_scale alone (transforms/pose/normalization/extractor.py:23-36 with hips_neck_extractor.py,
 *      bbox_extractor.py, hips_neck_bbox_fallback_extractor.py): the values Normalizer.shift / Normalizer.scale report
 *      after a call (normalizer.py:43-49; read by data/base/mixins/dataset/projection_2d_mixin.py:229-230).
 *      sample (M,J,c); shift out (M,dim), scale out (M); extractor 0 = hips/neck (dim 2 or 3), 1 = bbox, 2 = hips_neck_bbox. */
int p2c_extractor_shift_scale(const float* sample, float* shift, float* scale, int64_t m, int32_t j, int32_t c,
                              int32_t dim, int32_t hips, int32_t neck, int32_t extractor, float near_zero,
                              void* stream);

/* ---- Evaluation metrics on the step's output buffers: MPJPE (metrics/mpjpe.py:29-45), MRPE (metrics/mrpe.py:38-76),
 *      PCK (metrics/pck.py:54-97 with utils/tensors.py:12-40).  pred / gt (M,J,c).  `partials`: caller-owned workspace
 *      of p2c_metric_workspace_floats() floats; `out`: two DOUBLES on the device, (sum of distances, number of
 *      distances) or (correct, total); accumulate != 0 adds to them (the metric's running state), 0 overwrites.
 *      MRPE: *_world (M,c) or NULL (zeros).  PCK: mask_src (M,J,c_mask) = targets['projection_2d'] or NULL (no mask),
 *      mask_hips < 0: hips not exempt; norm_kind 0 = hips-neck distance of gt, 1 = bounding-box diagonal of gt.     */
int64_t p2c_metric_workspace_floats(void);
int p2c_mpjpe_sum(const float* pred, const float* gt, int64_t m, int32_t j, int32_t c, float* partials,
                  double* out, int32_t accumulate, void* stream);
int p2c_mrpe_sum(const float* pred, const float* gt, const float* pred_world, const float* gt_world,
                 int64_t m, int32_t j, int32_t c, int32_t hips, float* partials, double* out,
                 int32_t accumulate, void* stream);
int p2c_pck_count(const float* pred, const float* gt, const float* mask_src, int64_t m, int32_t j,
                  int32_t c, int32_t c_mask, int32_t mask_hips, int32_t hips, int32_t neck,
                  int32_t norm_kind, float threshold, float near_zero, float* partials, double* out,
                  int32_t accumulate, void* stream);

/* ---- The random draws of the synthetic-target generator (Carla2D3DIterableDataset.generate_batch,
 *      data/carla/datasets/carla_2d3d_dataset.py:145-183) on the device: per frame `changes_per_frame` distinct bones get
 *      Euler changes uniform in +-max_change (rad), world yaw changes uniform in +-max_initial_world_change (frame 0) /
 *      +-max_world_change (later frames), one skeleton type in [0, n_types) per clip.  Counter-based (Philox4x32-10):
 *      stateless, `offset` = the batch counter.  euler (N,L,B,3), world_euler (N,L,3), skeleton_index (N) int32 or NULL. */
int p2c_draw_sparse_changes(uint64_t seed, uint64_t offset, int32_t n_clips, int32_t clip_len, int32_t n_bones,
                            int32_t changes_per_frame, float max_change, float max_world_change,
                            float max_initial_world_change, int32_t n_types, float* euler, float* world_euler,
                            int32_t* skeleton_index, void* stream);

const char* p2c_version(void);
/* test hook: re-read the P2C_* debug switches from the environment (they are read once per process otherwise) */
void p2c_debug_reload_env(void);

#ifdef __cplusplus
}
#endif
#endif /* P2C_B200_H */
