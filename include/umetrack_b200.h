/*
 * umetrack_b200 -- C ABI of the B200 (sm_100a) implementation of UmeTrack's per-frame
 * multi-view inference path.
 *
 * The reference (facebookresearch/UmeTrack) is pure Python and has no native / FFI boundary;
 * its drop-in surface is the Python API in lib/tracker and lib/models.  This library sits
 * underneath the Python mirror of that API (package `umetrack_b200`), which binds it with
 * ctypes.  Each entry point below cites the reference function(s) it replaces
 * (file:line relative to the reference checkout).
 *
 * Conventions
 *   - every function returns 0 on success and a non-zero code on failure; the message is
 *     available from umt_last_error() (thread-local);
 *   - unless a parameter says "host", pointers are DEVICE pointers owned by the caller;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *   - no function synchronises the device; nothing is copied to the host;
 *   - except umt_model_create / umt_model_destroy, no function allocates memory.
 */
#ifndef UMETRACK_B200_H
#define UMETRACK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UMT_OK 0
#define UMT_ERR_INVALID 1
#define UMT_ERR_CUDA 2
#define UMT_ERR_UNSUPPORTED 3

#define UMT_CROP 96            /* crop edge length (lib/models/model_loader.py:56) */
#define UMT_RESAMPLE_PARAMS 48 /* doubles per crop for umt_resample_crops */
#define UMT_N_JOINTS 22
#define UMT_N_LANDMARKS 21

/* numeric mode of the encoder convolutions */
#define UMT_MODE_FP32_EQUIV 0 /* split-bf16 (hi+lo) operands, 3 tensor-core products, fp32 accumulate */
#define UMT_MODE_BF16 1       /* single bf16 product, fp32 accumulate */
#define UMT_MODE_FP16 2       /* single fp16 product (11-bit significands), fp32 accumulate; activations and
                                 BN-folded weights must stay inside the half range (|x| < 65504) */

/* regression head */
#define UMT_HEAD_KNOWN_SKELETON 0   /* _regressor_k, regress_pose_use_skeleton */
#define UMT_HEAD_UNKNOWN_SKELETON 1 /* _regressor_u, regress_pose_pred_skel_scale */

const char* umt_last_error(void);
int umt_abi_version(void);

/* ------------------------------------------------------------------------------------------
 * Stage (a): fisheye -> pinhole crop resampling.
 * Replaces `_warp_image` (lib/tracker/tracker.py:54-80): the float64 camera chain
 * window_to_eye / eye_to_world / world_to_eye / eye_to_window (lib/common/camera.py:257-281,
 * 68-85, 120-141), the float32 cast of the map, and cv2.remap(uint8, INTER_LINEAR,
 * BORDER_CONSTANT 0) in its 1/32-pixel fixed-point form; plus the `/255` conversion of
 * `_make_inputs` (tracker.py:211).
 *
 * params[n_crops][48] (float64): src fx fy cx cy | k1 k2 k3 k4 p1 p2 k5 k6 |
 *   src camera_to_world (16, row major, mm) | crop fx fy cx cy | crop camera_to_world (16).
 * Source pixel (x,y) of crop i is src_base[src_offset[i] + y*src_pitch + x]; this expresses
 * the reference's strided view frame[:, cam, :] of a [H, 4, W] frame.
 * Any of out_u8 / out_f32 / out_idx may be NULL.  out_idx receives (ix, iy, fx, fy) per
 * pixel: integer tap origin and 5-bit fractions, the quantities that must be bit-exact.
 */
int umt_resample_crops(const uint8_t* src_base, const int64_t* src_offset, int src_pitch,
                       int src_w, int src_h, const double* params, int n_crops,
                       uint8_t* out_u8 /* [n,96,96] */, float* out_f32 /* [n,96,96] */,
                       int32_t* out_idx /* [n,96,96,4] */, void* stream);

/* Batched crop-camera generation (float64), one (frame, hand) item per thread.  Replaces the host
 * loop HandTracker.gen_crop_cameras (lib/tracker/tracker.py:101-139) and below it
 * rank_hand_visibility_in_cameras / gen_crop_cameras_from_pose (lib/tracker/perspective_crop.py:54-162,
 * with sort_camera_index=True and at most 2 views), gen_crop_parameters_from_points
 * (lib/common/crop.py:15-82) and make_look_at_matrix (lib/common/affine.py:34-66).
 * crop_points [n,n_points,3] float32: FK landmarks of the given, neutral and open pose (the first 21
 * are tested for visibility); frame_idx [n] selects the frame's cameras in cam_to_world
 * [n_frames,n_cam,16] (mm); cam_intr [n_cam,12] = fx fy cx cy k1 k2 k3 k4 p1 p2 k5 k6; valid [n] may
 * be NULL.  Outputs per item: sel_count [n] (0..2 cameras chosen), sel_cam [n,2], and per chosen
 * camera the parameter row of umt_resample_crops (resample_params [n,2,48]) and the model's
 * float32 intrinsics [n,2,3,3] / extrinsics [n,2,4,4] (world -> crop eye, metres).  *error_flag is
 * set to 1 where the reference raises ValueError("Unable to create crop camera"), 2 on a
 * singular camera matrix. */
int umt_crop_cameras(const float* crop_points, const int32_t* frame_idx, const uint8_t* hand_is_right,
                     const uint8_t* valid, const double* cam_intr, const double* cam_to_world,
                     const double* cam_angles_deg, int n_cam, int cam_w, int cam_h, int n, int n_points,
                     int min_vis, double focal_multiplier, int32_t* sel_count, int32_t* sel_cam,
                     double* resample_params, float* intr_out, float* extr_out, int32_t* error_flag,
                     void* stream);

/* torch_data variant of stage (a): pinhole -> pinhole homography resampling of float32 images.
 * Replaces `_resample_images_batched` (lib/batched_dataset/data_transform.py:78-144) and, with
 * divide_by_255 != 0, the `/ 255` of `_perspective_crop_images` (:279).  src [n,src_h,src_w]
 * float32; xfs [n,4,4] float32 = K_orig * E_orig * E_new^-1 * K_new^-1 (`_compute_resample_matrix`,
 * :57-75) taking (u, v, 1) of the new image to the original; coordinates are evaluated in float64
 * like numpy's float32 x int32 promotion; pixels whose source falls outside 0 <= x < w-1,
 * 0 <= y < h-1 are 0.  out [n,out_h,out_w], out_w a multiple of 4. */
int umt_resample_homography(const float* src, int src_w, int src_h, const float* xfs, int n,
                            int out_w, int out_h, int divide_by_255, float* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Stage (b): per-view encoder, multi-view fusion, temporal memory, pose regression.
 * Replaces UmeTrackModel.regress_pose_use_skeleton / regress_pose_pred_skel_scale
 * (lib/models/umetrack_model.py:188-242) and everything they call: the ResNet backbone
 * (backbone_resnet.py:14-158, model_utils.py:107-138), FTL + fusion
 * (feature_extractor.py:61-141, model_utils.py:57-104,141-192), SimpleConvRNN
 * (temporal.py:51-139), SkeletonEncoder (skeleton_encoder.py:28-53), PoseRegressor and its
 * decoders (regressor.py:76-186, model_utils.py:17-54) and _recover_wrist_xfs_in_world
 * (umetrack_model.py:77-90).
 */
typedef struct umt_model umt_model;

/* Number / name / element count of the weight tensors umt_model_create expects, in order:
 * the reference's flat state_dict (lib/models/model_loader.py:84-87) without the
 * `num_batches_tracked` counters. */
int umt_model_num_tensors(void);
const char* umt_model_tensor_name(int i);
int64_t umt_model_tensor_numel(int i);

/* host_tensors: HOST pointers to float32 arrays, one per tensor above.  BatchNorm is folded,
 * weights are split into bf16 hi/lo pairs, packed for the tensor-core kernels and uploaded to
 * the current device (the only device allocations this library makes). */
int umt_model_create(const float* const* host_tensors, int n_tensors, umt_model** out);
void umt_model_destroy(umt_model* m);

/* Scratch memory (encoder activations of one chunk of `chunk_images` crops, per-batch tail
 * buffers).  The caller allocates `bytes` on the device (256-byte aligned) and binds it; binding
 * clears it once (the kernels rely on zero halo positions inside it).  max_images / max_samples
 * bound n_img / n_samples of later calls; batches larger than chunk_images run the encoder in
 * several passes over the same scratch. */
int umt_model_workspace_bytes(const umt_model* m, int chunk_images, int max_images,
                              int max_samples, size_t* bytes);
int umt_model_bind_workspace(umt_model* m, void* workspace, size_t bytes, int chunk_images,
                             int max_images, int max_samples, void* stream);

/* One forward pass over n_samples hand-frames built from n_img crops.
 * Inputs as in InputFrameData / InputFrameDesc / InputSkeletonData
 * (lib/models/umetrack_model.py:21-74): images [n_img,96,96] in [0,1]; intrinsics [n_img,3,3];
 * extrinsics [n_img,4,4] (world -> crop eye, metres); sample_range [n_samples,2] (1 or 2
 * views per sample); memory_idx, hand_idx [n_samples]; use_memory [n_samples] (bytes).
 * skel_axes / skel_rest [n_skel,22,3] with n_skel in {1, n_samples} (known-skeleton head
 * only).  mem_state [n_slots,18,6,6] and prev_extr [n_slots,4,4] are the recurrent state of
 * SimpleConvRNN (temporal.py:41-42), updated in place.
 * Outputs: joint_angles [n_samples,22]; wrist [n_samples,4,4] in world space (right hands
 * un-mirrored); scale [n_samples] (unknown-skeleton head, else may be NULL); sigma
 * [n_samples,21]. */
int umt_model_forward(umt_model* m, int mode, int head, const float* images,
                      const float* intrinsics, const float* extrinsics,
                      const int64_t* sample_range, const int64_t* memory_idx,
                      const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                      int n_samples, const float* skel_axes, const float* skel_rest, int n_skel,
                      float* mem_state, float* prev_extr, int n_slots, float* out_joint_angles,
                      float* out_wrist, float* out_scale, float* out_sigma, void* stream);

/* umt_model_forward on uint8 crops [n_img,96,96] (16-byte aligned), as cv2.remap leaves them: the
 * `crop.astype(np.float32) / 255.0` of HandTracker._make_inputs (lib/tracker/tracker.py:211) runs on
 * the device (IEEE division, identical values), so a host feeding crops uploads one byte per pixel
 * instead of four. */
int umt_model_forward_u8(umt_model* m, int mode, int head, const uint8_t* images_u8,
                         const float* intrinsics, const float* extrinsics,
                         const int64_t* sample_range, const int64_t* memory_idx,
                         const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                         int n_samples, const float* skel_axes, const float* skel_rest, int n_skel,
                         float* mem_state, float* prev_extr, int n_slots, float* out_joint_angles,
                         float* out_wrist, float* out_scale, float* out_sigma, void* stream);

/* Everything after the backbone: same arguments as umt_model_forward with the per-crop
 * `features` [n_img,72,6,6] (from umt_model_encode) in place of the images.  The backbone has no
 * temporal dependency, so a whole recording can be encoded in one batch and this entry point
 * stepped over its frames (fusion, lib/models/feature_extractor.py:96-141; SimpleConvRNN,
 * temporal.py:93-139; regressor and decoders, regressor.py:163-186). */
int umt_model_tail(umt_model* m, int head, const float* features, const float* intrinsics,
                   const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                   const uint8_t* use_memory, const int64_t* hand_idx, int n_img, int n_samples,
                   const float* skel_axes, const float* skel_rest, int n_skel, float* mem_state,
                   float* prev_extr, int n_slots, float* out_joint_angles, float* out_wrist,
                   float* out_scale, float* out_sigma, void* stream);

/* umt_model_tail stepped over a whole time scan in one call: the descriptor arrays (sample_range,
 * memory_idx, use_memory, hand_idx, and the outputs) hold the rows of ALL steps back to back; step s
 * covers rows [step_first[s], step_first[s] + step_rows[s]) (HOST int arrays; empty steps are skipped)
 * and runs after step s-1 on the same stream, so the recurrent state carries over exactly as in a
 * per-frame loop (lib/tracker/tracker.py:141-168 called once per frame).  per_row_skeletons != 0:
 * skel_axes / skel_rest are [rows,22,3]; else one skeleton [1,22,3] for every row.  Saves the host
 * round trip per frame: the scan of a recording is launch bound. */
int umt_model_tail_scan(umt_model* m, int head, const float* features, const float* intrinsics,
                        const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                        const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                        const int* step_first, const int* step_rows, int n_steps, const float* skel_axes,
                        const float* skel_rest, int per_row_skeletons, float* mem_state, float* prev_extr,
                        int n_slots, float* out_joint_angles, float* out_wrist, float* out_scale,
                        float* out_sigma, void* stream);

/* Backbone only: images [n_img,96,96] -> features [n_img,72,6,6]
 * (`_image_backbone`, lib/models/model_utils.py:107-138).  Uses the bound workspace. */
int umt_model_encode(umt_model* m, int mode, const float* images, int n_img, float* features,
                     void* stream);

/* umt_model_encode on uint8 crops (see umt_model_forward_u8). */
int umt_model_encode_u8(umt_model* m, int mode, const uint8_t* images_u8, int n_img, float* features,
                        void* stream);

/* Backbone for callers that cannot allocate device memory.  [No counterpart in the reference,
 * which is Python and lets torch allocate.]  Same result as umt_model_encode for n_img <= 4
 * crops (one tracker frame: two hands x two views), but it needs no bound workspace and makes
 * no device allocation: its scratch is a 16 MiB `static __device__` array inside this library,
 * resident on every GPU that loads the library until the process exits.  Meant for hosts that
 * created the model at start-up and later run where cudaMalloc is off limits (frozen memory
 * budget, a runtime that owns the allocator, a loop that cannot afford cudaMalloc's device-wide
 * synchronisation).  n_img > 4 returns UMT_ERR_INVALID.  The first call on a device resolves the
 * array's device address: make it at start-up with that device current and before any stream
 * capture.  Calls share the one array and clear it first, so calls whose GPU work may overlap
 * (other streams or threads) race: serialise them. */
int umt_model_encode_static(umt_model* m, int mode, const float* images, int n_img, float* features,
                            void* stream);

/* Number of kernels the last umt_model_forward / umt_model_encode call launched. */
int umt_model_last_launch_count(const umt_model* m);

/* Sticky status bits that kernels set on the device since the last call of this function; reads
 * them into *host_status (HOST pointer), clears them, and SYNCHRONISES `stream` (with
 * umt_model_profile_layers the only entry points that do).  The reference raises a Python exception
 * (or grows its memory bank, lib/models/temporal.py:101-123) where this library, which never reads
 * device data on the host, can only record the condition. */
#define UMT_STATUS_MEMORY_INDEX 1u  /* a memory_idx outside [0, n_slots): that sample ran without memory */
#define UMT_STATUS_FP16_RANGE 2u    /* UMT_MODE_FP16: an activation left the half range (result has inf) */
int umt_model_status(umt_model* m, unsigned* host_status, void* stream);

/* Measurement hook: one encoder pass over n_img crops (one chunk) with a CUDA event between
 * consecutive launches; out_ms is a HOST float[26]: stem, the 24 convolutions in execution order,
 * projection.  Synchronises the stream (this is the only entry point that does). */
int umt_model_profile_layers(umt_model* m, int mode, const float* images, int n_img, float* out_ms,
                             void* stream);

/* Test hook: run the encoder on `images` [n_img,96,96] up to convolution `layer` (0..23,
 * execution order), then evaluate that layer twice on the same inputs: with plain CUDA-core
 * arithmetic into out_simt and with the tensor-core kernel into out_umma, both as dense float32
 * [n_img, C, H, W] (no halo), so the tensor-core path can be checked layer by layer.  Either
 * output may be NULL.  out_shape is a HOST int[3] receiving C, H, W. */
int umt_model_debug_layer(umt_model* m, int mode, int layer, const float* images, int n_img,
                          float* out_simt, float* out_umma, int* out_shape, void* stream);

/* Test hook (allocates and synchronises): the convolution kernels rely on rows of the bound
 * workspace that are never written and must read as zero -- the halo positions of the padded raster
 * (which implement the zero padding of backbone_resnet.py:38,49 and separate neighbouring crops), the
 * guard rows around every plane and the padding beyond the workspace's capacity.  Counts the 16-byte
 * rows among them that are not zero into *host_nonzero_rows (HOST pointer). */
int umt_model_debug_halo_check(umt_model* m, long long* host_nonzero_rows, void* stream);

/* ------------------------------------------------------------------------------------------
 * Stage (c): forward kinematics + skinning, and metric reductions.
 * umt_fk_landmarks replaces skin_landmarks (lib/common/hand_skinning.py:17-209) including the
 * right-hand column flip of landmarks_from_hand_pose (lib/tracker/perspective_crop.py:40-51).
 * Hand model arrays are float32: axes/rest [22,3], lm_rest [21,3], lm_idx [21,3] (stored as
 * float like the reference's torch.Tensor(list)), lm_w [21,3].  mirror [n] (bytes) may be
 * NULL.  units follow the inputs (mm in the tracker).
 */
int umt_fk_landmarks(const float* axes, const float* rest, const float* lm_rest,
                     const float* lm_idx, const float* lm_w, const float* joint_angles /* [n,22] */,
                     const float* wrist /* [n,4,4] */, const uint8_t* mirror, int n,
                     float* out /* [n,21,3] */, void* stream);

/* sums[0] += sum over valid i of mean_21 |gt_i - pred_i|_2 ; sums[1] += number of valid i.
 * (run_eval_known_skeleton.py:88-89,117-119).  pred/gt [n,21,3] float32; sums float64[2]. */
int umt_mpjpe_accumulate(const float* pred, const float* gt, const uint8_t* valid, int n,
                         double* sums, void* stream);

/* Keypoint acceleration (load_eval.py:29-44): pts [n_seq, T, 21, 3], valid [n_seq, T];
 * sums[0] += sum of mean_21 |p[t] + p[t+2] - 2 p[t+1]| over t with three consecutive valid
 * frames; sums[1] += their count. */
int umt_accel_accumulate(const float* pts, const uint8_t* valid, int n_seq, int T, double* sums,
                         void* stream);

/* Per-hand-frame keypoint error err[i] = mean_21 |gt_i - pred_i|_2 in float64, evaluated in the
 * order numpy uses for np.linalg.norm(d, axis=-1).mean(axis=-1) (run_eval_known_skeleton.py:88-89,
 * load_eval.py:33-34: ((dx^2+dy^2)+dz^2, eight partial sums over the 21 landmarks), so the values
 * are bit-identical to the reference's for float32-valued keypoints.  pred/gt [n,21,3] float32. */
int umt_keypoint_errors(const float* pred, const float* gt, int n, double* err, void* stream);

/* PCK counts (lib/common/metric_utils.py:36-88 as used by load_eval.py:66-72):
 * counts[k] += number of i with valid[i] (NULL = all) and err[i] <= thresholds[k].
 * thresholds: ascending float64 [n_thresholds <= 1024]; counts float64 [n_thresholds]
 * (integers, exact below 2^53; float64 so that they ride in the same all-reduce as the sums). */
int umt_pck_accumulate(const double* err, const uint8_t* valid, long long n, const double* thresholds,
                       int n_thresholds, double* counts, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* UMETRACK_B200_H */
