"""Per-frame tracker: the reference's `HandTracker` API with the three stages on the GPU.

Names and behaviour follow lib/tracker/tracker.py:33-291, lib/tracker/perspective_crop.py:19-162
and lib/tracker/tracking_result.py:14-30: tracker I/O is in millimetres, model I/O in metres;
the model predicts left hands, right-hand crops are x-mirrored by the crop camera and the
result un-mirrored; an empty `crop_cameras` resets the history and returns an empty result.

What runs where: for the usual rig (FishEye62 views of one size) crop-camera generation runs on the
device (`umt_crop_cameras`, float64) and only the chosen cameras' parameters come back to build the
PinholePlaneCameraModel objects the API returns; any other camera model takes the host construction
(`crop_camera.py`).  The landmark FK, the crop resampling (`_warp_image`), the network and the result
FK run in libumetrack_b200.so.  A frame goes through persistent pinned / device buffers
(`_FrameBuffers`): two uploads, two reads, one CUDA-graph replay, no allocation.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, NamedTuple, Optional, Tuple

import numpy as np
import torch

from . import native, nvtx
from .cameras import CameraModel, PinholePlaneCameraModel, resample_params
from .crop_camera import gen_crop_parameters_from_points
from .hand_model import HandModel, NUM_HANDS, NUM_JOINTS_PER_HAND, RIGHT_HAND_INDEX, scaled_hand_model
from .model import InputFrameData, InputFrameDesc, InputSkeletonData, RegressorOutput
from .skinning import fk_landmarks_device

MM_TO_M = 0.001
M_TO_MM = 1000.0
CONFIDENCE_THRESHOLD = 0.5
MAX_VIEW_NUM = 2


class SingleHandPose(NamedTuple):
    """joint angles (22 DoF) + root-to-world wrist transform (mm)."""

    joint_angles: np.ndarray = np.zeros(NUM_JOINTS_PER_HAND, dtype=np.float32)
    wrist_xform: np.ndarray = np.eye(4, dtype=np.float32)
    hand_confidence: float = 1.0


class TrackingResult(NamedTuple):
    hand_poses: Dict[int, SingleHandPose] = {}
    num_views: Dict[int, int] = {}
    predicted_scales: Dict[int, float] = {}


@dataclass
class ViewData:
    image: np.ndarray
    camera: CameraModel
    camera_angle: float


@dataclass
class InputFrame:
    views: List[ViewData]


@dataclass
class HandTrackerOpts:
    num_crop_points: int = 63
    enable_memory: bool = True
    use_stored_pose_for_crop: bool = True
    hand_ratio_in_crop: float = 0.95
    min_required_vis_landmarks: int = 19


def neutral_joint_angles(hand_model: HandModel, lower_factor: float = 0.5) -> torch.Tensor:
    lim = hand_model.joint_limits
    assert lim is not None
    return lim[..., 0] * lower_factor + lim[..., 1] * (1 - lower_factor)


def landmarks_from_poses(hand_model: HandModel, joint_angles: np.ndarray, wrists: np.ndarray,
                         hand_indices) -> np.ndarray:
    """Batched `landmarks_from_hand_pose`: [n,22], [n,4,4], n hand indices -> [n,21,3] (host)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    ja = torch.from_numpy(np.ascontiguousarray(joint_angles, dtype=np.float32)).to(dev)
    wr = torch.from_numpy(np.ascontiguousarray(wrists, dtype=np.float32)).to(dev)
    mirror = torch.tensor([1 if h == RIGHT_HAND_INDEX else 0 for h in hand_indices], dtype=torch.uint8, device=dev)
    return fk_landmarks_device(hand_model, ja, wr, mirror).cpu().numpy()


def landmarks_from_hand_pose(hand_model: HandModel, hand_pose: SingleHandPose, hand_idx: int) -> np.ndarray:
    """World-space landmarks [21,3] of one pose (lib/tracker/perspective_crop.py:40-51)."""
    return landmarks_from_poses(hand_model, np.asarray(hand_pose.joint_angles)[None],
                                np.asarray(hand_pose.wrist_xform)[None], [hand_idx])[0]


def _visible_count(cam: CameraModel, pts_world: np.ndarray) -> int:
    eye = cam.world_to_eye(pts_world)
    win = cam.eye_to_window(eye)
    ok = ((win[..., 0] >= 0) & (win[..., 0] <= cam.width - 1) & (win[..., 1] >= 0)
          & (win[..., 1] <= cam.height - 1) & (eye[..., 2] > 0))
    return int(ok.sum())


def gen_crop_cameras_from_pose(cameras, camera_angles, hand_model: HandModel, hand_pose: SingleHandPose,
                               hand_idx: int, num_crop_points: int, new_image_size,
                               max_view_num: Optional[int] = None, sort_camera_index: bool = False,
                               focal_multiplier: float = 0.95, mirror_right_hand: bool = True,
                               min_required_vis_landmarks: int = 19) -> Dict[int, PinholePlaneCameraModel]:
    """Crop cameras for one hand (lib/tracker/perspective_crop.py:54-162): crop points are the
    landmarks of the given, the neutral and the open pose (one FK launch for all three);
    cameras seeing at least `min_required_vis_landmarks` landmarks are eligible, ranked by
    visibility (or by index when `sort_camera_index`), at most `max_view_num` are used."""
    assert num_crop_points in (21, 42, 63)
    n_pose = num_crop_points // 21
    ja = np.zeros((3, NUM_JOINTS_PER_HAND), dtype=np.float32)
    ja[0] = np.asarray(hand_pose.joint_angles, dtype=np.float32)
    ja[1] = neutral_joint_angles(hand_model).numpy()
    wr = np.repeat(np.asarray(hand_pose.wrist_xform)[None], 3, axis=0)
    lms = landmarks_from_poses(hand_model, ja[:n_pose], wr[:n_pose], [hand_idx] * n_pose)
    crop_points = lms.reshape(-1, 3)
    counts = [_visible_count(c, lms[0]) for c in cameras]
    eligible = [i for i, n in enumerate(counts) if n >= min_required_vis_landmarks]
    eligible.sort(reverse=True, key=lambda i: counts[i])
    if sort_camera_index:
        eligible = sorted(eligible)
    out: Dict[int, PinholePlaneCameraModel] = {}
    for cam_idx in eligible:
        out[cam_idx] = gen_crop_parameters_from_points(
            cameras[cam_idx], crop_points, new_image_size,
            mirror_img_x=(mirror_right_hand and hand_idx == 1),
            camera_angle=camera_angles[cam_idx], focal_multiplier=focal_multiplier)
        if len(out) == max_view_num:
            break
    return out


def warp_views_device(views: List[ViewData], pairs: List[Tuple[int, PinholePlaneCameraModel]],
                      device: torch.device, want_u8: bool = False):
    """GPU `_warp_image` for a list of (view index, crop camera) pairs of one frame: returns the
    float32 crops [n,96,96] in [0,1] (and the uint8 crops if asked) on `device`."""
    used = sorted({v for v, _ in pairs})
    slot = {v: k for k, v in enumerate(used)}
    h, w = views[used[0]].image.shape
    stack = np.empty((len(used), h, w), dtype=np.uint8)
    for v in used:
        if views[v].image.shape != (h, w):
            raise ValueError("all views of a frame must have the same size")
        stack[slot[v]] = views[v].image
    src = torch.from_numpy(stack).to(device)
    params = torch.from_numpy(np.stack([resample_params(views[v].camera, cam) for v, cam in pairs])).to(device)
    offsets = torch.tensor([slot[v] * h * w for v, _ in pairs], dtype=torch.int64, device=device)
    n = len(pairs)
    out_f = torch.empty((n, 96, 96), dtype=torch.float32, device=device)
    out_u = torch.empty((n, 96, 96), dtype=torch.uint8, device=device) if want_u8 else None
    with torch.cuda.device(device):
        native.check(native.lib().umt_resample_crops(
            native.ptr(src), native.ptr(offsets), w, w, h, native.ptr(params), n, native.ptr(out_u),
            native.ptr(out_f), 0, native.current_stream_ptr()), "umt_resample_crops")
    return (out_f, out_u) if want_u8 else out_f


_NP2T = {np.float64: torch.float64, np.float32: torch.float32, np.int64: torch.int64, np.int32: torch.int32,
         np.uint8: torch.uint8, np.bool_: torch.bool}


class _Section:
    """One pinned host buffer, one device buffer of the same layout, and typed views of both."""

    def __init__(self, device: torch.device, spec):
        off, self.offsets = 0, {}
        for name, dt, shape in spec:
            self.offsets[name] = off
            off += (int(np.prod(shape)) * np.dtype(dt).itemsize + 15) & ~15
        self.nbytes = off
        self.host = torch.zeros(off, dtype=torch.uint8).pin_memory()
        self.dev = torch.zeros(off, dtype=torch.uint8, device=device)
        raw = self.host.numpy()
        self.h, self.d = {}, {}
        for name, dt, shape in spec:
            o, nb = self.offsets[name], int(np.prod(shape)) * np.dtype(dt).itemsize
            self.h[name] = raw[o:o + nb].view(dt).reshape(shape)
            self.d[name] = self.dev[o:o + nb].view(_NP2T[dt]).view(shape)

    def upload(self, nbytes: Optional[int] = None):
        n = self.nbytes if nbytes is None else nbytes
        self.dev[:n].copy_(self.host[:n], non_blocking=True)

    def download(self):
        self.host.copy_(self.dev, non_blocking=True)


class _FrameBuffers:
    """Persistent buffers of the per-frame path for one rig (n_cam views of h x w pixels): every
    frame rewrites them in place, so a frame costs two host->device copies, two device->host copies
    and no allocation, and the CUDA graph of the forward pass reads its inputs where they are
    (InputFrameDesc.persistent_inputs)."""

    def __init__(self, device: torch.device, n_cam: int, h: int, w: int, n_pose: int):
        H, C = NUM_HANDS, NUM_HANDS * MAX_VIEW_NUM
        self.key = (n_cam, h, w, n_pose)
        self.crop_in = _Section(device, [
            ("intr", np.float64, (n_cam, 12)), ("c2w", np.float64, (n_cam, 16)), ("angles", np.float64, (n_cam,)),
            ("poses", np.float32, (H, n_pose, 22)), ("wrists", np.float32, (H, n_pose, 16)),
            ("mirror", np.uint8, (H, n_pose)), ("is_right", np.uint8, (H,)), ("valid", np.uint8, (H,)),
            ("frame_idx", np.int32, (H,))])
        self.crop_in.h["valid"][:] = 1
        self.crop_out = _Section(device, [
            ("err", np.int32, (4,)), ("sel_count", np.int32, (H,)), ("sel_cam", np.int32, (H, MAX_VIEW_NUM)),
            ("params", np.float64, (H, MAX_VIEW_NUM, 48)), ("intr", np.float32, (H, MAX_VIEW_NUM, 9)),
            ("extr", np.float32, (H, MAX_VIEW_NUM, 16))])
        self.points = torch.zeros((H, n_pose * 21, 3), dtype=torch.float32, device=device)
        self.frame_in = _Section(device, [
            ("intr", np.float32, (C, 3, 3)), ("extr", np.float32, (C, 4, 4)), ("range", np.int64, (H, 2)),
            ("hand", np.int64, (H,)), ("use_memory", np.bool_, (H,)), ("offsets", np.int64, (C,)),
            ("params", np.float64, (C, 48)), ("images", np.uint8, (n_cam, h, w))])
        self.crops = torch.zeros((C, 96, 96), dtype=torch.float32, device=device)


class HandTracker:
    def __init__(self, model, opts: HandTrackerOpts) -> None:
        if not torch.cuda.device_count():
            raise native.NativeError("HandTracker: no CUDA device (this implementation is GPU only)")
        self._device = torch.device("cuda", torch.cuda.current_device())
        self._model = model
        self._model.to(self._device)
        self._input_size = np.array(self._model.getInputImageSizes())
        self._num_crop_points = opts.num_crop_points
        self._enable_memory = opts.enable_memory
        self._hand_ratio_in_crop = opts.hand_ratio_in_crop
        self._min_required_vis_landmarks = opts.min_required_vis_landmarks
        self._valid_tracking_history = np.zeros(2, dtype=bool)
        self._device_crop_cameras = True      # False: always build crop cameras with numpy on the host
        self.use_cuda_graphs = True           # replay the per-frame forward pass as one CUDA graph (same results)
        self._skel_cache = None
        self._last_hands = None
        self._fb: Optional[_FrameBuffers] = None      # persistent buffers of the per-frame path
        self._crop_cache = None                       # device-side results of the last gen_crop_cameras
        self._fb_model_key = self._fb_intr_key = None
        self._stream_obj = None
        self._result_pin: Optional[torch.Tensor] = None   # pinned copy of the model's packed outputs

    def _sync(self):
        """Wait for the current stream.  The Stream object is kept between frames (building one costs
        ~20 us) and refreshed when the raw handle of the current stream changes."""
        st = self._stream_obj
        if st is None or st.cuda_stream != native.current_stream_ptr():
            st = self._stream_obj = torch.cuda.current_stream()
        st.synchronize()

    def _graph_scope(self):
        import contextlib
        scope = getattr(self._model, "frame_graph_scope", None)
        return scope() if (scope is not None and self.use_cuda_graphs) else contextlib.nullcontext()

    def reset_history(self) -> None:
        self._valid_tracking_history[:] = False
        self._skel_cache = None

    def gen_crop_cameras(self, cameras, camera_angles, hand_model: HandModel,
                         gt_tracking: Dict[int, SingleHandPose], min_num_crops: int):
        """Crop cameras of one frame (lib/tracker/tracker.py:101-139).  For the usual rig (FishEye62
        cameras of one size) all hands go through the float64 device kernel umt_crop_cameras in one
        launch and one device->host read (it agrees with the numpy construction below to ~1e-13);
        any other camera model takes the numpy path."""
        hands = [(h, p) for h, p in (gt_tracking or {}).items() if p.hand_confidence >= CONFIDENCE_THRESHOLD]
        if hands and self._device_crop_cameras and all(
                type(c).__name__ == "Fisheye62CameraModel" and c.width == cameras[0].width
                and c.height == cameras[0].height for c in cameras) and tuple(self._input_size) == (96, 96):
            with nvtx.stage("crop_cameras"):
                return self._gen_crop_cameras_device(cameras, camera_angles, hand_model, hands, min_num_crops)
        out: Dict[int, Dict[int, PinholePlaneCameraModel]] = {}
        for hand_idx, pose in (gt_tracking or {}).items():
            if pose.hand_confidence < CONFIDENCE_THRESHOLD:
                continue
            per_hand = gen_crop_cameras_from_pose(
                cameras, camera_angles, hand_model, pose, hand_idx, self._num_crop_points,
                self._input_size, max_view_num=MAX_VIEW_NUM, sort_camera_index=True,
                focal_multiplier=self._hand_ratio_in_crop, mirror_right_hand=True,
                min_required_vis_landmarks=self._min_required_vis_landmarks)
            if per_hand and len(per_hand) >= min_num_crops:
                out[hand_idx] = per_hand
        return out

    @staticmethod
    def _crop_signature(crop_cameras) -> tuple:
        """Hands, chosen cameras and the identity of the crop-camera objects: the persistent path is only
        taken while the dict still holds exactly what gen_crop_cameras put there."""
        return tuple((h, tuple((k, id(c)) for k, c in per.items())) for h, per in crop_cameras.items())

    def _buffers(self, n_cam: int, h: int, w: int) -> _FrameBuffers:
        n_pose = self._num_crop_points // 21
        if self._fb is None or self._fb.key != (n_cam, h, w, n_pose):
            self._fb = _FrameBuffers(self._device, n_cam, h, w, n_pose)
            self._fb_model_key = self._fb_intr_key = None
        return self._fb

    def _gen_crop_cameras_device(self, cameras, camera_angles, hand_model: HandModel, hands, min_num_crops: int):
        """All hands of the frame in two launches (landmark FK of the given / neutral / open pose, then
        umt_crop_cameras), one upload and one read, through the persistent buffers.  The device-side
        rows (resampling parameters, model intrinsics / extrinsics) are kept for track_frame."""
        from .batched import check_crop_camera_error
        from .skinning import _device_model

        n, n_cam = len(hands), len(cameras)
        fb = self._buffers(n_cam, int(cameras[0].height), int(cameras[0].width))
        n_pose = fb.key[3]
        ci, co = fb.crop_in, fb.crop_out
        hv = ci.h
        lim = hand_model.joint_limits
        mkey = (lim.data_ptr(), lim._version)
        if self._fb_model_key != mkey:          # pose slots 1 (neutral) and 2 (open = zeros) change with the hand model only
            hv["poses"][:] = 0
            if n_pose > 1:
                hv["poses"][:, 1] = neutral_joint_angles(hand_model).numpy()
            self._fb_model_key = mkey
        ikey = tuple((c.f, c.c, c.distort_coeffs) for c in cameras)      # compared by value: a few microseconds
        if ikey != self._fb_intr_key:
            hv["intr"][:] = [[*c.f, *c.c, *c.distort_coeffs] for c in cameras]
            self._fb_intr_key = ikey
        hv["c2w"][:] = [np.asarray(c.camera_to_world_xf, dtype=np.float64).reshape(16) for c in cameras]
        hv["angles"][:] = camera_angles
        for i, (h, p) in enumerate(hands):
            hv["poses"][i, 0] = p.joint_angles
            hv["wrists"][i] = np.asarray(p.wrist_xform, dtype=np.float32).reshape(16)
            hv["is_right"][i] = hv["mirror"][i] = 1 if h == RIGHT_HAND_INDEX else 0
        ci.upload()
        lib = native.lib()
        stream = native.current_stream_ptr()
        dv = ci.d
        axes, rest, lm_rest, lm_idx, lm_w = _device_model(hand_model, self._device)
        with torch.cuda.device(self._device):
            native.check(lib.umt_fk_landmarks(
                native.ptr(axes), native.ptr(rest), native.ptr(lm_rest), native.ptr(lm_idx), native.ptr(lm_w),
                native.ptr(dv["poses"]), native.ptr(dv["wrists"]), native.ptr(dv["mirror"]), n * n_pose,
                native.ptr(fb.points), stream), "umt_fk_landmarks")
            od = co.d
            native.check(lib.umt_crop_cameras(
                native.ptr(fb.points), native.ptr(dv["frame_idx"]), native.ptr(dv["is_right"]), native.ptr(dv["valid"]),
                native.ptr(dv["intr"]), native.ptr(dv["c2w"]), native.ptr(dv["angles"]), n_cam,
                int(cameras[0].width), int(cameras[0].height), n, n_pose * 21, self._min_required_vis_landmarks,
                float(self._hand_ratio_in_crop), native.ptr(od["sel_count"]), native.ptr(od["sel_cam"]),
                native.ptr(od["params"]), native.ptr(od["intr"]), native.ptr(od["extr"]), native.ptr(od["err"]),
                stream), "umt_crop_cameras")
        co.download()
        self._sync()
        oh = co.h
        if int(oh["err"][0]):        # the flag is only ever set by the kernel: clear it for the next frame
            code = int(oh["err"][0])
            co.dev.zero_()
            check_crop_camera_error(code)
        cnt, sel = oh["sel_count"][:n].copy(), oh["sel_cam"][:n].copy()
        rows = {"params": oh["params"][:n].copy(), "intr": oh["intr"][:n].copy(), "extr": oh["extr"][:n].copy()}
        out: Dict[int, Dict[int, PinholePlaneCameraModel]] = {}
        index = {}
        for i, (hand_idx, _) in enumerate(hands):
            par = rows["params"][i, :, 28:48]
            per_hand = {int(sel[i, k]): PinholePlaneCameraModel(
                width=int(self._input_size[0]), height=int(self._input_size[1]), f=par[k, 0:2].copy(),
                c=par[k, 2:4].copy(), distort_coeffs=(), camera_to_world_xf=par[k, 4:20].reshape(4, 4).copy())
                for k in range(int(cnt[i]))}
            if per_hand and len(per_hand) >= min_num_crops:
                out[hand_idx] = per_hand
                index[hand_idx] = i
        self._crop_cache = {"out": out, "index": index, "rows": rows, "views": [id(c) for c in cameras],
                            "sig": self._crop_signature(out)}
        return out

    def track_frame(self, sample: InputFrame, hand_model: HandModel, crop_cameras) -> TrackingResult:
        if not crop_cameras:
            self.reset_history()
            return TrackingResult()
        with nvtx.stage("inputs+resample"):
            data, desc, skel = self._make_inputs(sample, hand_model, crop_cameras)
        # one CUDA-graph replay per frame; the graph's static outputs are read before returning
        m = self._model
        prev = getattr(m, "_frame_graphs", None)
        if prev is not None and self.use_cuda_graphs:
            m._frame_graphs = True
        try:
            with nvtx.stage("forward"):
                out = m.regress_pose_use_skeleton(data, desc, skel)
            with nvtx.stage("result"):
                return self._gen_tracking_result(out, self._last_hands, crop_cameras)
        finally:
            if prev is not None:
                m._frame_graphs = prev

    def track_frame_and_calibrate_scale(self, sample: InputFrame, crop_cameras) -> TrackingResult:
        if not crop_cameras:
            self.reset_history()
            return TrackingResult()
        data, desc, _ = self._make_inputs(sample, None, crop_cameras)
        with self._graph_scope():
            out = self._model.regress_pose_pred_skel_scale(data, desc)
            return self._gen_tracking_result(out, self._last_hands, crop_cameras)

    def _skeleton(self, hand_model_mm: Optional[HandModel]):
        if hand_model_mm is None:
            return None
        # the metre-scaled skeleton is rebuilt when the hand model's tensors change (also in place)
        key = tuple((t.data_ptr(), t._version) for t in (hand_model_mm.joint_rotation_axes,
                                                        hand_model_mm.joint_rest_positions))
        if self._skel_cache is None or self._skel_cache[0] != key:
            hm = scaled_hand_model(hand_model_mm, MM_TO_M)
            self._skel_cache = (key, InputSkeletonData(
                joint_rotation_axes=hm.joint_rotation_axes.float().to(self._device).contiguous(),
                joint_rest_positions=hm.joint_rest_positions.float().to(self._device).contiguous()))
        return self._skel_cache[1]

    def _make_inputs_persistent(self, sample: InputFrame, hand_model_mm: Optional[HandModel], crop_cameras, cache):
        """`_make_inputs` when `crop_cameras` is what the last gen_crop_cameras call returned: the
        model's camera matrices and the resampling rows are the ones umt_crop_cameras wrote (the
        numbers the host construction would give, lib/tracker/tracker.py:188-236), everything goes
        up in one copy, and the crops are written into the buffer the CUDA graph reads."""
        fb, rows = self._fb, cache["rows"]
        fi = fb.frame_in
        hv = fi.h
        views = sample.views
        h, w = fb.key[1], fb.key[2]
        slot: Dict[int, int] = {}
        n = bs = 0
        hands = []
        for hand_idx, per_hand in crop_cameras.items():
            i, start = cache["index"][hand_idx], n
            for k, cam_idx in enumerate(per_hand.keys()):
                if cam_idx not in slot:
                    img = views[cam_idx].image
                    if img.shape != (h, w):
                        raise ValueError("all views of a frame must have the same size")
                    slot[cam_idx] = len(slot)
                    hv["images"][slot[cam_idx]] = img
                hv["intr"][n] = rows["intr"][i, k].reshape(3, 3)
                hv["extr"][n] = rows["extr"][i, k].reshape(4, 4)
                hv["params"][n] = rows["params"][i, k]
                hv["offsets"][n] = slot[cam_idx] * h * w
                n += 1
            hv["range"][bs] = (start, n)
            hv["hand"][bs] = hand_idx
            hv["use_memory"][bs] = self._valid_tracking_history[hand_idx]
            hands.append(hand_idx)
            bs += 1
        fi.upload(fi.offsets["images"] + len(slot) * h * w)
        dv = fi.d
        with torch.cuda.device(self._device):
            native.check(native.lib().umt_resample_crops(
                native.ptr(dv["images"]), native.ptr(dv["offsets"]), w, w, h, native.ptr(dv["params"]), n, 0,
                native.ptr(fb.crops), 0, native.current_stream_ptr()), "umt_resample_crops")
        data = InputFrameData(left_images=fb.crops[:n], intrinsics=dv["intr"][:n], extrinsics_xf=dv["extr"][:n])
        desc = InputFrameDesc(sample_range=dv["range"][:bs], memory_idx=dv["hand"][:bs],
                              use_memory=dv["use_memory"][:bs], hand_idx=dv["hand"][:bs],
                              n_memory_slots=NUM_HANDS, persistent_inputs=True)
        self._last_hands = np.array(hands)
        return data, desc, self._skeleton(hand_model_mm)

    def _make_inputs(self, sample: InputFrame, hand_model_mm: Optional[HandModel], crop_cameras):
        cache = self._crop_cache
        if (cache is not None and cache["out"] is crop_cameras and self._fb is not None
                and cache["sig"] == self._crop_signature(crop_cameras)
                and len(sample.views) == self._fb.key[0]
                and all(isinstance(v.image, np.ndarray) and v.image.dtype == np.uint8 for v in sample.views)):
            return self._make_inputs_persistent(sample, hand_model_mm, crop_cameras, cache)
        pairs, intr, extr, ranges, hands = [], [], [], [], []
        for hand_idx, per_hand in crop_cameras.items():
            start = len(pairs)
            for cam_idx, crop_cam in per_hand.items():
                pairs.append((cam_idx, crop_cam))
                intr.append(crop_cam.uv_to_window_matrix())
                w2e = np.linalg.inv(crop_cam.camera_to_world_xf)
                w2e[:3, 3] *= MM_TO_M
                extr.append(w2e)
            if len(pairs) > start:
                hands.append(hand_idx)
                ranges.append([start, len(pairs)])
        hands_np = np.array(hands)
        dev = self._device
        n, bs = len(pairs), len(hands)
        # three uploads per frame instead of seven: camera matrices, integer descriptors, memory flags
        cam = np.concatenate([np.stack(intr).reshape(-1), np.stack(extr).reshape(-1)]).astype(np.float32)
        cam_d = torch.from_numpy(cam).to(dev)
        ints = np.concatenate([np.asarray(ranges, dtype=np.int64).reshape(-1), hands_np.astype(np.int64)])
        ints_d = torch.from_numpy(ints).to(dev)
        data = InputFrameData(
            left_images=warp_views_device(sample.views, pairs, dev),
            intrinsics=cam_d[:9 * n].view(n, 3, 3),
            extrinsics_xf=cam_d[9 * n:].view(n, 4, 4))
        desc = InputFrameDesc(
            sample_range=ints_d[:2 * bs].view(bs, 2),
            memory_idx=ints_d[2 * bs:],
            use_memory=torch.from_numpy(self._valid_tracking_history[hands_np]).to(dev),
            hand_idx=ints_d[2 * bs:],
            n_memory_slots=NUM_HANDS)        # memory_idx is a hand index: the bank always has both slots
        self._last_hands = hands_np
        return data, desc, self._skeleton(hand_model_mm)

    def _gen_tracking_result(self, out: RegressorOutput, hand_indices: np.ndarray, crop_cameras) -> TrackingResult:
        n = out.joint_angles.shape[0]
        packed = getattr(self._model, "_last_packed", None)
        if packed is not None and packed[0].data_ptr() == out.joint_angles.data_ptr():
            # outputs of a CUDA-graph replay: slices of one buffer, read with one copy into pinned memory
            buf, (_, o_wr, _, o_sc) = packed
            pin = self._result_pin
            if pin is None or pin.numel() < buf.numel():
                pin = self._result_pin = torch.empty(max(buf.numel(), 256), dtype=torch.float32).pin_memory()
            pin[:buf.numel()].copy_(buf, non_blocking=True)
            self._sync()
            back = pin.numpy()
            ja = back[:n * 22].reshape(n, 22).copy()
            wr = back[o_wr:o_wr + n * 16].reshape(n, 4, 4).copy()
            scales = None if out.skel_scales is None else back[o_sc:o_sc + n].copy()
        else:
            parts = [out.joint_angles.reshape(n, 22), out.wrist_xfs.reshape(n, 16)]
            if out.skel_scales is not None:
                parts.append(out.skel_scales.reshape(n, 1))
            back = torch.cat(parts, dim=1).to("cpu").numpy()            # one device->host read per frame
            ja = np.ascontiguousarray(back[:, :22])
            wr = np.ascontiguousarray(back[:, 22:38]).reshape(n, 4, 4)
            scales = None if out.skel_scales is None else np.ascontiguousarray(back[:, 38])
        wr[..., :3, 3] *= M_TO_MM
        poses, n_views, pred_scales = {}, {}, {}
        for i, h in enumerate(hand_indices):
            h = int(h)
            poses[h] = SingleHandPose(joint_angles=ja[i], wrist_xform=wr[i], hand_confidence=1.0)
            n_views[h] = len(crop_cameras[h])
            if scales is not None:
                pred_scales[h] = scales[i]
        for h in range(NUM_HANDS):
            self._valid_tracking_history[h] = h in poses
        return TrackingResult(hand_poses=poses, num_views=n_views, predicted_scales=pred_scales)
