"""Batched known-skeleton tracking of whole recordings on one GPU.

The reference tracks a recording frame by frame (run_eval_known_skeleton.py:48-102): crop
cameras from the ground-truth pose, two crops per hand with cv2.remap, one model call per frame.
Only the ConvRNN memory is sequential (lib/models/temporal.py), and the crop cameras never depend
on earlier predictions, so this driver reorders the work without changing its result:

  1. crop cameras for every (frame, hand) of every recording at once  (umt_crop_cameras, float64)
  2. all crops of all recordings in one resampling launch              (umt_resample_crops)
  3. the backbone over all crops in large chunks                       (umt_model_encode)
  4. a time scan over frames; each step runs fusion + temporal memory + regressor for the hands of
     ALL recordings that are present at that frame index               (umt_model_tail)
  5. FK of predictions and labels, MPJPE / acceleration sums           (umt_fk_landmarks, metrics)

History semantics follow HandTracker (lib/tracker/tracker.py:141-150,278-285): a hand uses its
memory iff it was tracked in the previous frame; a frame without any hand resets both.
"""
from __future__ import annotations

import collections
import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import native, nvtx
from .cameras import FISHEYE62_FIELDS
from .hand_model import HandModel, load_hand_model_from_dict, scaled_hand_model
from .ingest import FrameStager
from .hand_tracker import CONFIDENCE_THRESHOLD, HandTrackerOpts, MM_TO_M, neutral_joint_angles
from .metrics import MetricSums, keypoint_errors_device
from .model import InputFrameDesc, InputSkeletonData
from .skinning import fk_landmarks_device


_NEUTRAL_CACHE: Dict[tuple, torch.Tensor] = {}


_LABEL_KEYS = ("joint_angles", "wrist_transforms", "hand_confidences", "camera_to_world_transforms", "camera_angles")
_LABEL_CACHE: "collections.OrderedDict" = collections.OrderedDict()


def label_arrays(labels: dict) -> dict:
    """The per-frame label lists of a recording (lib/tracker/video_pose_data.py:75-93: nested Python lists
    out of the JSON) as numpy arrays.  Converted once per set of list objects (the cache keeps the lists
    alive, so their identities stay valid) -- list -> array conversion was a fifth of the batched driver's
    host time.  Replacing a list in the dict (e.g. new confidences) gives a new entry; edits INSIDE a list
    that was already used are not seen (treat label lists as immutable)."""
    src = tuple(labels[k] for k in _LABEL_KEYS)
    key = tuple(id(v) for v in src)
    hit = _LABEL_CACHE.get(key)
    if hit is not None:
        _LABEL_CACHE.move_to_end(key)
        return hit[1]
    arr = {"joint_angles": np.asarray(src[0], dtype=np.float32), "wrist_transforms": np.asarray(src[1], dtype=np.float32),
           "hand_confidences": np.asarray(src[2], dtype=np.float64),
           "camera_to_world_transforms": np.asarray(src[3], dtype=np.float64),
           "camera_angles": np.asarray(src[4], dtype=np.float64)}
    _LABEL_CACHE[key] = (src, arr)
    while len(_LABEL_CACHE) > 256:
        _LABEL_CACHE.popitem(last=False)
    return arr


def crop_cameras_raw(hand_model: HandModel, ja: torch.Tensor, wr: torch.Tensor, is_right: torch.Tensor,
                     valid: torch.Tensor, frame_idx: torch.Tensor, intr: torch.Tensor, c2w: torch.Tensor,
                     angles: torch.Tensor, width: int, height: int,
                     opts: HandTrackerOpts = HandTrackerOpts()):
    """umt_crop_cameras for n (frame, hand) items already on the device: ja [n,22] f32, wr [n,4,4] f32
    (mm), is_right / valid [n] uint8, frame_idx [n] int32 into c2w [T,n_cam,4,4] f64, intr [n_cam,12]
    f64 (fx fy cx cy k1 k2 k3 k4 p1 p2 k5 k6), angles [n_cam] f64 (degrees).  Returns flat device
    tensors sel_count [n] i32, sel_cam [n,2] i32, resample_params [n,2,48] f64, intrinsics
    [n,2,3,3] f32, extrinsics [n,2,4,4] f32 and the error flag [1] i32 (checked by the caller)."""
    device = ja.device
    n = ja.shape[0]
    n_cam = intr.shape[0]
    # three poses per item: given, neutral, open (perspective_crop.py:89-115)
    lim = hand_model.joint_limits
    key = (lim.data_ptr(), lim._version, str(device))
    neutral = _NEUTRAL_CACHE.get(key)
    if neutral is None:                       # one upload per hand model, not per frame
        if len(_NEUTRAL_CACHE) > 64:
            _NEUTRAL_CACHE.clear()
        neutral = _NEUTRAL_CACHE[key] = neutral_joint_angles(hand_model).to(device=device, dtype=torch.float32)
    poses = torch.stack([ja, neutral.expand(n, 22), torch.zeros((n, 22), device=device)], dim=1)
    wrists = wr.reshape(n, 1, 4, 4).expand(n, 3, 4, 4)
    n_pose = opts.num_crop_points // 21
    pts = fk_landmarks_device(hand_model, poses[:, :n_pose].reshape(-1, 22).contiguous(),
                              wrists[:, :n_pose].reshape(-1, 4, 4).contiguous(),
                              is_right.repeat_interleave(n_pose)).reshape(n, n_pose * 21, 3).contiguous()
    sel_count = torch.zeros(n, dtype=torch.int32, device=device)
    sel_cam = torch.zeros((n, 2), dtype=torch.int32, device=device)
    params = torch.zeros((n, 2, 48), dtype=torch.float64, device=device)
    k_out = torch.zeros((n, 2, 3, 3), dtype=torch.float32, device=device)
    e_out = torch.zeros((n, 2, 4, 4), dtype=torch.float32, device=device)
    err = torch.zeros(1, dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        native.check(native.lib().umt_crop_cameras(
            native.ptr(pts), native.ptr(frame_idx), native.ptr(is_right), native.ptr(valid), native.ptr(intr),
            native.ptr(c2w), native.ptr(angles), n_cam, int(width), int(height), n,
            n_pose * 21, opts.min_required_vis_landmarks, float(opts.hand_ratio_in_crop), native.ptr(sel_count),
            native.ptr(sel_cam), native.ptr(params), native.ptr(k_out), native.ptr(e_out), native.ptr(err),
            native.current_stream_ptr()), "umt_crop_cameras")
    return {"sel_count": sel_count, "sel_cam": sel_cam, "resample_params": params, "intrinsics": k_out,
            "extrinsics": e_out, "err": err}


def check_crop_camera_error(code: int) -> None:
    if code == 1:
        raise ValueError("Unable to create crop camera")          # lib/common/crop.py:25-26
    if code:
        raise native.NativeError("umt_crop_cameras: singular camera matrix")


def crop_cameras_device(hand_model: HandModel, labels: dict, device: torch.device,
                        opts: HandTrackerOpts = HandTrackerOpts(), min_num_crops: int = 1):
    """Crop cameras for every (frame, hand) of one recording.  Returns a dict of device tensors:
    sel_count [T,2], sel_cam [T,2,2], resample_params [T,2,2,48] f64, intrinsics [T,2,2,3,3],
    extrinsics [T,2,2,4,4]; hands with fewer than `min_num_crops` views get sel_count 0."""
    cams = labels["cameras"]
    arr = label_arrays(labels)
    ja = torch.from_numpy(arr["joint_angles"]).to(device)               # [T,2,22] (float32 like perspective_crop.py:32-36)
    wr = torch.from_numpy(arr["wrist_transforms"]).to(device)           # [T,2,4,4]
    conf = arr["hand_confidences"]                                      # [T,2]
    T = ja.shape[0]
    n = T * 2
    mirror = torch.tensor([0, 1], dtype=torch.uint8, device=device).repeat(T)                          # hand index
    valid = torch.from_numpy(((conf > 0) & (conf >= CONFIDENCE_THRESHOLD)).reshape(n).astype(np.uint8)).to(device)
    frame_idx = torch.arange(T, dtype=torch.int32, device=device).repeat_interleave(2)
    intr = torch.tensor([[c["fx"], c["fy"], c["cx"], c["cy"]] + [c[k] for k in FISHEYE62_FIELDS] for c in cams],
                        dtype=torch.float64, device=device)
    c2w = torch.from_numpy(arr["camera_to_world_transforms"]).to(device).contiguous()
    angles = torch.from_numpy(arr["camera_angles"]).to(device)
    r = crop_cameras_raw(hand_model, ja.reshape(n, 22), wr.reshape(n, 4, 4), mirror, valid, frame_idx, intr, c2w,
                         angles, int(cams[0]["ImageSizeX"]), int(cams[0]["ImageSizeY"]), opts)
    check_crop_camera_error(int(r["err"].item()))
    sel_count = torch.where(r["sel_count"] < min_num_crops, torch.zeros_like(r["sel_count"]), r["sel_count"])
    return {"sel_count": sel_count.reshape(T, 2), "sel_cam": r["sel_cam"].reshape(T, 2, 2),
            "resample_params": r["resample_params"].reshape(T, 2, 2, 48),
            "intrinsics": r["intrinsics"].reshape(T, 2, 2, 3, 3), "extrinsics": r["extrinsics"].reshape(T, 2, 2, 4, 4)}


def track_recordings_batched(model, recordings: Sequence[Tuple[dict, np.ndarray]],
                             sums: Optional[MetricSums] = None, opts: HandTrackerOpts = HandTrackerOpts(),
                             chunk_frames: int = 32):
    """Known-skeleton tracking of several recordings [(labels, frames uint8 [T,H,n_cam*W]), ...] on
    the current CUDA device.  `frames` is a numpy array or a torch tensor; a tensor in pinned host
    memory is uploaded asynchronously (the upload of the raw frames is the largest single cost of
    this driver), a tensor already on the device is used in place.  Returns one dict per recording with the arrays the reference pickles
    (tracked_keypoints, gt_keypoints [2,T,21,3] float64 mm, valid_tracking [2,T]) and
    per_frame_mean_error; metric sums are accumulated on the GPU if `sums` is given."""
    if not recordings:
        return []
    dev = torch.device("cuda", torch.cuda.current_device())
    model.to(dev)
    model.reset_memory()
    lib = native.lib()
    R = len(recordings)
    hand_models = [load_hand_model_from_dict(lab["hand_model"]) for lab, _ in recordings]
    nvtx.push("crop_cameras")
    # ---- 1. crop cameras of every recording (small device work + the host's item lists) ----------------
    plans, counts = [], []
    for (lab, frames), hm in zip(recordings, hand_models):
        cc = crop_cameras_device(hm, lab, dev, opts, min_num_crops=1)
        T, H, W4 = frames.shape
        W = W4 // len(lab["cameras"])
        cnt = cc["sel_count"].cpu().numpy()                      # [T,2]
        sel = cc["sel_cam"].cpu().numpy()                        # [T,2,2]
        # crops are stored frame major, hand minor, view minor-most
        t_i, h_i, k_i = np.nonzero(np.arange(sel.shape[2])[None, None, :] < cnt[:, :, None])
        plan = {"t_i": t_i, "H": H, "W": W, "W4": W4}
        if len(t_i):
            thk = torch.from_numpy(np.stack([t_i, h_i, k_i])).to(dev)
            ti, hi, ki = thk[0], thk[1], thk[2]
            plan["params"] = cc["resample_params"][ti, hi, ki].contiguous()
            # offset of each item's source view inside its staging chunk
            plan["offs"] = torch.from_numpy((t_i % chunk_frames).astype(np.int64) * (H * W4)
                                            + sel[t_i, h_i, k_i].astype(np.int64) * W).to(dev)
            plan["intr"] = cc["intrinsics"][ti, hi, ki]
            plan["extr"] = cc["extrinsics"][ti, hi, ki]
        plans.append(plan)
        counts.append(cnt)
    base = np.cumsum([0] + [len(pl["t_i"]) for pl in plans])
    nvtx.pop()
    nvtx.push("resample")
    # ---- 2. resampling: frames go up in chunks through the staging ring (the upload of chunk k+1
    # overlaps the resampling of chunk k; the host only enqueues here, so uploads run back to back
    # across recordings); items are frame-major, so a chunk's items are contiguous ---------------------
    all_crops = torch.empty((int(base[-1]), 96, 96), dtype=torch.float32, device=dev)
    stager = FrameStager(dev, chunk_frames=chunk_frames)
    for r, ((lab, frames), pl) in enumerate(zip(recordings, plans)):
        t_i = pl["t_i"]
        if not len(t_i):
            continue
        for t0, t1, chunk, slot in stager.chunks(frames):
            i0, i1 = (int(v) for v in np.searchsorted(t_i, [t0, t1]))
            if i1 > i0:
                native.check(lib.umt_resample_crops(
                    chunk.data_ptr(), pl["offs"].data_ptr() + 8 * i0, pl["W4"], pl["W"], pl["H"],
                    pl["params"].data_ptr() + 8 * 48 * i0, i1 - i0, 0,
                    all_crops.data_ptr() + 4 * 9216 * (int(base[r]) + i0), 0, native.current_stream_ptr()),
                    "umt_resample_crops")
            stager.release(slot)
    nvtx.pop()
    nvtx.push("encoder")
    # ---- 3. backbone over every crop of every recording ---------------------------------------------
    feats = model.encode(all_crops) if len(all_crops) else torch.empty((0, 72, 6, 6), device=dev)
    with_items = [pl for pl in plans if len(pl["t_i"])]
    all_intr = torch.cat([pl["intr"] for pl in with_items], dim=0) if with_items else torch.empty((0, 3, 3), device=dev)
    all_extr = torch.cat([pl["extr"] for pl in with_items], dim=0) if with_items else torch.empty((0, 4, 4), device=dev)
    # crop index ranges per (recording, frame, hand): crops are stored frame major, hand minor
    start = []
    for r in range(R):
        flat = counts[r].reshape(-1).astype(np.int64)
        start.append((int(base[r]) + np.concatenate([[0], np.cumsum(flat)[:-1]])).reshape(counts[r].shape))
    nvtx.pop()
    nvtx.push("time_scan")
    # ---- 4. time scan: hands of all recordings present at frame t form one batch ---------------------
    # All per-step descriptors are built on the host first (numpy), uploaded once, and the scan itself
    # is one C call per frame index with offset pointers (UmeTrackModel.tail_scan).
    T_max = max(c.shape[0] for c in counts)
    off = np.cumsum([0] + [2 * c.shape[0] for c in counts])      # output slot of (r, h, t): off[r] + h*T_r + t
    valid = [(c > 0).T.copy() for c in counts]                   # [2, T_r]
    # One row per (frame, recording, hand) with crops, ordered by frame index, then recording, then hand.
    # A hand uses its memory iff it was tracked in the previous frame: after a frame the history is "hand in
    # present" (tracker.py:278-285), and a frame without hands clears it (tracker.py:147-150) -- both are
    # `present[t - 1, h]`.  (numpy instead of a Python loop over T x R: 25 ms per 32 x 256 frames.)
    parts = []
    for r in range(R):
        c = counts[r].astype(np.int64)
        T_r = c.shape[0]
        pres = c > 0
        prev = np.zeros_like(pres)
        prev[1:] = pres[:-1]
        t_i, h_i = np.nonzero(pres)
        s0 = start[r][t_i, h_i].astype(np.int64)
        parts.append(np.stack([t_i, np.full_like(t_i, r), h_i, s0, s0 + c[t_i, h_i], prev[t_i, h_i].astype(np.int64),
                               int(off[r]) + h_i * T_r + t_i], axis=1))
    rows = np.concatenate(parts).astype(np.int64) if parts else np.zeros((0, 7), dtype=np.int64)
    rows = rows[np.lexsort((rows[:, 2], rows[:, 1], rows[:, 0]))]
    per_t = np.bincount(rows[:, 0], minlength=T_max)
    first_t = np.concatenate([[0], np.cumsum(per_t)[:-1]])
    step_rows = list(zip(first_t.tolist(), per_t.tolist()))
    rows_sr, rows_mem, rows_use = rows[:, 3:5], rows[:, 1] * 2 + rows[:, 2], rows[:, 5].astype(np.bool_)
    rows_hand, rows_rec, rows_slot = rows[:, 2], rows[:, 1], rows[:, 6]
    ja_all = torch.zeros((int(off[-1]), 22), dtype=torch.float32, device=dev)
    wr_all = torch.eye(4, device=dev).repeat(int(off[-1]), 1, 1)
    if len(rows_mem):
        skel_axes = torch.stack([hm.joint_rotation_axes.float() for hm in hand_models]).to(dev)
        skel_rest = torch.stack([scaled_hand_model(hm, MM_TO_M).joint_rest_positions.float() for hm in hand_models]).to(dev)
        def dev_array(rows, dtype):      # numpy first: torch.tensor on a long Python list is ~10x slower
            return torch.from_numpy(np.ascontiguousarray(rows, dtype=dtype)).to(dev)

        ridx = dev_array(rows_rec, np.int64)
        ja, wr, _ = model.tail_scan(
            feats, all_intr, all_extr, step_rows,
            dev_array(rows_sr, np.int64), dev_array(rows_mem, np.int64),
            dev_array(rows_use, np.bool_), dev_array(rows_hand, np.int64),
            skel_axes[ridx].contiguous(), skel_rest[ridx].contiguous(), n_slots=2 * R)
        slot = dev_array(rows_slot, np.int64)
        wr = wr.clone()
        wr[:, :3, 3] *= 1000.0                   # m -> mm (tracker.py:258)
        ja_all[slot] = ja
        wr_all[slot] = wr
    out_ja = [ja_all[int(off[r]):int(off[r + 1])] for r in range(R)]
    out_wr = [wr_all[int(off[r]):int(off[r + 1])] for r in range(R)]
    nvtx.pop()
    nvtx.push("fk+metrics")
    # ---- 5. FK + metrics -----------------------------------------------------------------------------
    results = []
    for r, ((lab, _), hm) in enumerate(zip(recordings, hand_models)):
        T = counts[r].shape[0]
        mirror = torch.from_numpy(np.repeat(np.array([0, 1], dtype=np.uint8), T)).to(dev)
        pred = fk_landmarks_device(hm, out_ja[r].reshape(-1, 22), out_wr[r].reshape(-1, 4, 4), mirror)
        arr = label_arrays(lab)
        gja = torch.from_numpy(arr["joint_angles"]).to(dev).permute(1, 0, 2)
        gwr = torch.from_numpy(arr["wrist_transforms"]).to(dev).permute(1, 0, 2, 3)
        gt = fk_landmarks_device(hm, gja.reshape(-1, 22).contiguous(), gwr.reshape(-1, 4, 4).contiguous(), mirror)
        v = torch.from_numpy(valid[r].reshape(-1)).to(dev)
        if sums is not None:
            sums.add_errors(pred, gt, v)
            sums.add_accelerations(pred.reshape(2, T, 21, 3), v.reshape(2, T), gt=False)
            sums.add_accelerations(gt.reshape(2, T, 21, 3), v.reshape(2, T), gt=True)
        tk = np.zeros((2, T, 21, 3))
        gk = np.zeros((2, T, 21, 3))
        pm, gm = pred.reshape(2, T, 21, 3).cpu().numpy(), gt.reshape(2, T, 21, 3).cpu().numpy()
        tk[valid[r]] = pm[valid[r]]
        gk[valid[r]] = gm[valid[r]]
        err = keypoint_errors_device(pred, gt)[v].cpu().numpy()      # hand-major, like (gt - tracked)[valid]
        results.append({"tracked_keypoints": tk, "gt_keypoints": gk, "valid_tracking": valid[r],
                        "per_frame_mean_error": err})
    nvtx.pop()
    return results
