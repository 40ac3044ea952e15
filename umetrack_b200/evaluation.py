"""Recording-level drivers: the per-frame loop of the reference's eval scripts, and sharding of
recordings across GPUs.

`track_recording` is the body of `_track_sequence` (run_eval_known_skeleton.py:48-102):
gen_crop_cameras -> track_frame -> landmarks for prediction and ground truth, MPJPE over valid
hand-frames.  `calibrate_scale` is pass 1 of run_eval_unknown_skeleton.py:49-78.  Recordings
are supplied as the label dict of lib/tracker/video_pose_data.py:75-93 plus raw uint8 frames
[T, H, n_cam*W] (video decode is outside the path).
"""
from __future__ import annotations

import copy as _copy
from typing import Dict, Iterator, Optional, Tuple

import numpy as np
import torch

from .cameras import read_camera_from_json
from .hand_model import HandModel, NUM_HANDS, NUM_LANDMARKS_PER_HAND, load_hand_model_from_dict, scaled_hand_model
from .hand_tracker import (HandTracker, HandTrackerOpts, InputFrame, SingleHandPose, ViewData,
                           landmarks_from_poses)
from .metrics import MetricSums, keypoint_errors_device


class RecordingStream:
    """Iterates (InputFrame, gt_tracking) like SyncedImagePoseStream.__iter__
    (lib/tracker/video_pose_data.py:106-151)."""

    def __init__(self, labels: dict, frames: np.ndarray):
        self.cameras = [read_camera_from_json(c) for c in labels["cameras"]]
        self.camera_angles = labels["camera_angles"]
        self.hand_model: HandModel = load_hand_model_from_dict(labels["hand_model"])
        self.joint_angles = np.array(labels["joint_angles"])
        self.wrist_transforms = np.array(labels["wrist_transforms"])
        self.hand_confidences = np.array(labels["hand_confidences"])
        self.camera_to_world = np.array(labels["camera_to_world_transforms"])
        self.frames = frames
        assert len(self.joint_angles) == len(frames)
        # every frame's cameras are `camera.copy(camera_to_world_xf=...)` of the same four cameras: the
        # copy's intrinsics are computed once here, a frame's camera is a shallow clone with its own pose
        self._camera_copies = [c.copy() for c in self.cameras]

    def __len__(self):
        return len(self.frames)

    def __iter__(self) -> Iterator[Tuple[InputFrame, Dict[int, SingleHandPose]]]:
        n_cam = len(self.cameras)
        for t, raw in enumerate(self.frames):
            gt = {}
            for h in range(2):
                if self.hand_confidences[t, h] > 0:
                    gt[h] = SingleHandPose(self.joint_angles[t, h], self.wrist_transforms[t, h],
                                           self.hand_confidences[t, h])
            mv = raw.reshape(raw.shape[0], n_cam, -1)
            if self.camera_to_world[t].sum() == 0:
                assert not gt, "Cameras are not tracked, expecting no ground truth tracking!"
            views = []
            for c in range(n_cam):
                cam = _copy.copy(self._camera_copies[c])
                cam.camera_to_world_xf = self.camera_to_world[t, c]
                views.append(ViewData(image=mv[:, c, :], camera=cam, camera_angle=self.camera_angles[c]))
            yield InputFrame(views=views), gt


def track_recording(model, labels: dict, frames: np.ndarray, hand_model: Optional[HandModel] = None,
                    sums: Optional[MetricSums] = None, tracker: Optional[HandTracker] = None):
    """Known-skeleton tracking of one recording.  Returns dict with tracked_keypoints,
    gt_keypoints [2,T,21,3] (float64, mm), valid_tracking [2,T] and per_frame_mean_error --
    the same arrays the reference pickles (run_eval_known_skeleton.py:92-100).  `hand_model`
    replaces the recording's own model for the PREDICTED keypoints only (the calibrated model of
    run_eval_unknown_skeleton.py:119-127); crop cameras and ground-truth keypoints always use the
    recording's model.  `tracker` continues an existing tracker (pass 2 after calibration)."""
    stream = RecordingStream(labels, frames)
    T = len(stream)
    pred_model = stream.hand_model if hand_model is None else hand_model
    tracked = np.zeros([NUM_HANDS, T, NUM_LANDMARKS_PER_HAND, 3])
    gtk = np.zeros([NUM_HANDS, T, NUM_LANDMARKS_PER_HAND, 3])
    valid = np.zeros([NUM_HANDS, T], dtype=bool)
    tracker = tracker or HandTracker(model, HandTrackerOpts())
    # The reference computes the landmarks of the tracked and the ground-truth pose inside the loop
    # (run_eval_known_skeleton.py:70-81).  They do not feed back into the tracking, so here the poses are
    # collected and both landmark sets come from one FK launch each after the loop (same kernel, same
    # per-pose arithmetic; two device->host reads per recording instead of two per frame).
    items, ja_p, wr_p, ja_g, wr_g = [], [], [], [], []
    for t, (frame, gt) in enumerate(stream):
        cc = tracker.gen_crop_cameras([v.camera for v in frame.views], stream.camera_angles,
                                      stream.hand_model, gt, min_num_crops=1)
        res = tracker.track_frame(frame, pred_model, cc)
        for h in sorted(res.hand_poses.keys()):
            items.append((h, t))
            ja_p.append(res.hand_poses[h].joint_angles); wr_p.append(res.hand_poses[h].wrist_xform)
            ja_g.append(gt[h].joint_angles); wr_g.append(gt[h].wrist_xform)
    if items:
        hands_all = [h for h, _ in items]
        lm_pred = landmarks_from_poses(pred_model, np.stack(ja_p), np.stack(wr_p), hands_all)
        lm_gt = landmarks_from_poses(stream.hand_model, np.stack(ja_g), np.stack(wr_g), hands_all)
        for k, (h, t) in enumerate(items):
            tracked[h, t], gtk[h, t], valid[h, t] = lm_pred[k], lm_gt[k], True
    dev = sums.device if sums is not None else torch.device("cuda", torch.cuda.current_device())
    p = torch.from_numpy(tracked.reshape(-1, 21, 3)).float().to(dev)
    g = torch.from_numpy(gtk.reshape(-1, 21, 3)).float().to(dev)
    v = torch.from_numpy(valid.reshape(-1)).to(dev)
    err = keypoint_errors_device(p, g)[v].cpu().numpy()              # run_eval_known_skeleton.py:88-89
    if sums is not None:
        sums.add_errors(p, g, v)
        sums.add_accelerations(p.reshape(2, T, 21, 3), v.reshape(2, T), gt=False)
        sums.add_accelerations(g.reshape(2, T, 21, 3), v.reshape(2, T), gt=True)
    return {"tracked_keypoints": tracked, "gt_keypoints": gtk, "valid_tracking": valid,
            "per_frame_mean_error": err}


def calibrate_scale(model, labels: dict, frames: np.ndarray, generic_hand_model: HandModel,
                    n_calibration_samples: int = 30, tracker: Optional[HandTracker] = None):
    """Pass 1 of the unknown-skeleton eval (run_eval_unknown_skeleton.py:49-78): mean predicted scale
    over the first `n_calibration_samples` two-view hand-frames (0 = all) -> scaled generic hand
    model.  Crop cameras come from the recording's own hand model, like the reference's."""
    stream = RecordingStream(labels, frames)
    tracker = tracker or HandTracker(model, HandTrackerOpts())
    scales = []
    for frame, gt in stream:
        cc = tracker.gen_crop_cameras([v.camera for v in frame.views], stream.camera_angles,
                                      stream.hand_model, gt, min_num_crops=2)
        res = tracker.track_frame_and_calibrate_scale(frame, cc)
        for h in res.hand_poses.keys():
            scales.append(res.predicted_scales[h])
        if n_calibration_samples != 0 and len(scales) >= n_calibration_samples:
            scales = scales[:n_calibration_samples]
            break
    assert len(scales) > 0, "No samples collected for scale calibration!"
    mean_scale = float(np.mean(scales))
    return scaled_hand_model(generic_hand_model, mean_scale), mean_scale, np.array(scales)


def track_recording_unknown_skeleton(model, labels: dict, frames: np.ndarray, generic_hand_model: HandModel,
                                     n_calibration_samples: int = 30, sums: Optional[MetricSums] = None):
    """The whole unknown-skeleton eval of one recording (run_eval_unknown_skeleton.py:82-131):
    calibration pass, history reset (the recurrent memory is not cleared: it is zeroed lazily by
    use_memory=False on the next frame), then the known-skeleton loop with the calibrated model for
    the predictions.  Returns (result dict of track_recording, mean_scale, scales)."""
    tracker = HandTracker(model, HandTrackerOpts())
    calibrated, mean_scale, scales = calibrate_scale(model, labels, frames, generic_hand_model,
                                                     n_calibration_samples, tracker=tracker)
    tracker.reset_history()
    res = track_recording(model, labels, frames, hand_model=calibrated, sums=sums, tracker=tracker)
    return res, mean_scale, scales
