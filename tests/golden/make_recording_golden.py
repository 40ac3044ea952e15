"""Recording-level golden vectors from the UNMODIFIED reference at the survey's sizes (SURVEY.md 8(d)
configs 2/3) plus the cases the first generator did not cover.

    python tests/golden/make_recording_golden.py [handset] [long] [hand]    # build container only

Sections (each writes its own .npz; default = all):
  handset  9-frame recording in which the set of tracked hands changes at equal size: hand 0 only
           (3 frames), hand 1 only (3 frames), both (3 frames).  The reference regrows its memory
           bank when hand 1 first appears alone (lib/models/temporal.py:101-123).
  long     one 256-frame two-hand 4-camera recording whose ground-truth hand model is the generic
           model x 1.07, in two frame variants (seeded noise, smooth): the known-skeleton loop
           (run_eval_known_skeleton.py:48-102) and the unknown-skeleton loop (30-sample scale
           calibration, history reset, pass 2 with the calibrated model;
           run_eval_unknown_skeleton.py:49-131), the reference's own crop cameras of every crop, a
           CRC32 per crop of the fixed-point map (sx, sy) = cvRound(32 * float32 map) that
           cv2.remap derives from `_warp_image`'s map (lib/tracker/tracker.py:54-80), the uint8 crops
           of 32 sampled crops, and the MPJPE / acceleration sums of load_eval.py:26-44.
  hand     scaled_hand_model / mirrored_hand_model (lib/common/hand.py:65-134) on batched models.
Inputs come from seeds (umetrack_b200.synthetic); only outputs are stored.  Shims as in
make_golden.py (so3_exp_map restated, `av` stubbed).
"""
import os
import sys
import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

LONG_T = 256
LONG_SEED = 7
LONG_MODEL_SCALE = 1.07
N_CALIBRATION = 30


def handset_presence():
    return np.array([[1, 0]] * 3 + [[0, 1]] * 3 + [[1, 1]] * 3, dtype=np.float64)


def long_hand_model_dict(hm_dict):
    """Ground-truth hand model of the long recording: the generic model, 7 % larger."""
    d = dict(hm_dict)
    d["joint_rest_positions"] = (np.array(hm_dict["joint_rest_positions"]) * LONG_MODEL_SCALE).tolist()
    d["landmark_rest_positions"] = (np.array(hm_dict["landmark_rest_positions"]) * LONG_MODEL_SCALE).tolist()
    return d


def map_crc(sw_f32):
    """CRC32 of cv2.remap's fixed-point coordinates of one float32 map [96,96,2]:
    int32 (sx, sy) = round-half-even(32 * map)."""
    s = np.rint(sw_f32.astype(np.float64) * 32.0).astype(np.int32)   # f32 * 32 is exact in f64; rint = cvRound
    return zlib.crc32(np.ascontiguousarray(s).tobytes())


class Ref:
    """The reference modules, imported once."""

    def __init__(self):
        import make_golden as mg

        mg.install_reference()
        self.mg = mg
        from lib.common.camera import read_camera_from_json
        from lib.common.hand import mirrored_hand_model, scaled_hand_model
        from lib.tracker import tracker as rtracker
        from lib.tracker.perspective_crop import landmarks_from_hand_pose
        from lib.tracker.tracking_result import SingleHandPose
        from lib.tracker.video_pose_data import load_hand_model_from_dict

        self.read_camera_from_json = read_camera_from_json
        self.scaled_hand_model, self.mirrored_hand_model = scaled_hand_model, mirrored_hand_model
        self.rtracker, self.landmarks_from_hand_pose = rtracker, landmarks_from_hand_pose
        self.SingleHandPose, self.load_hand_model_from_dict = SingleHandPose, load_hand_model_from_dict


class Stream:
    """(InputFrame, gt_tracking) per frame like SyncedImagePoseStream.__iter__
    (lib/tracker/video_pose_data.py:106-151), fed from arrays instead of PyAV."""

    def __init__(self, ref, labels, frames):
        self.ref, self.labels, self.frames = ref, labels, frames
        self.cams = [ref.read_camera_from_json(c) for c in labels["cameras"]]
        self.cam_xf = np.array(labels["camera_to_world_transforms"])
        self.ja = np.array(labels["joint_angles"])
        self.wr = np.array(labels["wrist_transforms"])
        self.conf = np.array(labels["hand_confidences"])
        self.hand_model = ref.load_hand_model_from_dict(labels["hand_model"])

    def __len__(self):
        return len(self.frames)

    def __iter__(self):
        rt = self.ref.rtracker
        for t in range(len(self.frames)):
            gt = {h: self.ref.SingleHandPose(self.ja[t, h], self.wr[t, h], self.conf[t, h])
                  for h in range(2) if self.conf[t, h] > 0}
            mv = self.frames[t].reshape(self.frames.shape[1], 4, -1)
            views = [rt.ViewData(image=mv[:, c, :], camera=self.cams[c].copy(camera_to_world_xf=self.cam_xf[t, c]),
                                 camera_angle=self.labels["camera_angles"][c]) for c in range(4)]
            yield rt.InputFrame(views=views), gt


def known_skeleton_pass(ref, tracker, stream, pred_model, collect_crops=None):
    """Body of _track_sequence (run_eval_known_skeleton.py:68-89 / run_eval_unknown_skeleton.py:110-131)."""
    T = len(stream)
    tracked = np.zeros((2, T, 21, 3))
    gtk = np.zeros((2, T, 21, 3))
    valid = np.zeros((2, T), dtype=bool)
    ja = np.zeros((2, T, 22), dtype=np.float32)
    wr = np.zeros((2, T, 4, 4), dtype=np.float32)
    for t, (frame, gt) in enumerate(stream):
        cc = tracker.gen_crop_cameras([v.camera for v in frame.views], stream.labels["camera_angles"],
                                      stream.hand_model, gt, 1)
        if collect_crops is not None:
            collect_crops(t, frame, cc)
        res = tracker.track_frame(frame, pred_model, cc)
        for h, pose in res.hand_poses.items():
            tracked[h, t] = ref.landmarks_from_hand_pose(pred_model, pose, h)
            gtk[h, t] = ref.landmarks_from_hand_pose(stream.hand_model, gt[h], h)
            valid[h, t] = True
            ja[h, t], wr[h, t] = pose.joint_angles, pose.wrist_xform
    err = np.linalg.norm((gtk - tracked)[valid], axis=-1).mean(axis=-1)
    return {"tracked": tracked, "gt": gtk, "valid": valid, "per_frame_error": err, "joint_angles": ja, "wrist": wr}


def section_handset(ref, hm_dict, sd):
    from umetrack_b200 import synthetic

    labels, frames = synthetic.make_recording(hm_dict, seed=5, n_frames=9, presence=handset_presence())
    st = Stream(ref, labels, frames)
    tracker = ref.rtracker.HandTracker(ref.mg.build_reference_model(sd), ref.rtracker.HandTrackerOpts())
    r = known_skeleton_pass(ref, tracker, st, st.hand_model)
    assert r["valid"].sum() == 12
    np.savez_compressed(os.path.join(HERE, "tracker_handset_golden.npz"), **r)
    print("handset: per-frame error", r["per_frame_error"])


def section_long(ref, hm_dict, sd):
    import load_eval
    from umetrack_b200 import synthetic

    generic = ref.load_hand_model_from_dict(hm_dict)
    out = {}
    for variant, smooth in (("noise", False), ("smooth", True)):
        labels, frames = synthetic.make_recording(long_hand_model_dict(hm_dict), seed=LONG_SEED, n_frames=LONG_T,
                                                  smooth_images=smooth)
        st = Stream(ref, labels, frames)
        rows, crcs, items = [], [], []

        def collect(t, frame, cc):
            for h, per in cc.items():
                for c, cam in per.items():
                    rows.append(np.concatenate([[t, h, c], cam.f, cam.c, cam.camera_to_world_xf.reshape(-1)]))
                    # the float32 map exactly as _warp_image builds it (tracker.py:61-78)
                    px, py = np.meshgrid(np.arange(96), np.arange(96))
                    w = np.column_stack((px.flatten(), py.flatten()))
                    src = frame.views[c].camera
                    e = src.world_to_eye(cam.eye_to_world(cam.window_to_eye(w)))
                    sw = src.eye_to_window(e)
                    sw[e[:, 2] < 0] = -1
                    crcs.append(map_crc(sw.astype(np.float32).reshape(96, 96, 2)))
                    items.append((src, cam, frame.views[c].image))

        # ---- known skeleton ----
        tracker = ref.rtracker.HandTracker(ref.mg.build_reference_model(sd), ref.rtracker.HandTrackerOpts())
        r = known_skeleton_pass(ref, tracker, st, st.hand_model, collect if variant == "noise" else None)
        m = load_eval._compute_metrics(r["gt"], r["tracked"], r["valid"])
        out[f"{variant}_k_tracked"] = r["tracked"].astype(np.float32)
        out[f"{variant}_k_valid"] = r["valid"]
        out[f"{variant}_k_mpjpe"] = np.float64(m.keypoint_errors.mean())
        out[f"{variant}_k_accel"] = np.float64(m.keypoint_accelerations.mean())
        out[f"{variant}_k_gt_accel"] = np.float64(m.gt_keypoint_accelerations.mean())
        if variant == "noise":
            out["gt"] = r["gt"].astype(np.float32)
            out["crop_rows"] = np.array(rows)
            out["map_crc"] = np.array(crcs, dtype=np.uint32)
            pick = np.linspace(0, len(items) - 1, 32).astype(int)
            out["crop_pick"] = pick
        # uint8 crops of the sampled items in this variant's frames
        pick = out["crop_pick"]
        if variant == "noise":
            out["noise_crops_u8"] = np.stack([ref.rtracker._warp_image(*items[i]) for i in pick])
            noise_items = items
        else:
            crops = []
            for i in pick:
                t, c = int(out["crop_rows"][i][0]), int(out["crop_rows"][i][2])
                img = frames[t].reshape(frames.shape[1], 4, -1)[:, c, :]
                crops.append(ref.rtracker._warp_image(noise_items[i][0], noise_items[i][1], img))
            out["smooth_crops_u8"] = np.stack(crops)
        print(f"long/{variant}: known MPJPE {out[f'{variant}_k_mpjpe']:.4f} accel {out[f'{variant}_k_accel']:.4f}"
              f" valid {int(r['valid'].sum())} crops {len(rows)}")
        # ---- unknown skeleton: calibration pass, reset, pass 2 (run_eval_unknown_skeleton.py:49-131) ----
        tracker = ref.rtracker.HandTracker(ref.mg.build_reference_model(sd), ref.rtracker.HandTrackerOpts())
        scales = []
        for frame, gt in st:
            cc = tracker.gen_crop_cameras([v.camera for v in frame.views], labels["camera_angles"], st.hand_model, gt, 2)
            res = tracker.track_frame_and_calibrate_scale(frame, cc)
            for h in res.hand_poses.keys():
                scales.append(res.predicted_scales[h])
            if len(scales) >= N_CALIBRATION:
                scales = scales[:N_CALIBRATION]
                break
        mean_scale = np.mean(scales)
        calibrated = ref.scaled_hand_model(generic, mean_scale)
        tracker.reset_history()
        r2 = known_skeleton_pass(ref, tracker, st, calibrated)
        m2 = load_eval._compute_metrics(r2["gt"], r2["tracked"], r2["valid"])
        out[f"{variant}_u_scales"] = np.array(scales, dtype=np.float64)
        out[f"{variant}_u_mean_scale"] = np.float64(mean_scale)
        out[f"{variant}_u_tracked"] = r2["tracked"].astype(np.float32)
        out[f"{variant}_u_valid"] = r2["valid"]
        out[f"{variant}_u_mpjpe"] = np.float64(m2.keypoint_errors.mean())
        out[f"{variant}_u_accel"] = np.float64(m2.keypoint_accelerations.mean())
        print(f"long/{variant}: mean scale {mean_scale:.6f}, unknown MPJPE {out[f'{variant}_u_mpjpe']:.4f}")
    np.savez_compressed(os.path.join(HERE, "tracker_long_golden.npz"), **out)


def section_hand(ref, hm_dict):
    hm = ref.load_hand_model_from_dict(hm_dict)
    batched = type(hm)(*[(v[None].repeat(3, *([1] * v.dim())) if isinstance(v, torch.Tensor) and v.dim() > 0 else v)
                         for v in hm])
    s1 = ref.scaled_hand_model(hm, 1.0375)
    s3 = ref.scaled_hand_model(batched, 0.91)
    mask = torch.tensor([True, False, True])
    m3 = ref.mirrored_hand_model(batched, mask)
    m1 = ref.mirrored_hand_model(type(hm)(*[(v[None] if isinstance(v, torch.Tensor) and v.dim() > 0 else v) for v in hm]),
                                 torch.tensor([True]))
    np.savez_compressed(
        os.path.join(HERE, "hand_model_golden.npz"),
        s1_rest=s1.joint_rest_positions.numpy(), s1_lm=s1.landmark_rest_positions.numpy(),
        s3_rest=s3.joint_rest_positions.numpy(), s3_lm=s3.landmark_rest_positions.numpy(),
        m3_axes=m3.joint_rotation_axes.numpy(), m3_rest=m3.joint_rest_positions.numpy(),
        m3_lm=m3.landmark_rest_positions.numpy(), m1_axes=m1.joint_rotation_axes.numpy(),
        m1_rest=m1.joint_rest_positions.numpy(), m1_lm=m1.landmark_rest_positions.numpy())
    print("hand: written")


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    from umetrack_b200 import synthetic

    sections = sys.argv[1:] or ["handset", "long", "hand"]
    ref = Ref()
    hm_dict = synthetic.load_hand_model_json(os.path.join(HERE, "generic_hand_model_nomesh.json"))
    sd = synthetic.make_state_dict(0)
    if "handset" in sections:
        section_handset(ref, hm_dict, sd)
    if "long" in sections:
        section_long(ref, hm_dict, sd)
    if "hand" in sections:
        section_hand(ref, hm_dict)


if __name__ == "__main__":
    main()
