"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference modules are imported as they are; the two third-party modules that are absent
here are stubbed before import: `pytorch3d.transforms.so3_exp_map` (restated, see
oracle/umetrack_oracle.py: parity unpinned at that one boundary) and `av` (video decode only;
frames are supplied as uint8 arrays).  Inputs are regenerated from seeds by
`umetrack_b200.synthetic`, so only outputs (and small parameter blocks) are stored.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("UMETRACK_REFERENCE", "/root/reference")


def install_reference():
    sys.path.insert(0, REF)
    from oracle.umetrack_oracle import so3_exp_map

    p3d = types.ModuleType("pytorch3d")
    tr = types.ModuleType("pytorch3d.transforms")
    tr.so3_exp_map = so3_exp_map
    p3d.transforms = tr
    sys.modules["pytorch3d"] = p3d
    sys.modules["pytorch3d.transforms"] = tr
    sys.modules["av"] = types.ModuleType("av")


def build_reference_model(sd):
    from lib.models import feature_extractor as fe, regressor as reg, skeleton_encoder as se
    from lib.models import temporal as tem
    from lib.models.model_opts import ModelOpts
    from lib.models.model_loader import _create_regressor
    from lib.models.umetrack_model import UmeTrackModel

    mo = ModelOpts()
    fx = fe.FeatureExtractor((96, 96), ModelOpts())
    model = UmeTrackModel(
        feature_extractor=fx,
        temporal=tem.create_temporal_model(mo, fx.output_feature_sizes),
        skeleton_encoder=se.SkeletonEncoder([mo.nSkeletonFeatureChannels, *fx.output_feature_sizes]),
        regressor_k=_create_regressor(mo, fx.output_feature_sizes, use_skel=True, predict_skel_scale=False),
        regressor_u=_create_regressor(mo, fx.output_feature_sizes, use_skel=False, predict_skel_scale=True),
    )
    model.load_state_dict(sd)  # strict: proves the synthetic key/shape list matches the reference
    model.eval()
    return model


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    install_reference()
    from umetrack_b200 import synthetic

    from lib.common.hand_skinning import skin_landmarks
    from lib.models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData
    from lib.tracker import tracker as rtracker
    from lib.tracker.perspective_crop import landmarks_from_hand_pose
    from lib.tracker.tracking_result import SingleHandPose
    from lib.tracker.video_pose_data import load_hand_model_from_dict
    from lib.common.camera import read_camera_from_json

    hm_dict = synthetic.load_hand_model_json(os.path.join(HERE, "generic_hand_model_nomesh.json"))
    hand_model = load_hand_model_from_dict(hm_dict)
    sd = synthetic.make_state_dict(0)

    # ---------------- model: two time steps on a mixed 1/2-view batch ----------------
    model = build_reference_model(sd)
    out = {}
    skel = InputSkeletonData(
        joint_rotation_axes=hand_model.joint_rotation_axes.float()[None],
        joint_rest_positions=(hand_model.joint_rest_positions * 1e-3).float()[None],
    )
    with torch.no_grad():
        for step, use_mem in enumerate((False, True)):
            b = synthetic.make_model_batch(6, seed=100 + step, single_view_every=3)
            b["use_memory"][:] = use_mem
            fd = InputFrameData(b["left_images"], b["intrinsics"], b["extrinsics_xf"])
            desc = InputFrameDesc(b["sample_range"], b["memory_idx"], b["use_memory"], b["hand_idx"])
            if step == 0:
                feats = model._feature_extractor._image_backbone(b["left_images"].unsqueeze(1))
                out["k_backbone_feats_step0"] = feats[:3].numpy()
            r = model.regress_pose_use_skeleton(fd, desc, skel)
            out[f"k_joint_angles_{step}"] = r.joint_angles.numpy()
            out[f"k_wrist_xfs_{step}"] = r.wrist_xfs.numpy()
            out[f"k_sigmas_{step}"] = r.landmark_uncertainty_sigmas.numpy()
        out["k_mem_after"] = model._temporal._mem_features.numpy().copy()
        # unknown-skeleton head on an all-two-view batch (fresh recurrent state)
        model2 = build_reference_model(sd)
        for step, use_mem in enumerate((False, True)):
            b = synthetic.make_model_batch(4, seed=200 + step)
            b["use_memory"][:] = use_mem
            fd = InputFrameData(b["left_images"], b["intrinsics"], b["extrinsics_xf"])
            desc = InputFrameDesc(b["sample_range"], b["memory_idx"], b["use_memory"], b["hand_idx"])
            r = model2.regress_pose_pred_skel_scale(fd, desc)
            out[f"u_joint_angles_{step}"] = r.joint_angles.numpy()
            out[f"u_wrist_xfs_{step}"] = r.wrist_xfs.numpy()
            out[f"u_scales_{step}"] = r.skel_scales.numpy()
            out[f"u_sigmas_{step}"] = r.landmark_uncertainty_sigmas.numpy()
    np.savez_compressed(os.path.join(HERE, "model_golden.npz"), **out)

    # ---------------- FK / skinning ----------------
    g = np.random.default_rng(5)
    n = 16
    lim = hand_model.joint_limits.numpy()
    ja = (lim[:, 0] + g.random((n, 22)) * (lim[:, 1] - lim[:, 0])).astype(np.float32)
    ja[0] = 0
    wr = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    for i in range(1, n):
        wr[i, :3, :3] = synthetic._random_rotation(g, 2.0)
        wr[i, :3, 3] = g.normal(0, 200, size=3)
    wr[n // 2:, :, 0] *= -1  # mirrored (right-hand style) transforms too
    hm_b = type(hand_model)(*[
        (v[None].expand(n, *v.shape) if isinstance(v, torch.Tensor) and v.dim() > 0 else v)
        for v in hand_model
    ])
    lms = skin_landmarks(hm_b, torch.from_numpy(ja), torch.from_numpy(wr)).numpy()
    np.savez_compressed(os.path.join(HERE, "fk_golden.npz"), joint_angles=ja, wrist=wr, landmarks=lms)

    # ---------------- tracker loop on a synthetic recording ----------------
    T = 12
    labels, frames = synthetic.make_recording(hm_dict, seed=3, n_frames=T)
    cams = [read_camera_from_json(c) for c in labels["cameras"]]
    cam_xf = np.array(labels["camera_to_world_transforms"])
    ja_gt = np.array(labels["joint_angles"])
    wr_gt = np.array(labels["wrist_transforms"])
    conf = np.array(labels["hand_confidences"])
    trk_model = build_reference_model(sd)
    tracker = rtracker.HandTracker(trk_model, rtracker.HandTrackerOpts())
    tracked = np.zeros((2, T, 21, 3))
    gtk = np.zeros((2, T, 21, 3))
    valid = np.zeros((2, T), dtype=bool)
    crop_rows, crops_u8, maps = [], [], []
    out_ja = np.zeros((2, T, 22), dtype=np.float32)
    out_wr = np.zeros((2, T, 4, 4), dtype=np.float32)
    for t in range(T):
        gt = {}
        for h in range(2):
            if conf[t, h] > 0:
                gt[h] = SingleHandPose(ja_gt[t, h], wr_gt[t, h], conf[t, h])
        mv = frames[t].reshape(frames.shape[1], 4, -1)
        views = [rtracker.ViewData(image=mv[:, c, :], camera=cams[c].copy(camera_to_world_xf=cam_xf[t, c]),
                                   camera_angle=labels["camera_angles"][c]) for c in range(4)]
        frame = rtracker.InputFrame(views=views)
        cc = tracker.gen_crop_cameras([v.camera for v in views], labels["camera_angles"], hand_model, gt, 1)
        for h, per in cc.items():
            for c, cam in per.items():
                crop_rows.append(np.concatenate([[t, h, c], cam.f, cam.c, cam.camera_to_world_xf.reshape(-1)]))
                if t < 2:
                    img = rtracker._warp_image(views[c].camera, cam, views[c].image)
                    crops_u8.append(img)
                    if t == 0:
                        # recompute the float32 map exactly as _warp_image does (tracker.py:61-78)
                        px, py = np.meshgrid(np.arange(96), np.arange(96))
                        w = np.column_stack((px.flatten(), py.flatten()))
                        e = views[c].camera.world_to_eye(cam.eye_to_world(cam.window_to_eye(w)))
                        sw = views[c].camera.eye_to_window(e)
                        sw[e[:, 2] < 0] = -1
                        maps.append(sw.astype(np.float32).reshape(96, 96, 2))
        res = tracker.track_frame(frame, hand_model, cc)
        for h, pose in res.hand_poses.items():
            tracked[h, t] = landmarks_from_hand_pose(hand_model, pose, h)
            gtk[h, t] = landmarks_from_hand_pose(hand_model, gt[h], h)
            valid[h, t] = True
            out_ja[h, t] = pose.joint_angles
            out_wr[h, t] = pose.wrist_xform
    err = np.linalg.norm((gtk - tracked)[valid], axis=-1).mean(axis=-1)
    np.savez_compressed(
        os.path.join(HERE, "tracker_golden.npz"),
        crop_rows=np.array(crop_rows), crops_u8=np.array(crops_u8), maps=np.array(maps),
        tracked=tracked, gt=gtk, valid=valid, per_frame_error=err,
        joint_angles=out_ja, wrist=out_wr,
    )

    # ---------------- unknown-skeleton calibration pass ----------------
    cal_model = build_reference_model(sd)
    tracker = rtracker.HandTracker(cal_model, rtracker.HandTrackerOpts())
    scales = []
    for t in range(T):
        gt = {h: SingleHandPose(ja_gt[t, h], wr_gt[t, h], conf[t, h]) for h in range(2) if conf[t, h] > 0}
        mv = frames[t].reshape(frames.shape[1], 4, -1)
        views = [rtracker.ViewData(image=mv[:, c, :], camera=cams[c].copy(camera_to_world_xf=cam_xf[t, c]),
                                   camera_angle=labels["camera_angles"][c]) for c in range(4)]
        cc = tracker.gen_crop_cameras([v.camera for v in views], labels["camera_angles"], hand_model, gt, 2)
        res = tracker.track_frame_and_calibrate_scale(rtracker.InputFrame(views=views), cc)
        for h in range(2):
            scales.append(res.predicted_scales.get(h, np.nan))
    np.savez_compressed(os.path.join(HERE, "calibration_golden.npz"), scales=np.array(scales, dtype=np.float64))
    # ---------------- torch_data variant: homography resampling (data_transform.py:57-144,216-281) ----
    from lib.batched_dataset import data_transform as dt

    g = np.random.default_rng(21)
    n_frames, n_views, hh, ww = 3, 2, 96, 128
    orig = (g.random((n_frames, n_views, hh, ww)) * 255).astype(np.float32)
    intr = np.zeros((n_frames, n_views, 3, 3), dtype=np.float32)
    extr = np.zeros((n_frames, n_views, 4, 4), dtype=np.float32)
    pts = np.zeros((n_frames, 21, 3), dtype=np.float32)
    for f in range(n_frames):
        centre = np.array([g.uniform(-40, 40), g.uniform(-30, 30), g.uniform(280, 400)])
        pts[f] = centre + g.normal(0, 35, size=(21, 3))
        for v in range(n_views):
            intr[f, v] = [[150 + 10 * v, 0, (ww - 1) / 2], [0, 150 + 10 * v, (hh - 1) / 2], [0, 0, 1]]
            c2w = np.eye(4)
            c2w[:3, :3] = synthetic._rot_y(0.1 * (1 - 2 * v)) @ synthetic._rot_x(0.05 * f)
            c2w[:3, 3] = [40.0 * (2 * v - 1), 3.0 * f, 0.0]
            extr[f, v] = np.linalg.inv(c2w)
    td = {}
    for hand in (0, 1):
        imgs, ex, kk = dt._perspective_crop_images(orig, extr, intr, pts, hand, (96, 96))
        td[f"crops_{hand}"] = imgs.astype(np.float32)
        td[f"extr_{hand}"] = ex
        td[f"intr_{hand}"] = kk
    # raw kernel check: transforms that throw part of the grid outside the source image
    xfs = np.tile(np.eye(4, dtype=np.float32), (4, 1, 1))
    xfs[0, 0, 3] = -7.25
    xfs[1, :2, :2] = [[1.3, 0.2], [-0.1, 0.9]]
    xfs[1, :2, 3] = [10.5, -3.0]
    xfs[2, 2, :] = [1e-3, -2e-3, 1.0, 0.0]          # perspective row (z depends on u, v)
    xfs[3, :2, 3] = [200.0, 200.0]                   # everything outside -> zeros
    src4 = (g.random((4, hh, ww)) * 255).astype(np.float32)
    out4 = np.zeros((4, 96, 96), dtype=np.float32)
    dt._resample_images_batched(src4, out4, xfs)
    td["raw_xfs"], td["raw_out"] = xfs, out4
    np.savez_compressed(os.path.join(HERE, "torchdata_golden.npz"), **td)
    print("golden written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
    print("tracker per-frame error (mm):", err[:4], "n crops", len(crop_rows))


if __name__ == "__main__":
    main()
