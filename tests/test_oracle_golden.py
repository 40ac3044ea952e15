"""The oracle (oracle/umetrack_oracle.py) against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import synthetic


def _hm_tensors(hm):
    return {k: torch.tensor(v) for k, v in hm.items() if isinstance(v, list)}


def test_warp_maps_and_crops_match_reference(golden_dir, hand_model_dict):
    g = np.load(os.path.join(golden_dir, "tracker_golden.npz"))
    labels, frames = synthetic.make_recording(hand_model_dict, seed=3, n_frames=12)
    rows = g["crop_rows"]
    k_map = 0
    k_crop = 0
    for row in rows:
        t, h, c = int(row[0]), int(row[1]), int(row[2])
        if t >= 2:
            break
        cam = labels["cameras"][c]
        src_xf = np.array(labels["camera_to_world_transforms"][t][c])
        coeffs = [cam[n] for n in ("k1", "k2", "k3", "k4", "p1", "p2", "k5", "k6")]
        f, cc, xf = row[3:5], row[5:7], row[7:].reshape(4, 4)
        mx, my = O.warp_map((cam["fx"], cam["fy"]), (cam["cx"], cam["cy"]), coeffs, src_xf, f, cc, xf)
        if t == 0:
            ref = g["maps"][k_map]
            k_map += 1
            assert np.array_equal(mx, ref[..., 0]) and np.array_equal(my, ref[..., 1])
        src = frames[t].reshape(frames.shape[1], 4, -1)[:, c, :]
        crop = O.remap_u8_bilinear(np.ascontiguousarray(src), mx, my)
        assert np.array_equal(crop, g["crops_u8"][k_crop])
        k_crop += 1
    assert k_crop == len(g["crops_u8"]) and k_map == len(g["maps"])


def test_remap_model_is_bit_exact_vs_cv2():
    cv2 = pytest.importorskip("cv2")
    g = np.random.default_rng(0)
    src = g.integers(0, 256, size=(120, 160), dtype=np.uint8)
    mx = g.uniform(-4, 164, size=(96, 96)).astype(np.float32)
    my = g.uniform(-4, 124, size=(96, 96)).astype(np.float32)
    mx[0, :8] = -1
    my[0, :8] = -1
    mx[1, :4] = 158.984375  # exact 1/64 ties exercise round-half-even
    my[1, :4] = 10.015625
    mx[2, :4] = 1e6
    ref = cv2.remap(src, mx, my, cv2.INTER_LINEAR)
    assert np.array_equal(O.remap_u8_bilinear(src, mx, my), ref)
    # strided source view like frame[:, cam, :] (video_pose_data.py:123-125)
    wide = g.integers(0, 256, size=(120, 4, 160), dtype=np.uint8)
    ref = cv2.remap(wide[:, 2, :], mx, my, cv2.INTER_LINEAR)
    assert np.array_equal(O.remap_u8_bilinear(np.ascontiguousarray(wide[:, 2, :]), mx, my), ref)


def test_model_matches_reference(golden_dir, hand_model_dict):
    g = np.load(os.path.join(golden_dir, "model_golden.npz"))
    sd = synthetic.make_state_dict(0)
    hm = _hm_tensors(hand_model_dict)
    ax = hm["joint_rotation_axes"].float()[None]
    rest = (hm["joint_rest_positions"] * 1e-3).float()[None]
    m = O.OracleModel(sd)
    for step, use_mem in enumerate((False, True)):
        b = synthetic.make_model_batch(6, seed=100 + step, single_view_every=3)
        b["use_memory"][:] = use_mem
        if step == 0:
            f = O.backbone_forward(sd, b["left_images"][:3])
            np.testing.assert_allclose(f.numpy(), g["k_backbone_feats_step0"], atol=2e-5)
        out = m.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"],
                                          b["sample_range"], b["memory_idx"], b["use_memory"],
                                          b["hand_idx"], ax, rest)
        np.testing.assert_allclose(out["joint_angles"].numpy(), g[f"k_joint_angles_{step}"], atol=1e-5)
        np.testing.assert_allclose(out["wrist_xfs"].numpy(), g[f"k_wrist_xfs_{step}"], atol=1e-5)
        np.testing.assert_allclose(out["landmark_uncertainty_sigmas"].numpy(), g[f"k_sigmas_{step}"], atol=1e-5)
    np.testing.assert_allclose(m.state.mem.numpy(), g["k_mem_after"], atol=1e-5)
    m2 = O.OracleModel(sd)
    for step, use_mem in enumerate((False, True)):
        b = synthetic.make_model_batch(4, seed=200 + step)
        b["use_memory"][:] = use_mem
        out = m2.regress_pose_pred_skel_scale(b["left_images"], b["intrinsics"], b["extrinsics_xf"],
                                              b["sample_range"], b["memory_idx"], b["use_memory"],
                                              b["hand_idx"])
        np.testing.assert_allclose(out["skel_scales"].numpy(), g[f"u_scales_{step}"], atol=1e-5)
        np.testing.assert_allclose(out["wrist_xfs"].numpy(), g[f"u_wrist_xfs_{step}"], atol=1e-5)
        np.testing.assert_allclose(out["joint_angles"].numpy(), g[f"u_joint_angles_{step}"], atol=1e-5)


def test_fk_matches_reference_and_rest_pose(golden_dir, hand_model_dict):
    g = np.load(os.path.join(golden_dir, "fk_golden.npz"))
    hm = _hm_tensors(hand_model_dict)
    out = O.skin_landmarks(hm["joint_rotation_axes"], hm["joint_rest_positions"],
                           hm["landmark_rest_positions"], hm["landmark_rest_bone_indices"],
                           hm["landmark_rest_bone_weights"], torch.from_numpy(g["joint_angles"]),
                           torch.from_numpy(g["wrist"]))
    np.testing.assert_allclose(out.numpy(), g["landmarks"], atol=2e-3)  # mm, |coords| up to ~700
    # known answer: rest pose + identity wrist returns the rest landmarks
    np.testing.assert_allclose(out[0].numpy(), hm["landmark_rest_positions"].numpy(), atol=1e-4)


def test_metrics_definitions(golden_dir):
    g = np.load(os.path.join(golden_dir, "tracker_golden.npz"))
    e = O.mpjpe_per_frame(g["gt"], g["tracked"], g["valid"])
    np.testing.assert_allclose(e, g["per_frame_error"], rtol=1e-12)
    a = O.accel_per_frame(g["tracked"], g["valid"])
    assert a.ndim == 1 and len(a) == int((g["valid"][:, :-2] & g["valid"][:, 1:-1] & g["valid"][:, 2:]).sum())


def test_homography_resampler_matches_reference(golden_dir):
    t = np.load(os.path.join(golden_dir, "torchdata_golden.npz"))
    g = np.random.default_rng(21)
    orig = (g.random((3, 2, 96, 128)) * 255).astype(np.float32)   # consume the stream like make_golden.py
    for f in range(3):
        g.uniform(-40, 40), g.uniform(-30, 30), g.uniform(280, 400)
        g.normal(0, 35, size=(21, 3))
    src4 = (g.random((4, 96, 128)) * 255).astype(np.float32)
    out = O.resample_homography(src4, t["raw_xfs"])
    np.testing.assert_allclose(out, t["raw_out"], rtol=0, atol=2e-5)
    assert (out[3] == 0).all()


def test_aggregate_metrics_oracle_matches_reference(golden_dir):
    """PCK curve, AUC and the load_eval summary against the reference's outputs."""
    from tests.helpers import aggregate_inputs

    g = np.load(os.path.join(golden_dir, "aggregate_golden.npz"))
    recs = aggregate_inputs()
    errs = []
    for i, r in enumerate(recs):
        e = O.mpjpe_per_frame(r["gt_keypoints"], r["tracked_keypoints"], r["valid_tracking"])
        assert np.array_equal(e, g[f"errors_{i}"])
        assert np.array_equal(O.accel_per_frame(r["tracked_keypoints"], r["valid_tracking"]), g[f"accel_{i}"])
        assert np.array_equal(O.accel_per_frame(r["gt_keypoints"], r["valid_tracking"]), g[f"gt_accel_{i}"])
        errs.append(e)
    errs = np.concatenate(errs)
    assert np.array_equal(O.PCK_THRESHOLDS, g["thresholds"])
    pck = O.pck_curve(errs, O.PCK_THRESHOLDS) * 100.0
    assert np.array_equal(pck, g["pck_curve"])
    assert float(O.normalized_auc(O.PCK_THRESHOLDS, pck)) == float(g["auc"])
    e2 = np.stack([errs[:60], errs[60:120]])
    mask = (np.arange(120).reshape(2, 60) % 3 != 0).astype(np.float64)
    assert np.array_equal(O.pck_curve(e2, O.PCK_THRESHOLDS, mask=mask, axis=0), g["pck_axis0_masked"])
    a = O.aggregate_metrics(recs)
    text = str(g["summary_text"])
    assert f"Tracked {a['n_valid']} out of {a['n_total']}, success rate: {a['success_rate']}%" in text
    assert f"Mean keypoint error: {a['mean_keypoint_error']}" in text
    assert f"AUC score: {a['auc']}" in text
    assert f"Mean keypoint accelerations: {a['mean_keypoint_acceleration']}" in text
    assert f"GT mean keypoint accelerations: {a['gt_mean_keypoint_acceleration']}" in text


def test_oracle_maps_match_reference_on_every_crop_of_the_256_frame_recording(golden_dir, hand_model_dict):
    """All 1018 crops of the survey-size recording: the CRC of the oracle's fixed-point map equals the
    CRC the unmodified reference produced (tests/golden/make_recording_golden.py long), and the
    sampled uint8 crops are identical in both frame variants."""
    import importlib.util
    import zlib

    spec = importlib.util.spec_from_file_location("make_recording_golden", os.path.join(golden_dir, "make_recording_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    g = np.load(os.path.join(golden_dir, "tracker_long_golden.npz"))
    rows, pick = g["crop_rows"], set(int(i) for i in g["crop_pick"])
    assert len(rows) == len(g["map_crc"]) == 1018
    recs = {v: synthetic.make_recording(mg.long_hand_model_dict(hand_model_dict), seed=mg.LONG_SEED, n_frames=mg.LONG_T,
                                        smooth_images=(v == "smooth")) for v in ("noise", "smooth")}
    labels = recs["noise"][0]
    crops = {"noise": [], "smooth": []}
    for k, row in enumerate(rows):
        t, c = int(row[0]), int(row[2])
        cam = labels["cameras"][c]
        src_xf = np.array(labels["camera_to_world_transforms"][t][c])
        coeffs = [cam[n] for n in ("k1", "k2", "k3", "k4", "p1", "p2", "k5", "k6")]
        mx, my = O.warp_map((cam["fx"], cam["fy"]), (cam["cx"], cam["cy"]), coeffs, src_xf, row[3:5], row[5:7],
                            row[7:].reshape(4, 4))
        assert mg.map_crc(np.stack([mx, my], axis=-1)) == int(g["map_crc"][k]), f"crop {k}"
        ix, iy, fx, fy = O.remap_fixed_point_indices(mx, my)
        s = np.stack([ix * 32 + fx, iy * 32 + fy], axis=-1).astype(np.int32)
        assert zlib.crc32(np.ascontiguousarray(s).tobytes()) == int(g["map_crc"][k])
        if k in pick:
            for v in crops:
                fr = recs[v][1]
                src = fr[t].reshape(fr.shape[1], 4, -1)[:, c, :]
                crops[v].append(O.remap_u8_bilinear(np.ascontiguousarray(src), mx, my))
    assert np.array_equal(np.stack(crops["noise"]), g["noise_crops_u8"])
    assert np.array_equal(np.stack(crops["smooth"]), g["smooth_crops_u8"])


def test_scaled_and_mirrored_hand_model_match_reference(golden_dir, hand_model_dict):
    """C4: scaled_hand_model / mirrored_hand_model against the unmodified reference
    (lib/common/hand.py:65-134), single and batched models.  Host code: runs without a GPU."""
    from umetrack_b200 import hand_model as hmod

    g = np.load(os.path.join(golden_dir, "hand_model_golden.npz"))
    hm = hmod.load_hand_model_from_dict(hand_model_dict)
    batched = type(hm)(*[(v[None].repeat(3, *([1] * v.dim())) if isinstance(v, torch.Tensor) and v.dim() > 0 else v)
                         for v in hm])
    s1, s3 = hmod.scaled_hand_model(hm, 1.0375), hmod.scaled_hand_model(batched, 0.91)
    assert np.array_equal(s1.joint_rest_positions.numpy(), g["s1_rest"])
    assert np.array_equal(s1.landmark_rest_positions.numpy(), g["s1_lm"])
    assert np.array_equal(s3.joint_rest_positions.numpy(), g["s3_rest"])
    assert np.array_equal(s3.landmark_rest_positions.numpy(), g["s3_lm"])
    assert s1.joint_rotation_axes is hm.joint_rotation_axes
    m3 = hmod.mirrored_hand_model(batched, torch.tensor([True, False, True]))
    one = type(hm)(*[(v[None] if isinstance(v, torch.Tensor) and v.dim() > 0 else v) for v in hm])
    m1 = hmod.mirrored_hand_model(one, torch.tensor([True]))
    for m, p in ((m3, "m3"), (m1, "m1")):
        assert np.array_equal(m.joint_rotation_axes.numpy(), g[p + "_axes"])
        assert np.array_equal(m.joint_rest_positions.numpy(), g[p + "_rest"])
        assert np.array_equal(m.landmark_rest_positions.numpy(), g[p + "_lm"])
    assert np.array_equal(batched.joint_rest_positions.numpy()[0], hm.joint_rest_positions.numpy())   # inputs untouched
