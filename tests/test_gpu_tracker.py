"""End-to-end tracker parity with the reference's per-frame loop (run_eval_known_skeleton.py:48-102)."""
import numpy as np
import pytest
import torch

from umetrack_b200 import hand_model as hmod
from umetrack_b200 import synthetic
from umetrack_b200.evaluation import calibrate_scale, track_recording
from umetrack_b200.hand_tracker import HandTracker, HandTrackerOpts, TrackingResult
from umetrack_b200.metrics import MetricSums
from umetrack_b200.model import UmeTrackModel
from tests.helpers import golden, hm_dict

pytestmark = pytest.mark.gpu


def test_known_skeleton_recording_matches_reference():
    g = golden("tracker_golden.npz")
    labels, frames = synthetic.make_recording(hm_dict(), seed=3, n_frames=12)
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    sums = MetricSums("cuda")
    res = track_recording(model, labels, frames, sums=sums)
    assert np.array_equal(res["valid_tracking"], g["valid"])        # includes the confidence dropouts
    d = np.abs(res["tracked_keypoints"] - g["tracked"])[g["valid"]]
    assert d.max() < 0.05, f"keypoints differ by {d.max()} mm"       # north_star: 0.05 mm
    assert np.abs(res["gt_keypoints"] - g["gt"]).max() < 0.05
    assert abs(res["per_frame_mean_error"].mean() - g["per_frame_error"].mean()) < 0.01   # MPJPE gate
    assert abs(sums.result()["mpjpe"] - g["per_frame_error"].mean()) < 0.01


def test_empty_crop_cameras_resets_history():
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    tr = HandTracker(model, HandTrackerOpts())
    tr._valid_tracking_history[:] = True
    out = tr.track_frame(None, None, {})
    assert out == TrackingResult() and not tr._valid_tracking_history.any()
    assert tr.gen_crop_cameras([], [], None, {}, 1) == {}


def test_unknown_skeleton_calibration_matches_reference():
    g = golden("calibration_golden.npz")["scales"]
    labels, frames = synthetic.make_recording(hm_dict(), seed=3, n_frames=12)
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    generic = hmod.load_hand_model_from_dict(hm_dict())
    scaled, mean_scale, scales = calibrate_scale(model, labels, frames, generic, n_calibration_samples=30)
    ref = g[~np.isnan(g)]
    assert len(scales) == len(ref)
    assert np.abs(scales - ref).max() < 1e-4
    assert abs(mean_scale - ref.mean()) < 1e-4
    assert torch.allclose(scaled.joint_rest_positions, generic.joint_rest_positions * mean_scale)


def test_known_skeleton_recording_in_fp16_mode_meets_the_gates():
    g = golden("tracker_golden.npz")
    labels, frames = synthetic.make_recording(hm_dict(), seed=3, n_frames=12)
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda").set_mode("fp16")
    res = track_recording(model, labels, frames)
    assert np.array_equal(res["valid_tracking"], g["valid"])
    d = np.abs(res["tracked_keypoints"] - g["tracked"])[g["valid"]]
    assert d.max() < 0.05, f"fp16 keypoints differ by {d.max()} mm"
    assert abs(res["per_frame_mean_error"].mean() - g["per_frame_error"].mean()) < 0.01


def test_device_and_host_crop_cameras_agree():
    """HandTracker.gen_crop_cameras: the float64 device kernel (default for FishEye62 rigs) against the
    numpy construction the reference uses -- same hands, same cameras in the same order, parameters
    equal to float64 round-off (so crop pixel indices cannot differ in practice)."""
    from umetrack_b200.evaluation import RecordingStream
    from umetrack_b200.hand_tracker import HandTracker, HandTrackerOpts

    labels, frames = synthetic.make_recording(hm_dict(), seed=3, n_frames=24)
    tr = HandTracker(UmeTrackModel(synthetic.make_state_dict(0)), HandTrackerOpts())
    st = RecordingStream(labels, frames)
    n = 0
    for frame, gt in st:
        cams = [v.camera for v in frame.views]
        tr._device_crop_cameras = True
        a = tr.gen_crop_cameras(cams, st.camera_angles, st.hand_model, gt, 1)
        tr._device_crop_cameras = False
        b = tr.gen_crop_cameras(cams, st.camera_angles, st.hand_model, gt, 1)
        assert list(a.keys()) == list(b.keys())
        for h in a:
            assert list(a[h].keys()) == list(b[h].keys())
            for c in a[h]:
                x, y = a[h][c], b[h][c]
                assert (x.width, x.height) == (y.width, y.height) == (96, 96)
                np.testing.assert_allclose(np.array(x.f), np.array(y.f), rtol=1e-11)
                np.testing.assert_allclose(np.array(x.c), np.array(y.c), rtol=1e-11)
                np.testing.assert_allclose(x.camera_to_world_xf, y.camera_to_world_xf, rtol=0, atol=1e-9)
                n += 1
    assert n > 60
    # min_num_crops = 2 drops single-view hands on both paths alike
    frame, gt = next(iter(st))
    cams = [v.camera for v in frame.views]
    tr._device_crop_cameras = True
    a = tr.gen_crop_cameras(cams, st.camera_angles, st.hand_model, gt, 2)
    tr._device_crop_cameras = False
    b = tr.gen_crop_cameras(cams, st.camera_angles, st.hand_model, gt, 2)
    assert list(a.keys()) == list(b.keys())
    assert tr.gen_crop_cameras(cams, st.camera_angles, st.hand_model, {}, 1) == {}


def test_hand_set_changing_at_equal_size_matches_reference():
    """Hand 0 alone for three frames, then hand 1 alone, then both: the memory index tensor keeps its
    size (and, with a caching allocator, its address) while its VALUE changes from [0] to [1].  The
    reference grows its memory bank when hand 1 first appears alone (lib/models/temporal.py:101-123,
    lib/tracker/tracker.py:278-285); golden from the unmodified reference
    (tests/golden/make_recording_golden.py handset)."""
    import importlib.util
    import os

    from tests.helpers import GOLDEN
    spec = importlib.util.spec_from_file_location("make_recording_golden", os.path.join(GOLDEN, "make_recording_golden.py"))
    mrg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mrg)
    g = golden("tracker_handset_golden.npz")
    labels, frames = synthetic.make_recording(hm_dict(), seed=5, n_frames=9, presence=mrg.handset_presence())
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    res = track_recording(model, labels, frames)
    assert np.array_equal(res["valid_tracking"], g["valid"]) and g["valid"].sum() == 12
    d = np.abs(res["tracked_keypoints"] - g["tracked"])[g["valid"]]
    assert d.max() < 0.05, f"keypoints differ by {d.max()} mm"
    assert np.abs(res["per_frame_mean_error"] - g["per_frame_error"]).max() < 0.01
    model.check_status()          # no memory index fell outside the bank
    assert len(model._mem_features) == 2


def test_persistent_frame_path_equals_the_general_path():
    """track_frame takes the persistent-buffer path (device-side camera rows, CUDA graph reading its
    inputs in place, packed outputs) when it is handed gen_crop_cameras' own dict, and the general path
    (camera matrices rebuilt on the host, separate uploads) for any other dict.  Same recording through
    both: poses agree to float32 round-off of the camera matrices, and both use the recurrent memory."""
    from umetrack_b200.evaluation import RecordingStream
    labels, frames = synthetic.make_recording(hm_dict(), seed=11, n_frames=10)
    sd = synthetic.make_state_dict(0)
    poses = {}
    for variant in ("persistent", "general", "general-no-graph"):
        model = UmeTrackModel(sd).to("cuda")
        tr = HandTracker(model, HandTrackerOpts())
        tr.use_cuda_graphs = variant != "general-no-graph"
        stream = RecordingStream(labels, frames)
        out = []
        for frame, gt in stream:
            cc = tr.gen_crop_cameras([v.camera for v in frame.views], stream.camera_angles, stream.hand_model, gt, 1)
            if variant != "persistent":
                cc = {h: dict(v) for h, v in cc.items()}          # equal content, not the tracker's own object
            res = tr.track_frame(frame, stream.hand_model, cc)
            took_fast = tr._crop_cache is not None and tr._crop_cache["out"] is cc
            assert took_fast == (variant == "persistent")
            out.append({h: (p.joint_angles.copy(), p.wrist_xform.copy()) for h, p in res.hand_poses.items()})
        poses[variant] = out
    n_checked = 0
    for a, b, c in zip(poses["persistent"], poses["general"], poses["general-no-graph"]):
        assert a.keys() == b.keys() == c.keys()
        for h in a:
            for other in (b, c):
                assert np.abs(a[h][0] - other[h][0]).max() < 2e-5            # joint angles (rad)
                assert np.abs(a[h][1] - other[h][1]).max() < 2e-3            # wrist transform (mm)
            n_checked += 1
    assert n_checked >= 10
