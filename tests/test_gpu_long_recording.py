"""Parity at the survey's sizes: one 256-frame, two-hand, 4-camera
recording with confidence dropouts and 1-view cases, in a seeded-noise and a smooth frame variant,
against vectors from the unmodified reference (tests/golden/make_recording_golden.py long).

Reference loops matched: run_eval_known_skeleton.py:48-102, run_eval_unknown_skeleton.py:49-131.
Gates (north_star): crop pixel indices bit-exact, crops identical as uint8 (=> within 1e-3),
keypoints within 0.05 mm, MPJPE / acceleration within 0.01, mean scale within 1e-4.
"""
import functools
import importlib.util
import os
import zlib

import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import hand_model as hmod
from umetrack_b200 import native, synthetic
from umetrack_b200.batched import crop_cameras_device, track_recordings_batched
from umetrack_b200.cameras import resample_params
from umetrack_b200.evaluation import RecordingStream, track_recording, track_recording_unknown_skeleton
from umetrack_b200.hand_tracker import HandTracker, HandTrackerOpts
from umetrack_b200.metrics import MetricSums
from umetrack_b200.model import UmeTrackModel
from tests.helpers import GOLDEN, crop_params_from_row, golden, hm_dict

pytestmark = pytest.mark.gpu


@functools.lru_cache(maxsize=None)
def _gen():
    spec = importlib.util.spec_from_file_location("make_recording_golden", os.path.join(GOLDEN, "make_recording_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@functools.lru_cache(maxsize=2)
def _recording(smooth):
    mg = _gen()
    return synthetic.make_recording(mg.long_hand_model_dict(hm_dict()), seed=mg.LONG_SEED, n_frames=mg.LONG_T,
                                    smooth_images=smooth)


def _check_tracked(res, g, prefix):
    valid = g[prefix + "_valid"]
    assert np.array_equal(res["valid_tracking"], valid)
    d = np.abs(res["tracked_keypoints"] - g[prefix + "_tracked"].astype(np.float64))[valid]
    assert d.max() < 0.05, f"{prefix}: keypoints differ from the reference by {d.max()} mm"
    assert np.abs(res["gt_keypoints"] - g["gt"].astype(np.float64))[valid].max() < 0.05
    return float(d.max())


@pytest.mark.parametrize("variant", ["noise", "smooth"])
def test_known_skeleton_256_frames_per_frame_tracker(variant):
    g = golden("tracker_long_golden.npz")
    labels, frames = _recording(variant == "smooth")
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    sums = MetricSums("cuda")
    res = track_recording(model, labels, frames, sums=sums)
    _check_tracked(res, g, f"{variant}_k")
    r = sums.result()
    assert abs(r["mpjpe"] - float(g[f"{variant}_k_mpjpe"])) < 0.01
    assert abs(r["mpjpa"] - float(g[f"{variant}_k_accel"])) < 0.01          # "MPJPA" of the README
    assert abs(r["gt_mpjpa"] - float(g[f"{variant}_k_gt_accel"])) < 0.01
    assert abs(res["per_frame_mean_error"].mean() - float(g[f"{variant}_k_mpjpe"])) < 0.01
    model.check_status()


@pytest.mark.parametrize("variant", ["noise", "smooth"])
def test_unknown_skeleton_256_frames_calibration_and_pass_two(variant):
    g = golden("tracker_long_golden.npz")
    labels, frames = _recording(variant == "smooth")
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    generic = hmod.load_hand_model_from_dict(hm_dict())
    sums = MetricSums("cuda")
    res, mean_scale, scales = track_recording_unknown_skeleton(model, labels, frames, generic, _gen().N_CALIBRATION, sums=sums)
    ref_scales = g[f"{variant}_u_scales"]
    assert len(scales) == len(ref_scales) == 30
    assert np.abs(scales - ref_scales).max() < 1e-4
    assert abs(mean_scale - float(g[f"{variant}_u_mean_scale"])) < 1e-4
    _check_tracked(res, g, f"{variant}_u")
    r = sums.result()
    assert abs(r["mpjpe"] - float(g[f"{variant}_u_mpjpe"])) < 0.01
    assert abs(r["mpjpa"] - float(g[f"{variant}_u_accel"])) < 0.01


def test_batched_driver_256_frames_matches_reference():
    g = golden("tracker_long_golden.npz")
    labels, frames = _recording(False)
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    sums = MetricSums("cuda")
    res = track_recordings_batched(model, [(labels, frames)], sums=sums)[0]
    _check_tracked(res, g, "noise_k")
    assert abs(sums.result()["mpjpe"] - float(g["noise_k_mpjpe"])) < 0.01
    assert abs(sums.result()["mpjpa"] - float(g["noise_k_accel"])) < 0.01
    model.check_status()


def _resample_with_indices(frames_dev, H, W4, W, offsets, params, want_u8=False):
    dev = frames_dev.device
    n = len(offsets)
    off = torch.tensor(offsets, dtype=torch.int64, device=dev)
    par = torch.from_numpy(np.asarray(params, dtype=np.float64).reshape(n, 48)).to(dev)
    oi = torch.empty((n, 96, 96, 4), dtype=torch.int32, device=dev)
    ou = torch.empty((n, 96, 96), dtype=torch.uint8, device=dev) if want_u8 else None
    native.check(native.lib().umt_resample_crops(frames_dev.data_ptr(), off.data_ptr(), W4, W, H, par.data_ptr(), n,
                                                 native.ptr(ou), 0, oi.data_ptr(), 0), "resample")
    torch.cuda.synchronize()
    return oi.cpu().numpy(), (ou.cpu().numpy() if want_u8 else None)


def _crc_of_indices(idx):
    """CRC32 of (sx, sy) = (32*ix + fx, 32*iy + fy), the fixed-point map cv2.remap works from."""
    s = np.stack([idx[..., 0] * 32 + idx[..., 2], idx[..., 1] * 32 + idx[..., 3]], axis=-1).astype(np.int32)
    return zlib.crc32(np.ascontiguousarray(s).tobytes())


@pytest.mark.parametrize("device_cameras", [True, False])
def test_end_to_end_crop_indices_with_our_own_crop_cameras(device_cameras):
    """north_star: crop pixel indices bit-exact.  The crop cameras here are the repo's own
    (HandTracker.gen_crop_cameras: float64 device kernel, or the host numpy construction), fed by the
    repo's own float32 FK on the GPU -- not the reference's parameters -- and the resulting
    (ix, iy, fx, fy) of every pixel of every crop of the 256-frame recording is compared with the
    indices the reference's own crop cameras give (lib/tracker/tracker.py:54-80,
    lib/tracker/perspective_crop.py:40-51,89-115): per crop against the CRC of the reference's
    fixed-point map, per pixel against the oracle evaluated on the reference's crop cameras."""
    g = golden("tracker_long_golden.npz")
    labels, frames = _recording(False)
    T, H, W4 = frames.shape
    W = W4 // 4
    dev = torch.device("cuda")
    tracker = HandTracker(UmeTrackModel(synthetic.make_state_dict(0)), HandTrackerOpts())
    tracker._device_crop_cameras = device_cameras
    stream = RecordingStream(labels, frames)
    keys, params, offsets = [], [], []
    for t, (frame, gt) in enumerate(stream):
        cc = tracker.gen_crop_cameras([v.camera for v in frame.views], stream.camera_angles, stream.hand_model, gt, 1)
        for h, per in cc.items():
            for c, cam in per.items():
                keys.append((t, h, c))
                params.append(resample_params(frame.views[c].camera, cam))
                offsets.append(t * H * W4 + c * W)
    rows = g["crop_rows"]
    assert keys == [(int(r[0]), int(r[1]), int(r[2])) for r in rows]       # same hands, same cameras, same order
    fr = torch.from_numpy(frames).to(dev)
    idx, crops = _resample_with_indices(fr, H, W4, W, offsets, params, want_u8=True)
    crc_mismatch = [k for k in range(len(keys)) if _crc_of_indices(idx[k]) != int(g["map_crc"][k])]
    # per-pixel count for the crops whose CRC differs (and a sample of the others as a self-check)
    mism = 0
    check = sorted(set(crc_mismatch) | set(range(0, len(keys), 97)))
    for k in check:
        t, h, c, sf, sc, co, sxf, cf, ccn, cxf = crop_params_from_row(labels, rows[k])
        mx, my = O.warp_map(sf, sc, co, sxf, cf, ccn, cxf)
        ref = np.stack(O.remap_fixed_point_indices(mx, my), axis=-1)
        assert _crc_of_indices(ref) == int(g["map_crc"][k])                 # oracle == reference on this crop
        mism += int((ref != idx[k]).any(axis=-1).sum())
    total = len(keys) * 96 * 96
    assert mism == 0 and not crc_mismatch, (
        f"{mism} of {total} crop pixel indices ({len(crc_mismatch)} of {len(keys)} crops) differ from the "
        f"reference's with {'device' if device_cameras else 'host'} crop cameras")
    # and the sampled uint8 crops are the reference's cv2.remap output
    pick = g["crop_pick"]
    assert np.array_equal(crops[pick], g["noise_crops_u8"])


def test_batched_device_crop_cameras_give_the_reference_indices_and_smooth_crops():
    """Same index gate for the batched driver's crop cameras (crop_cameras_device), and the sampled
    crops of the smooth frame variant."""
    g = golden("tracker_long_golden.npz")
    labels, frames = _recording(True)
    T, H, W4 = frames.shape
    W = W4 // 4
    dev = torch.device("cuda")
    cc = crop_cameras_device(hmod.load_hand_model_from_dict(labels["hand_model"]), labels, dev)
    cnt, sel, par = cc["sel_count"].cpu().numpy(), cc["sel_cam"].cpu().numpy(), cc["resample_params"].cpu().numpy()
    rows = g["crop_rows"]
    keys, params, offsets = [], [], []
    for t in range(T):
        for h in range(2):
            for k in range(int(cnt[t, h])):
                keys.append((t, h, int(sel[t, h, k])))
                params.append(par[t, h, k])
                offsets.append(t * H * W4 + int(sel[t, h, k]) * W)
    assert keys == [(int(r[0]), int(r[1]), int(r[2])) for r in rows]
    idx, crops = _resample_with_indices(torch.from_numpy(frames).to(dev), H, W4, W, offsets, params, want_u8=True)
    bad = [k for k in range(len(keys)) if _crc_of_indices(idx[k]) != int(g["map_crc"][k])]
    assert not bad, f"{len(bad)} of {len(keys)} crops have index differences"
    assert np.array_equal(crops[g["crop_pick"]], g["smooth_crops_u8"])
