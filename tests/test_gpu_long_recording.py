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


def test_kernel_crop_indices_bit_exact_on_every_reference_crop_camera():
    """north_star: crop pixel indices bit-exact.  Kernel level, full size: the resample kernel is fed
    the crop cameras the unmodified reference built for all 1018 crops of the 256-frame recording
    (lib/tracker/tracker.py:101-139) and must reproduce the fixed-point map cv2.remap derives from
    `_warp_image`'s float32 map (tracker.py:54-80) for every one of the 9.38 M pixels -- checked
    against the CRC the reference itself produced -- and the reference's uint8 crops."""
    g = golden("tracker_long_golden.npz")
    rows = g["crop_rows"]
    for variant, smooth in (("noise", False), ("smooth", True)):
        labels, frames = _recording(smooth)
        T, H, W4 = frames.shape
        W = W4 // 4
        params, offsets = [], []
        for r in rows:
            t, h, c, sf, sc, co, sxf, cf, ccn, cxf = crop_params_from_row(labels, r)
            params.append(np.concatenate([sf, sc, co, sxf.reshape(-1), cf, ccn, cxf.reshape(-1)]))
            offsets.append(t * H * W4 + c * W)
        idx, crops = _resample_with_indices(torch.from_numpy(frames).cuda(), H, W4, W, offsets, params, want_u8=True)
        if variant == "noise":
            bad = [k for k in range(len(rows)) if _crc_of_indices(idx[k]) != int(g["map_crc"][k])]
            assert not bad, f"{len(bad)} of {len(rows)} crops have index differences at kernel level"
        assert np.array_equal(crops[g["crop_pick"]], g[f"{variant}_crops_u8"])


# What the end-to-end index count can and cannot be.  The crop cameras are a float64 function of 63
# float32 crop points, which the reference gets from torch-CPU float32 forward kinematics
# (lib/common/hand_skinning.py:17-67): sin / cos there come from Intel MKL's vector math library
# (closed source, <1 ulp but not correctly rounded: 5 % of its values differ from the correctly
# rounded result), after pytorch3d's so3_exp_map (absent here: parity unpinned at that boundary).
# Neither can be reproduced bit for bit on the GPU, so some crop points differ in their last bit,
# the crop camera moves by ~1e-7 relative, and a pixel whose source coordinate sits that close to
# a 1/64-pixel rounding boundary lands one 1/32-pixel step aside.  Measured on this recording:
# 2148 of 9,381,888 pixels (2.3e-4), every one of them by exactly one step.  The gate below holds
# that level; the bit-exact gates are the kernel-level test above (reference crop cameras in,
# all 9.38 M indices equal) and the float64 crop-camera construction (1e-13 against numpy).
END_TO_END_INDEX_SLIP_FRACTION = 5e-4


@pytest.mark.parametrize("device_cameras", [True, False])
def test_end_to_end_crop_indices_with_our_own_crop_cameras(device_cameras):
    """The repo's own crop cameras (HandTracker.gen_crop_cameras: float64 device kernel, or the host
    numpy construction) on the repo's own float32 GPU forward kinematics, then the resample kernel:
    (ix, iy, fx, fy) of every pixel of every crop of the 256-frame recording against the indices the
    reference's own crop cameras give (lib/tracker/tracker.py:54-80,
    lib/tracker/perspective_crop.py:40-51,89-115)."""
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
    idx, _ = _resample_with_indices(torch.from_numpy(frames).to(dev), H, W4, W, offsets, params)
    crc_mismatch = [k for k in range(len(keys)) if _crc_of_indices(idx[k]) != int(g["map_crc"][k])]
    mism, worst = 0, 0
    for k in crc_mismatch:
        t, h, c, sf, sc, co, sxf, cf, ccn, cxf = crop_params_from_row(labels, rows[k])
        mx, my = O.warp_map(sf, sc, co, sxf, cf, ccn, cxf)
        ref = np.stack(O.remap_fixed_point_indices(mx, my), axis=-1)
        assert _crc_of_indices(ref) == int(g["map_crc"][k])                 # oracle == reference on this crop
        d = np.stack([(idx[k][..., 0] * 32 + idx[k][..., 2]) - (ref[..., 0] * 32 + ref[..., 2]),
                      (idx[k][..., 1] * 32 + idx[k][..., 3]) - (ref[..., 1] * 32 + ref[..., 3])], axis=-1)
        mism += int((d != 0).any(axis=-1).sum())
        worst = max(worst, int(np.abs(d).max()))
    total = len(keys) * 96 * 96
    print(f"end-to-end crop indices ({'device' if device_cameras else 'host'} crop cameras): {mism} of {total} pixels "
          f"({len(crc_mismatch)} of {len(keys)} crops) differ, largest step {worst}/32 px")
    assert worst <= 1 and mism <= END_TO_END_INDEX_SLIP_FRACTION * total, (mism, total, worst)


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
    # same end-to-end bound as above (float32 FK of the crop points is the only difference); the
    # sampled smooth crops then differ in at most a few pixels by one grey level step of the slip
    sx = idx[..., 0].astype(np.int64) * 32 + idx[..., 2]
    sy = idx[..., 1].astype(np.int64) * 32 + idx[..., 3]
    slipped = 0
    for k in range(len(keys)):
        if _crc_of_indices(idx[k]) != int(g["map_crc"][k]):
            t, h, c, sf, sc, co, sxf, cf, ccn, cxf = crop_params_from_row(labels, rows[k])
            ref = np.stack(O.remap_fixed_point_indices(*O.warp_map(sf, sc, co, sxf, cf, ccn, cxf)), axis=-1)
            dx = sx[k] - (ref[..., 0] * 32 + ref[..., 2])
            dy = sy[k] - (ref[..., 1] * 32 + ref[..., 3])
            assert max(np.abs(dx).max(), np.abs(dy).max()) <= 1
            slipped += int(((dx != 0) | (dy != 0)).sum())
    assert slipped <= END_TO_END_INDEX_SLIP_FRACTION * len(keys) * 9216
    diff = crops[g["crop_pick"]].astype(np.int32) - g["smooth_crops_u8"].astype(np.int32)
    assert (diff != 0).mean() <= 2 * END_TO_END_INDEX_SLIP_FRACTION and np.abs(diff).max() <= 8
