"""torch_data variant of stage (a): homography resampling against the reference's numpy resampler."""
import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import native, synthetic
from umetrack_b200.torch_data import perspective_crop_images, resample_images_device
from tests.helpers import golden

pytestmark = pytest.mark.gpu
CROP_TOL = 1e-3   # north_star: resampled crops within 1e-3 abs (on the [0,1] scale)


def _inputs():
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
    src4 = (g.random((4, hh, ww)) * 255).astype(np.float32)
    return orig, extr, intr, pts, src4


def test_perspective_crops_match_reference():
    t = golden("torchdata_golden.npz")
    orig, extr, intr, pts, _ = _inputs()
    for hand in (0, 1):
        crops, ex, kk = perspective_crop_images(orig, extr, intr, pts, hand, (96, 96))
        np.testing.assert_allclose(ex, t[f"extr_{hand}"], atol=1e-5)
        np.testing.assert_allclose(kk, t[f"intr_{hand}"], rtol=1e-6)
        assert np.abs(crops - t[f"crops_{hand}"]).max() <= CROP_TOL
        assert crops.max() <= 1.0 and crops.min() >= 0.0 and crops.max() > 0.5


def test_raw_kernel_matches_reference_and_oracle_including_outside_pixels():
    t = golden("torchdata_golden.npz")
    src4 = _inputs()[4]
    out = resample_images_device(torch.from_numpy(src4).cuda(), torch.from_numpy(t["raw_xfs"]).cuda()).cpu().numpy()
    ref = t["raw_out"]
    assert np.abs(out - ref).max() <= 255 * 1e-6          # 0..255 scale here; float32 rounding only
    assert (out[3] == 0).all() and (ref[3] == 0).all()    # everything outside -> zeros
    assert ((out == 0) == (ref == 0)).mean() > 0.9999     # same inside/outside decisions
    orc = O.resample_homography(src4, t["raw_xfs"])
    assert np.abs(out - orc).max() <= 255 * 1e-6


def test_homography_argument_validation():
    lib = native.lib()
    assert lib.umt_resample_homography(0, 8, 8, 0, 0, 96, 96, 0, 0, 0) == 0     # n = 0
    assert lib.umt_resample_homography(0, 8, 8, 0, 1, 97, 96, 0, 0, 0) != 0     # out_w % 4
    assert lib.umt_resample_homography(0, 1, 8, 0, 1, 96, 96, 0, 0, 0) != 0     # degenerate source


def test_sequence_folder_end_to_end_against_oracle():
    """idx/bin folder -> parse -> GPU crops -> model stepped over the sequence
    (run_inference_torch_data.py:39-133), against the oracle fed the same prepared inputs."""
    import os

    from tests.helpers import GOLDEN
    from umetrack_b200.model import UmeTrackModel
    from umetrack_b200.torch_data_reader import SequenceFolder, eval_sequences, parse_raw_buffers, prepare_model_inputs

    ds = SequenceFolder(os.path.join(GOLDEN, "torchdata_seq"))
    sd = synthetic.make_state_dict(0)
    model = UmeTrackModel(sd).to("cuda")
    for i in range(len(ds)):
        raw = ds[i]
        inp = prepare_model_inputs(parse_raw_buffers(**raw), (96, 96))
        seq = raw["mono"].shape[0]
        assert inp["left_images"].shape == (seq, 2, 96, 96) and float(inp["left_images"].max()) <= 1.0
        # crops: same as calling the crop function directly on the (scaled) sample
        s2 = parse_raw_buffers(**ds[i])
        s2.scaled(0.001)
        crops, ex, kk = perspective_crop_images(s2.images.astype(np.float32), s2.extrinsics, s2.intrinsics,
                                                s2.enclosing_points, int(s2.hand[0]), (96, 96))
        assert np.array_equal(crops, inp["left_images"].numpy()) and np.array_equal(ex, inp["extrinsics_xf"].numpy())
        err_mm, ja, wr = eval_sequences(model, [inp])
        # oracle: same stepping on the CPU
        orc = O.OracleModel(sd)
        hm = inp["left_hand_model"]
        kp_o, kp_g = [], []
        for t in range(seq):
            ref = orc.regress_pose_use_skeleton(
                inp["left_images"][t], inp["intrinsics"][t], inp["extrinsics_xf"][t], torch.tensor([[0, 2]]),
                torch.tensor([0]), torch.tensor([t > 0]), inp["hand_idx"][t:t + 1].long(),
                hm.joint_rotation_axes[t][None], hm.joint_rest_positions[t][None])
            un = [hm.joint_rotation_axes[t].clone(), hm.joint_rest_positions[t].clone(), hm.landmark_rest_positions[t].clone()]
            if int(inp["hand_idx"][t]) == 1:      # un-mirror (lib/common/hand.py:101-134)
                un[0][:, 1:] *= -1
                un[1][:, 0] *= -1
                un[2][:, 0] *= -1
            args = (un[0], un[1], un[2], hm.landmark_rest_bone_indices[t], hm.landmark_rest_bone_weights[t])
            kp_o.append(O.skin_landmarks(*args, ref["joint_angles"], ref["wrist_xfs"]))
            kp_g.append(O.skin_landmarks(*args, inp["joint_angles"][t:t + 1], inp["wrist_xfs"][t:t + 1]))
            d_ja = float((ja[0, t].cpu() - ref["joint_angles"][0]).abs().max())
            assert d_ja < 2e-4, (i, t, d_ja)
        kp_o, kp_g = torch.cat(kp_o), torch.cat(kp_g)
        want = float((kp_g - kp_o).norm(dim=-1).mean() * 1000)
        assert abs(float(err_mm[0]) - want) < 0.01, (float(err_mm[0]), want)
