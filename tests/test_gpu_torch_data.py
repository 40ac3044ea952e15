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
