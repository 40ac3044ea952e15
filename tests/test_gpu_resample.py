"""Stage (a) parity: crop indices bit-exact, uint8 crops identical (=> float crops within 1e-3)."""
import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import native, synthetic
from tests.helpers import crop_params_from_row, golden, hm_dict, pack_resample_params

pytestmark = pytest.mark.gpu


def run_resample(src_u8, offsets, pitch, w, h, params):
    dev = torch.device("cuda")
    n = len(params)
    src = torch.from_numpy(np.ascontiguousarray(src_u8)).to(dev)
    off = torch.tensor(offsets, dtype=torch.int64, device=dev)
    par = torch.from_numpy(np.asarray(params, dtype=np.float64).reshape(n, 48)).to(dev)
    ou = torch.full((n, 96, 96), 77, dtype=torch.uint8, device=dev)
    of = torch.empty((n, 96, 96), dtype=torch.float32, device=dev)
    oi = torch.empty((n, 96, 96, 4), dtype=torch.int32, device=dev)
    native.check(native.lib().umt_resample_crops(src.data_ptr(), off.data_ptr(), pitch, w, h, par.data_ptr(), n,
                                                 ou.data_ptr(), of.data_ptr(), oi.data_ptr(), 0), "resample")
    torch.cuda.synchronize()
    return ou.cpu().numpy(), of.cpu().numpy(), oi.cpu().numpy()


def test_crops_match_reference_golden():
    g = golden("tracker_golden.npz")
    labels, frames = synthetic.make_recording(hm_dict(), seed=3, n_frames=12)
    T, H, W4 = frames.shape
    W = W4 // 4
    rows = [r for r in g["crop_rows"] if int(r[0]) < 2]
    params, offsets, maps = [], [], []
    for r in rows:
        t, h, c, sf, sc, co, sxf, cf, cc, cxf = crop_params_from_row(labels, r)
        params.append(pack_resample_params(sf, sc, co, sxf, cf, cc, cxf))
        offsets.append(t * H * W4 + c * W)      # strided view frame[:, cam, :] of [H, 4, W]
        maps.append(O.warp_map(sf, sc, co, sxf, cf, cc, cxf))
    ou, of, oi = run_resample(frames, offsets, W4, W, H, params)
    assert np.array_equal(ou, g["crops_u8"]), "uint8 crops differ from the reference's cv2.remap output"
    assert np.array_equal(of, ou.astype(np.float32) / 255.0)
    assert np.abs(of - g["crops_u8"].astype(np.float32) / 255.0).max() <= 1e-3
    mism = 0
    for k, (mx, my) in enumerate(maps):
        ix, iy, fx, fy = O.remap_fixed_point_indices(mx, my)
        ref = np.stack([ix, iy, fx, fy], axis=-1)
        mism += int((ref != oi[k]).any(axis=-1).sum())
    assert mism == 0, f"{mism} crop pixel indices differ"


def test_edge_cases_behind_camera_borders_and_large_window():
    g = np.random.default_rng(7)
    H, W = 120, 200
    src = g.integers(0, 256, size=(2, H, W), dtype=np.uint8)
    coeffs = [1e-2, -3e-3, 5e-4, 0, 2e-4, -1e-4, 0, 0]
    eye = np.eye(4)
    cases = []
    # (crop focal, crop rotation about y, image index)
    for f, yaw, im in [(60.0, 0.0, 0), (25.0, 0.4, 1), (8.0, 0.0, 0), (40.0, 2.6, 1), (300.0, -0.2, 0), (12.0, 1.2, 1)]:
        xf = np.eye(4)
        xf[:3, :3] = synthetic._rot_y(yaw) @ synthetic._rot_x(0.1 * yaw)
        cases.append((f, xf, im))
    params, offsets, refs = [], [], []
    for f, xf, im in cases:
        sf, sc = (70.0, 71.0), (99.3, 60.2)
        params.append(pack_resample_params(sf, sc, coeffs, eye, (f, f), (47.5, 47.5), xf))
        offsets.append(im * H * W)
        mx, my = O.warp_map(sf, sc, coeffs, eye, (f, f), (47.5, 47.5), xf)
        refs.append((O.remap_u8_bilinear(src[im], mx, my), O.remap_fixed_point_indices(mx, my)))
    ou, of, oi = run_resample(src, offsets, W, W, H, params)
    for k, (crop, (ix, iy, fx, fy)) in enumerate(refs):
        assert np.array_equal(ou[k], crop), f"case {k}"
        assert np.array_equal(oi[k], np.stack([ix, iy, fx, fy], axis=-1)), f"case {k} indices"
    # case 3 looks backwards: every pixel is masked to -1 -> zero crop (tracker.py:71-73)
    assert ou[3].max() == 0


def test_empty_and_invalid_arguments():
    lib = native.lib()
    assert lib.umt_resample_crops(0, 0, 16, 16, 16, 0, 0, 0, 0, 0, 0) == 0          # n = 0 is a no-op
    assert lib.umt_resample_crops(0, 0, 8, 16, 16, 0, 1, 0, 0, 0, 0) != 0           # pitch < width
    assert b"geometry" in lib.umt_last_error()
