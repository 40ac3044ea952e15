"""Camera models used on the host side of the path (numpy float64).

Same public surface as the reference's lib/common/camera.py:164-355 for what the tracker
touches: `PinholePlaneCameraModel`, `Fisheye62CameraModel`, `read_camera_from_json`, and the
methods world_to_eye / eye_to_world / eye_to_window / window_to_eye / crop / copy /
uv_to_window_matrix.  The per-pixel use of these maps (`_warp_image`) runs on the GPU
(csrc/resample.cu); here they are only evaluated on a handful of landmarks per frame to choose
and build crop cameras, which must stay float64 to keep crop pixel indices bit-exact.
"""
from __future__ import annotations

import json
from typing import Sequence, Tuple

import numpy as np

FISHEYE62_FIELDS = ("k1", "k2", "k3", "k4", "p1", "p2", "k5", "k6")


def _unit(v: np.ndarray, eps: float = 5.43e-20) -> np.ndarray:
    return v / np.maximum(eps, (v * v).sum(axis=-1, keepdims=True) ** 0.5)


class CameraModel:
    """width/height in pixels, focal f=(fx,fy), centre c=(cx,cy), camera_to_world_xf 4x4 (mm)."""

    model_name = "base"

    def __init__(self, width, height, f, c, distort_coeffs=(), camera_to_world_xf=None):
        self.width = width
        self.height = height
        self.f = tuple(np.broadcast_to(f, 2))
        self.c = tuple(c)
        self.distort_coeffs = tuple(float(x) for x in distort_coeffs)
        self.camera_to_world_xf = np.eye(4) if camera_to_world_xf is None else camera_to_world_xf

    def __repr__(self):
        return f"{type(self).__name__}({self.width}x{self.height}, f={self.f} c={self.c})"

    # -- rigid part -------------------------------------------------------------------------
    def world_to_eye(self, v):
        m = self.camera_to_world_xf
        d = np.asarray(v) - m[:3, 3]
        return (d.reshape(-1, 3) @ m[:3, :3]).reshape(d.shape)

    def eye_to_world(self, v):
        m = self.camera_to_world_xf
        v = np.asarray(v)
        return (v.reshape(-1, 3) @ m[:3, :3].T).reshape(v.shape) + m[:3, 3]

    # -- projection part (overridden) ---------------------------------------------------------
    def project(self, v):
        raise NotImplementedError

    def unproject(self, p):
        raise NotImplementedError

    def distort(self, p):
        return p

    def eye_to_window(self, v):
        return self.distort(self.project(np.asarray(v))) * self.f + self.c

    def window_to_eye(self, w):
        assert not self.distort_coeffs, "Only unprojection for NoDistortion camera is supported"
        return self.unproject((np.asarray(w) - self.c) / self.f)

    # -- derived cameras ----------------------------------------------------------------------
    def crop(self, src_x, src_y, target_width, target_height, scale=1, camera_to_world_xf=None):
        return type(self)(
            target_width, target_height, np.asarray(self.f) * scale,
            (np.array(self.c) - (src_x, src_y) + 0.5) * scale - 0.5, self.distort_coeffs,
            self.camera_to_world_xf if camera_to_world_xf is None else camera_to_world_xf)

    def copy(self, camera_to_world_xf=None):
        return self.crop(0, 0, self.width, self.height, camera_to_world_xf=camera_to_world_xf)


class PinholePlaneCameraModel(CameraModel):
    model_name = "PinholePlane"

    def project(self, v):
        return v[..., :2] / v[..., 2, None]

    def unproject(self, p):
        p = np.asarray(p)
        return _unit(np.concatenate([p, np.ones(p.shape[:-1] + (1,), dtype=p.dtype)], axis=-1))

    def uv_to_window_matrix(self):
        return np.array([[self.f[0], 0, self.c[0]], [0, self.f[1], self.c[1]], [0, 0, 1]])


class Fisheye62CameraModel(CameraModel):
    """Equidistant (arctan) projection with 6 radial + 2 tangential distortion coefficients,
    ordered k1 k2 k3 k4 p1 p2 k5 k6 (lib/common/camera.py:79-85,105-141)."""

    model_name = "FishEye62"

    def project(self, v, eps: float = 2.0**-128):
        x, y, z = v[..., 0], v[..., 1], v[..., 2]
        r = np.sqrt(x * x + y * y)
        s = np.arctan2(r, z) / np.maximum(r, eps)
        return np.stack((x * s, y * s), axis=-1)

    def distort(self, p):
        k1, k2, k3, k4, p1, p2, k5, k6 = self.distort_coeffs
        r2 = np.clip((p * p).sum(axis=-1, keepdims=True), -np.pi**2, np.pi**2)
        r4 = r2 * r2
        r6 = r2 * r4
        radial = 1 + k1 * r2 + k2 * r4 + k3 * r6 + k4 * (r4 * r4) + k5 * (r4 * r6) + k6 * (r6 * r6)
        uv = p * radial
        x, y = uv[..., 0], uv[..., 1]
        xy, x2, y2 = x * y, x * x, y * y
        rr = x2 + y2
        return np.stack((x + (2 * p2 * xy + p1 * (rr + 2 * x2)),
                         y + (2 * p1 * xy + p2 * (rr + 2 * y2))), axis=-1)


def read_camera_from_json(js):
    """Same schema as lib/common/camera.py:334-355 (ImageSizeX/Y, DistortionModel, fx fy cx cy,
    and k1..k6,p1,p2 for FishEye62)."""
    if isinstance(js, str):
        js = json.loads(js)
    js = js.get("Camera", js)
    kind = js["DistortionModel"]
    if kind == "PinholePlane":
        cls, coeffs = PinholePlaneCameraModel, ()
    elif kind == "FishEye62":
        cls, coeffs = Fisheye62CameraModel, [js[n] for n in FISHEYE62_FIELDS]
    else:
        raise ValueError(f"unsupported DistortionModel {kind!r}")
    return cls(js["ImageSizeX"], js["ImageSizeY"], (js["fx"], js["fy"]), (js["cx"], js["cy"]), coeffs)


def resample_params(src: Fisheye62CameraModel, crop: PinholePlaneCameraModel) -> np.ndarray:
    """The 48 float64 values `umt_resample_crops` takes for one (source view, crop camera) pair."""
    if not isinstance(src, Fisheye62CameraModel):
        raise TypeError("the GPU resampler implements the FishEye62 source model only")
    return np.concatenate([
        np.asarray(src.f, dtype=np.float64), np.asarray(src.c, dtype=np.float64),
        np.asarray(src.distort_coeffs, dtype=np.float64),
        np.asarray(src.camera_to_world_xf, dtype=np.float64).reshape(-1),
        np.asarray(crop.f, dtype=np.float64), np.asarray(crop.c, dtype=np.float64),
        np.asarray(crop.camera_to_world_xf, dtype=np.float64).reshape(-1)])
