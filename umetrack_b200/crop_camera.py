"""Crop-camera construction (host, float64): look-at rotation towards the hand, optional
x-mirroring for right hands, focal length from the bounding landmarks.

Equivalent of the reference's lib/common/crop.py:15-82 and lib/common/affine.py:34-66; raises the
same `ValueError("Unable to create crop camera", fx_fy)` (crop.py:25-26).
"""
from __future__ import annotations

from typing import Tuple

import numpy as np
from scipy.spatial.transform import Rotation

from .cameras import CameraModel, PinholePlaneCameraModel, _unit


def _apply(m: np.ndarray, pts: np.ndarray) -> np.ndarray:
    return (pts.reshape(-1, 3) @ m[:3, :3].T).reshape(pts.shape) + m[:3, 3]


def rotation_between(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Rotation taking direction a to direction b (Rodrigues form of affine.py:34-44)."""
    a, b = _unit(a), _unit(b)
    v = np.cross(a, b)
    s = np.linalg.norm(v)
    c = np.dot(a, b)
    k = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=v.dtype)
    return np.eye(3, 3) + k + np.matmul(k, k) * (1 - c) / (max(s * s, 1e-15))


def look_at(world_to_eye: np.ndarray, target: np.ndarray, camera_angle: float = 0) -> np.ndarray:
    """World->eye transform of a camera at the same position whose z axis passes through
    `target`, rolled by `camera_angle` degrees about z (affine.py:47-66)."""
    local = _apply(world_to_eye, target)
    zdir = local / np.linalg.norm(local)
    delta = rotation_between(np.array([0, 0, 1], dtype=target.dtype), zdir)
    eye_to_world = np.linalg.inv(world_to_eye)
    out = eye_to_world.copy()
    out[0:3, 0:3] = eye_to_world[0:3, 0:3] @ delta
    roll = Rotation.from_euler("z", camera_angle, degrees=True).as_matrix()
    out[0:3, 0:3] = out[0:3, 0:3] @ roll
    return np.linalg.inv(out)


def intrinsics_from_points(pts_eye: np.ndarray, image_w: int, image_h: int,
                           min_focal: float = 5) -> Tuple[np.ndarray, np.ndarray]:
    ndc = pts_eye[..., 0:2] / pts_eye[..., 2:]
    size = np.array([image_w, image_h], dtype=pts_eye.dtype)
    cx_cy = (size - 1) / 2
    fx_fy = cx_cy / np.absolute(ndc).max()
    if np.any(pts_eye[..., 2:] < 0.0001) or np.any(fx_fy < min_focal):
        raise ValueError("Unable to create crop camera", fx_fy)
    return fx_fy, cx_cy


def gen_crop_parameters_from_points(camera_orig: CameraModel, pts_world, new_image_size,
                                    mirror_img_x: bool, camera_angle: float = 0,
                                    focal_multiplier: float = 0.95) -> PinholePlaneCameraModel:
    """Pinhole camera at `camera_orig`'s position that keeps every point of `pts_world` inside a
    `new_image_size` image (crop.py:31-82)."""
    w2e = np.linalg.inv(camera_orig.camera_to_world_xf)
    centre = (pts_world.min(axis=0) + pts_world.max(axis=0)) / 2.0
    new_w2e = look_at(w2e, centre, camera_angle)
    if mirror_img_x:
        flip = np.eye(4, dtype=np.float32)
        flip[0, 0] = -1
        new_w2e = flip @ new_w2e
    fx_fy, cx_cy = intrinsics_from_points(_apply(new_w2e, pts_world), new_image_size[0], new_image_size[1])
    return PinholePlaneCameraModel(width=new_image_size[0], height=new_image_size[1],
                                   f=focal_multiplier * fx_fy, c=cx_cy, distort_coeffs=(),
                                   camera_to_world_xf=np.linalg.inv(new_w2e))
