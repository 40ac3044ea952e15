"""torch_data preprocessing on the GPU: perspective hand crops from pinhole views.

Mirrors the crop part of the reference's lib/batched_dataset/data_transform.py:
`_compute_resample_matrix` (:57-75), `_gen_crop_matrices` (:147-213) and
`_perspective_crop_images` (:216-281).  The crop cameras are built on the host in float64 (a few
4x4 products per view); the per-pixel resampling (`_resample_images_batched`, :78-144) runs in
`umt_resample_homography`.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
import torch

from . import native
from .cameras import PinholePlaneCameraModel
from .crop_camera import gen_crop_parameters_from_points


def compute_resample_matrix(camera_orig: PinholePlaneCameraModel, camera_new: PinholePlaneCameraModel) -> np.ndarray:
    """float32 4x4 taking a pixel (u, v, 1) of `camera_new` to a pixel of `camera_orig`."""
    k_inv_new = np.eye(4, 4)
    k_inv_new[0:3, 0:3] = np.linalg.inv(camera_new.uv_to_window_matrix())
    k_orig = np.eye(4, 4)
    k_orig[0:3, 0:3] = camera_orig.uv_to_window_matrix()
    m = k_orig @ np.linalg.inv(camera_orig.camera_to_world_xf) @ camera_new.camera_to_world_xf @ k_inv_new
    return m.astype(np.float32)


def gen_crop_matrices(orig_extrinsics: np.ndarray, orig_intrinsics: np.ndarray, crop_points: np.ndarray,
                      mirror_image: bool, crop_size: Tuple[int, int], focal_multiplier: float = 0.95):
    """Per view: (new world->eye [n,4,4], new intrinsics [n,3,3], resample matrices [n,4,4]), float32."""
    n_views = orig_extrinsics.shape[0]
    extr = np.empty([n_views, 4, 4], dtype=np.float32)
    intr = np.empty([n_views, 3, 3], dtype=np.float32)
    res = np.empty([n_views, 4, 4], dtype=np.float32)
    for v in range(n_views):
        k = orig_intrinsics[v]
        cam_orig = PinholePlaneCameraModel(width=crop_size[0], height=crop_size[1], f=(k[0, 0], k[1, 1]),
                                           c=(k[0, 2], k[1, 2]), distort_coeffs=(),
                                           camera_to_world_xf=np.linalg.inv(orig_extrinsics[v]))
        cam_new = gen_crop_parameters_from_points(cam_orig, crop_points, crop_size, mirror_image,
                                                  camera_angle=0, focal_multiplier=focal_multiplier)
        extr[v] = np.linalg.inv(cam_new.camera_to_world_xf)
        intr[v] = cam_new.uv_to_window_matrix()
        res[v] = compute_resample_matrix(cam_orig, cam_new)
    return extr, intr, res


def resample_images_device(images: torch.Tensor, resample_xfs: torch.Tensor, crop_size: Tuple[int, int] = (96, 96),
                           divide_by_255: bool = False) -> torch.Tensor:
    """images [n,h,w] float32 CUDA, resample_xfs [n,4,4] float32 CUDA -> [n,crop_h,crop_w] (CUDA)."""
    if not images.is_cuda:
        raise native.NativeError("resample_images_device needs CUDA tensors (no CPU fallback)")
    n, h, w = images.shape
    src = images.contiguous().float()
    xfs = resample_xfs.to(images.device, torch.float32).contiguous()
    out = torch.empty((n, crop_size[1], crop_size[0]), dtype=torch.float32, device=images.device)
    native.check(native.lib().umt_resample_homography(native.ptr(src), w, h, native.ptr(xfs), n, crop_size[0],
                                                      crop_size[1], 1 if divide_by_255 else 0, native.ptr(out),
                                                      native.current_stream_ptr()), "umt_resample_homography")
    return out


def perspective_crop_images(orig_images: np.ndarray, orig_extrinsics: np.ndarray, orig_intrinsics: np.ndarray,
                            crop_points: np.ndarray, hand_idx: int, crop_size: Tuple[int, int]) -> List[np.ndarray]:
    """[frames, views, H, W] images -> [crops in [0,1], new extrinsics, new intrinsics] as the reference
    returns them (host arrays); the resampling itself runs on the current CUDA device."""
    n_frames, n_views = orig_images.shape[:2]
    extr = np.empty([n_frames, n_views, 4, 4], dtype=np.float32)
    intr = np.empty([n_frames, n_views, 3, 3], dtype=np.float32)
    res = np.empty([n_frames, n_views, 4, 4], dtype=np.float32)
    for f in range(n_frames):
        extr[f], intr[f], res[f] = gen_crop_matrices(orig_extrinsics[f], orig_intrinsics[f], crop_points[f],
                                                     hand_idx == 1, crop_size)
    dev = torch.device("cuda", torch.cuda.current_device())
    imgs = torch.from_numpy(np.ascontiguousarray(orig_images, dtype=np.float32)).to(dev)
    out = resample_images_device(imgs.reshape(-1, *imgs.shape[2:]), torch.from_numpy(res.reshape(-1, 4, 4)).to(dev),
                                 crop_size, divide_by_255=True)
    return [out.reshape(n_frames, n_views, crop_size[1], crop_size[0]).cpu().numpy(), extr, intr]
