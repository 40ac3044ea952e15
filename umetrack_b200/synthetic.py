"""Seeded synthetic weights, camera rigs, recordings and model batches.

Neither `pretrained_weights.torch` nor the `UmeTrack_data` recordings ship with the reference
checkout, so parity and throughput are established on synthetic inputs in the reference's
own formats: a flat `state_dict` with the reference's 252 keys (lib/models/model_loader.py:84-87)
and recordings in the label schema of lib/tracker/video_pose_data.py:75-93.  Everything here
depends only on numpy's PCG64 stream, so the same seed gives the same bytes on any machine.
"""
from __future__ import annotations

import json
import math
import zlib
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import arch


def _rng_for(seed: int, key: str) -> np.random.Generator:
    return np.random.default_rng([seed, zlib.crc32(key.encode())])


def make_state_dict(seed: int = 0) -> Dict[str, torch.Tensor]:
    """Random weights with non-trivial BatchNorm statistics (so BN folding is exercised)."""
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, kind in arch.state_dict_spec():
        g = _rng_for(seed, key)
        if kind == "conv":
            fan = shape[0] * shape[2] * shape[3]
            v = g.normal(0.0, math.sqrt(2.0 / fan), size=shape)
        elif kind == "linear":
            v = g.normal(0.0, 1.0 / math.sqrt(shape[1]), size=shape)
        elif kind == "bias":
            v = g.normal(0.0, 0.05, size=shape)
        elif kind == "bn_gamma":
            v = g.uniform(0.6, 1.4, size=shape)
        elif kind == "bn_beta":
            v = g.normal(0.0, 0.1, size=shape)
        elif kind == "bn_mean":
            v = g.normal(0.0, 0.1, size=shape)
        elif kind == "bn_var":
            v = g.uniform(0.5, 1.5, size=shape)
        elif kind == "bn_count":
            sd[key] = torch.tensor(100, dtype=torch.int64)
            continue
        else:  # pragma: no cover
            raise AssertionError(kind)
        sd[key] = torch.from_numpy(np.asarray(v * _conditioning(key, kind), dtype=np.float32))
    _shape_outputs(sd, seed)
    return sd


def _conditioning(key: str, kind: str) -> float:
    """Fixed gains that keep activations O(1) through the 30-conv stack with random weights, so
    that absolute-tolerance parity gates (mm, radians) are meaningful on synthetic weights."""
    if kind == "bn_gamma" and key.endswith("bn2.weight"):
        return 0.45          # damp the residual branch
    if kind == "conv" and key == arch.BACKBONE + "1.weight":
        return 0.35          # 256 -> 72 projection
    if kind == "conv" and key.startswith(arch.FUSION):
        return 0.7
    if kind == "conv" and key.startswith(arch.TEMPORAL):
        return 0.7
    return 1.0


def _shape_outputs(sd: Dict[str, torch.Tensor], seed: int) -> None:
    """Give the two regression heads plausible output statistics: finger angles of a few tenths
    of a radian, seven wrist points near the rigid template 0.35 m in front of camera 0 with
    feature-driven variation of a few centimetres, log-scale near 0."""
    tmpl = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, -1, 0], [-1, 0, -1],
                     [0, -1, -1]], dtype=np.float64)
    nrm = np.linalg.norm(tmpl, axis=1, keepdims=True)
    tmpl = np.where(nrm > 0, tmpl / np.maximum(nrm, 1e-12) * 0.1, tmpl)
    for name, nout in (("_regressor_k", arch.REG_K_OUT), ("_regressor_u", arch.REG_U_OUT)):
        g = _rng_for(seed, name + ".shape")
        w = sd[name + "._pose_regression_layers.2.weight"].numpy().astype(np.float64)
        b = np.zeros(nout)
        gain = np.full(nout, 0.05)
        gain[0:20] = 0.25
        b[0:20] = g.uniform(-0.2, 0.6, size=20)
        gain[20:41] = 0.03
        rot = _random_rotation(g, 0.8)
        b[20:41] = (tmpl @ rot.T + np.array([0.02, 0.03, 0.35])).reshape(-1)
        if nout == arch.REG_U_OUT:
            gain[41] = 0.05
            b[41] = 0.02
        w = w / (np.abs(w).mean() * math.sqrt(w.shape[1]) + 1e-12) * gain[:, None, None, None]
        sd[name + "._pose_regression_layers.2.weight"] = torch.from_numpy(w.astype(np.float32))
        sd[name + "._pose_regression_layers.2.bias"] = torch.from_numpy(b.astype(np.float32))


def _rot_x(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)


def _rot_y(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def _rot_z(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def make_rig(seed: int = 0, width: int = 640, height: int = 480):
    """Four FishEye62 cameras on a headset: returns (camera_json_list, camera_angles,
    camera_to_world [4,4,4] in mm).  World frame: x right, y down, z forward."""
    g = np.random.default_rng([seed, 7])
    cams, angles, xfs = [], [], []
    yaws = np.deg2rad([-38.0, -12.0, 12.0, 38.0])
    pitch = np.deg2rad([12.0, 4.0, 4.0, 12.0])
    xs = [-60.0, -22.0, 22.0, 60.0]
    for i in range(4):
        cams.append({
            "ImageSizeX": width, "ImageSizeY": height, "DistortionModel": "FishEye62",
            "fx": float(240.0 + g.uniform(-4, 4)), "fy": float(240.0 + g.uniform(-4, 4)),
            "cx": float((width - 1) / 2 + g.uniform(-3, 3)),
            "cy": float((height - 1) / 2 + g.uniform(-3, 3)),
            "k1": float(g.normal(0, 2e-2)), "k2": float(g.normal(0, 5e-3)),
            "k3": float(g.normal(0, 1e-3)), "k4": float(g.normal(0, 2e-4)),
            "p1": float(g.normal(0, 3e-4)), "p2": float(g.normal(0, 3e-4)),
            "k5": float(g.normal(0, 5e-5)), "k6": float(g.normal(0, 1e-5)),
        })
        roll = float(g.uniform(-8, 8)) + (15.0 if i in (0, 3) else 0.0) * (1 if i == 3 else -1)
        angles.append(roll)
        m = np.eye(4)
        m[:3, :3] = _rot_y(yaws[i]) @ _rot_x(pitch[i]) @ _rot_z(np.deg2rad(roll))
        m[:3, 3] = [xs[i], float(g.uniform(-5, 5)), float(g.uniform(-5, 5))]
        xfs.append(m)
    return cams, angles, np.stack(xfs)


def _random_rotation(g: np.random.Generator, max_angle: float) -> np.ndarray:
    axis = g.normal(size=3)
    axis /= np.linalg.norm(axis)
    ang = g.uniform(-max_angle, max_angle)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * K @ K


def load_hand_model_json(path: str) -> dict:
    with open(path, "rb") as f:
        return json.load(f)


def make_recording(hand_model_dict: dict, seed: int = 0, n_frames: int = 32,
                   width: int = 640, height: int = 480, smooth_images: bool = False,
                   dropouts: bool = True, with_frames: bool = True, presence=None):
    """Synthetic recording: (labels_dict in the reference JSON schema, frames uint8
    [T, H, 4*W]).  Both hands move on smooth trajectories 250-450 mm in front of the rig;
    joint angles stay inside `joint_limits`; a few confidence dropouts exercise the tracker's
    history reset; frames are seeded noise (optionally box-blurred).  `presence` ([T,2] of 0/1),
    if given, replaces the hand confidences (e.g. one hand only for a stretch of frames); it does
    not change any other generated value."""
    g = np.random.default_rng([seed, 11])
    cams, angles, cam_xf = make_rig(seed, width, height)
    limits = np.asarray(hand_model_dict["joint_limits"], dtype=np.float64)
    T = n_frames
    tt = np.arange(T)[:, None] / max(T - 1, 1)
    joint_angles = np.zeros((T, 2, 22))
    wrist = np.zeros((T, 2, 4, 4))
    conf = np.ones((T, 2))
    for h in range(2):
        ph = g.uniform(0, 2 * np.pi, size=(1, 22))
        fr = g.uniform(0.5, 2.0, size=(1, 22))
        u = 0.5 + 0.45 * np.sin(2 * np.pi * fr * tt + ph)
        ja = limits[None, :, 0] + u * (limits[None, :, 1] - limits[None, :, 0])
        ja[:, 20:] = 0.0
        joint_angles[:, h] = ja
        base = np.array([-110.0 if h == 0 else 110.0, 60.0, 340.0])
        amp = g.uniform(15, 45, size=3)
        php = g.uniform(0, 2 * np.pi, size=3)
        r0 = _random_rotation(g, 0.6)
        for t in range(T):
            pos = base + amp * np.sin(2 * np.pi * t / max(T, 2) + php)
            rot = r0 @ _rot_y(0.3 * math.sin(2 * np.pi * t / max(T, 2) + h)) @ _rot_x(
                0.25 * math.cos(2 * np.pi * t / max(T, 2)))
            m = np.eye(4)
            m[:3, :3] = rot
            m[:3, 3] = pos
            if h == 1:
                m[:, 0] *= -1  # right-hand labels carry the mirrored column (perspective_crop.py:48-49)
            wrist[t, h] = m
    if dropouts and T >= 12:
        conf[T // 3, 0] = 0.0
        conf[T // 2: T // 2 + 2, 1] = 0.0
    if presence is not None:
        conf = np.asarray(presence, dtype=np.float64).reshape(T, 2).copy()
    frames = g.integers(0, 256, size=(T, height, 4 * width), dtype=np.uint8) if with_frames else None
    if smooth_images and frames is not None:
        f = frames.astype(np.float32)
        for _ in range(2):
            f = (f + np.roll(f, 1, 1) + np.roll(f, -1, 1) + np.roll(f, 1, 2) + np.roll(f, -1, 2)) / 5
        frames = np.clip(np.rint(f * 2 - 127), 0, 255).astype(np.uint8)
    labels = {
        "cameras": cams,
        "camera_angles": angles,
        "hand_model": hand_model_dict,
        "joint_angles": joint_angles.tolist(),
        "wrist_transforms": wrist.tolist(),
        "hand_confidences": conf.tolist(),
        "camera_to_world_transforms": np.broadcast_to(cam_xf, (T, 4, 4, 4)).tolist(),
    }
    return labels, frames


def make_model_batch(n_samples: int, seed: int = 1234, single_view_every: int = 0,
                     device: str = "cpu"):
    """BASELINE config 4 batch: `n_samples` hand-frames, two 96x96 crops each (optionally every
    k-th sample single-view).  Returns a dict of tensors matching InputFrameData /
    InputFrameDesc (lib/models/umetrack_model.py:21-74)."""
    g = np.random.default_rng([seed, 13])
    views = [1 if (single_view_every and (i % single_view_every == single_view_every - 1)) else 2
             for i in range(n_samples)]
    n_img = int(sum(views))
    images = g.random((n_img, 96, 96), dtype=np.float32)
    intr = np.zeros((n_img, 3, 3), dtype=np.float32)
    extr = np.zeros((n_img, 4, 4), dtype=np.float32)
    rng_rows = []
    k = 0
    for i, nv in enumerate(views):
        start = k
        centre = np.array([g.uniform(-0.1, 0.1), g.uniform(-0.05, 0.1), g.uniform(0.28, 0.42)])
        for v in range(nv):
            f = g.uniform(120, 220)
            intr[k] = [[f, 0, 47.5], [0, f, 47.5], [0, 0, 1]]
            cam_pos = np.array([(-0.04 if v == 0 else 0.04) + g.uniform(-0.01, 0.01),
                                g.uniform(-0.01, 0.01), g.uniform(-0.01, 0.01)])
            z = centre - cam_pos
            z /= np.linalg.norm(z)
            x = np.cross(np.array([0.0, 1.0, 0.0]), z)
            x /= np.linalg.norm(x)
            y = np.cross(z, x)
            c2w = np.eye(4)
            c2w[:3, :3] = np.stack([x, y, z], axis=1) @ _rot_z(g.uniform(-0.3, 0.3))
            c2w[:3, 3] = cam_pos
            extr[k] = np.linalg.inv(c2w).astype(np.float32)
            k += 1
        rng_rows.append([start, k])
    batch = {
        "left_images": torch.from_numpy(images),
        "intrinsics": torch.from_numpy(intr),
        "extrinsics_xf": torch.from_numpy(extr),
        "sample_range": torch.tensor(rng_rows, dtype=torch.int64),
        "memory_idx": torch.arange(n_samples, dtype=torch.int64),
        "use_memory": torch.ones(n_samples, dtype=torch.bool),
        "hand_idx": torch.from_numpy((np.arange(n_samples) % 2).astype(np.int64)),
    }
    return {k_: v.to(device) for k_, v in batch.items()}
