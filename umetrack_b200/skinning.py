"""Forward kinematics + skinning of the 21 hand landmarks on the GPU.

`skin_landmarks` keeps the signature of the reference's lib/common/hand_skinning.py:189-209 and
runs `umt_fk_landmarks` (csrc/fk_metrics.cu).  The clamped-angle exponential map of
pytorch3d.so3_exp_map (hand_skinning.py:46) is built into the kernel.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import native
from .hand_model import HandModel, NUM_JOINTS_PER_HAND, NUM_LANDMARKS_PER_HAND

_model_cache: Dict[tuple, tuple] = {}


def _device_model(hand_model: HandModel, device: torch.device):
    """float32 device copies of the five arrays the kernel needs, cached per model object."""
    key = (id(hand_model.joint_rotation_axes), id(hand_model.joint_rest_positions),
           id(hand_model.landmark_rest_positions), str(device))
    hit = _model_cache.get(key)
    if hit is not None and hit[0] is hand_model.joint_rotation_axes:
        return hit[1]
    arrs = tuple(
        t.detach().to(device=device, dtype=torch.float32).contiguous()
        for t in (hand_model.joint_rotation_axes, hand_model.joint_rest_positions,
                  hand_model.landmark_rest_positions, hand_model.landmark_rest_bone_indices,
                  hand_model.landmark_rest_bone_weights))
    if len(_model_cache) > 64:
        _model_cache.clear()
    _model_cache[key] = (hand_model.joint_rotation_axes, arrs)
    return arrs


def fk_landmarks_device(hand_model: HandModel, joint_angles: torch.Tensor, wrist: torch.Tensor,
                        mirror: Optional[torch.Tensor] = None) -> torch.Tensor:
    """joint_angles [n,22], wrist [n,4,4] float32 CUDA tensors -> landmarks [n,21,3] (CUDA).
    `mirror` [n] uint8/bool flips column 0 of the wrist transform first (right hands)."""
    if not joint_angles.is_cuda:
        raise native.NativeError("fk_landmarks_device needs CUDA tensors (no CPU fallback)")
    n = joint_angles.shape[0]
    axes, rest, lm_rest, lm_idx, lm_w = _device_model(hand_model, joint_angles.device)
    if axes.dim() != 2:
        raise ValueError("fk_landmarks_device takes one (unbatched) hand model")
    ja = joint_angles.to(torch.float32).contiguous()
    wr = wrist.to(torch.float32).contiguous()
    mi = None if mirror is None else mirror.to(torch.uint8).contiguous()
    out = torch.empty((n, NUM_LANDMARKS_PER_HAND, 3), dtype=torch.float32, device=ja.device)
    native.check(native.lib().umt_fk_landmarks(
        native.ptr(axes), native.ptr(rest), native.ptr(lm_rest), native.ptr(lm_idx), native.ptr(lm_w),
        native.ptr(ja), native.ptr(wr), native.ptr(mi), n, native.ptr(out),
        native.current_stream_ptr()), "umt_fk_landmarks")
    return out


def skin_landmarks(hand_model: HandModel, joint_angles: torch.Tensor,
                   wrist_transforms: torch.Tensor) -> torch.Tensor:
    """[..., 22] joint angles and [..., 4, 4] wrist transforms -> [..., 21, 3] landmarks.

    Like the reference, the hand model's leading dims must equal those of `joint_angles`
    (or the model is unbatched and shared by every pose).  Inputs may live on the host; the
    work is always done on the current CUDA device and the result returned on the input's device.
    """
    lead = tuple(joint_angles.shape[:-1])
    src_device = joint_angles.device
    dev = src_device if joint_angles.is_cuda else torch.device("cuda", torch.cuda.current_device())
    ja = joint_angles.reshape(-1, NUM_JOINTS_PER_HAND).to(dev, torch.float32)
    wr = wrist_transforms.reshape(-1, 4, 4).to(dev, torch.float32)
    model_lead = tuple(hand_model.joint_rest_positions.shape[:-2])
    if model_lead == ():
        out = fk_landmarks_device(hand_model, ja, wr)
    else:
        assert model_lead == lead, f"Leading dimensions do not match, got {lead} and {model_lead}"
        n = ja.shape[0]
        flat = [t.reshape(n, *t.shape[len(lead):]) for t in (
            hand_model.joint_rotation_axes, hand_model.joint_rest_positions,
            hand_model.landmark_rest_positions, hand_model.landmark_rest_bone_indices,
            hand_model.landmark_rest_bone_weights)]
        out = torch.empty((n, NUM_LANDMARKS_PER_HAND, 3), dtype=torch.float32, device=dev)
        for i in range(n):  # one launch per distinct model (rare path)
            one = hand_model._replace(
                joint_rotation_axes=flat[0][i], joint_rest_positions=flat[1][i],
                landmark_rest_positions=flat[2][i], landmark_rest_bone_indices=flat[3][i],
                landmark_rest_bone_weights=flat[4][i])
            out[i:i + 1] = fk_landmarks_device(one, ja[i:i + 1], wr[i:i + 1])
    return out.reshape(*lead, NUM_LANDMARKS_PER_HAND, 3).to(src_device)
