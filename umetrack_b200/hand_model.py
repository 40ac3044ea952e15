"""Hand model container and its two transforms.

Mirrors the public names of the reference's lib/common/hand.py:12-134 (HandModel fields,
scaled_hand_model, mirrored_hand_model, the index constants) and
`load_hand_model_from_dict` of lib/tracker/video_pose_data.py:63-72, so the same hand-model
JSON (e.g. dataset/generic_hand_model.json) loads unchanged.
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

NUM_HANDS = 2
NUM_LANDMARKS_PER_HAND = 21
NUM_FINGERTIPS_PER_HAND = 5
NUM_JOINTS_PER_HAND = 22
LEFT_HAND_INDEX = 0
RIGHT_HAND_INDEX = 1
NUM_DIGITS = 5
NUM_JOINT_FRAMES = 17  # root, wrist, three frames per finger
DOF_PER_FINGER = 4


class HandModel(NamedTuple):
    joint_rotation_axes: torch.Tensor
    joint_rest_positions: torch.Tensor
    joint_frame_index: torch.Tensor
    joint_parent: torch.Tensor
    joint_first_child: torch.Tensor
    joint_next_sibling: torch.Tensor
    landmark_rest_positions: torch.Tensor
    landmark_rest_bone_weights: torch.Tensor
    landmark_rest_bone_indices: torch.Tensor
    hand_scale: Optional[torch.Tensor]
    mesh_vertices: Optional[torch.Tensor] = None
    mesh_triangles: Optional[torch.Tensor] = None
    dense_bone_weights: Optional[torch.Tensor] = None
    joint_limits: Optional[torch.Tensor] = None


def load_hand_model_from_dict(d: dict) -> HandModel:
    """Lists become float32 tensors (torch.Tensor(list)), everything else is kept as is."""
    return HandModel(**{k: (torch.Tensor(v) if isinstance(v, list) else v) for k, v in d.items()})


def scaled_hand_model(hand: HandModel, multiplier) -> HandModel:
    """Scale rest positions, landmarks and mesh; axes, weights and indices are unchanged
    (lib/common/hand.py:65-98).  `multiplier` may be a float or a tensor of the leading dims."""
    lead = hand.joint_rest_positions.shape[:-2]
    m = torch.ones(lead, dtype=hand.joint_rest_positions.dtype,
                   device=hand.joint_rest_positions.device) * multiplier
    m = m[..., None, None]
    return hand._replace(
        joint_rest_positions=hand.joint_rest_positions * m,
        landmark_rest_positions=hand.landmark_rest_positions * m,
        mesh_vertices=None if hand.mesh_vertices is None else hand.mesh_vertices * m,
    )


def mirrored_hand_model(hand: HandModel, to_mirror: torch.Tensor) -> HandModel:
    """Left<->right mirroring of the entries selected by the boolean mask `to_mirror`
    (lib/common/hand.py:101-134): rotation axes flip y,z; rest/landmark positions flip x.
    The mesh is returned untouched, as in the reference (its `*= 1` at :116 is a no-op)."""
    mask = to_mirror.reshape(-1)
    nlead = len(to_mirror.shape)

    def flipped(t: torch.Tensor, sl) -> torch.Tensor:
        out = t.clone()
        flat = out.reshape(-1, *t.shape[nlead:]) if nlead else out.reshape(1, *t.shape)
        sel = flat[mask]
        sel[..., sl] *= -1
        flat[mask] = sel
        return flat.reshape(t.shape)

    return hand._replace(
        joint_rotation_axes=flipped(hand.joint_rotation_axes, slice(1, None)),
        joint_rest_positions=flipped(hand.joint_rest_positions, slice(0, 1)),
        landmark_rest_positions=flipped(hand.landmark_rest_positions, slice(0, 1)),
        mesh_vertices=None if hand.mesh_vertices is None else hand.mesh_vertices.clone(),
    )
