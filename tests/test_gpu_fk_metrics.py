"""Stage (c) parity: FK/skinning vs the reference's golden landmarks, metric sums vs numpy."""
import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import hand_model as hmod
from umetrack_b200.metrics import MetricSums
from umetrack_b200.skinning import fk_landmarks_device, skin_landmarks
from tests.helpers import golden, hm_dict, hm_tensors

pytestmark = pytest.mark.gpu
KEYPOINT_TOL_MM = 0.05  # north_star: keypoints within 0.05 mm


def test_fk_matches_reference_golden():
    g = golden("fk_golden.npz")
    model = hmod.load_hand_model_from_dict(hm_dict())
    out = skin_landmarks(model, torch.from_numpy(g["joint_angles"]), torch.from_numpy(g["wrist"]))
    assert out.shape == (16, 21, 3) and out.device.type == "cpu"
    assert np.abs(out.numpy() - g["landmarks"]).max() < 2e-3  # float32 at |x| ~ 700 mm
    assert np.abs(out.numpy() - g["landmarks"]).max() < KEYPOINT_TOL_MM
    # known answer: rest pose + identity wrist -> rest landmarks
    assert np.abs(out[0].numpy() - model.landmark_rest_positions.numpy()).max() < 1e-4


def test_fk_batched_shapes_mirror_and_large_batch():
    model = hmod.load_hand_model_from_dict(hm_dict())
    t = hm_tensors()
    g = np.random.default_rng(3)
    n = 4099
    ja = torch.from_numpy(g.uniform(-0.5, 1.2, size=(n, 22)).astype(np.float32))
    wr = torch.eye(4).repeat(n, 1, 1)
    wr[:, :3, 3] = torch.from_numpy(g.normal(0, 150, size=(n, 3)).astype(np.float32))
    mirror = torch.from_numpy((np.arange(n) % 2).astype(np.uint8))
    out = fk_landmarks_device(model, ja.cuda(), wr.cuda(), mirror.cuda()).cpu()
    wr_m = wr.clone()
    wr_m[mirror.bool(), :, 0] *= -1
    ref = O.skin_landmarks(t["joint_rotation_axes"], t["joint_rest_positions"], t["landmark_rest_positions"],
                           t["landmark_rest_bone_indices"], t["landmark_rest_bone_weights"], ja, wr_m)
    assert (out - ref).abs().max() < 2e-3
    # leading dims [2, 5] with a batched model, as the reference allows
    lead = skin_landmarks(hmod.HandModel(*[(v[None, None].expand(2, 5, *v.shape) if isinstance(v, torch.Tensor) and v.dim() > 0 else v)
                                           for v in model]), ja[:10].reshape(2, 5, 22), wr[:10].reshape(2, 5, 4, 4))
    assert lead.shape == (2, 5, 21, 3)
    assert (lead.reshape(10, 21, 3) - O.skin_landmarks(t["joint_rotation_axes"], t["joint_rest_positions"],
            t["landmark_rest_positions"], t["landmark_rest_bone_indices"], t["landmark_rest_bone_weights"],
            ja[:10], wr[:10])).abs().max() < 2e-3


def test_metric_sums_match_numpy():
    g = golden("tracker_golden.npz")
    gt, tr, valid = g["gt"], g["tracked"], g["valid"]
    s = MetricSums("cuda")
    p = torch.from_numpy(tr.reshape(-1, 21, 3)).float().cuda()
    q = torch.from_numpy(gt.reshape(-1, 21, 3)).float().cuda()
    v = torch.from_numpy(valid.reshape(-1)).cuda()
    s.add_errors(p, q, v)
    T = gt.shape[1]
    s.add_accelerations(p.reshape(2, T, 21, 3), v.reshape(2, T))
    s.add_accelerations(q.reshape(2, T, 21, 3), v.reshape(2, T), gt=True)
    r = s.result()
    # the reference reduces float64 copies of float32 landmarks
    e = O.mpjpe_per_frame(gt.astype(np.float32).astype(np.float64), tr.astype(np.float32).astype(np.float64), valid)
    a = O.accel_per_frame(tr.astype(np.float32).astype(np.float64), valid)
    ag = O.accel_per_frame(gt.astype(np.float32).astype(np.float64), valid)
    assert abs(r["mpjpe"] - e.mean()) < 1e-9 * max(1.0, e.mean()) and r["n_hand_frames"] == len(e)
    assert abs(r["mpjpa"] - a.mean()) < 1e-9 * max(1.0, a.mean()) and r["n_accel"] == len(a)
    assert abs(r["gt_mpjpa"] - ag.mean()) < 1e-9 * max(1.0, ag.mean())
    assert abs(r["mpjpe"] - g["per_frame_error"].mean()) < 0.01  # north_star: MPJPE within 0.01


def test_metric_large_and_empty():
    g = np.random.default_rng(0)
    n = 200_003
    p = torch.from_numpy(g.normal(0, 30, size=(n, 21, 3)).astype(np.float32)).cuda()
    q = torch.from_numpy(g.normal(0, 30, size=(n, 21, 3)).astype(np.float32)).cuda()
    s = MetricSums("cuda")
    s.add_errors(p, q)
    ref = torch.linalg.norm((q.double() - p.double()), dim=-1).mean(dim=-1).mean().item()
    assert abs(s.result()["mpjpe"] - ref) < 1e-9 * ref
    s2 = MetricSums("cuda")
    s2.add_errors(p[:0], q[:0])
    assert s2.result()["n_hand_frames"] == 0
