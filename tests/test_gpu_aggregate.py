"""Aggregate metrics on the GPU against the reference's golden outputs
(tests/golden/make_aggregate_golden.py) and the oracle: per-frame errors bit-exact, PCK counts
bit-exact (integers), AUC / means to float64 round-off, result files readable both ways."""
import contextlib
import io
import os
import pickle

import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from tests.helpers import aggregate_inputs, golden
from umetrack_b200 import aggregate
from umetrack_b200.metrics import PCK_THRESHOLDS, MetricSums, keypoint_errors_device

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda().reshape(-1, 21, 3)


def test_keypoint_errors_bit_exact_vs_reference():
    g = golden("aggregate_golden.npz")
    for i, r in enumerate(aggregate_inputs()):
        err = keypoint_errors_device(_dev(r["tracked_keypoints"]), _dev(r["gt_keypoints"])).cpu().numpy()
        assert np.array_equal(err[r["valid_tracking"].reshape(-1)], g[f"errors_{i}"])


def test_pck_curve_and_auc_match_reference():
    g = golden("aggregate_golden.npz")
    errs = np.concatenate([g[f"errors_{i}"] for i in range(3)])
    pck = aggregate.PCK_curve(errs, PCK_THRESHOLDS) * 100.0
    assert np.array_equal(pck, g["pck_curve"])                      # integer counts / n: exact
    assert float(aggregate.normalized_AUC(PCK_THRESHOLDS, pck)) == float(g["auc"])
    e2 = np.stack([errs[:60], errs[60:120]])
    mask = (np.arange(120).reshape(2, 60) % 3 != 0).astype(np.float64)
    assert np.array_equal(aggregate.PCK_curve(e2, PCK_THRESHOLDS, mask=mask, axis=0), g["pck_axis0_masked"])
    # empty mask row -> zeros (the reference's _safe_div); NaN errors are never counted
    z = aggregate.PCK_curve(np.array([[1.0, 2.0], [3.0, np.nan]]), PCK_THRESHOLDS,
                            mask=np.array([[0.0, 0.0], [1.0, 1.0]]), axis=0)
    assert np.all(z[0] == 0) and z[1, -1] == 0.5
    with pytest.raises(AssertionError):
        aggregate.PCK_curve(np.zeros(3), PCK_THRESHOLDS, mask=np.zeros(4))


def test_aggregate_directory_matches_reference_summary(tmp_path, capsys):
    g = golden("aggregate_golden.npz")
    recs = aggregate_inputs()
    for i, r in enumerate(recs):
        aggregate.save_eval_result(str(tmp_path / f"user{i}" / f"rec{i}.npy"), r)
    with open(tmp_path / "user0" / "rec0.npy", "rb") as fp:        # the reference's pickle layout
        back = pickle.load(fp)
    assert set(back) == {"tracked_keypoints", "gt_keypoints", "valid_tracking"}
    assert back["tracked_keypoints"].dtype == np.float64 and back["valid_tracking"].dtype == bool
    out = aggregate.aggregate_metrics(str(tmp_path))
    ref = O.aggregate_metrics(recs)
    assert out["n_valid"] == ref["n_valid"] and out["n_total"] == ref["n_total"]
    assert out["success_rate"] == ref["success_rate"]
    assert np.array_equal(out["pck_curve"], ref["pck_curve"].reshape(-1))
    assert abs(out["auc"] - ref["auc"]) < 1e-9
    assert abs(out["mean_keypoint_error"] - ref["mean_keypoint_error"]) < 1e-9       # gate: 0.01
    assert abs(out["mean_keypoint_acceleration"] - ref["mean_keypoint_acceleration"]) < 1e-6
    assert abs(out["gt_mean_keypoint_acceleration"] - ref["gt_mean_keypoint_acceleration"]) < 1e-6
    printed = capsys.readouterr().out.splitlines()
    want = str(g["summary_text"]).splitlines()
    assert printed[0] == want[0]                                     # "Tracked n out of m, success rate: x%"
    assert printed[2] == want[2]                                     # AUC line, digit for digit
    assert aggregate.aggregate_metrics(str(tmp_path / "nothing_here"), verbose=False) is None


def test_pck_full_size_properties():
    """1M hand-frames: counts are monotone in the threshold, the last count is the number of
    errors <= 50 mm, and two half-batches add up to the whole (linearity of the accumulation)."""
    gen = torch.Generator(device="cuda").manual_seed(5)
    n = 1 << 20
    gt = torch.rand((n, 21, 3), device="cuda", generator=gen) * 100
    pred = gt + torch.randn((n, 21, 3), device="cuda", generator=gen) * 20
    valid = torch.rand(n, device="cuda", generator=gen) > 0.1
    s = MetricSums("cuda")
    s.add_errors(pred, gt, valid)
    a, b = MetricSums("cuda"), MetricSums("cuda")
    a.add_errors(pred[: n // 2], gt[: n // 2], valid[: n // 2])
    b.add_errors(pred[n // 2:], gt[n // 2:], valid[n // 2:])
    c = s.sums[8:].cpu().numpy()
    assert np.all(np.diff(c) >= 0)
    err = keypoint_errors_device(pred, gt)
    assert c[-1] == float(((err <= 50.0) & valid).sum())
    assert np.array_equal(c, (a.sums[8:] + b.sums[8:]).cpu().numpy())
    assert s.sums[6].item() == n and s.sums[1].item() == float(valid.sum())
