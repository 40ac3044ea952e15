"""Aggregate metrics over eval results: the reference's
`load_eval.py:47-89` and `lib/common/metric_utils.py`, with the reductions on the GPU.

* result files are the reference's own format: a pickle (named *.npy) of
  {"tracked_keypoints", "gt_keypoints": float64 [2,T,21,3] mm, "valid_tracking": bool [2,T]}
  (run_eval_known_skeleton.py:92-100), so `load_eval.py` can read what `save_eval_result`
  writes and `aggregate_metrics` can read what the reference scripts wrote;
* per-hand-frame errors, the PCK counts over the 101 thresholds and the acceleration sums are
  CUDA kernels (umt_keypoint_errors, umt_pck_accumulate, umt_accel_accumulate); only the 101-point
  trapezoid of the AUC is host arithmetic.  Keypoints are float32 values (the reference computes
  landmarks in float32, lib/tracker/perspective_crop.py:32-36), so the upload is lossless.
"""
from __future__ import annotations

import fnmatch
import os
import pickle
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np
import torch

from . import native
from .metrics import (MAX_LANDMARK_ERROR_MM, PCK_THRESHOLDS, MetricSums, keypoint_errors_device,  # noqa: F401
                      normalized_AUC)


def _cuda_device(device=None) -> torch.device:
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise native.NativeError("aggregate metrics run on a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def PCK_curve(errors, thresholds: np.ndarray, mask: Optional[np.ndarray] = None,
              axis: Optional[int] = None, device=None) -> np.ndarray:
    """Fraction of (masked) errors <= each threshold (lib/common/metric_utils.py:48-88): one
    curve [len(thresholds)] over everything, or [n, len(thresholds)] with one per element along
    `axis`.  `mask` is 0/1.  Rows whose
    mask is empty give 0, like the reference's _safe_div."""
    dev = _cuda_device(device)
    e = np.asarray(errors, dtype=np.float64)
    m = None if mask is None else np.asarray(mask)
    if m is not None and m.shape != e.shape:
        raise AssertionError("mask and errors must have the same shape")   # metric_utils.py:24
    if axis is None:
        e2 = e.reshape(1, -1)
        m2 = None if m is None else m.reshape(1, -1)
    else:
        n = e.shape[axis]
        e2 = np.moveaxis(e, axis, 0).reshape(n, -1)
        m2 = None if m is None else np.moveaxis(m, axis, 0).reshape(n, -1)
    thr = torch.from_numpy(np.ascontiguousarray(thresholds, dtype=np.float64)).to(dev)
    if thr.numel() > 1 and bool((thr[1:] < thr[:-1]).any()):
        raise ValueError("thresholds must be ascending")
    rows, cols = e2.shape
    err = torch.from_numpy(np.ascontiguousarray(e2)).to(dev)
    val = None if m2 is None else torch.from_numpy(np.ascontiguousarray(m2 != 0)).to(dev).to(torch.uint8)
    counts = torch.zeros((rows, thr.numel()), dtype=torch.float64, device=dev)
    lib = native.lib()
    with torch.cuda.device(dev):
        for r in range(rows):
            native.check(lib.umt_pck_accumulate(err[r].data_ptr(), 0 if val is None else val[r].data_ptr(), cols,
                                                thr.data_ptr(), thr.numel(), counts[r].data_ptr(),
                                                native.current_stream_ptr()), "umt_pck_accumulate")
    den = np.full(rows, float(cols)) if m2 is None else m2.astype(np.float64).sum(axis=-1)
    c = counts.cpu().numpy()
    out = np.zeros_like(c)
    ok = den >= 1e-6
    out[ok] = c[ok] / den[ok, None]
    return out[0] if axis is None else out


@dataclass
class Metrics:
    """What load_eval.py:20-24 keeps per recording, with the accelerations already reduced."""
    keypoint_errors: np.ndarray          # float64 [n_valid]: per valid (hand, frame), hand-major
    accel_sum: float                     # sum over valid triples of the mean-21 acceleration norm
    gt_accel_sum: float
    n_accel: int


def compute_metrics(gt_keypoints: np.ndarray, tracked_keypoints: np.ndarray, valid_tracking: np.ndarray,
                    device=None, sums: Optional[MetricSums] = None) -> Metrics:
    """load_eval.py:27-44 for one recording ([2,T,21,3] float64 mm, valid [2,T]).  If `sums` is
    given the recording is also added to it (errors, PCK counts, accelerations)."""
    dev = _cuda_device(device if sums is None else sums.device)
    T = valid_tracking.shape[1]
    g = torch.from_numpy(np.ascontiguousarray(gt_keypoints, dtype=np.float32)).to(dev).reshape(-1, 21, 3)
    p = torch.from_numpy(np.ascontiguousarray(tracked_keypoints, dtype=np.float32)).to(dev).reshape(-1, 21, 3)
    v = torch.from_numpy(np.ascontiguousarray(valid_tracking)).to(dev).reshape(-1)
    with torch.cuda.device(dev):
        err = keypoint_errors_device(p, g)[v].cpu().numpy()
        local = MetricSums(dev)
        for s in ([local] if sums is None else [local, sums]):
            if s is not local:
                s.add_errors(p, g, v)
            s.add_accelerations(p.reshape(2, T, 21, 3), v.reshape(2, T), gt=False)
            s.add_accelerations(g.reshape(2, T, 21, 3), v.reshape(2, T), gt=True)
        acc = local.sums[:6].cpu().tolist()
    return Metrics(keypoint_errors=err, accel_sum=acc[2], gt_accel_sum=acc[4], n_accel=int(acc[3]))


def save_eval_result(output_path: str, result: Dict[str, np.ndarray]) -> None:
    """Write one recording's result exactly as run_eval_known_skeleton.py:90-100 does (a pickle of
    the three arrays; the reference names these files *.npy)."""
    d = os.path.dirname(output_path)
    if d and not os.path.exists(d):
        os.makedirs(d)
    with open(output_path, "wb") as fp:
        pickle.dump({"tracked_keypoints": np.asarray(result["tracked_keypoints"], dtype=np.float64),
                     "gt_keypoints": np.asarray(result["gt_keypoints"], dtype=np.float64),
                     "valid_tracking": np.asarray(result["valid_tracking"], dtype=bool)}, fp)


def aggregate_metrics(output_dir: str, device=None, verbose: bool = True) -> Optional[Dict[str, object]]:
    """load_eval.py:47-89: walk `output_dir` for *.npy pickles, reduce them on the GPU and print
    the same five lines.  Returns the numbers (None if no file was found, where the reference
    prints nothing)."""
    dev = _cuda_device(device)
    sums = MetricSums(dev)
    n_files = 0
    for cur_dir, _, filenames in os.walk(output_dir):
        for fname in fnmatch.filter(filenames, "*.npy"):
            with open(os.path.join(cur_dir, fname), "rb") as fp:
                data = pickle.load(fp)
            compute_metrics(data["gt_keypoints"], data["tracked_keypoints"], data["valid_tracking"], sums=sums)
            n_files += 1
    if n_files == 0:
        return None
    r = sums.result()
    out = {"n_valid": int(r["n_hand_frames"]), "n_total": int(r["n_slots"]), "success_rate": r["success_rate"],
           "mean_keypoint_error": r["mpjpe"], "auc": r["auc"], "pck_curve": r["pck_curve"],
           "mean_keypoint_acceleration": r["mpjpa"], "gt_mean_keypoint_acceleration": r["gt_mpjpa"]}
    if verbose:
        print(f"  Tracked {out['n_valid']} out of {out['n_total']}, success rate: {out['success_rate']}%")
        print(f"  Mean keypoint error: {out['mean_keypoint_error']}")
        print(f"  AUC score: {out['auc']}")
        print(f"  Mean keypoint accelerations: {out['mean_keypoint_acceleration']}")
        print(f"  GT mean keypoint accelerations: {out['gt_mean_keypoint_acceleration']}")
    return out
