"""MPJPE / keypoint-acceleration ("MPJPA") sums on the GPU and their reduction across ranks.

Definitions follow the reference scripts: per valid (hand, frame) the mean over 21 landmarks of
the L2 error, then the global mean over all valid hand-frames
(run_eval_known_skeleton.py:88-89,117-119); acceleration |p[t] + p[t+2] - 2 p[t+1]| averaged over
landmarks on triples of consecutive valid frames (load_eval.py:29-44).  Each rank keeps ONE float64
vector: [sum_err, n_err, sum_acc_pred, n_acc, sum_acc_gt, n_acc_gt, n_slots, 0] followed by the
101 PCK counts (errors <= threshold, lib/common/metric_utils.py:19-20,36-88); `all_reduce()` adds
it across the process group (the only collective on the path -- NCCL on GPUs, gloo in CPU
tests).  `result()` derives what load_eval.py:47-89 prints: success rate, mean error, PCK curve
(percent), normalised AUC, mean accelerations.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from . import native


MAX_LANDMARK_ERROR_MM = 50                                   # lib/common/metric_utils.py:19
PCK_THRESHOLDS = np.linspace(0, MAX_LANDMARK_ERROR_MM, 101)  # lib/common/metric_utils.py:20
_N_SCALARS = 8


def keypoint_errors_device(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """Per-hand-frame mean-over-21 L2 error, float64 [n], for [n,21,3] float32 CUDA tensors
    (run_eval_known_skeleton.py:88-89), bit-identical to numpy's evaluation order."""
    p, g = pred.contiguous().float(), gt.contiguous().float()
    n = p.shape[0]
    err = torch.empty(n, dtype=torch.float64, device=p.device)
    native.check(native.lib().umt_keypoint_errors(native.ptr(p), native.ptr(g), n, native.ptr(err),
                                                  native.current_stream_ptr()), "umt_keypoint_errors")
    return err


def normalized_AUC(x: np.ndarray, y: np.ndarray, y_max: float = 1.0) -> np.ndarray:
    """Trapezoid area under curves sharing the x axis, over (x[-1]-x[0])*y_max
    (lib/common/metric_utils.py:91-112).  101 numbers: host arithmetic."""
    shape = y.shape[:-1]
    y2 = y.reshape(-1, y.shape[-1])
    auc = ((x[1:] - x[:-1]).reshape(1, -1) * ((y2[..., 1:] + y2[..., :-1]) * 0.5)).sum(axis=-1)
    return (auc / ((x[-1] - x[0]) * y_max)).reshape(shape)


class MetricSums:
    def __init__(self, device: Optional[torch.device] = None):
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.sums = torch.zeros(_N_SCALARS + len(PCK_THRESHOLDS), dtype=torch.float64, device=self.device)
        self._thr = None

    # ---- accumulation (CUDA kernels) ---------------------------------------------------------
    def add_errors(self, pred: torch.Tensor, gt: torch.Tensor, valid: Optional[torch.Tensor] = None):
        """pred/gt [n,21,3] float32 CUDA tensors (mm); valid [n] bool/uint8 or None."""
        self._need_cuda()
        n = pred.shape[0]
        p, g = pred.contiguous().float(), gt.contiguous().float()
        v = None if valid is None else valid.to(torch.uint8).contiguous()
        native.check(native.lib().umt_mpjpe_accumulate(native.ptr(p), native.ptr(g), native.ptr(v), n,
                                                       self.sums.data_ptr(), native.current_stream_ptr()),
                     "umt_mpjpe_accumulate")
        # PCK counts + number of (hand, frame) slots seen (success rate, load_eval.py:73-78)
        if self._thr is None:
            self._thr = torch.from_numpy(PCK_THRESHOLDS).to(self.device)
        err = keypoint_errors_device(p, g)
        native.check(native.lib().umt_pck_accumulate(native.ptr(err), native.ptr(v), n, native.ptr(self._thr),
                                                     len(PCK_THRESHOLDS), self.sums.data_ptr() + 8 * _N_SCALARS,
                                                     native.current_stream_ptr()), "umt_pck_accumulate")
        self.sums[6] += n

    def add_accelerations(self, pts: torch.Tensor, valid: torch.Tensor, gt: bool = False):
        """pts [n_seq,T,21,3] float32 CUDA, valid [n_seq,T]."""
        self._need_cuda()
        n_seq, T = pts.shape[0], pts.shape[1]
        p = pts.contiguous().float()
        v = valid.to(torch.uint8).contiguous()
        off = 4 if gt else 2
        native.check(native.lib().umt_accel_accumulate(native.ptr(p), native.ptr(v), n_seq, T,
                                                       self.sums.data_ptr() + 8 * off,
                                                       native.current_stream_ptr()), "umt_accel_accumulate")

    def _need_cuda(self):
        if self.device.type != "cuda":
            raise native.NativeError("MetricSums accumulates on a CUDA device (no CPU fallback)")

    # ---- reduction -----------------------------------------------------------------------------
    def all_reduce(self, group=None):
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.sums, op=dist.ReduceOp.SUM, group=group)
        return self

    def result(self) -> Dict[str, float]:
        s = self.sums.detach().cpu().tolist()
        div = lambda a, b: (a / b) if b > 0 else float("nan")
        pck = np.array([div(c, s[1]) for c in s[_N_SCALARS:]]) * 100.0     # percent, as load_eval.py:66-71
        return {"mpjpe": div(s[0], s[1]), "n_hand_frames": s[1], "mpjpa": div(s[2], s[3]),
                "gt_mpjpa": div(s[4], s[5]), "n_accel": s[3], "n_slots": s[6],
                "success_rate": div(s[1], s[6]) * 100.0, "pck_curve": pck,
                "auc": float(normalized_AUC(PCK_THRESHOLDS, pck)) if s[1] > 0 else float("nan")}


def shard_indices(n_items: int, rank: int, world_size: int, costs=None):
    """Recording indices owned by `rank`: round-robin, or longest-processing-time-first when
    per-item `costs` (e.g. frame counts) are given.  Every item is owned by exactly one rank."""
    if costs is None:
        return list(range(rank, n_items, world_size))
    order = sorted(range(n_items), key=lambda i: (-costs[i], i))
    load = [0.0] * world_size
    owner = {}
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        owner[i] = r
        load[r] += costs[i]
    return sorted(i for i, r in owner.items() if r == rank)
