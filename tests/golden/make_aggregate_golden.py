"""Golden vectors for the aggregate metrics from the UNMODIFIED
reference: load_eval.py:_compute_metrics / aggregate_metrics and lib/common/metric_utils.py.

    python tests/golden/make_aggregate_golden.py        # build container only (needs /root/reference)

Inputs are regenerated from the seed by `aggregate_inputs` below (imported by the tests), so only
the reference's outputs are stored in aggregate_golden.npz.
"""
import contextlib
import io
import os
import pickle
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("UMETRACK_REFERENCE", "/root/reference")


def aggregate_inputs(seed=7, n_rec=3):
    """Synthetic eval results shaped like the reference's pickles: float32-valued keypoints in
    float64 arrays [2,T,21,3] (mm), errors spread over 0..80 mm so every PCK bin is exercised,
    some errors exactly on a threshold, invalid frames zeroed like the reference leaves them."""
    rng = np.random.default_rng(seed)
    out = []
    for r in range(n_rec):
        T = 40 + 17 * r
        gt = rng.uniform(-300, 300, size=(2, T, 21, 3)).astype(np.float32)
        mag = rng.uniform(0, 80, size=(2, T, 1, 1)).astype(np.float32)
        d = rng.normal(size=(2, T, 21, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=-1, keepdims=True)
        tracked = (gt + d * mag).astype(np.float32)
        # a few frames whose 21 errors are all exactly 12.5 / 50 mm (on a threshold)
        tracked[0, 3] = gt[0, 3] + np.array([12.5, 0, 0], dtype=np.float32)
        tracked[1, 5] = gt[1, 5] + np.array([0, 50.0, 0], dtype=np.float32)
        valid = rng.uniform(size=(2, T)) > 0.15
        valid[0, 3] = valid[1, 5] = True
        gt64, tr64 = gt.astype(np.float64), tracked.astype(np.float64)
        gt64[~valid] = 0.0
        tr64[~valid] = 0.0
        out.append({"tracked_keypoints": tr64, "gt_keypoints": gt64, "valid_tracking": valid})
    return out


def main():
    sys.path.insert(0, REF)
    import load_eval
    from lib.common import metric_utils

    recs = aggregate_inputs()
    store = {}
    errs = []
    for i, r in enumerate(recs):
        m = load_eval._compute_metrics(r["gt_keypoints"], r["tracked_keypoints"], r["valid_tracking"])
        store[f"errors_{i}"] = m.keypoint_errors
        store[f"accel_{i}"] = m.keypoint_accelerations
        store[f"gt_accel_{i}"] = m.gt_keypoint_accelerations
        errs.append(m.keypoint_errors)
    errs = np.concatenate(errs)
    pck = metric_utils.PCK_curve(errs, metric_utils.PCK_THRESHOLDS) * 100.0
    store["thresholds"] = metric_utils.PCK_THRESHOLDS
    store["pck_curve"] = pck
    store["auc"] = np.float64(metric_utils.normalized_AUC(metric_utils.PCK_THRESHOLDS, pck))
    # masked / per-axis variant of PCK_curve
    e2 = np.stack([errs[:60], errs[60:120]])
    mask = (np.arange(120).reshape(2, 60) % 3 != 0).astype(np.float64)
    store["pck_axis0_masked"] = metric_utils.PCK_curve(e2, metric_utils.PCK_THRESHOLDS, mask=mask, axis=0)
    # the printed summary of aggregate_metrics over a directory of pickles
    with tempfile.TemporaryDirectory() as d:
        for i, r in enumerate(recs):
            os.makedirs(os.path.join(d, f"user{i}"), exist_ok=True)
            with open(os.path.join(d, f"user{i}", f"rec{i}.npy"), "wb") as fp:
                pickle.dump(r, fp)
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            load_eval.aggregate_metrics(d)
    store["summary_text"] = np.array(buf.getvalue())
    np.savez_compressed(os.path.join(HERE, "aggregate_golden.npz"), **store)
    print(buf.getvalue())
    print("auc", store["auc"], "pck[::20]", pck.reshape(-1)[::20])


if __name__ == "__main__":
    main()
