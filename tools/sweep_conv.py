"""Tuning aid (GPU box): sweep tile/occupancy/grouping configurations of the tensor-core conv
kernel and print the algorithmic TFLOP/s per output-width class and numeric mode.

    python tools/sweep_conv.py "4,2,3,2,2" "4,1,3,2,2" ...      # config = S,occ,G,cluster,nbuf
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from umetrack_b200 import synthetic  # noqa: E402
from umetrack_b200.model import UmeTrackModel  # noqa: E402

configs = sys.argv[1:] or ["4,2,3,2,2"]
n = int(os.environ.get("SWEEP_CROPS", "2048"))
sd = synthetic.make_state_dict(0)
imgs = torch.rand((n, 96, 96), device="cuda")
flops = np.array(bench.conv_flops_per_crop(), dtype=np.float64) * n
cls = [32] * 4 + [64] * 6 + [128] * 10 + [256] * 4
for cfg in configs:
    for c in (32, 64, 128, 256):
        os.environ[f"UMT_CONV_CFG{c}"] = cfg
    model = UmeTrackModel(sd).to("cuda")
    row = [f"cfg {cfg:<12}"]
    for mode in ("fp32", "bf16", "fp16"):
        model.set_mode(mode)
        try:
            model.profile_layers(imgs)
            t = np.mean([model.profile_layers(imgs) for _ in range(3)], axis=0)[1:25]
        except Exception as e:  # noqa: BLE001
            row.append(f"{mode}: failed ({str(e)[-60:]})")
            continue
        parts = []
        for c in (32, 64, 128, 256):
            idx = [i for i, k in enumerate(cls) if k == c]
            parts.append(f"{c}:{flops[idx].sum() / (t[idx].sum() * 1e-3) / 1e12:6.0f}")
        row.append(f"{mode} [{' '.join(parts)}] all:{flops.sum() / (t.sum() * 1e-3) / 1e12:6.1f} ms:{t.sum():.3f}")
    print(" | ".join(row), flush=True)
    del model
    torch.cuda.empty_cache()
