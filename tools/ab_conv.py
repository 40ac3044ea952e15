"""A/B of conv configurations inside one process (GPU box): the per-class UMT_CONV_CFG<cout>
strings are read when a model is built, so two models with different settings can be timed
alternately, which cancels clock / power drift between runs.

    python tools/ab_conv.py "128=2,2,5,2,2,0" "128=2,2,9,2,2,0" [reps]
"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from umetrack_b200 import synthetic  # noqa: E402
from umetrack_b200.model import UmeTrackModel  # noqa: E402

specs = sys.argv[1:3]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 12
mode = os.environ.get("AB_MODE", "fp16")
sd = synthetic.make_state_dict(0)
imgs = torch.rand((2048, 96, 96), device="cuda")
models = []
for spec in specs:
    for c in (32, 64, 128, 256):
        os.environ.pop(f"UMT_CONV_CFG{c}", None)
    for part in spec.split(";"):
        if part:
            c, cfg = part.split("=")
            os.environ[f"UMT_CONV_CFG{c}"] = cfg
    m = UmeTrackModel(sd).to("cuda"); m.set_mode(mode); m.profile_layers(imgs)
    models.append(m)
flops = np.array(bench.conv_flops_per_crop(), dtype=np.float64) * 2048
cls = np.array([32] * 4 + [64] * 6 + [128] * 10 + [256] * 4)
t = [[], []]
for r in range(reps):
    for k in (0, 1) if r % 2 == 0 else (1, 0):
        t[k].append(np.array(models[k].profile_layers(imgs))[1:25])
for k in (0, 1):
    med = np.median(np.array(t[k]), axis=0)
    parts = [f"{c}:{flops[cls == c].sum() / (med[cls == c].sum() * 1e-3) / 1e12:6.0f}" for c in (32, 64, 128, 256)]
    print(f"{specs[k]:<40} [{' '.join(parts)}] conv ms {med.sum():.3f}")
