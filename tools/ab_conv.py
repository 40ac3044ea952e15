"""A/B of conv configurations inside one process (GPU box): the per-class UMT_CONV_CFG<cout>
strings are read when a model is built, so two models with different settings can be timed
alternately, which cancels clock / power drift between runs.

    python tools/ab_conv.py "128=2,2,5,2,2,0" "128=2,2,9,2,2,0" [reps]

Parts of the form NAME=VALUE with a non-numeric NAME (e.g. UMT_CONV_NA=2) are per-launch knobs:
they are set in the environment around every timed call of that variant.
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
models, launch_env = [], []
for spec in specs:
    for c in (32, 64, 128, 256):
        os.environ.pop(f"UMT_CONV_CFG{c}", None)
    env = {}
    for part in spec.split(";"):
        if part:
            c, cfg = part.split("=")
            if c.isdigit():
                os.environ[f"UMT_CONV_CFG{c}"] = cfg
            else:
                env[c] = cfg
    launch_env.append(env)
    os.environ.update(env)
    m = UmeTrackModel(sd).to("cuda"); m.set_mode(mode); m.profile_layers(imgs)
    models.append(m)
    for k in env:
        os.environ.pop(k, None)
flops = np.array(bench.conv_flops_per_crop(), dtype=np.float64) * 2048
cls = np.array([32] * 4 + [64] * 6 + [128] * 10 + [256] * 4)
t = [[], []]
for r in range(reps):
    for k in (0, 1) if r % 2 == 0 else (1, 0):
        os.environ.update(launch_env[k])
        t[k].append(np.array(models[k].profile_layers(imgs))[1:25])
        for name in launch_env[k]:
            os.environ.pop(name, None)
for k in (0, 1):
    med = np.median(np.array(t[k]), axis=0)
    parts = [f"{c}:{flops[cls == c].sum() / (med[cls == c].sum() * 1e-3) / 1e12:6.0f}" for c in (32, 64, 128, 256)]
    print(f"{specs[k]:<40} [{' '.join(parts)}] conv ms {med.sum():.3f}")
