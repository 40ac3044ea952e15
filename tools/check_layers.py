"""Diagnostic (GPU box): per-layer error of the tensor-core convs vs the CUDA-core check kernel and
the oracle, in both numeric modes.  Prints a table; exits non-zero if the fp32-equivalent mode is off."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import umetrack_oracle as O  # noqa: E402  (diagnostic tool, not product)
from umetrack_b200 import synthetic  # noqa: E402
from umetrack_b200.model import UmeTrackModel  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 9
sd = synthetic.make_state_dict(0)
imgs = synthetic.make_model_batch((n + 1) // 2, seed=100)["left_images"][:n]
taps = {}
feats = O.backbone_forward(sd, imgs, taps)
names = []
for li, nb in enumerate((2, 3, 5, 2)):
    for b in range(nb):
        names += [None, f"layer{li + 1}.{b}"]
model = UmeTrackModel(sd).to("cuda")
bad = False
for mode in ("fp32", "bf16", "fp16"):
    model.set_mode(mode)
    for layer in range(24):
        simt, umma = model.debug_layer(layer, imgs)
        torch.cuda.synchronize()
        simt, umma = simt.cpu(), umma.cpu()
        sc = float(simt.abs().max()) + 1e-9
        e1 = float((simt - umma).abs().max()) / sc
        msg = f"{mode} layer {layer:2d} shape {tuple(umma.shape)} |simt|max {sc:.3e} umma-vs-simt {e1:.2e}"
        if names[layer]:
            ref = taps[names[layer]]
            e2 = float((umma - ref).abs().max()) / (float(ref.abs().max()) + 1e-9)
            e3 = float((simt - ref).abs().max()) / (float(ref.abs().max()) + 1e-9)
            msg += f"  umma-vs-oracle {e2:.2e} simt-vs-oracle {e3:.2e}"
            if mode == "fp32" and e2 > 1e-3:
                bad = True
        if mode == "fp32" and e1 > 1e-4:
            bad = True
        print(msg, flush=True)
    out = model.encode(imgs).cpu()
    print(f"{mode} features rel err {float((out - feats).abs().max()) / float(feats.abs().max()):.3e}", flush=True)
sys.exit(1 if bad else 0)
