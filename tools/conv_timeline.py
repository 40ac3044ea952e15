"""Diagnostic (GPU box): where do the roles of the conv kernel wait?  Runs the encoder up to layer L
with the per-CTA cycle counters switched on for that one launch and prints the averages."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from umetrack_b200 import native, synthetic  # noqa: E402
from umetrack_b200.model import UmeTrackModel  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16"
layers = [int(x) for x in sys.argv[2:]] or [1, 7, 13, 22]
n = int(os.environ.get("SWEEP_CROPS", "2048"))
lib = native.lib()
lib.umt_debug_set_conv_timeline.argtypes = [ctypes.c_void_p]
lib.umt_debug_set_conv_timeline.restype = None
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
model.set_mode(mode)
imgs = torch.rand((n, 96, 96), device="cuda")
model.encode(imgs)
names = ["A:wait_empty", "A:total", "B:wait_empty", "M:wait_acc", "M:wait_a", "M:wait_b", "M:total", "M:tiles",
         "E:wait_full", "E:total", "g0", "g1", "sm"]
for layer in layers:
    stage = 0 if layer < 4 else 1 if layer < 10 else 2 if layer < 20 else 3
    c, h = {0: (32, 48), 1: (64, 24), 2: (128, 12), 3: (256, 6)}[stage]
    dense = torch.empty((n, c, h, h), dtype=torch.float32, device="cuda")
    buf = torch.zeros((1024, 16), dtype=torch.int64, device="cuda")
    out = (ctypes.c_int * 3)()
    model._ensure_workspace(n, 8)
    lib.umt_debug_set_conv_timeline(buf.data_ptr())
    # debug_layer runs stem + convs 0..layer-1, then `layer` itself: every launch overwrites the
    # counters, the last writer is `layer`
    native.check(lib.umt_model_debug_layer(model._handle, model._mode, layer, imgs.data_ptr(), n, None,
                                           dense.data_ptr(), out, 0), "dbg")
    torch.cuda.synchronize()
    lib.umt_debug_set_conv_timeline(None)
    b = buf.cpu().double()
    used = b[:, 6] > 0
    m = b[used].mean(dim=0)
    g0, g1, sm = b[used][:, 10], b[used][:, 11], b[used][:, 12]
    t0 = g0.min()
    per_sm = {}
    for a_, e_, s_ in zip(g0.tolist(), g1.tolist(), sm.tolist()):
        per_sm.setdefault(int(s_), []).append(((a_ - t0) / 1e3, (e_ - t0) / 1e3))
    overlap = sum(1 for v in per_sm.values() if len(v) >= 2 and sorted(v)[1][0] < sorted(v)[0][1])
    print(f"   wall: first start 0, last start {(g0.max() - t0) / 1e3:.1f} us, last end {(g1.max() - t0) / 1e3:.1f} us, "
          f"mean CTA life {((g1 - g0).mean()) / 1e3:.1f} us, SMs used {len(per_sm)}, SMs with 2 overlapping CTAs {overlap}, "
          f"clk/us {(m[6] / ((g1 - g0).mean() / 1e3)):.0f}")
    print(f"layer {layer} ({mode}) C={out[0]} HW={out[1]} ctas={int(used.sum())} tiles/cta={m[7]:.1f}: " +
          " ".join(f"{nm}={m[i]:.0f}" for i, nm in enumerate(names) if i not in (7, 10, 11, 12)), flush=True)
