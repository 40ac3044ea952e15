"""Where a tracker frame's forward pass spends its time on the device (GPU box): global-timer stamps of the
first CTA start and the last CTA end of every conv_umma_kernel launch of one 4-crop pass (plain launches).

    UMT_CONV_DBG_STRIDE=4096 python tools/frame_timeline.py [mode]
"""
import ctypes, os, sys
os.environ.setdefault("UMT_CONV_DBG_STRIDE", "4096")
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from umetrack_b200 import native, synthetic
from umetrack_b200.model import UmeTrackModel

stride = int(os.environ["UMT_CONV_DBG_STRIDE"])
lib = native.lib()
lib.umt_debug_set_conv_timeline.argtypes = [ctypes.c_void_p]
lib.umt_debug_set_conv_timeline.restype = None
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
model.set_mode(sys.argv[1] if len(sys.argv) > 1 else "fp32")
x = torch.rand((4, 96, 96), device="cuda")
for _ in range(3):
    model.encode(x)
torch.cuda.synchronize()
buf = torch.zeros((64 * stride,), dtype=torch.int64, device="cuda")
lib.umt_debug_set_conv_timeline(buf.data_ptr())
model.encode(x)
torch.cuda.synchronize()
lib.umt_debug_set_conv_timeline(None)
b = buf.cpu().view(64, stride // 16, 16).double()
t0 = None
prev_end = None
for i in range(64):
    used = b[i][:, 6] > 0
    if not used.any():
        continue
    g0, g1 = b[i][used][:, 10], b[i][used][:, 11]
    if t0 is None:
        t0 = g0.min()
    gap = "" if prev_end is None else f"  gap after previous timed launch {(g0.min() - prev_end) / 1e3:6.1f} us"
    m = b[i][used].mean(dim=0)
    roles = (f"  [cycles/CTA: A total {m[1]:.0f} (wait for a free slot {m[0]:.0f}); MMA total {m[6]:.0f}, waits: acc {m[3]:.0f} "
             f"A {m[4]:.0f} B {m[5]:.0f}; epilogue total {m[9]:.0f}, wait for the accumulator {m[8]:.0f}]")
    print(f"launch {i:2d}: ctas {int(used.sum()):3d}  first start {(g0.min() - t0) / 1e3:7.1f} us  last start "
          f"{(g0.max() - t0) / 1e3:7.1f}  last end {(g1.max() - t0) / 1e3:7.1f}  life {(g1 - g0).mean() / 1e3:5.1f} us{gap}")
    print(roles)
    prev_end = g1.max()
