"""Time umt_resample_crops on the crops of a 512-frame synthetic rig (GPU box)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from umetrack_b200 import hand_model as hmod, native, synthetic  # noqa: E402
from umetrack_b200.batched import crop_cameras_device  # noqa: E402
from tests.helpers import hm_dict  # noqa: E402

dev = torch.device("cuda")
T, H, W = 512, 480, 640
hd = hm_dict()
labels, _ = synthetic.make_recording(hd, seed=11, n_frames=T, dropouts=False, with_frames=False)
cc = crop_cameras_device(hmod.load_hand_model_from_dict(hd), labels, dev)
cnt, sel = cc["sel_count"].cpu().numpy(), cc["sel_cam"].cpu().numpy()
item = [(t, h, k) for t in range(T) for h in range(2) for k in range(int(cnt[t, h]))]
ti, hi, ki = (torch.tensor([i[j] for i in item], device=dev) for j in range(3))
params = cc["resample_params"][ti, hi, ki].contiguous()
offs = torch.tensor([t * H * 4 * W + int(sel[t, h, k]) * W for t, h, k in item], dtype=torch.int64, device=dev)
frames = torch.randint(0, 256, (T, H, 4 * W), dtype=torch.uint8, device=dev)
crops = torch.empty((len(item), 96, 96), dtype=torch.float32, device=dev)
lib = native.lib()
run = lambda: native.check(lib.umt_resample_crops(frames.data_ptr(), offs.data_ptr(), 4 * W, W, H, params.data_ptr(),
                                                  len(item), 0, crops.data_ptr(), 0, native.current_stream_ptr()), "resample")
run(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"{len(item)} crops, {ms:.3f} ms, {len(item) / ms / 1e3:.2f} M crops/s, checksum {float(crops.sum()):.1f}")
