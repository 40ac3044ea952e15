"""Serialised stage breakdown of the batched driver on recordings whose frames are already in HBM (the
`recordings_sharded` stage of bench.py).  GPU box:  python tools/profile_sharded.py [n_recordings] [n_frames] [mode]"""
import os, sys, time
os.environ["UMETRACK_STAGE_TIMES"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from umetrack_b200 import nvtx, synthetic
from umetrack_b200.batched import label_arrays, track_recordings_batched
from umetrack_b200.metrics import MetricSums
from umetrack_b200.model import UmeTrackModel
from tests.helpers import hm_dict

R = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda:0")
model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
model.set_mode(sys.argv[3] if len(sys.argv) > 3 else "fp32")
hm = hm_dict()
recs = []
for r in range(R):
    lab = synthetic.make_recording(hm, seed=r, n_frames=T, with_frames=False)[0]
    label_arrays(lab)
    g = torch.Generator(device=dev).manual_seed(7000 + r)
    recs.append((lab, torch.randint(0, 256, (T, 480, 2560), dtype=torch.uint8, device=dev, generator=g)))
track_recordings_batched(model, recs, sums=MetricSums(dev))
nvtx.stage_times.clear()
torch.cuda.synchronize()
t0 = time.perf_counter()
res = track_recordings_batched(model, recs, sums=MetricSums(dev))
torch.cuda.synchronize()
wall = time.perf_counter() - t0
n = int(sum(x["valid_tracking"].sum() for x in res))
print(f"{R} recordings x {T} frames, {n} hand-frames: {wall * 1e3:.1f} ms with a synchronisation after every stage "
      f"({n / wall:.0f} hand-frames/s)")
for k, v in nvtx.stage_times.items():
    print(f"  {k:14s} {v * 1e3:8.1f} ms  {100 * v / wall:5.1f} %")
