"""Host profile of track_recordings_batched (where does the wall time go?).  Needs a B200.
Usage: python tools/profile_batched.py [n_recordings] [n_frames]"""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from umetrack_b200 import synthetic
from umetrack_b200.batched import track_recordings_batched
from umetrack_b200.model import UmeTrackModel
from tests.helpers import hm_dict

R = int(sys.argv[1]) if len(sys.argv) > 1 else 8
T = int(sys.argv[2]) if len(sys.argv) > 2 else 48
dev = torch.device("cuda:0")
model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
model.set_mode("fp16")
hm = hm_dict()
recs = [synthetic.make_recording(hm, seed=100 + r, n_frames=T) for r in range(R)]
recs = [(lab, torch.from_numpy(fr).pin_memory()) for lab, fr in recs]     # decoded frames in pinned memory, like bench.py
track_recordings_batched(model, recs[:2])
torch.cuda.synchronize()
for _ in range(2):
    t0 = time.perf_counter(); track_recordings_batched(model, recs); torch.cuda.synchronize()
    print("wall", time.perf_counter() - t0)
pr = cProfile.Profile(); pr.enable()
track_recordings_batched(model, recs); torch.cuda.synchronize()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(35); print(s.getvalue())
