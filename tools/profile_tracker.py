"""Host profile of the per-frame tracker loop (evaluation.track_recording).  Needs a B200."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from umetrack_b200 import synthetic
from umetrack_b200.evaluation import track_recording
from umetrack_b200.model import UmeTrackModel
from tests.helpers import hm_dict

model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
lab, fr = synthetic.make_recording(hm_dict(), seed=100, n_frames=48)
track_recording(model, lab, fr)
torch.cuda.synchronize()
t0 = time.perf_counter(); track_recording(model, lab, fr); torch.cuda.synchronize()
print("wall per frame ms", (time.perf_counter() - t0) / 48 * 1e3)
pr = cProfile.Profile(); pr.enable()
track_recording(model, lab, fr); torch.cuda.synchronize()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue())
