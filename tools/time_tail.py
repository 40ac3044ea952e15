"""Device time of one tail step (fusion + temporal + regressor + head) on small batches, fused
one-CTA-per-sample kernel and its cluster variant against the tensor-core tail (GPU box):  python tools/time_tail.py [bs ...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from umetrack_b200 import hand_model as hmod, synthetic
from umetrack_b200.model import InputFrameDesc, InputSkeletonData, UmeTrackModel
from tests.helpers import hm_dict

hm = hmod.load_hand_model_from_dict(hm_dict())
skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].cuda(), (hm.joint_rest_positions * 1e-3).float()[None].cuda())
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
for bs in [int(a) for a in sys.argv[1:]] or [2, 16, 64, 256]:
    b = synthetic.make_model_batch(bs, seed=1)
    d = {k: v.cuda() for k, v in b.items()}
    feats = model.encode(d["left_images"])
    desc = InputFrameDesc(d["sample_range"], d["memory_idx"], d["use_memory"], d["hand_idx"], n_memory_slots=bs)
    n_steps = 200
    ax = skel.joint_rotation_axes.expand(bs, 22, 3).contiguous()
    rs = skel.joint_rest_positions.expand(bs, 22, 3).contiguous()
    for name, limit, cl in (("tensor ", "0", "0"), ("fused  ", "4096", "0"), ("cluster", "4096", "4096")):
        os.environ["UMT_TAIL_FUSED_MAX"] = limit
        os.environ["UMT_TAIL_CLUSTER_MAX"] = cl

        def scan():   # one C call stepping the same rows n_steps times: no Python between the steps
            model.tail_scan(feats, d["intrinsics"], d["extrinsics_xf"], [(0, bs)] * n_steps, d["sample_range"],
                            d["memory_idx"], d["use_memory"], d["hand_idx"], ax, rs, n_slots=bs)
        scan()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        scan()
        e1.record(); torch.cuda.synchronize()
        print(f"bs {bs:4d} {name}: {e0.elapsed_time(e1) / n_steps * 1e3:8.1f} us per step", flush=True)
