"""Device time of the per-frame forward pass (CUDA graph replay, 1-4 crops) and of its parts.  GPU box."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from umetrack_b200 import hand_model as hmod, synthetic
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel
from tests.helpers import hm_dict

hm = hmod.load_hand_model_from_dict(hm_dict())
skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].cuda(), (hm.joint_rest_positions * 1e-3).float()[None].cuda())
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
mode = sys.argv[1] if len(sys.argv) > 1 else "fp32"
model.set_mode(mode)


def timed(fn, n=200):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for bs in (1, 2):
    b = synthetic.make_model_batch(bs, seed=1)
    d = {k: v.cuda() for k, v in b.items()}
    d["use_memory"][:] = True
    fd = InputFrameData(d["left_images"], d["intrinsics"], d["extrinsics_xf"])
    desc = InputFrameDesc(d["sample_range"], d["memory_idx"], d["use_memory"], d["hand_idx"], n_memory_slots=2,
                          persistent_inputs=True)
    with model.frame_graph_scope():
        t_graph = timed(lambda: model.regress_pose_use_skeleton(fd, desc, skel))
    t_plain = timed(lambda: model.regress_pose_use_skeleton(fd, desc, skel), 50)
    t_enc = timed(lambda: model.encode(d["left_images"]), 50)
    print(f"{mode} hands {bs} crops {len(d['left_images'])}: graph replay {t_graph:7.1f} us, plain launches {t_plain:7.1f} us, "
          f"encoder alone (plain) {t_enc:7.1f} us", flush=True)
