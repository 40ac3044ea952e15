"""One forward pass on two hand-frames (4 crops), plain launches: the subject of an ncu launch list."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from umetrack_b200 import hand_model as hmod, synthetic
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel
from tests.helpers import hm_dict

hm = hmod.load_hand_model_from_dict(hm_dict())
skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].cuda(), (hm.joint_rest_positions * 1e-3).float()[None].cuda())
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
model.set_mode(sys.argv[1] if len(sys.argv) > 1 else "fp32")
d = {k: v.cuda() for k, v in synthetic.make_model_batch(2, seed=1).items()}
fd = InputFrameData(d["left_images"], d["intrinsics"], d["extrinsics_xf"])
desc = InputFrameDesc(d["sample_range"], d["memory_idx"], d["use_memory"], d["hand_idx"], n_memory_slots=2)
for _ in range(3):
    model.regress_pose_use_skeleton(fd, desc, skel)
torch.cuda.synchronize()
