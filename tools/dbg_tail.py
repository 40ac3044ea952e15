import os, sys
sys.path.insert(0, "/root/repo")
import torch
from umetrack_b200 import hand_model as hmod, synthetic
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel
from tests.helpers import hm_dict
hm = hmod.load_hand_model_from_dict(hm_dict())
skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].cuda(), (hm.joint_rest_positions * 1e-3).float()[None].cuda())
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
b = synthetic.make_model_batch(4, seed=1); d = {k: v.cuda() for k, v in b.items()}
feats = model.encode(d["left_images"])
desc = InputFrameDesc(d["sample_range"], d["memory_idx"], d["use_memory"], d["hand_idx"], n_memory_slots=4)
for i in range(3):
    if i == 2: os.environ["UMT_TAIL_DBG"] = "1"
    model.regress_pose_from_features(feats, d["intrinsics"], d["extrinsics_xf"], desc, skel)
    torch.cuda.synchronize()
