"""Small end-to-end pass for compute-sanitizer (memcheck): every kernel of the library once, tiny sizes."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from umetrack_b200 import hand_model as hmod  # noqa: E402
from umetrack_b200 import synthetic  # noqa: E402
from umetrack_b200.batched import track_recordings_batched  # noqa: E402
from umetrack_b200.metrics import MetricSums  # noqa: E402
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel  # noqa: E402
from umetrack_b200.torch_data import resample_images_device  # noqa: E402

hm_dict = synthetic.load_hand_model_json(os.path.join(ROOT, "tests", "golden", "generic_hand_model_nomesh.json"))
hm = hmod.load_hand_model_from_dict(hm_dict)
model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
for mode in ("fp32", "fp16", "bf16"):
    model.set_mode(mode)
    b = synthetic.make_model_batch(3, seed=1, single_view_every=3)
    out = model.regress_pose_use_skeleton(
        InputFrameData(b["left_images"].cuda(), b["intrinsics"].cuda(), b["extrinsics_xf"].cuda()),
        InputFrameDesc(b["sample_range"].cuda(), b["memory_idx"].cuda(), b["use_memory"].cuda(), b["hand_idx"].cuda()),
        InputSkeletonData(hm.joint_rotation_axes.float()[None].cuda(), (hm.joint_rest_positions * 1e-3).float()[None].cuda()))
    torch.cuda.synchronize()
    print(mode, "forward ok", float(out.joint_angles.abs().max()))
model.set_mode("fp32")
b = synthetic.make_model_batch(2, seed=2)
out = model.regress_pose_pred_skel_scale(
    InputFrameData(b["left_images"].cuda(), b["intrinsics"].cuda(), b["extrinsics_xf"].cuda()),
    InputFrameDesc(b["sample_range"].cuda(), b["memory_idx"].cuda(), b["use_memory"].cuda(), b["hand_idx"].cuda()))
rec = synthetic.make_recording(hm_dict, seed=3, n_frames=3)
sums = MetricSums("cuda")
res = track_recordings_batched(model, [rec], sums=sums)
print("recording ok", sums.result())
x = resample_images_device(torch.rand(2, 40, 60).cuda() * 255, torch.eye(4).repeat(2, 1, 1).cuda())
torch.cuda.synchronize()
print("all ok")
