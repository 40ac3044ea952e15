"""Where does the end-to-end step lose time against the device-resident step?  Needs a B200.

Times, interleaved and repeated so that clock drift under the power cap hits all variants alike:
  f32   float32 crops resident in HBM (bench.py's `value` leg)
  u8    uint8 crops resident in HBM (conversion on the device, no copies)
  u8+c  uint8 crops, host->device copy of the next step on a copy stream, keypoints copied back
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from umetrack_b200 import synthetic
from umetrack_b200 import hand_model as hmod
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel
from umetrack_b200.skinning import fk_landmarks_device

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
mode = sys.argv[1] if len(sys.argv) > 1 else "fp32"
B, K = 2048, 10
dev = torch.device("cuda", 0)
hm = hmod.load_hand_model_from_dict(
    synthetic.load_hand_model_json(os.path.join(ROOT, "tests", "golden", "generic_hand_model_nomesh.json")))
model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
model.set_mode(mode)
host = synthetic.make_model_batch(B, seed=1234)
u8 = (host["left_images"] * 255.0).round().to(torch.uint8)
host["left_images"] = u8.float() / 255.0
host["use_memory"][:] = True
d = {k: v.to(dev) for k, v in host.items()}
d_u8 = u8.to(dev)
pinned = {k: (u8 if k == "left_images" else v).pin_memory() for k, v in host.items()}
skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].to(dev),
                         (hm.joint_rest_positions * 1e-3).float()[None].to(dev))
mirror = (d["hand_idx"] == 1).to(torch.uint8)


def step(images, dd):
    out = model.regress_pose_use_skeleton(
        InputFrameData(images, dd["intrinsics"], dd["extrinsics_xf"]),
        InputFrameDesc(dd["sample_range"], dd["memory_idx"], dd["use_memory"], dd["hand_idx"], n_memory_slots=B), skel)
    wr = out.wrist_xfs.clone()
    wr[:, :3, 3] *= 1000.0
    return fk_landmarks_device(hm, out.joint_angles, wr, mirror)


out_host = torch.empty((B, 21, 3), dtype=torch.float32).pin_memory()
copy_stream = torch.cuda.Stream(device=dev)
slots = [{k: torch.empty_like(v, device=dev) for k, v in pinned.items()} for _ in range(2)]
ready = [torch.cuda.Event(), torch.cuda.Event()]
consumed = [torch.cuda.Event(), torch.cuda.Event()]
state = {"i": 0}


def stage(slot):
    with torch.cuda.stream(copy_stream):
        copy_stream.wait_event(consumed[slot])
        for k, v in pinned.items():
            slots[slot][k].copy_(v, non_blocking=True)
        ready[slot].record(copy_stream)


def e2e(copy_in=True, copy_out=True):
    slot = state["i"] & 1
    state["i"] += 1
    if copy_in:
        stage(slot ^ 1)
        torch.cuda.current_stream().wait_event(ready[slot])
    dd = slots[slot]
    kp = step(dd["left_images"], dd)
    consumed[slot].record()
    if copy_out:
        out_host.copy_(kp, non_blocking=True)


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K


for ev in consumed:
    ev.record()
stage(0); stage(1)
torch.cuda.synchronize()
variants = {
    "f32": lambda: step(d["left_images"], d),
    "u8": lambda: step(d_u8, d),
    "u8 slots, no copies": lambda: e2e(False, False),
    "u8 + d2h": lambda: e2e(False, True),
    "u8 + h2d": lambda: e2e(True, False),
    "u8 + h2d + d2h": lambda: e2e(True, True),
}
for fn in variants.values():
    for _ in range(3):
        fn()
res = {k: [] for k in variants}
for rep in range(4):
    for k, fn in variants.items():
        res[k].append(timed(fn))
for k, v in res.items():
    print(f"{mode} {k:22s} ms/step " + " ".join(f"{x:.3f}" for x in v) + f"   min {min(v):.3f}")
