"""Stage (b) parity, part 2: full model outputs against vectors from the unmodified reference."""
import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import hand_model as hmod
from umetrack_b200 import synthetic
from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel, load_pretrained_model
from umetrack_b200.skinning import fk_landmarks_device
from tests.helpers import golden, hm_dict

pytestmark = pytest.mark.gpu
KEYPOINT_TOL_MM = 0.05


def _inputs(b, dev="cuda"):
    fd = InputFrameData(b["left_images"].to(dev), b["intrinsics"].to(dev), b["extrinsics_xf"].to(dev))
    desc = InputFrameDesc(b["sample_range"].to(dev), b["memory_idx"].to(dev), b["use_memory"].to(dev), b["hand_idx"].to(dev))
    return fd, desc


def _keypoints_mm(model_hm, ja, wr_m):
    wr = wr_m.clone().float().cuda()
    wr[:, :3, 3] *= 1000.0
    return fk_landmarks_device(model_hm, ja.float().cuda(), wr).cpu()


def test_known_skeleton_head_two_steps_mixed_views(tmp_path):
    g = golden("model_golden.npz")
    sd = synthetic.make_state_dict(0)
    path = tmp_path / "pretrained_weights.torch"
    torch.save(sd, str(path))
    model = load_pretrained_model(str(path))      # same file format as the reference
    model.eval()
    model.to("cuda")
    hm = hmod.load_hand_model_from_dict(hm_dict())
    skel = InputSkeletonData(hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None])
    for step, use_mem in enumerate((False, True)):
        b = synthetic.make_model_batch(6, seed=100 + step, single_view_every=3)
        b["use_memory"][:] = use_mem
        fd, desc = _inputs(b)
        out = model.regress_pose_use_skeleton(fd, desc, skel)
        ja, wr, sg = out.joint_angles.cpu(), out.wrist_xfs.cpu(), out.landmark_uncertainty_sigmas.cpu()
        assert out.skel_scales is None
        assert np.abs(ja.numpy() - g[f"k_joint_angles_{step}"]).max() < 2e-4
        assert np.abs(wr.numpy() - g[f"k_wrist_xfs_{step}"]).max() < 2e-5          # metres
        assert np.abs(sg.numpy() - g[f"k_sigmas_{step}"]).max() < 2e-4
        kp = _keypoints_mm(hm, ja, wr)
        kp_ref = _keypoints_mm(hm, torch.from_numpy(g[f"k_joint_angles_{step}"]), torch.from_numpy(g[f"k_wrist_xfs_{step}"]))
        assert float((kp - kp_ref).abs().max()) < KEYPOINT_TOL_MM
    assert np.abs(model._mem_features.cpu().numpy() - g["k_mem_after"]).max() < 2e-4


def test_unknown_skeleton_head_and_single_view_rejection():
    g = golden("model_golden.npz")
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    for step, use_mem in enumerate((False, True)):
        b = synthetic.make_model_batch(4, seed=200 + step)
        b["use_memory"][:] = use_mem
        out = model.regress_pose_pred_skel_scale(*_inputs(b))
        assert np.abs(out.skel_scales.cpu().numpy() - g[f"u_scales_{step}"]).max() < 2e-5
        assert np.abs(out.wrist_xfs.cpu().numpy() - g[f"u_wrist_xfs_{step}"]).max() < 2e-5
        assert np.abs(out.joint_angles.cpu().numpy() - g[f"u_joint_angles_{step}"]).max() < 2e-4
    b = synthetic.make_model_batch(3, seed=5, single_view_every=3)
    with pytest.raises(AssertionError):
        model.regress_pose_pred_skel_scale(*_inputs(b))


def test_large_batch_against_oracle_and_memory_slots():
    """64 hand-frames, permuted memory slots, three steps; compared with the oracle on CPU."""
    sd = synthetic.make_state_dict(0)
    model = UmeTrackModel(sd, chunk_images=64).to("cuda")   # 128 crops -> 2 encoder chunks
    orc = O.OracleModel(sd)
    hm = hmod.load_hand_model_from_dict(hm_dict())
    ax, rest = hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None]
    perm = torch.from_numpy(np.random.default_rng(1).permutation(64))
    for step in range(3):
        b = synthetic.make_model_batch(64, seed=300 + step)
        b["memory_idx"] = perm.clone()
        b["use_memory"][:] = step > 0
        if step == 2:
            b["use_memory"][::5] = False   # some hands lose tracking: their memory is zeroed
        out = model.regress_pose_use_skeleton(*_inputs(b), InputSkeletonData(ax, rest))
        ref = orc.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"], b["sample_range"],
                                            b["memory_idx"], b["use_memory"], b["hand_idx"], ax, rest)
        kp = _keypoints_mm(hm, out.joint_angles.cpu(), out.wrist_xfs.cpu())
        kp_ref = _keypoints_mm(hm, ref["joint_angles"], ref["wrist_xfs"])
        assert float((kp - kp_ref).abs().max()) < KEYPOINT_TOL_MM, f"step {step}"
    assert float((model._mem_features.cpu() - orc.state.mem).abs().max()) < 5e-4


def test_errors_are_loud():
    sd = synthetic.make_state_dict(0)
    bad = dict(sd)
    bad.pop("_regressor_u._pose_regression_layers.2.bias")
    with pytest.raises(RuntimeError, match="Missing key"):
        UmeTrackModel(bad)
    bad = dict(sd)
    bad["_temporal._temporal_module.0.weight"] = torch.zeros(90, 91, 1, 1)
    with pytest.raises(RuntimeError, match="size mismatch"):
        UmeTrackModel(bad)
    m = UmeTrackModel(sd)
    with pytest.raises(RuntimeError):
        m.to("cpu")
    with pytest.raises(RuntimeError):
        m.encode(torch.zeros(1, 96, 96))   # no .to('cuda') yet


@pytest.mark.parametrize("mode,tol_mm", [("fp16", KEYPOINT_TOL_MM)])
def test_sixteen_bit_encoder_mode_meets_the_keypoint_gate(mode, tol_mm):
    """BASELINE configs[3]: 16-bit encoder vs fp32 tolerance check.  fp16 operands (fp32 accumulate)
    must keep the keypoints inside the same 0.05 mm gate as the fp32-equivalent mode."""
    sd = synthetic.make_state_dict(0)
    model = UmeTrackModel(sd).to("cuda").set_mode(mode)
    orc = O.OracleModel(sd)
    hm = hmod.load_hand_model_from_dict(hm_dict())
    ax, rest = hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None]
    worst = 0.0
    for step in range(3):
        b = synthetic.make_model_batch(48, seed=400 + step, single_view_every=5)
        b["use_memory"][:] = step > 0
        out = model.regress_pose_use_skeleton(*_inputs(b), InputSkeletonData(ax, rest))
        ref = orc.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"], b["sample_range"],
                                            b["memory_idx"], b["use_memory"], b["hand_idx"], ax, rest)
        kp = _keypoints_mm(hm, out.joint_angles.cpu(), out.wrist_xfs.cpu())
        kp_ref = _keypoints_mm(hm, ref["joint_angles"], ref["wrist_xfs"])
        worst = max(worst, float((kp - kp_ref).abs().max()))
    assert worst < tol_mm, f"{mode}: keypoints differ from the oracle by {worst} mm"


def test_memory_bank_follows_the_index_values_not_the_tensor_identity():
    """[0] then [1] in a same-sized index tensor (the stale-cache case): the bank must hold both slots
    and slot 1's memory must be written and used, as the reference's regrowth does
    (lib/models/temporal.py:101-123).  Checked against the oracle."""
    sd = synthetic.make_state_dict(0)
    model = UmeTrackModel(sd).to("cuda")
    orc = O.OracleModel(sd)
    hm = hmod.load_hand_model_from_dict(hm_dict())
    ax, rest = hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None]
    idx_dev = torch.zeros(1, dtype=torch.int64, device="cuda")       # one tensor, mutated in place
    for step, (slot, use) in enumerate([(0, False), (0, True), (1, False), (1, True), (0, True)]):
        b = synthetic.make_model_batch(1, seed=500 + step)
        b["memory_idx"][:] = slot
        b["use_memory"][:] = use
        fd, desc = _inputs(b)
        idx_dev.fill_(slot)
        desc.memory_idx = idx_dev
        out = model.regress_pose_use_skeleton(fd, desc, InputSkeletonData(ax, rest))
        ref = orc.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"], b["sample_range"],
                                            b["memory_idx"], b["use_memory"], b["hand_idx"], ax, rest)
        kp = _keypoints_mm(hm, out.joint_angles.cpu(), out.wrist_xfs.cpu())
        kp_ref = _keypoints_mm(hm, ref["joint_angles"], ref["wrist_xfs"])
        assert float((kp - kp_ref).abs().max()) < KEYPOINT_TOL_MM, f"step {step}"
    assert len(model._mem_features) == 2
    assert float((model._mem_features.cpu() - orc.state.mem).abs().max()) < 5e-4
    model.check_status()


def test_out_of_range_memory_index_is_reported():
    from umetrack_b200 import native
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    hm = hmod.load_hand_model_from_dict(hm_dict())
    skel = InputSkeletonData(hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None])
    b = synthetic.make_model_batch(2, seed=9)
    fd, desc = _inputs(b)
    desc.n_memory_slots = 2
    model.regress_pose_use_skeleton(fd, desc, skel)
    model.check_status()                       # indices 0, 1 fit two slots
    desc.memory_idx = torch.tensor([0, 2], device="cuda")
    model.regress_pose_use_skeleton(fd, desc, skel)      # broken promise: index 2 with two slots
    with pytest.raises(native.NativeError, match="memory_idx"):
        model.check_status()
    model.check_status()                       # the flag is cleared by the read


def test_uint8_crops_give_the_same_result_as_their_float_conversion():
    """`crop.astype(np.float32) / 255.0` (lib/tracker/tracker.py:211) on the device: identical values,
    hence bit-identical outputs, for ragged counts over two chunks."""
    model = UmeTrackModel(synthetic.make_state_dict(0), chunk_images=16).to("cuda")
    hm = hmod.load_hand_model_from_dict(hm_dict())
    skel = InputSkeletonData(hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None])
    b = synthetic.make_model_batch(11, seed=77, single_view_every=4)
    u8 = (b["left_images"] * 255).round().to(torch.uint8)
    as_float = torch.from_numpy(u8.numpy().astype(np.float32) / 255.0)
    assert torch.equal(model.encode(u8.cuda()), model.encode(as_float.cuda()))
    b["use_memory"][:] = False
    outs = []
    for imgs in (u8, as_float):
        b["left_images"] = imgs
        fd, desc = _inputs(b)
        outs.append(model.regress_pose_use_skeleton(fd, desc, skel))
    assert torch.equal(outs[0].joint_angles, outs[1].joint_angles) and torch.equal(outs[0].wrist_xfs, outs[1].wrist_xfs)


def test_fp16_mode_refuses_weights_and_reports_activations_outside_the_half_range():
    """UMT_MODE_FP16 keeps activations and BN-folded weights in IEEE half: a weight beyond 65504 is
    refused when the mode is used, an activation beyond it is reported through the status word (the
    fp32-equivalent and bf16 modes are unaffected)."""
    from umetrack_b200 import native
    imgs = synthetic.make_model_batch(2, seed=3)["left_images"].cuda()
    sd = synthetic.make_state_dict(0)
    big_w = dict(sd)
    key = "_feature_extractor._image_backbone.0._layers.1.0.conv1.weight"
    big_w[key] = sd[key] * 1e7
    m = UmeTrackModel(big_w).to("cuda")
    m.encode(imgs)                                   # fp32-equivalent mode still runs
    with pytest.raises(native.NativeError, match="half range"):
        m.set_mode("fp16").encode(imgs)
    big_a = dict(sd)
    key = "_feature_extractor._image_backbone.0._layers.0.1.bias"
    big_a[key] = sd[key] + 1e5                       # BatchNorm shift: weights stay small, the stem's outputs do not
    m = UmeTrackModel(big_a).to("cuda").set_mode("fp16")
    m.encode(imgs)
    with pytest.raises(native.NativeError, match="half range"):
        m.check_status()
    m.set_mode("bf16").encode(imgs)
    m.check_status()                                 # bf16 has the float32 exponent range
    ok = UmeTrackModel(sd).to("cuda").set_mode("fp16")
    ok.encode(imgs)
    ok.check_status()


def test_fused_small_batch_tail_and_tensor_core_tail_agree(monkeypatch):
    """Batches up to 128 samples run the tail as one CTA per sample (csrc/tail_fused.cu, float32 FMAs);
    larger ones through the tensor-core convolutions of csrc/tail.cu.  Same inputs, same recurrent state:
    both within the keypoint gate of the oracle and within 1e-3 mm of each other, over three steps with
    mixed 1/2-view samples and memory resets."""
    sd = synthetic.make_state_dict(0)
    hm = hmod.load_hand_model_from_dict(hm_dict())
    ax, rest = hm.joint_rotation_axes.float()[None], (hm.joint_rest_positions * 1e-3).float()[None]
    results = {}
    for name, limit, cluster in (("cluster", "128", "4096"), ("fused", "128", "0"), ("tensor", "0", "0")):
        monkeypatch.setenv("UMT_TAIL_FUSED_MAX", limit)
        monkeypatch.setenv("UMT_TAIL_CLUSTER_MAX", cluster)     # a sample per CTA cluster / per CTA
        model = UmeTrackModel(sd).to("cuda")
        orc = O.OracleModel(sd)
        kps, launches = [], []
        for step in range(3):
            b = synthetic.make_model_batch(24, seed=600 + step, single_view_every=4)
            b["use_memory"][:] = step > 0
            if step == 2:
                b["use_memory"][::3] = False
            out = model.regress_pose_use_skeleton(*_inputs(b), InputSkeletonData(ax, rest))
            launches.append(model.last_launch_count)
            ref = orc.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"], b["sample_range"],
                                                b["memory_idx"], b["use_memory"], b["hand_idx"], ax, rest)
            kp = _keypoints_mm(hm, out.joint_angles.cpu(), out.wrist_xfs.cpu())
            kp_ref = _keypoints_mm(hm, ref["joint_angles"], ref["wrist_xfs"])
            assert float((kp - kp_ref).abs().max()) < KEYPOINT_TOL_MM, f"{name} step {step}"
            kps.append(kp)
        assert float((model._mem_features.cpu() - orc.state.mem).abs().max()) < 5e-4
        results[name] = (torch.stack(kps), launches)
        # unknown-skeleton head on the same path
        g = golden("model_golden.npz")
        model2 = UmeTrackModel(sd).to("cuda")
        for step, use_mem in enumerate((False, True)):
            b = synthetic.make_model_batch(4, seed=200 + step)
            b["use_memory"][:] = use_mem
            out = model2.regress_pose_pred_skel_scale(*_inputs(b))
            assert np.abs(out.skel_scales.cpu().numpy() - g[f"u_scales_{step}"]).max() < 2e-5
            assert np.abs(out.wrist_xfs.cpu().numpy() - g[f"u_wrist_xfs_{step}"]).max() < 2e-5
    assert float((results["fused"][0] - results["tensor"][0]).abs().max()) < 1e-3
    # the cluster variant deals the same per-element FMA chains to eight CTAs: bit-identical keypoints
    assert torch.equal(results["cluster"][0], results["fused"][0])
    assert results["cluster"][1] == results["fused"][1]
    assert results["fused"][1][0] == 26 + 2 and results["tensor"][1][0] == 26 + 15     # launches per forward pass
