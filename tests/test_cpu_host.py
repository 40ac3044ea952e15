"""CPU-side checks: the C ABI exports what the header declares, host geometry code agrees with
the reference's golden crop cameras, sharding and the gloo reduction of metric sums."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import arch, native, synthetic
from umetrack_b200.cameras import Fisheye62CameraModel, PinholePlaneCameraModel, read_camera_from_json, resample_params
from umetrack_b200.crop_camera import gen_crop_parameters_from_points
from umetrack_b200.metrics import MetricSums, shard_indices
from tests.helpers import golden, hm_dict, hm_tensors

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "umetrack_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(umt_[a-z0-9_]+)\s*\(", header))
    assert declared == set(native.SIGNATURES), declared ^ set(native.SIGNATURES)
    lib = native.lib()   # loading needs no GPU
    for name in declared:
        assert hasattr(lib, name)
    assert lib.umt_abi_version() == 1


def test_weight_registry_matches_state_dict_spec():
    lib = native.lib()
    names = [lib.umt_model_tensor_name(i).decode() for i in range(lib.umt_model_num_tensors())]
    spec = [(k, s) for k, s, kind in arch.state_dict_spec() if kind != "bn_count"]
    assert names == [k for k, _ in spec]
    assert [lib.umt_model_tensor_numel(i) for i in range(len(names))] == [int(np.prod(s)) for _, s in spec]
    assert len(arch.state_dict_spec()) == 252
    assert sum(int(np.prod(s)) for _, s, kind in arch.state_dict_spec() if kind != "bn_count") == 4_259_371


def test_argument_validation_without_gpu():
    lib = native.lib()
    assert lib.umt_model_create(None, 3, ctypes.byref(ctypes.c_void_p())) != 0
    assert b"expected 213 tensors" in lib.umt_last_error()
    assert lib.umt_fk_landmarks(0, 0, 0, 0, 0, 0, 0, 0, -1, 0, 0) != 0
    assert lib.umt_fk_landmarks(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0) == 0      # empty batch is fine
    assert lib.umt_mpjpe_accumulate(0, 0, 0, 0, 0, 0) == 0


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "umetrack_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("no oracle", ""), fn


def test_cameras_match_oracle_projection():
    labels, _ = synthetic.make_recording(hm_dict(), seed=3, n_frames=2)
    cam = read_camera_from_json(labels["cameras"][1]).copy(np.array(labels["camera_to_world_transforms"][0][1]))
    assert isinstance(cam, Fisheye62CameraModel) and cam.width == 640
    g = np.random.default_rng(0)
    pts = g.normal(0, 100, size=(50, 3)) + [0, 0, 350]
    eye = cam.world_to_eye(pts)
    win = cam.eye_to_window(eye)
    c = labels["cameras"][1]
    ref = O.fisheye62_distort(O.arctan_project(eye), [c[k] for k in ("k1", "k2", "k3", "k4", "p1", "p2", "k5", "k6")])
    ref = ref * [c["fx"], c["fy"]] + [c["cx"], c["cy"]]
    np.testing.assert_allclose(win, ref, rtol=0, atol=1e-10)
    np.testing.assert_allclose(cam.eye_to_world(eye), pts, atol=1e-9)
    pin = PinholePlaneCameraModel(96, 96, (100.0, 101.0), (47.5, 47.5))
    w = np.array([[3.0, 90.0], [47.5, 47.5]])
    np.testing.assert_allclose(pin.eye_to_window(pin.window_to_eye(w)), w, atol=1e-10)
    with pytest.raises(AssertionError):
        cam.window_to_eye(w)
    assert resample_params(cam, pin).shape == (48,)


def test_crop_cameras_match_reference_golden():
    """gen_crop_parameters_from_points on the reference's crop points reproduces its cameras."""
    g = golden("tracker_golden.npz")
    hm = hm_dict()
    t = hm_tensors(hm)
    labels, _ = synthetic.make_recording(hm, seed=3, n_frames=12)
    lim = t["joint_limits"]
    neutral = (lim[:, 0] * 0.5 + lim[:, 1] * 0.5).numpy()
    checked = 0
    for row in g["crop_rows"][:12]:
        ti, h, c = int(row[0]), int(row[1]), int(row[2])
        ja = np.array(labels["joint_angles"][ti][h])
        wr = np.array(labels["wrist_transforms"][ti][h])
        pts = np.concatenate([O.landmarks_from_hand_pose(t, a, wr, h) for a in (ja, neutral, np.zeros(22, np.float32))])
        cam = read_camera_from_json(labels["cameras"][c]).copy(np.array(labels["camera_to_world_transforms"][ti][c]))
        crop = gen_crop_parameters_from_points(cam, pts, (96, 96), mirror_img_x=(h == 1),
                                               camera_angle=labels["camera_angles"][c], focal_multiplier=0.95)
        np.testing.assert_allclose(np.concatenate([crop.f, crop.c]), row[3:7], rtol=1e-9)
        np.testing.assert_allclose(crop.camera_to_world_xf.reshape(-1), row[7:], atol=1e-9)
        checked += 1
    assert checked == 12
    with pytest.raises(ValueError, match="Unable to create crop camera"):
        wide = cam.eye_to_world(np.array([[-1000.0, 0, 1.0], [1000.0, 0, 1.0], [0, 0, 1.0]]))
        gen_crop_parameters_from_points(cam, wide, (96, 96), False)   # focal would be < 5 px


def test_shard_indices_partition():
    for n, w in [(0, 2), (5, 2), (64, 8), (7, 4)]:
        parts = [shard_indices(n, r, w) for r in range(w)]
        assert sorted(sum(parts, [])) == list(range(n))
    costs = [100, 1, 1, 1, 50, 49, 2, 2]
    parts = [shard_indices(len(costs), r, 2, costs) for r in range(2)]
    assert sorted(sum(parts, [])) == list(range(len(costs)))
    loads = [sum(costs[i] for i in p) for p in parts]
    assert abs(loads[0] - loads[1]) <= 6


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from umetrack_b200.metrics import MetricSums, shard_indices
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
errs = [float(i) * 0.5 for i in range(11)]           # per-recording error sums, 3 hand-frames each
mine = shard_indices(len(errs), rank, world)
s = MetricSums("cpu")
for i in mine:
    s.sums[0] += errs[i]; s.sums[1] += 3
s.all_reduce()
r = s.result()
assert abs(r["mpjpe"] - sum(errs) / (3 * len(errs))) < 1e-12, r
assert r["n_hand_frames"] == 3 * len(errs)
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_metric_sums_all_reduce_gloo_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29641", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=120)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)


def test_torch_data_reader_matches_reference_parser(tmp_path):
    """the committed idx/bin files, as parsed by the reference when the goldens
    were made (tests/golden/make_torchdata_reader_golden.py), vs our reader; plus error cases."""
    from umetrack_b200.torch_data_reader import (MAGIC, SequenceFolder, TorchIdx, find_sequence_folders,
                                                  parse_raw_buffers, write_torch_idx_bin)

    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = np.load(os.path.join(gdir, "torchdata_reader_golden.npz"))
    folder = os.path.join(gdir, "torchdata_seq")
    assert find_sequence_folders(gdir) == [folder]
    ds = SequenceFolder(folder)
    assert len(ds) == int(g["n"])
    mono_idx = TorchIdx(os.path.join(folder, "mono.torch.idx"))
    assert mono_idx.is_uniform == bool(g["mono_is_uniform"]) and mono_idx.shape is None
    with pytest.raises(ValueError):
        mono_idx.item_shape()
    for i in range(len(ds)):
        item = ds[i]
        assert item["mono"].dtype == np.uint8 and np.array_equal(item["mono"], g[f"mono_{i}"])
        assert tuple(g[f"mono_dims_{i}"]) == mono_idx.item_shape(i)
        assert int(g[f"mono_offset_{i}"]) == mono_idx.byte_offset(i)
        s = parse_raw_buffers(**item)
        for f in ("extrinsics", "intrinsics", "enclosing_points", "hand", "wrist", "joint_angles",
                  "solved_wrist_xfs", "solved_joint_angles", "pinch"):
            v = getattr(s, f)
            assert v.dtype == np.float32 and np.array_equal(v, g[f"{f}_{i}"]), f
        assert np.array_equal(s.hand_model.joint_rest_positions.numpy(), g[f"hm_rest_{i}"])
        assert float(s.hand_model.hand_scale) == float(g[f"hm_scale_{i}"])
        s.scaled(0.001)
        assert np.array_equal(s.extrinsics, g[f"scaled_extrinsics_{i}"])
        assert np.array_equal(s.hand_model.joint_rest_positions.numpy(), g[f"scaled_hm_rest_{i}"])
    uni = TorchIdx(os.path.join(folder, "uniform.torch.idx"))
    assert uni.shape == tuple(g["uniform_shape"]) and np.array_equal(uni.read_bin(), g["uniform_data"])
    assert uni.item_shape() == (2, 3) and uni.data_size_bytes() == 4 * 6 * 4
    # round trip + the reference's error behaviour
    p = str(tmp_path / "x.torch.idx")
    write_torch_idx_bin(p, [np.arange(6, dtype=np.int16).reshape(2, 3), np.arange(4, dtype=np.int16)], dtype="int16")
    idx = TorchIdx(p)
    assert not idx.is_uniform and np.array_equal(idx.read_item(1), np.arange(4))
    with pytest.raises(IndexError):
        idx.read_item(2)
    raw = np.fromfile(p, dtype=np.int64)
    for pos, val, exc in ((0, 123, ValueError), (1, 7, ValueError), (2, 99, KeyError), (3, 4, ValueError)):
        bad = raw.copy()
        bad[pos] = val
        with pytest.raises(exc):
            TorchIdx(p, buffer=bad.tobytes())
    with pytest.raises(ValueError):
        TorchIdx(p, buffer=raw.tobytes()[:-3])
    with pytest.raises(ValueError):
        idx.view_frame(0, b"123")
    assert raw[0] == MAGIC and raw[1] == 1
    with pytest.raises(ValueError):
        SequenceFolder(str(tmp_path), fields=("mono",))                       # missing field
    write_torch_idx_bin(str(tmp_path / "a.torch.idx"), [np.zeros(2, dtype=np.uint8)], dtype="uint8")
    with pytest.raises(ValueError):
        SequenceFolder(str(tmp_path), fields=("x", "a"))                      # 2 items vs 1
    with pytest.raises(TypeError):
        parse_raw_buffers(np.zeros(1), {"unknown_field": [1.0]})


def test_torch_idx_round_trip_fuzz(tmp_path):
    """Property test: whatever write_torch_idx_bin writes, TorchIdx reads back item by item and
    (for uniform shapes) as one array -- ragged ranks, empty items and every tensor dtype included."""
    from hypothesis import HealthCheck, given, settings, strategies as st

    from umetrack_b200.torch_data_reader import TorchIdx, write_torch_idx_bin

    dtypes = ["uint8", "int8", "int16", "int32", "int64", "float32", "float64"]
    shape = st.lists(st.integers(0, 4), min_size=0, max_size=3).map(tuple)
    counter = [0]

    @settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
    @given(dt=st.sampled_from(dtypes), shapes=st.lists(shape, min_size=1, max_size=5), seed=st.integers(0, 2**31 - 1),
           uniform=st.booleans())
    def run(dt, shapes, seed, uniform):
        if uniform:
            shapes = [shapes[0]] * len(shapes)
        rng = np.random.default_rng(seed)
        items = [(rng.integers(-100, 100, size=s) if "int" in dt else rng.normal(size=s)).astype(dt) for s in shapes]
        if dt == "uint8":
            items = [np.abs(a).astype(dt) for a in items]
        counter[0] += 1
        p = str(tmp_path / f"f{counter[0]}.torch.idx")
        write_torch_idx_bin(p, items, dtype=dt)
        idx = TorchIdx(p)
        assert len(idx) == len(items) and idx.dtype == np.dtype(dt)
        for i, a in enumerate(items):
            got = idx.read_item(i)
            assert got.shape == a.shape and got.dtype == a.dtype and np.array_equal(got, a)
        whole = idx.read_bin()
        if idx.is_uniform:
            assert np.array_equal(whole, np.stack(items))
        else:
            assert all(np.array_equal(w, a) for w, a in zip(whole, items))
        assert idx.data_size_bytes() == sum(a.nbytes for a in items)

    run()
    # msgpack items
    recs = [{"a": [1.5, 2.5], "b": {"c": "x"}}, {}, {"k": list(range(50))}]
    p = str(tmp_path / "m.torch.idx")
    write_torch_idx_bin(p, recs)
    idx = TorchIdx(p)
    assert idx.is_msgpack and not idx.is_uniform and [idx.read_item(i) for i in range(3)] == recs


def test_nvtx_ranges_are_off_by_default_and_cost_nothing():
    from umetrack_b200 import nvtx
    assert not nvtx.enabled()
    with nvtx.stage("anything"):
        pass
    nvtx.push("anything")
    nvtx.pop()
    assert nvtx.stage("a") is nvtx.stage("b")          # the shared null context, no allocation per call
