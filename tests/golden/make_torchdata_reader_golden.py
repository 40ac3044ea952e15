"""Fixtures for the torch_data reader.

    python tests/golden/make_torchdata_reader_golden.py      # build container only (needs /root/reference)

Writes a tiny dataset folder tests/golden/torchdata_seq/{mono,labels}.torch.{idx,bin} with
umetrack_b200.torch_data_reader.write_torch_idx_bin (the reference ships no writer), then parses
the files with the UNMODIFIED reference (lib/data_utils/idxbinfile.py:TorchIdx,
lib/batched_dataset/sample.py:parse_raw_buffers) and stores what IT read in
torchdata_reader_golden.npz.  The tests check that our reader returns the same from the same files.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("UMETRACK_REFERENCE", "/root/reference")
FOLDER = os.path.join(HERE, "torchdata_seq")


def make_samples(n=3, seed=11):
    """Seeded synthetic sequences in the labels schema of lib/batched_dataset/sample.py:16-40."""
    from umetrack_b200 import synthetic

    hm = synthetic.load_hand_model_json(os.path.join(HERE, "generic_hand_model_nomesh.json"))
    hm_lists = {k: v for k, v in hm.items() if isinstance(v, list)}
    g = np.random.default_rng(seed)
    monos, labels = [], []
    for i in range(n):
        seq, views, hh, ww = 2 + i, 2, 48, 64          # ragged sequence lengths
        monos.append((g.random((seq, views, hh, ww)) * 255).astype(np.uint8))
        intr = np.zeros((seq, views, 3, 3), dtype=np.float32)
        extr = np.zeros((seq, views, 4, 4), dtype=np.float32)
        pts = np.zeros((seq, 21, 3), dtype=np.float32)
        for f in range(seq):
            centre = np.array([g.uniform(-40, 40), g.uniform(-30, 30), g.uniform(280, 400)])
            pts[f] = centre + g.normal(0, 35, size=(21, 3))
            for v in range(views):
                intr[f, v] = [[75 + 5 * v, 0, (ww - 1) / 2], [0, 75 + 5 * v, (hh - 1) / 2], [0, 0, 1]]
                c2w = np.eye(4)
                c2w[:3, :3] = synthetic._rot_y(0.1 * (1 - 2 * v)) @ synthetic._rot_x(0.05 * f)
                c2w[:3, 3] = [40.0 * (2 * v - 1), 3.0 * f, 0.0]
                extr[f, v] = np.linalg.inv(c2w)
        wrist = np.tile(np.eye(4, dtype=np.float32), (seq, 1, 1))
        wrist[:, :3, 3] = g.uniform(-50, 50, size=(seq, 3))
        scaled = dict(hm_lists)
        scaled["joint_rest_positions"] = (np.asarray(hm_lists["joint_rest_positions"]) * 1.07).tolist()
        scaled["hand_scale"] = 1.07
        generic = dict(hm_lists)
        generic["hand_scale"] = 1.0
        labels.append({
            "extrinsics": extr.tolist(), "intrinsics": intr.tolist(), "enclosing_points": pts.tolist(),
            "hand": [float(i % 2)] * seq, "hand_model": scaled, "wrist": wrist.tolist(),
            "joint_angles": g.uniform(-0.3, 0.6, size=(seq, 22)).astype(np.float32).tolist(),
            "solved_wrist_xfs": wrist.tolist(),
            "solved_joint_angles": g.uniform(-0.3, 0.6, size=(seq, 22)).astype(np.float32).tolist(),
            "generic_hand_model": generic, "pinch": g.uniform(0, 1, size=(seq, 4)).astype(np.float32).tolist()})
    return monos, labels


def main():
    from umetrack_b200.torch_data_reader import write_torch_idx_bin

    monos, labels = make_samples()
    write_torch_idx_bin(os.path.join(FOLDER, "mono.torch.idx"), monos, dtype="uint8")
    write_torch_idx_bin(os.path.join(FOLDER, "labels.torch.idx"), labels)
    # a uniform float32 field as well (exercises `shape` / read_bin)
    write_torch_idx_bin(os.path.join(FOLDER, "uniform.torch.idx"),
                        [np.full((2, 3), i, dtype=np.float32) for i in range(4)], dtype="float32")

    sys.path.insert(0, REF)
    from lib.batched_dataset.sample import parse_raw_buffers
    from lib.data_utils.idxbinfile import TorchIdx

    store = {}
    mono_idx = TorchIdx(os.path.join(FOLDER, "mono.torch.idx"))
    lab_idx = TorchIdx(os.path.join(FOLDER, "labels.torch.idx"))
    uni_idx = TorchIdx(os.path.join(FOLDER, "uniform.torch.idx"))
    assert len(mono_idx) == len(lab_idx) == len(monos)
    store["n"] = np.int64(len(mono_idx))
    store["uniform_shape"] = np.array(uni_idx.shape)
    store["uniform_data"] = np.array(uni_idx.read_bin())
    store["mono_is_uniform"] = np.bool_(mono_idx.is_uniform)
    # (the reference's read_bin() only works for uniform fields; per-item views are its
    # view_buffer_at over the whole file's bytes)
    mono_buf = open(mono_idx.bin_path, "rb").read()
    lab_buf = open(lab_idx.bin_path, "rb").read()
    mono_items = [mono_idx.view_buffer_at(i, mono_buf) for i in range(len(mono_idx))]
    lab_items = [lab_idx.view_buffer_at(i, lab_buf) for i in range(len(lab_idx))]
    for i in range(len(mono_idx)):
        store[f"mono_{i}"] = np.array(mono_items[i])
        store[f"mono_dims_{i}"] = np.array(mono_idx.item_shape(i))
        store[f"mono_offset_{i}"] = np.int64(mono_idx.byte_offset(i))
        s = parse_raw_buffers(np.array(mono_items[i]), lab_items[i])
        for f in ("extrinsics", "intrinsics", "enclosing_points", "hand", "wrist", "joint_angles",
                  "solved_wrist_xfs", "solved_joint_angles", "pinch"):
            v = getattr(s, f)
            assert v.dtype == np.float32
            store[f"{f}_{i}"] = v.copy()
        store[f"hm_rest_{i}"] = s.hand_model.joint_rest_positions.numpy()
        store[f"hm_scale_{i}"] = np.float32(s.hand_model.hand_scale)
        s.scaled(0.001)
        store[f"scaled_extrinsics_{i}"] = s.extrinsics
        store[f"scaled_hm_rest_{i}"] = s.hand_model.joint_rest_positions.numpy()
    np.savez_compressed(os.path.join(HERE, "torchdata_reader_golden.npz"), **store)
    print("wrote", FOLDER, sorted(os.listdir(FOLDER)), "n =", len(mono_idx))


if __name__ == "__main__":
    main()
