"""Reader for the reference's torch_data sequence files.

A dataset folder holds one `<field>.torch.idx` / `<field>.torch.bin` pair per field
(run_inference_torch_data.py:143-144 asks for "mono" and "labels"): the .idx file is an int64
array describing how the .bin file splits into items (lib/data_utils/idxbinfile.py:115-130)

    [0] magic 0x584449544E54 ("TNTIDX")   [1] version 1 (version 0: magic 0)
    [2] type code: 1 uint8, 2 int8, 3 int16, 4 int32, 5 int64, 6 float32, 7 float64, 8 msgpack object
    [3] itemsize    [4] N items    [5] S = total number of dims
    then N+1 dim offsets (into `sizes`), N+1 data offsets (units of itemsize), S sizes

"mono" items are uint8 image stacks [seq, n_views, H, W]; "labels" items are msgpack dicts
(lib/batched_dataset/sample.py:42-53).  `parse_raw_buffers` types them like the reference; the
crops are then made on the GPU by `torch_data.perspective_crop_images` (umt_resample_homography)
in `prepare_model_inputs`.  The reference ships no writer; `write_torch_idx_bin` exists for tests
and tooling and produces files the reference's `TorchIdx` parses (checked when the goldens are
generated, tests/golden/make_torchdata_reader_golden.py).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

MAGIC = 0x584449544E54
_CODE_TO_DTYPE = {1: "uint8", 2: "int8", 3: "int16", 4: "int32", 5: "int64", 6: "float32", 7: "float64", 8: "object"}
_DTYPE_TO_CODE = {v: k for k, v in _CODE_TO_DTYPE.items()}
_MSGPACK_CODE = 8


def _bin_path_for_idx(idx_path: str) -> str:
    suffix = ".torch.idx"
    assert idx_path.endswith(suffix)                       # idxbinfile.py:41
    return idx_path[: -len(suffix)] + ".torch.bin"


class TorchIdx:
    """Parsed `.torch.idx` file (lib/data_utils/idxbinfile.py:48-186): `len()` items, `dims[i]`,
    `dtype`, `itemsize`, `is_uniform`/`shape`, `byte_offset(i)`, and `read_item(i)` /
    `read_bin()` on the matching `.torch.bin`.  Raises the reference's errors: ValueError for a
    bad magic number, an unsupported version, a byte length that is not a multiple of 8 or an
    itemsize that contradicts the dtype; KeyError for an unknown type code."""

    def __init__(self, path: str, bin_path: Optional[str] = None, buffer: Optional[bytes] = None):
        if bin_path is None:
            bin_path = _bin_path_for_idx(path)
        if buffer is None:
            with open(path, "rb") as fp:
                buffer = fp.read()
        self.source, self.bin_path = path, bin_path
        if len(buffer) % 8 != 0:
            raise ValueError(f".torch.idx data has invalid length: {len(buffer)}%8 != 0")
        data = np.frombuffer(buffer, dtype=np.int64)
        if data[1] == 0:
            if data[0] != 0:
                raise ValueError(f"bad magic number in .torch.idx file {path}")
        elif data[1] == 1:
            if data[0] != MAGIC:
                raise ValueError(f"bad magic number in .torch.idx file {path}")
        else:
            raise ValueError(f"unsupported version {data[1]} in .torch.idx file {path}")
        code = int(data[2])
        self.itemsize = int(data[3])
        if code not in _CODE_TO_DTYPE:
            raise KeyError(f"unrecognized type {code}")
        self.is_msgpack = code == _MSGPACK_CODE
        self.dtype = np.dtype(_CODE_TO_DTYPE[code])
        if not self.is_msgpack and self.dtype.itemsize != self.itemsize:
            raise ValueError(f"item size {self.itemsize} not compatible with "
                             f"dtype {self.dtype.name}.itemsize={self.dtype.itemsize}")
        n, s = int(data[4]), int(data[5])
        ofs = 6
        dim_offsets = data[ofs: ofs + n + 1]
        ofs += n + 1
        self._data_offsets = data[ofs: ofs + n + 1].copy()
        ofs += n + 1
        sizes = data[ofs: ofs + s]
        self.dims: List[Tuple[int, ...]] = [tuple(int(v) for v in sizes[dim_offsets[i]: dim_offsets[i + 1]])
                                            for i in range(n)]
        self.is_uniform = (not self.is_msgpack and n > 0 and all(d == self.dims[0] for d in self.dims))
        self.shape = (n, *self.dims[0]) if self.is_uniform else None

    def __len__(self) -> int:
        return len(self.dims)

    def __repr__(self) -> str:
        if self.is_uniform:
            return f"TorchIdx(shape={self.shape}, dtype={self.dtype}, source={self.source})"
        return f"TorchIdx(N={len(self)}, dtype={self.dtype}, is_uniform=False, source={self.source})"

    def byte_offset(self, i: int) -> int:
        """Byte offset of item i in the .bin file; i == -1 or len() gives the end."""
        if i == -1:
            i = len(self)
        return int(self._data_offsets[i]) * self.itemsize

    def data_size_bytes(self) -> int:
        return self.byte_offset(-1) - self.byte_offset(0)

    def item_shape(self, i: Optional[int] = None) -> Tuple[int, ...]:
        if i is None:
            if self.shape is None:
                raise ValueError(f"Dataset does not have uniform shape: {self.source}")
            return self.shape[1:]
        return self.dims[i]

    def view_frame(self, index: int, buffer) -> Union[np.ndarray, Dict[str, Any]]:
        """One item from exactly its bytes (idxbinfile.py:271-288)."""
        want = self.byte_offset(index + 1) - self.byte_offset(index)
        got = memoryview(buffer).nbytes
        if want != got:
            raise ValueError(f"expected {want} bytes but got {got} for {self.bin_path}[{index}:{index + 1}]")
        if self.is_msgpack:
            import msgpack

            return msgpack.unpackb(buffer)
        return np.ndarray(shape=self.dims[index], dtype=self.dtype, buffer=buffer)

    def read_item(self, index: int) -> Union[np.ndarray, Dict[str, Any]]:
        if not 0 <= index < len(self):
            raise IndexError(index)
        s0, s1 = self.byte_offset(index), self.byte_offset(index + 1)
        with open(self.bin_path, "rb") as fp:
            fp.seek(s0)
            buf = fp.read(s1 - s0)
        return self.view_frame(index, buf)

    def read_bin(self):
        """The whole .bin file: one ndarray [N, ...] if uniform, else a list of items."""
        with open(self.bin_path, "rb") as fp:
            fp.seek(self.byte_offset(0))
            buf = fp.read(self.data_size_bytes())
        if self.shape is not None:
            return np.ndarray(shape=self.shape, dtype=self.dtype, buffer=buf)
        base = self.byte_offset(0)
        return [self.view_frame(i, buf[self.byte_offset(i) - base: self.byte_offset(i + 1) - base])
                for i in range(len(self))]


def write_torch_idx_bin(idx_path: str, items: Sequence[Any], dtype: Optional[str] = None) -> None:
    """Write `items` (ndarrays of one dtype, or msgpack-able dicts when dtype is None/"object") as
    a version-1 .torch.idx / .torch.bin pair."""
    bin_path = _bin_path_for_idx(idx_path)
    msgpack_items = dtype in (None, "object")
    if msgpack_items:
        import msgpack

        blobs = [msgpack.packb(it) for it in items]
        code, itemsize = _MSGPACK_CODE, 1
        dims = [(len(b),) for b in blobs]
    else:
        dt = np.dtype(dtype)
        arrs = [np.asarray(it, dtype=dt, order="C") for it in items]      # (ascontiguousarray would make 0-d items 1-d)
        blobs = [a.tobytes() for a in arrs]
        code, itemsize = _DTYPE_TO_CODE[dt.name], dt.itemsize
        dims = [a.shape for a in arrs]
    dim_offsets = np.cumsum([0] + [len(d) for d in dims])
    data_offsets = np.cumsum([0] + [len(b) // itemsize for b in blobs])
    sizes = [int(v) for d in dims for v in d]
    header = [MAGIC, 1, code, itemsize, len(items), len(sizes)]
    idx = np.array(header + dim_offsets.tolist() + data_offsets.tolist() + sizes, dtype=np.int64)
    os.makedirs(os.path.dirname(os.path.abspath(idx_path)), exist_ok=True)
    with open(idx_path, "wb") as fp:
        fp.write(idx.tobytes())
    with open(bin_path, "wb") as fp:
        for b in blobs:
            fp.write(b)


class SequenceFolder:
    """One dataset folder: item i -> {field: value} for the requested fields
    (lib/data_utils/async_dataset.py:245-300).  ValueError if a field's files are missing or the
    fields disagree on the number of items."""

    def __init__(self, folder: str, fields: Sequence[str] = ("mono", "labels")):
        self.folder, self._idx = folder, {}
        self.n = None
        for field in fields:
            path = os.path.join(folder, field + ".torch.idx")
            if not os.path.exists(path):
                raise ValueError(f"required field {field} not found in folder {folder}, "
                                 f"which only has {sorted(os.listdir(folder))}")
            idx = TorchIdx(path)
            if self.n is not None and self.n != len(idx):
                raise ValueError(f"folder has mismatched field sizes: {field}")
            self.n = len(idx)
            self._idx[field] = idx

    def __len__(self) -> int:
        return int(self.n or 0)

    def __getitem__(self, i: int) -> Dict[str, Any]:
        return {field: idx.read_item(i) for field, idx in self._idx.items()}


def find_sequence_folders(root: str, fields: Sequence[str] = ("mono", "labels")) -> List[str]:
    """Folders under `root` that hold every `<field>.torch.idx` (the reference's find_dataset walk)."""
    out = []
    for cur, _, files in sorted(os.walk(root)):
        if all((f + ".torch.idx") in files for f in fields):
            out.append(cur)
    return out


@dataclass
class RawSample:
    """lib/batched_dataset/sample.py:16-40 (units as stored: mm)."""
    images: np.ndarray
    extrinsics: np.ndarray
    intrinsics: np.ndarray
    enclosing_points: np.ndarray
    hand: np.ndarray
    hand_model: Any
    wrist: np.ndarray
    joint_angles: np.ndarray
    solved_wrist_xfs: np.ndarray
    solved_joint_angles: np.ndarray
    generic_hand_model: Any
    pinch: np.ndarray

    def scaled(self, factor: float):
        from .hand_model import scaled_hand_model

        self.extrinsics[..., :3, 3] *= factor
        self.enclosing_points *= factor
        self.wrist[..., :3, 3] *= factor
        self.solved_wrist_xfs[..., :3, 3] *= factor
        self.hand_model = scaled_hand_model(self.hand_model, factor)
        self.generic_hand_model = scaled_hand_model(self.generic_hand_model, factor)


def parse_raw_buffers(mono: np.ndarray, labels: Dict[str, Any]) -> RawSample:
    """lib/batched_dataset/sample.py:42-53: every label field becomes a float32 array, except the
    two hand models, which become HandModel tuples of tensors.  Unknown or missing fields raise
    TypeError from the dataclass constructor, as in the reference."""
    import torch

    from .hand_model import HandModel

    typed = {}
    for field, value in labels.items():
        if "hand_model" in field:
            typed[field] = HandModel(**{k: torch.tensor(v) for k, v in value.items()})
        else:
            typed[field] = np.array(value, dtype=np.float32)
    return RawSample(images=mono, **typed)


def prepare_model_inputs(sample: RawSample, crop_size: Tuple[int, int] = (96, 96)) -> Dict[str, Any]:
    """The part of prepare_inputs_targets (lib/batched_dataset/data_transform.py:284-381) that the
    inference script consumes: mm -> m, perspective hand crops on the GPU, left-hand (mirrored)
    skeletons repeated over the sequence.  Returns torch tensors: left_images [seq,views,h,w] in
    [0,1], intrinsics [seq,views,3,3], extrinsics_xf [seq,views,4,4], hand_idx [seq], joint_angles /
    wrist_xfs (GT) and solved_* [seq,...], left_hand_model / left_generic_hand_model (HandModel
    with a leading seq dimension)."""
    import torch

    from .hand_model import mirrored_hand_model
    from .torch_data import perspective_crop_images

    sample.scaled(0.001)
    images = sample.images.astype(np.float32)
    seq = images.shape[0]
    th = lambda a: torch.from_numpy(np.ascontiguousarray(a)).float()    # noqa: E731
    rep = lambda hm: type(hm)(*[None if t is None else t.unsqueeze(0).expand(seq, *t.shape) for t in hm])  # noqa: E731
    to_mirror = th(sample.hand) == 1
    left_images, extr, intr = perspective_crop_images(images, sample.extrinsics, sample.intrinsics,
                                                      sample.enclosing_points, int(sample.hand[0]), crop_size)
    return {"left_images": th(left_images), "intrinsics": th(intr), "extrinsics_xf": th(extr),
            "hand_idx": th(sample.hand), "joint_angles": th(sample.joint_angles), "wrist_xfs": th(sample.wrist),
            "solved_joint_angles": th(sample.solved_joint_angles), "solved_wrist_xfs": th(sample.solved_wrist_xfs),
            "left_hand_model": mirrored_hand_model(rep(sample.hand_model), to_mirror),
            "left_generic_hand_model": mirrored_hand_model(rep(sample.generic_hand_model), to_mirror),
            "pinch": th(sample.pinch)}


def eval_sequences(model, inputs: Sequence[Dict[str, Any]], seq_mode: str = "multiv", use_skel: bool = True):
    """run_inference_torch_data.py:39-133 for a batch of prepared sequences of equal length: step
    the frames through the model (memory slot = sequence index, memory used from frame 1 on), skin
    predicted and ground-truth landmarks with the un-mirrored GT skeleton and return the
    per-sequence keypoint error in mm (mean over frames and 21 landmarks) as a float32 tensor,
    plus the predicted joint angles [bs,seq,22] and wrist transforms [bs,seq,4,4] (metres)."""
    import torch

    from .hand_model import mirrored_hand_model
    from .model import InputFrameData, InputFrameDesc, InputSkeletonData
    from .skinning import skin_landmarks

    if seq_mode not in ("multiv", "singlev"):
        raise ValueError(f"Unknown sequence mode: {seq_mode}")
    nv = 2 if seq_mode == "multiv" else 1
    assert use_skel or seq_mode == "multiv", "Skeleton scale prediction requires multiv data"
    dev = torch.device("cuda", torch.cuda.current_device())
    bs, seq_len = len(inputs), int(inputs[0]["left_images"].shape[0])
    stack = lambda key: torch.stack([d[key] for d in inputs]).to(dev)    # noqa: E731
    images, intr, extr, hand_idx = stack("left_images"), stack("intrinsics"), stack("extrinsics_xf"), stack("hand_idx")
    left_hm = type(inputs[0]["left_hand_model"])(*[
        None if inputs[0]["left_hand_model"][k] is None else torch.stack([d["left_hand_model"][k] for d in inputs]).to(dev)
        for k in range(len(inputs[0]["left_hand_model"]))])
    sample_range = torch.tensor([(i * nv, (i + 1) * nv) for i in range(bs)], device=dev).long()
    memory_idx = torch.arange(0, bs, device=dev).long()
    model.reset_memory()
    ja, wr = [], []
    for t in range(seq_len):
        use_memory = torch.full((bs,), t > 0, device=dev, dtype=torch.bool)
        fd = InputFrameData(left_images=torch.flatten(images[:, t, 0:nv], 0, 1).contiguous(),
                            intrinsics=torch.flatten(intr[:, t, 0:nv], 0, 1).contiguous(),
                            extrinsics_xf=torch.flatten(extr[:, t, 0:nv], 0, 1).contiguous())
        desc = InputFrameDesc(hand_idx=hand_idx[:, t].long(), sample_range=sample_range, memory_idx=memory_idx,
                              use_memory=use_memory, n_memory_slots=bs)
        if use_skel:
            out = model.regress_pose_use_skeleton(fd, desc, InputSkeletonData(
                joint_rotation_axes=left_hm.joint_rotation_axes[:, t].contiguous(),
                joint_rest_positions=left_hm.joint_rest_positions[:, t].contiguous()))
        else:
            out = model.regress_pose_pred_skel_scale(fd, desc)
        ja.append(out.joint_angles)
        wr.append(out.wrist_xfs)
    ja, wr = torch.stack(ja, dim=1), torch.stack(wr, dim=1)
    hand_model = mirrored_hand_model(left_hm, hand_idx == 1)
    gt = skin_landmarks(hand_model, stack("joint_angles"), stack("wrist_xfs"))
    pred = skin_landmarks(hand_model, ja, wr)
    err_mm = (gt - pred).norm(dim=-1).mean(dim=(1, 2)) * 1000
    return err_mm, ja, wr
