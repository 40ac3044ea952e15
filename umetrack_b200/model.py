"""`UmeTrackModel` on the B200: the reference's model API over the CUDA library.

Public names and semantics follow lib/models/umetrack_model.py:21-242, lib/models/regressor.py:124-129
and lib/models/model_loader.py:53-88:

    model = load_pretrained_model("pretrained_models/pretrained_weights.torch")
    model.eval(); model.to("cuda")
    out = model.regress_pose_use_skeleton(InputFrameData(...), InputFrameDesc(...), InputSkeletonData(...))
    out = model.regress_pose_pred_skel_scale(InputFrameData(...), InputFrameDesc(...))

The weight file is the same flat `state_dict`; it is validated strictly (every key, every shape).
All arithmetic runs in `libumetrack_b200.so`; calling the model without a CUDA device raises.
Like the reference, the model owns the recurrent ConvRNN state (lib/models/temporal.py:41-42)
and mutates it on every call; it is not re-entrant.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import arch, native
from .arch import ModelOpts


@dataclass
class InputFrameData:
    """left_images [n_images,h,w]; intrinsics [n_images,3,3]; extrinsics_xf [n_images,4,4]."""

    left_images: torch.Tensor
    intrinsics: torch.Tensor
    extrinsics_xf: torch.Tensor


@dataclass
class InputFrameDesc:
    """sample_range [bs,2] (start,end image index, 1 or 2 views); memory_idx [bs]; use_memory
    [bs] bool; hand_idx [bs] (1 = right hand).

    `n_memory_slots` is an optional extension: the caller's promise that every memory_idx is below
    it.  The reference reads `max(memory_idx) + 1` on the host at every call to grow its memory bank
    (lib/models/temporal.py:101-123); with the promise the bank is sized without that device->host
    read.  A broken promise does not corrupt memory: the sample runs without recurrent memory and
    UMT_STATUS_MEMORY_INDEX is raised by `UmeTrackModel.check_status()`.

    `persistent_inputs` is a second extension for per-frame callers: the promise that the image,
    camera and descriptor tensors of this call are the same device buffers at every call of this
    shape (the caller rewrites them in place between calls).  A CUDA-graph replay then reads them
    where they are instead of copying nine small tensors into buffers of its own."""

    sample_range: torch.Tensor
    memory_idx: torch.Tensor
    use_memory: torch.Tensor
    hand_idx: torch.Tensor
    n_memory_slots: Optional[int] = None
    persistent_inputs: bool = False


@dataclass
class InputSkeletonData:
    """joint_rotation_axes / joint_rest_positions [bs or 1, 22, 3] (metres)."""

    joint_rotation_axes: torch.Tensor
    joint_rest_positions: torch.Tensor


@dataclass
class RegressorOutput:
    joint_angles: torch.Tensor
    wrist_xfs: torch.Tensor
    skel_scales: Optional[torch.Tensor] = None
    landmark_uncertainty_sigmas: Optional[torch.Tensor] = None


_MODES = {"fp32": native.MODE_FP32_EQUIV, "fp32_equiv": native.MODE_FP32_EQUIV, "bf16": native.MODE_BF16,
          "fp16": native.MODE_FP16}


def validate_state_dict(state_dict: Dict[str, torch.Tensor]) -> None:
    """Strict key/shape check, the equivalent of load_state_dict(strict=True)
    (lib/models/model_loader.py:87)."""
    spec = arch.state_dict_spec()
    want = {k for k, _, _ in spec}
    have = set(state_dict.keys())
    missing, unexpected = sorted(want - have), sorted(have - want)
    if missing or unexpected:
        raise RuntimeError(
            "Error(s) in loading state_dict for UmeTrackModel: "
            f"Missing key(s): {missing[:8]}{'...' if len(missing) > 8 else ''} "
            f"Unexpected key(s): {unexpected[:8]}{'...' if len(unexpected) > 8 else ''}")
    for key, shape, _ in spec:
        if tuple(state_dict[key].shape) != tuple(shape):
            raise RuntimeError(f"size mismatch for {key}: checkpoint {tuple(state_dict[key].shape)}, "
                               f"model {tuple(shape)}")


class UmeTrackModel:
    """Inference-only model bound to one CUDA device."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], opts: ModelOpts = ModelOpts(),
                 chunk_images: int = 4096):
        arch.check_supported(opts)
        validate_state_dict(state_dict)
        self._opts = opts
        self._sd = {k: v.detach().to("cpu") for k, v in state_dict.items()}
        self._handle: Optional[ctypes.c_void_p] = None
        self._device: Optional[torch.device] = None
        self._mode = native.MODE_FP32_EQUIV
        self._chunk_cap = int(chunk_images)
        self._ws: Optional[torch.Tensor] = None
        self._ws_shape = (0, 0, 0)  # (chunk, max_images, max_samples)
        # recurrent state (lib/models/temporal.py:41-42)
        self._mem_features = torch.empty(0)
        self._prev_extrinsics = torch.empty(0)
        self.last_launch_count = 0
        self._frame_graphs = False
        self._graphs: Dict[tuple, dict] = {}
        self._last_packed = None     # (buffer, offsets) of the last graph replay's outputs, else None

    # ---- nn.Module-like surface the callers use ----------------------------------------------
    def eval(self):
        return self

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return dict(self._sd)

    def getInputImageSizes(self) -> Tuple[int, int]:
        return (arch.CROP_SIZE, arch.CROP_SIZE)

    def set_mode(self, mode: str):
        """'fp32' (default; split-bf16, three tensor-core products, fp32-equivalent results) or
        'bf16' (single product, ~7e-3 relative feature error) or 'fp16' (single product, 11-bit
        significands: ~1e-3 feature error, needs activations/weights inside the half range)."""
        self._mode = _MODES[mode]
        return self

    def to(self, device):
        device = torch.device(device)
        if device.type != "cuda":
            raise native.NativeError(
                f"UmeTrackModel.to({device}): this implementation runs on CUDA (sm_100a) only; "
                "there is no CPU path")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if self._handle is not None and self._device == device:
            return self
        self._release()
        lib = native.lib()
        n = lib.umt_model_num_tensors()
        arrays = []
        ptrs = (ctypes.c_void_p * n)()
        for i in range(n):
            name = lib.umt_model_tensor_name(i).decode()
            t = self._sd[name].to(torch.float32).contiguous()
            if t.numel() != lib.umt_model_tensor_numel(i):
                raise RuntimeError(f"size mismatch for {name}")
            arrays.append(t)
            ptrs[i] = t.data_ptr()
        out = ctypes.c_void_p()
        with torch.cuda.device(device):
            native.check(lib.umt_model_create(ptrs, n, ctypes.byref(out)), "umt_model_create")
        self._handle = out
        self._device = device
        self._ws = None
        self._ws_shape = (0, 0, 0)
        return self

    def cuda(self, index: Optional[int] = None):
        return self.to(torch.device("cuda", torch.cuda.current_device() if index is None else index))

    def enable_frame_graphs(self, on: bool = True):
        """Latency mode for per-frame callers (HandTracker): a forward pass on at most 8 crops is
        captured once per shape into a CUDA graph and replayed -- one launch instead of ~45 and no
        per-call host work in the C library.  The returned tensors are the graph's static outputs:
        they are overwritten by the next call of the same shape, so read them before calling again
        (HandTracker copies them to the host straight away)."""
        self._frame_graphs = bool(on)
        if not on:
            self._graphs.clear()
        return self

    def frame_graph_scope(self):
        """Context manager: frame graphs on for the calls inside (used by HandTracker, which reads the
        static outputs before it returns), previous setting restored afterwards."""
        import contextlib

        @contextlib.contextmanager
        def scope():
            prev, self._frame_graphs = self._frame_graphs, True
            try:
                yield self
            finally:
                self._frame_graphs = prev
        return scope()

    def reset_memory(self):
        self._mem_features = torch.empty(0)
        self._prev_extrinsics = torch.empty(0)
        self._graphs.clear()

    def _release(self):
        if self._handle is not None:
            native.lib().umt_model_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    # ---- plumbing --------------------------------------------------------------------------
    def _require_handle(self):
        if self._handle is None:
            raise native.NativeError("UmeTrackModel: call .to('cuda') before inference (CUDA only)")

    def _ensure_workspace(self, n_img: int, n_samples: int):
        chunk, mi, ms = self._ws_shape
        if self._ws is not None and n_img <= mi and n_samples <= ms and (n_img <= chunk or chunk == self._chunk_cap):
            return
        new_chunk = min(self._chunk_cap, max(chunk, _round_up(n_img, 8)))
        new_mi = max(mi, _round_up(n_img, 8))
        new_ms = max(ms, _round_up(n_samples, 8))
        lib = native.lib()
        need = ctypes.c_size_t()
        native.check(lib.umt_model_workspace_bytes(self._handle, new_chunk, new_mi, new_ms, ctypes.byref(need)),
                     "umt_model_workspace_bytes")
        self._ws = None
        self._ws = torch.empty(need.value + 256, dtype=torch.uint8, device=self._device)
        base = (self._ws.data_ptr() + 255) & ~255
        native.check(lib.umt_model_bind_workspace(self._handle, base, need.value, new_chunk, new_mi, new_ms,
                                                  native.current_stream_ptr()), "umt_model_bind_workspace")
        self._ws_shape = (new_chunk, new_mi, new_ms)

    def _grow_memory(self, required: int):
        """Lazily grown memory bank, zero initialised (lib/models/temporal.py:101-123)."""
        if len(self._mem_features) >= required and self._mem_features.device == self._device:
            return
        mem = torch.zeros(required, arch.N_MEMORY, arch.FEAT_HW, arch.FEAT_HW, dtype=torch.float32,
                          device=self._device)
        ext = torch.zeros(required, 4, 4, dtype=torch.float32, device=self._device)
        old = len(self._mem_features)
        if old:
            mem[:old] = self._mem_features.to(self._device)
            ext[:old] = self._prev_extrinsics.to(self._device)
        self._mem_features, self._prev_extrinsics = mem, ext

    def _prep(self, t: torch.Tensor, dtype) -> torch.Tensor:
        return t.to(device=self._device, dtype=dtype).contiguous()

    def _forward(self, head: int, frame_data: InputFrameData, frame_desc: InputFrameDesc,
                 skel: Optional[InputSkeletonData], features: Optional[torch.Tensor] = None) -> RegressorOutput:
        self._require_handle()
        if features is not None:   # backbone already run (encode()): left_images is not used
            images = self._prep(features, torch.float32)
            if images.dim() != 4 or tuple(images.shape[1:]) != (arch.N_IMAGE_FEATURES, arch.FEAT_HW, arch.FEAT_HW):
                raise ValueError(f"features must be [n,72,6,6], got {tuple(images.shape)}")
        else:
            # uint8 crops (as cv2.remap leaves them) are an accepted extension: the `/ 255` of
            # lib/tracker/tracker.py:211 then runs on the device (umt_model_forward_u8)
            u8 = frame_data.left_images.dtype == torch.uint8
            images = self._prep(frame_data.left_images, torch.uint8 if u8 else torch.float32)
            if images.dim() != 3 or tuple(images.shape[1:]) != (arch.CROP_SIZE, arch.CROP_SIZE):
                raise ValueError(f"left_images must be [n,96,96], got {tuple(images.shape)}")
        intr = self._prep(frame_data.intrinsics, torch.float32)
        extr = self._prep(frame_data.extrinsics_xf, torch.float32)
        srange = self._prep(frame_desc.sample_range, torch.int64)
        mem_idx = self._prep(frame_desc.memory_idx, torch.int64)
        use_mem = self._prep(frame_desc.use_memory, torch.bool)
        hand_idx = self._prep(frame_desc.hand_idx, torch.int64)
        n_img, bs = images.shape[0], srange.shape[0]
        if bs == 0:
            raise ValueError("empty batch")
        self._ensure_workspace(8 if features is not None else n_img, bs)
        # memory bank size: like the reference, max(memory_idx) + 1 read on the host at every call
        # (temporal.py:101) -- from the caller's tensor if that is still on the host, which costs
        # nothing -- unless the caller states the slot count (InputFrameDesc.n_memory_slots)
        if frame_desc.n_memory_slots is not None:
            need = int(frame_desc.n_memory_slots)
        elif frame_desc.memory_idx.device.type == "cpu":
            need = int(frame_desc.memory_idx.max()) + 1
        else:
            need = int(mem_idx.max().item()) + 1
        if need < 1:
            raise ValueError("memory_idx must be non-negative")
        self._grow_memory(need)
        dev = self._device
        if (self._frame_graphs and features is None and n_img <= 8 and (skel is not None or head == native.HEAD_UNKNOWN)):
            return self._forward_graph(head, images, intr, extr, srange, mem_idx, use_mem, hand_idx, skel,
                                       persistent=bool(frame_desc.persistent_inputs))
        self._last_packed = None
        ja = torch.empty((bs, 22), dtype=torch.float32, device=dev)
        wr = torch.empty((bs, 4, 4), dtype=torch.float32, device=dev)
        sg = torch.empty((bs, 21), dtype=torch.float32, device=dev)
        sc = torch.empty((bs,), dtype=torch.float32, device=dev) if head == native.HEAD_UNKNOWN else None
        axes = rest = None
        n_skel = 0
        if head == native.HEAD_KNOWN:
            axes = self._prep(skel.joint_rotation_axes, torch.float32).reshape(-1, 22, 3)
            rest = self._prep(skel.joint_rest_positions, torch.float32).reshape(-1, 22, 3)
            n_skel = axes.shape[0]
        lib = native.lib()
        with torch.cuda.device(dev):
            if features is not None:
                native.check(lib.umt_model_tail(
                    self._handle, head, native.ptr(images), native.ptr(intr), native.ptr(extr),
                    native.ptr(srange), native.ptr(mem_idx), native.ptr(use_mem), native.ptr(hand_idx),
                    n_img, bs, native.ptr(axes), native.ptr(rest), n_skel,
                    native.ptr(self._mem_features), native.ptr(self._prev_extrinsics),
                    len(self._mem_features), native.ptr(ja), native.ptr(wr), native.ptr(sc), native.ptr(sg),
                    native.current_stream_ptr()), "umt_model_tail")
            else:
                fwd = lib.umt_model_forward_u8 if images.dtype == torch.uint8 else lib.umt_model_forward
                native.check(fwd(
                    self._handle, self._mode, head, native.ptr(images), native.ptr(intr), native.ptr(extr),
                    native.ptr(srange), native.ptr(mem_idx), native.ptr(use_mem), native.ptr(hand_idx),
                    n_img, bs, native.ptr(axes), native.ptr(rest), n_skel,
                    native.ptr(self._mem_features), native.ptr(self._prev_extrinsics),
                    len(self._mem_features), native.ptr(ja), native.ptr(wr), native.ptr(sc), native.ptr(sg),
                    native.current_stream_ptr()), "umt_model_forward")
        self.last_launch_count = lib.umt_model_last_launch_count(self._handle)
        return RegressorOutput(joint_angles=ja, wrist_xfs=wr, skel_scales=sc, landmark_uncertainty_sigmas=sg)

    def _forward_graph(self, head, images, intr, extr, srange, mem_idx, use_mem, hand_idx, skel,
                       persistent: bool = False) -> RegressorOutput:
        """Replay (or capture) the CUDA graph of umt_model_forward for this shape.  The outputs are
        slices of one buffer (`_last_packed`), so that a per-frame caller reads them with one copy."""
        dev = self._device
        n_img, bs = images.shape[0], srange.shape[0]
        axes = rest = None
        if head == native.HEAD_KNOWN:
            axes = self._prep(skel.joint_rotation_axes, torch.float32).reshape(-1, 22, 3)
            rest = self._prep(skel.joint_rest_positions, torch.float32).reshape(-1, 22, 3)
        n_skel = 0 if axes is None else axes.shape[0]
        ins = [images, intr, extr, srange, mem_idx, use_mem, hand_idx] + ([axes, rest] if axes is not None else [])
        key = (head, self._mode, n_img, bs, n_skel, images.dtype, self._ws.data_ptr(), self._mem_features.data_ptr(),
               len(self._mem_features), tuple(t.data_ptr() for t in ins) if persistent else None)
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) > 16:
                self._graphs.clear()
            st = ins if persistent else [t.clone() for t in ins]
            r4 = lambda v: (v + 3) & ~3          # every output starts on a 16-byte boundary
            o_wr, o_sg, o_sc = r4(bs * 22), r4(bs * 22) + bs * 16, r4(r4(bs * 22) + bs * 16 + bs * 21)
            pack = torch.empty(o_sc + r4(bs), dtype=torch.float32, device=dev)
            out = {"ja": pack[:bs * 22].view(bs, 22), "wr": pack[o_wr:o_wr + bs * 16].view(bs, 4, 4),
                   "sg": pack[o_sg:o_sg + bs * 21].view(bs, 21),
                   "sc": pack[o_sc:o_sc + bs] if head == native.HEAD_UNKNOWN else None,
                   "pack": pack, "offsets": (0, o_wr, o_sg, o_sc)}
            lib = native.lib()
            fwd = lib.umt_model_forward_u8 if images.dtype == torch.uint8 else lib.umt_model_forward

            def call():
                native.check(fwd(
                    self._handle, self._mode, head, native.ptr(st[0]), native.ptr(st[1]), native.ptr(st[2]),
                    native.ptr(st[3]), native.ptr(st[4]), native.ptr(st[5]), native.ptr(st[6]), n_img, bs,
                    native.ptr(st[7]) if n_skel else 0, native.ptr(st[8]) if n_skel else 0, n_skel,
                    native.ptr(self._mem_features), native.ptr(self._prev_extrinsics), len(self._mem_features),
                    native.ptr(out["ja"]), native.ptr(out["wr"]), native.ptr(out["sc"]), native.ptr(out["sg"]),
                    native.current_stream_ptr()), "umt_model_forward")

            # one plain call first (per-device kernel attributes, driver entry points), on a copy of the
            # recurrent state so that the capture below is this call's one and only state update
            mem, ext = self._mem_features.clone(), self._prev_extrinsics.clone()
            with torch.cuda.device(dev):
                call()
                self._mem_features.copy_(mem)
                self._prev_extrinsics.copy_(ext)
                torch.cuda.synchronize(dev)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    call()
            g = {"graph": graph, "st": st, "out": out, "launches": native.lib().umt_model_last_launch_count(self._handle)}
            self._graphs[key] = g
        elif not persistent:
            for dst, src in zip(g["st"], ins):
                dst.copy_(src)
        g["graph"].replay()
        self.last_launch_count = g["launches"]
        o = g["out"]
        self._last_packed = (o["pack"], o["offsets"])
        return RegressorOutput(joint_angles=o["ja"], wrist_xfs=o["wr"], skel_scales=o["sc"],
                               landmark_uncertainty_sigmas=o["sg"])

    # ---- the two entry points of the reference ------------------------------------------------
    def regress_pose_use_skeleton(self, frame_data: InputFrameData, frame_desc: InputFrameDesc,
                                  skel_data: InputSkeletonData) -> RegressorOutput:
        return self._forward(native.HEAD_KNOWN, frame_data, frame_desc, skel_data)

    def regress_pose_pred_skel_scale(self, frame_data: InputFrameData,
                                     frame_desc: InputFrameDesc) -> RegressorOutput:
        r = frame_desc.sample_range
        assert bool(((r[:, 1] - r[:, 0]) != 1).all()), \
            "Unsupported: found single-view samples when calibration scale"
        return self._forward(native.HEAD_UNKNOWN, frame_data, frame_desc, None)

    def regress_pose_from_features(self, features: torch.Tensor, intrinsics: torch.Tensor, extrinsics_xf: torch.Tensor,
                                   frame_desc: InputFrameDesc, skel_data: Optional[InputSkeletonData]) -> RegressorOutput:
        """Fusion + temporal memory + regression on backbone features computed earlier with
        `encode()` ([n,72,6,6]); `skel_data=None` selects the unknown-skeleton head.  The backbone
        has no temporal dependency, so recordings can be encoded as one batch and stepped here."""
        fd = InputFrameData(features, intrinsics, extrinsics_xf)
        head = native.HEAD_KNOWN if skel_data is not None else native.HEAD_UNKNOWN
        return self._forward(head, fd, frame_desc, skel_data, features=features)

    def tail_scan(self, features: torch.Tensor, intrinsics: torch.Tensor, extrinsics_xf: torch.Tensor,
                  step_rows, sample_range: torch.Tensor, memory_idx: torch.Tensor, use_memory: torch.Tensor,
                  hand_idx: torch.Tensor, skel_axes: torch.Tensor, skel_rest: torch.Tensor, n_slots: int):
        """Time scan for batched recordings: `step_rows` is a list of (first_row, n_rows); the
        descriptor tensors hold the rows of ALL steps back to back (already on the device), so a step
        is one C call with offset pointers -- no tensor construction, no host sync.  Uses the
        known-skeleton head with one skeleton per row.  Returns (joint_angles [rows,22], wrist_xfs
        [rows,4,4] in metres, sigmas [rows,21]) in row order."""
        self._require_handle()
        dev = self._device
        feats = self._prep(features, torch.float32)
        intr = self._prep(intrinsics, torch.float32)
        extr = self._prep(extrinsics_xf, torch.float32)
        sr = self._prep(sample_range, torch.int64)
        mi = self._prep(memory_idx, torch.int64)
        um = self._prep(use_memory, torch.bool)
        hi = self._prep(hand_idx, torch.int64)
        ax = self._prep(skel_axes, torch.float32)
        rs = self._prep(skel_rest, torch.float32)
        rows = sr.shape[0]
        max_rows = max((n for _, n in step_rows), default=0)
        self._ensure_workspace(8, max(max_rows, 8))
        self._grow_memory(n_slots)
        ja = torch.empty((rows, 22), dtype=torch.float32, device=dev)
        wr = torch.empty((rows, 4, 4), dtype=torch.float32, device=dev)
        sg = torch.empty((rows, 21), dtype=torch.float32, device=dev)
        lib = native.lib()
        first = (ctypes.c_int * len(step_rows))(*[int(f) for f, _ in step_rows])
        count = (ctypes.c_int * len(step_rows))(*[int(n) for _, n in step_rows])
        with torch.cuda.device(dev):
            native.check(lib.umt_model_tail_scan(
                self._handle, native.HEAD_KNOWN, feats.data_ptr(), intr.data_ptr(), extr.data_ptr(), sr.data_ptr(),
                mi.data_ptr(), um.data_ptr(), hi.data_ptr(), feats.shape[0], first, count, len(step_rows),
                ax.data_ptr(), rs.data_ptr(), 1, self._mem_features.data_ptr(), self._prev_extrinsics.data_ptr(),
                len(self._mem_features), ja.data_ptr(), wr.data_ptr(), 0, sg.data_ptr(),
                native.current_stream_ptr()), "umt_model_tail_scan")
        self.last_launch_count = lib.umt_model_last_launch_count(self._handle)
        return ja, wr, sg

    def check_status(self) -> None:
        """Raise if a kernel recorded a condition since the last check: a memory index outside the
        bank (only possible with a wrong `n_memory_slots` promise or a too small `n_slots`), or an
        fp16-mode activation outside the half range.  Synchronises the current stream."""
        self._require_handle()
        st = ctypes.c_uint(0)
        with torch.cuda.device(self._device):
            native.check(native.lib().umt_model_status(self._handle, ctypes.byref(st), native.current_stream_ptr()),
                         "umt_model_status")
        if st.value & native.STATUS_MEMORY_INDEX:
            raise native.NativeError("a memory_idx was outside the memory bank: that sample ran without "
                                     "recurrent memory (n_memory_slots / n_slots too small)")
        if st.value & native.STATUS_FP16_RANGE:
            raise native.NativeError("fp16 mode: an activation left the half range (|x| > 65504); "
                                     "use set_mode('fp32') or 'bf16' for these weights")

    # ---- extras ----------------------------------------------------------------------------
    def encode(self, images: torch.Tensor) -> torch.Tensor:
        """Backbone only: [n,96,96] (float32 in [0,1], or uint8) -> [n,72,6,6] (the reference's
        `_image_backbone`)."""
        self._require_handle()
        u8 = images.dtype == torch.uint8
        images = self._prep(images, torch.uint8 if u8 else torch.float32)
        n = images.shape[0]
        self._ensure_workspace(n, 8)
        out = torch.empty((n, arch.N_IMAGE_FEATURES, arch.FEAT_HW, arch.FEAT_HW), dtype=torch.float32,
                          device=self._device)
        with torch.cuda.device(self._device):
            enc = native.lib().umt_model_encode_u8 if u8 else native.lib().umt_model_encode
            native.check(enc(self._handle, self._mode, native.ptr(images), n, native.ptr(out),
                             native.current_stream_ptr()), "umt_model_encode")
        self.last_launch_count = native.lib().umt_model_last_launch_count(self._handle)
        return out

    def profile_layers(self, images: torch.Tensor):
        """Per-launch device times (ms) of one encoder pass: [stem, conv0..conv23, projection]."""
        self._require_handle()
        images = self._prep(images, torch.float32)
        n = images.shape[0]
        self._ensure_workspace(n, 8)
        out = (ctypes.c_float * 26)()
        with torch.cuda.device(self._device):
            native.check(native.lib().umt_model_profile_layers(self._handle, self._mode, native.ptr(images), n,
                                                               out, native.current_stream_ptr()),
                         "umt_model_profile_layers")
        return [float(x) for x in out]

    def debug_layer(self, layer: int, images: torch.Tensor):
        """(cuda-core result, tensor-core result) of backbone convolution `layer`, dense NCHW."""
        self._require_handle()
        images = self._prep(images, torch.float32)
        n = images.shape[0]
        self._ensure_workspace(n, 8)
        shape = (ctypes.c_int * 3)()
        lib = native.lib()
        # first call only for the shape
        cmax = 256 * 48 * 48  # upper bound is loose; sizes are tiny in tests
        hw = {0: (32, 48), 1: (64, 24), 2: (128, 12), 3: (256, 6)}
        stage = 0 if layer < 4 else 1 if layer < 10 else 2 if layer < 20 else 3
        c, h = hw[stage]
        a = torch.zeros((n, c, h, h), dtype=torch.float32, device=self._device)
        b = torch.zeros((n, c, h, h), dtype=torch.float32, device=self._device)
        with torch.cuda.device(self._device):
            native.check(lib.umt_model_debug_layer(self._handle, self._mode, layer, native.ptr(images), n,
                                                   native.ptr(a), native.ptr(b), shape,
                                                   native.current_stream_ptr()), "umt_model_debug_layer")
        assert (shape[0], shape[1], shape[2]) == (c, h, h), (tuple(shape), (c, h, h))
        return a, b


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def load_pretrained_model(model_path: str) -> UmeTrackModel:
    """Read the reference's weight file (a flat state_dict saved with torch.save) and build the
    model (lib/models/model_loader.py:53-88).  Move it to the GPU with `.to("cuda")`."""
    with open(model_path, "rb") as fp:
        state_dict = torch.load(fp, map_location="cpu")
    return UmeTrackModel(state_dict)
