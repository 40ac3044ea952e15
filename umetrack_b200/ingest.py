"""Recording ingest: labels JSON -> dict, and staging of the decoded
frames to the GPU.

The reference reads a recording as `<name>.mp4` + `<name>.json` (lib/tracker/video_pose_data.py:
63-151): the JSON holds cameras, camera_angles, hand_model, joint_angles [T,2,22],
wrist_transforms [T,2,4,4] (mm), hand_confidences [T,2] and camera_to_world_transforms
[T,n_cam,4,4]; every video frame is one uint8 image [H, n_cam*W] with the views side by side
(:123-125,144).  Video decode (PyAV) stays on the host and is not part of this package: frames
arrive here as a uint8 array / tensor [T, H, n_cam*W] or as an iterator of such chunks.

`FrameStager` moves them to the device in chunks of frames through a small ring of pinned host
buffers and device buffers on a dedicated copy stream, so that (a) the upload of chunk k+1
overlaps the resampling of chunk k, and (b) device memory holds `n_slots` chunks instead of the
whole video (1.2 MB per 4-view 640x480 frame).  Crops (36 KB per view) are what stays resident.
"""
from __future__ import annotations

import json
from typing import Iterator, Tuple, Union

import numpy as np
import torch

from . import native

Frames = Union[np.ndarray, torch.Tensor]

_LABEL_KEYS = ("cameras", "camera_angles", "hand_model", "joint_angles", "wrist_transforms",
               "hand_confidences", "camera_to_world_transforms")


def load_recording_labels(json_path: str) -> dict:
    """Read `<recording>.json` (lib/tracker/video_pose_data.py:75-93).  Raises KeyError naming the
    first missing field, like the reference's direct indexing does."""
    with open(json_path, "rb") as fp:
        labels = json.load(fp)
    for k in _LABEL_KEYS:
        if k not in labels:
            raise KeyError(k)
    return labels


class FrameStager:
    """Chunked, double-buffered host->device staging of uint8 frames [T, H, W4].

        stager = FrameStager(device, chunk_frames=32, n_slots=3)
        for t0, t1, dev_chunk, slot in stager.chunks(frames):
            ... launch kernels reading dev_chunk[: t1 - t0] on the current stream ...
            stager.release(slot)

    `frames` may be a numpy array (copied into the pinned ring by the host, then uploaded), a
    pinned torch tensor (uploaded straight from it), or a CUDA tensor (yielded in place, no copy).
    The current stream waits for each chunk's upload before the caller's kernels run; the copy
    stream waits for `release` before a slot is overwritten."""

    def __init__(self, device, chunk_frames: int = 32, n_slots: int = 3):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise native.NativeError("FrameStager stages frames to a CUDA device")
        if chunk_frames < 1 or n_slots < 2:
            raise ValueError("chunk_frames >= 1 and n_slots >= 2 required")
        self.chunk_frames, self.n_slots = int(chunk_frames), int(n_slots)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._shape = None
        self._host, self._dev, self._ready, self._free = [], [], [], []
        self.bytes_uploaded = 0

    def _ensure(self, H: int, W4: int, need_host: bool):
        if self._shape != (H, W4):
            self._shape = (H, W4)
            self._host, self._dev = [], []
            for _ in range(self.n_slots):
                self._dev.append(torch.empty((self.chunk_frames, H, W4), dtype=torch.uint8, device=self.device))
            self._ready = [torch.cuda.Event() for _ in range(self.n_slots)]
            self._free = [None] * self.n_slots
        if need_host and not self._host:
            self._host = [torch.empty((self.chunk_frames, H, W4), dtype=torch.uint8).pin_memory()
                          for _ in range(self.n_slots)]

    def _upload(self, slot: int, src: Frames, t0: int, t1: int):
        n = t1 - t0
        if isinstance(src, np.ndarray):
            if self._free[slot] is not None:
                self._ready[slot].synchronize()                    # the previous upload has left this pinned buffer
            self._host[slot][:n].copy_(torch.from_numpy(np.ascontiguousarray(src[t0:t1])))
            host = self._host[slot][:n]
        else:
            host = src[t0:t1]
        with torch.cuda.stream(self.copy_stream):
            if self._free[slot] is not None:
                self.copy_stream.wait_event(self._free[slot])      # consumer finished with the old contents
            self._dev[slot][:n].copy_(host, non_blocking=True)
            self._ready[slot].record(self.copy_stream)
        self.bytes_uploaded += n * self._shape[0] * self._shape[1]

    def chunks(self, frames: Frames) -> Iterator[Tuple[int, int, torch.Tensor, int]]:
        if frames.ndim != 3 or frames.dtype not in (np.uint8, torch.uint8):
            raise ValueError("frames must be uint8 [T, H, n_cam*W]")
        T, H, W4 = int(frames.shape[0]), int(frames.shape[1]), int(frames.shape[2])
        cf = self.chunk_frames
        spans = [(t0, min(T, t0 + cf)) for t0 in range(0, T, cf)]
        if isinstance(frames, torch.Tensor) and frames.is_cuda:
            fr = frames.contiguous()
            for t0, t1 in spans:
                yield t0, t1, fr[t0:t1], -1
            return
        if isinstance(frames, torch.Tensor):
            frames = frames.contiguous()
            if not frames.is_pinned():
                frames = frames.numpy()                            # pageable: go through the pinned ring
        self._ensure(H, W4, need_host=isinstance(frames, np.ndarray))
        ahead = 0                                                  # chunks whose upload has been issued
        for k, (t0, t1) in enumerate(spans):
            # chunk a reuses the slot of chunk a - n_slots, which the caller has released iff it is < k
            while ahead < len(spans) and ahead < k + self.n_slots:
                a0, a1 = spans[ahead]
                self._upload(ahead % self.n_slots, frames, a0, a1)
                ahead += 1
            slot = k % self.n_slots
            torch.cuda.current_stream(self.device).wait_event(self._ready[slot])
            yield t0, t1, self._dev[slot][: t1 - t0], slot

    def release(self, slot: int):
        """The kernels reading this slot have been launched on the current stream."""
        if slot < 0:
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._free[slot] = ev
