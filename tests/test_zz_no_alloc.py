"""umt_model_encode_static -- the entry point for hosts that may not allocate device memory.

A tracker embedded in someone else's runtime, or a per-frame loop running after the application
froze its GPU memory budget, can still create the model at start-up but cannot cudaMalloc (or let
torch allocate) a workspace later.  For them the library carries its own 16 MiB scratch array and
encodes up to 4 crops (one frame) out of it.  This test drives that entry point through the C ABI
and requires the same features as the regular, workspace-backed umt_model_encode -- twice in a row,
because the second call starts from the first call's leftovers in the shared array.

Collected last on purpose (file name): a failure here must not hide the other parity tests.
"""
import ctypes

import pytest
import torch

from umetrack_b200 import native, synthetic
from umetrack_b200.model import UmeTrackModel

pytestmark = pytest.mark.gpu


def _out_like(images):
    return torch.full((images.shape[0], 72, 6, 6), float("nan"), dtype=torch.float32, device=images.device)


def _encode_static(model, images, mode=native.MODE_FP32_EQUIV, out=None):
    out = _out_like(images) if out is None else out      # outputs belong to the caller
    rc = native.lib().umt_model_encode_static(model._handle, mode, images.data_ptr(), images.shape[0],
                                              out.data_ptr(), native.current_stream_ptr())
    return rc, out


def test_static_scratch_encode_matches_workspace_encode_twice():
    dev = torch.device("cuda", torch.cuda.current_device())
    model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
    imgs = synthetic.make_model_batch(4, seed=77)["left_images"].to(dev)      # 8 crops
    first, second = imgs[:4].contiguous(), imgs[4:7].contiguous()            # 4 crops, then 3 (ragged)
    ref_first = model.encode(first)          # existing entry point, caller-bound workspace
    ref_second = model.encode(second)
    out_first, out_second = _out_like(first), _out_like(second)
    # start-up call, as the header asks: it resolves the array's device address (and lets the
    # driver load the 16 MiB array), so the calls below are the steady state a real host sees
    assert _encode_static(model, first[:1].contiguous())[0] == 0
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0])
    rc1, got_first = _encode_static(model, first, out=out_first)
    rc2, got_second = _encode_static(model, second, out=out_second)    # reuses the array the first call dirtied
    torch.cuda.synchronize()
    assert rc1 == 0 and rc2 == 0, native.lib().umt_last_error()
    # the launch count of the call is reported on the caller's handle: stem + 24 convolutions + projection
    assert native.lib().umt_model_last_launch_count(model._handle) == 26
    # neither torch's allocator nor the driver handed out device memory during the two calls
    assert torch.cuda.memory_allocated(dev) == before[0]
    assert torch.cuda.mem_get_info(dev)[0] >= before[1]
    # same kernels, same operands: identical up to nothing; keep the encoder tests' tolerance
    for got, ref in ((got_first, ref_first), (got_second, ref_second)):
        assert not torch.isnan(got).any()
        assert float((got - ref).abs().max()) <= 3e-4 * float(ref.abs().max())
    assert torch.equal(got_first, ref_first) and torch.equal(got_second, ref_second)
    # bf16 mode goes through the same scratch
    model.set_mode("bf16")
    ref_b = model.encode(first)
    model.set_mode("fp32")
    rc3, got_b = _encode_static(model, first, native.MODE_BF16)
    torch.cuda.synchronize()
    assert rc3 == 0 and torch.equal(got_b, ref_b)


def test_static_scratch_rejects_more_than_its_capacity():
    dev = torch.device("cuda", torch.cuda.current_device())
    model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
    imgs = torch.zeros((5, 96, 96), dtype=torch.float32, device=dev)
    rc, _ = _encode_static(model, imgs)
    assert rc != 0
    msg = native.lib().umt_last_error().decode()
    assert "exceed the static scratch capacity of 4" in msg
    rc0, _ = _encode_static(model, imgs[:0])
    assert rc0 == 0        # empty batch is a no-op


def test_static_scratch_rejects_a_call_from_another_device():
    """The scratch address is resolved for the current device while the weights stay where the model
    was created; mixing the two must be an argument error, not an illegal address."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda:0")
    with torch.cuda.device(1):
        imgs = torch.zeros((1, 96, 96), dtype=torch.float32, device="cuda:1")
        rc, _ = _encode_static(model, imgs)
    assert rc == 1 and "lives on device 0" in native.lib().umt_last_error().decode()
