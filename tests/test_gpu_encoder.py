"""Stage (b) parity, part 1: the tensor-core backbone against the oracle, layer by layer."""
import numpy as np
import pytest
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import synthetic
from umetrack_b200.model import UmeTrackModel
from tests.helpers import golden

pytestmark = pytest.mark.gpu

# execution-order index of the conv that ends each block -> oracle tap name
BLOCK_END = {}
_i = 0
for _li, _n in enumerate((2, 3, 5, 2)):
    for _b in range(_n):
        BLOCK_END[_i + 1] = f"layer{_li + 1}.{_b}"
        _i += 2


@pytest.fixture(scope="module")
def setup():
    sd = synthetic.make_state_dict(0)
    model = UmeTrackModel(sd).to("cuda")
    imgs = synthetic.make_model_batch(5, seed=100)["left_images"][:9]   # 9 crops: ragged vs tile sizes
    taps = {}
    with torch.no_grad():
        feats = O.backbone_forward(sd, imgs, taps)
    return sd, model, imgs, taps, feats


@pytest.mark.parametrize("layer", list(range(24)))
def test_each_conv_tensor_core_vs_cuda_core_and_oracle(setup, layer):
    sd, model, imgs, taps, _ = setup
    simt, umma = model.debug_layer(layer, imgs)
    simt, umma = simt.cpu(), umma.cpu()
    scale = float(simt.abs().max()) + 1e-6
    # same operands, same program: only accumulation order and the dropped lo*lo term differ
    assert float((simt - umma).abs().max()) <= 2e-5 * scale, f"layer {layer}: tensor-core != cuda-core"
    if layer in BLOCK_END:
        ref = taps[BLOCK_END[layer]]
        assert umma.shape == ref.shape
        assert float((umma - ref).abs().max()) <= 2e-4 * (float(ref.abs().max()) + 1e-6), \
            f"layer {layer} ({BLOCK_END[layer]}) differs from the oracle"


def test_backbone_features_vs_reference_golden(setup):
    sd, model, imgs, taps, feats = setup
    out = model.encode(imgs).cpu()
    assert out.shape == (9, 72, 6, 6)
    assert float((out - feats).abs().max()) <= 3e-4 * float(feats.abs().max())
    g = golden("model_golden.npz")["k_backbone_feats_step0"]
    b = synthetic.make_model_batch(6, seed=100, single_view_every=3)
    got = model.encode(b["left_images"][:3]).cpu().numpy()
    assert np.abs(got - g).max() <= 3e-4 * np.abs(g).max()
    assert model.last_launch_count == 26   # stem + 24 convs + projection


def test_chunking_and_batch_independence(setup):
    sd, _, imgs, _, feats = setup
    small = UmeTrackModel(sd, chunk_images=8).to("cuda")   # 9 images -> two encoder passes
    out = small.encode(imgs).cpu()
    assert float((out - feats).abs().max()) <= 3e-4 * float(feats.abs().max())
    one = small.encode(imgs[4:5]).cpu()
    assert torch.equal(one[0], out[4])


def test_reduced_precision_modes_report_their_error(setup):
    sd, model, imgs, _, feats = setup
    rel = {}
    for mode in ("bf16", "fp16"):
        model.set_mode(mode)
        try:
            out = model.encode(imgs).cpu()
        finally:
            model.set_mode("fp32")
        rel[mode] = float((out - feats).abs().max()) / float(feats.abs().max())
    assert 1e-4 < rel["bf16"] < 8e-2, rel   # bf16 really is lower precision, but not garbage
    assert 1e-5 < rel["fp16"] < rel["bf16"] / 3, rel   # 11-bit vs 8-bit significands


def test_halo_and_guard_rows_stay_zero_after_ragged_passes(setup):
    """The zero padding of every 3x3 convolution (lib/models/backbone_resnet.py:38,49) and the
    separation of neighbouring crops rest on rows of the workspace that no kernel may ever write.
    After passes at several ragged crop counts, in every numeric mode and over two chunks, all of
    them must still be zero (there is no compute-sanitizer on this pool to catch a stray store)."""
    import ctypes

    from umetrack_b200 import native
    sd, _, imgs, _, feats = setup
    model = UmeTrackModel(sd, chunk_images=16).to("cuda")
    big = synthetic.make_model_batch(12, seed=321)["left_images"]        # 24 crops
    for mode in ("fp32", "bf16", "fp16"):
        model.set_mode(mode)
        for n in (24, 16, 9, 1, 23, 5):                                  # 24 and 23 use two passes over the 16-crop chunk
            model.encode(big[:n])
    model.set_mode("fp32")
    out = model.encode(imgs).cpu()                                       # and the results are still right
    assert float((out - feats).abs().max()) <= 3e-4 * float(feats.abs().max())
    bad = ctypes.c_longlong(-1)
    native.check(native.lib().umt_model_debug_halo_check(model._handle, ctypes.byref(bad), native.current_stream_ptr()),
                 "umt_model_debug_halo_check")
    assert bad.value == 0, f"{bad.value} halo / guard rows of the activation workspace are no longer zero"


@pytest.mark.parametrize("mode", ["fp32", "fp16"])
def test_small_batch_launch_variants_are_bit_identical(setup, monkeypatch, mode):
    """A tracker frame (2-4 crops) runs the 128 / 256-channel layers as (tile, 64-channel slice) CTAs and
    launches the convolutions as programmatic dependents.  Dependent launches change no arithmetic: bit-identical
    features, pass after pass (an overlap bug would show up intermittently).  Slices change no arithmetic in the
    16-bit modes either; in the split mode a 64-channel slice uses the merged form of the three products
    (two accumulators added in the epilogue) where the whole layer uses one accumulator, i.e. the same products
    summed in another order (and rounded to hi + lo again after each of the 14 layers): equal to ~1e-5."""
    sd, _, imgs, _, _ = setup
    model = UmeTrackModel(sd).to("cuda").set_mode(mode)
    x = imgs[:4].cuda()
    outs = {}
    for name, slice_rows, overlap in (("default", None, None), ("no slices", "0", None), ("no overlap", None, "0"),
                                      ("neither", "0", "0")):
        for key, val in (("UMT_CONV_SLICE_ROWS", slice_rows), ("UMT_CONV_OVERLAP_MAX", overlap)):
            if val is None:
                monkeypatch.delenv(key, raising=False)
            else:
                monkeypatch.setenv(key, val)
        a = model.encode(x).clone()
        for _ in range(5):
            assert torch.equal(model.encode(x), a), f"{name}: not reproducible"
        outs[name] = a.cpu()
    model.check_status()
    assert torch.equal(outs["default"], outs["no overlap"]), f"dependent launches changed the result ({mode})"
    assert torch.equal(outs["no slices"], outs["neither"]), f"dependent launches changed the result ({mode})"
    if mode == "fp16":
        assert torch.equal(outs["default"], outs["no slices"])
    else:
        scale = float(outs["default"].abs().max())
        assert float((outs["default"] - outs["no slices"]).abs().max()) <= 5e-5 * scale   # (features vs oracle: 3e-4)
