"""Static description of the one architecture this library runs.

The reference builds its network from `ModelOpts` defaults (lib/models/model_opts.py:10-39,
lib/models/model_loader.py:53-88).  The kernels here are specialised to exactly that
configuration; `state_dict_spec()` lists every tensor of the reference's flat `state_dict`
(252 entries) with its shape so weight files can be validated strictly, the way
`load_state_dict(strict=True)` does in the reference (model_loader.py:84-87).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

CROP_SIZE = 96
FEAT_HW = 6
FEAT_PIX = FEAT_HW * FEAT_HW
N_IMAGE_FEATURES = 72
N_SKELETON_FEATURES = 4
N_MEMORY = 18
N_TEMPORAL = N_IMAGE_FEATURES + N_MEMORY
N_VIEWS = 2
N_RIGID_PTS = 7
BN_EPS = 1e-5
CANONICAL_FOCAL = 200.0

STAGE_BLOCKS = (2, 3, 5, 2)
STAGE_PLANES = (32, 64, 128, 256)
STAGE_HW = (48, 24, 12, 6)

REG_K_CH, REG_K_OUT = N_IMAGE_FEATURES + N_SKELETON_FEATURES, 62
REG_U_CH, REG_U_OUT = N_IMAGE_FEATURES, 63

BACKBONE = "_feature_extractor._image_backbone."
FUSION = "_feature_extractor._multi_view_fusion."
TEMPORAL = "_temporal._temporal_module."
SKELETON = "_skeleton_enc._layers."


@dataclass(frozen=True)
class ModelOpts:
    """Same fields and defaults as lib/models/model_opts.py:10-39."""

    network: str = "resnet_layers_2352-f32"
    nImageFeatureChannels: int = 72
    nSkeletonFeatureChannels: int = 4
    nTemporalMemoryChannels: int = 18
    useUnscaledAsCanonical: bool = False
    nMultiViewFusionBlocks: int = 2
    nTemporalBlocks: int = 3
    nPoseRegressionBlocks: int = 2
    spatialFTLRatio: float = 1.0
    temporalFTLRatio: float = 1.0
    nWristRigidPts: int = 7


def check_supported(opts: ModelOpts) -> None:
    """The CUDA path is compiled for the default options only; anything else fails loudly."""
    if opts != ModelOpts():
        raise NotImplementedError(
            f"umetrack_b200 kernels are specialised to ModelOpts() defaults, got {opts}")


def _bn(prefix: str, c: int) -> List[Tuple[str, Tuple[int, ...], str]]:
    return [
        (prefix + ".weight", (c,), "bn_gamma"),
        (prefix + ".bias", (c,), "bn_beta"),
        (prefix + ".running_mean", (c,), "bn_mean"),
        (prefix + ".running_var", (c,), "bn_var"),
        (prefix + ".num_batches_tracked", (), "bn_count"),
    ]


def _block(prefix: str, cin: int, cout: int, downsample: bool):
    out = [(prefix + ".conv1.weight", (cout, cin, 3, 3), "conv")]
    out += _bn(prefix + ".bn1", cout)
    out += [(prefix + ".conv2.weight", (cout, cout, 3, 3), "conv")]
    out += _bn(prefix + ".bn2", cout)
    if downsample:
        out += [(prefix + ".downsample.0.weight", (cout, cin, 1, 1), "conv")]
        out += _bn(prefix + ".downsample.1", cout)
    return out


def state_dict_spec() -> List[Tuple[str, Tuple[int, ...], str]]:
    """(key, shape, kind) for all 252 entries, in the reference's registration order."""
    spec: List[Tuple[str, Tuple[int, ...], str]] = []
    b = BACKBONE + "0._layers."
    spec += [(b + "0.0.weight", (32, 1, 3, 3), "conv"), (b + "0.0.bias", (32,), "bias")]
    spec += _bn(b + "0.1", 32)
    cin = 32
    for li, (nblk, cout) in enumerate(zip(STAGE_BLOCKS, STAGE_PLANES)):
        for k in range(nblk):
            spec += _block(f"{b}{li + 1}.{k}", cin, cout, k == 0 and li > 0)
            cin = cout
    spec += [(BACKBONE + "1.weight", (72, 256, 1, 1), "conv"), (BACKBONE + "1.bias", (72,), "bias")]
    spec += [(FUSION + "0.weight", (108, 144, 1, 1), "conv"), (FUSION + "0.bias", (108,), "bias")]
    spec += _bn(FUSION + "1", 108)
    spec += [(FUSION + "3.weight", (72, 108, 1, 1), "conv"), (FUSION + "3.bias", (72,), "bias")]
    spec += _bn(FUSION + "4", 72)
    spec += [(FUSION + "6.weight", (72, 72, 1, 1), "conv"), (FUSION + "6.bias", (72,), "bias")]
    for i in (0, 2, 4):
        spec += [(f"{TEMPORAL}{i}.weight", (90, 90, 1, 1), "conv"),
                 (f"{TEMPORAL}{i}.bias", (90,), "bias")]
    spec += [(SKELETON + "0.weight", (144, 132), "linear"), (SKELETON + "0.bias", (144,), "bias")]
    spec += _bn(SKELETON + "2", 4)
    for name, ch, nout in (("_regressor_k", REG_K_CH, REG_K_OUT), ("_regressor_u", REG_U_CH, REG_U_OUT)):
        p = name + "._pose_regression_layers."
        spec += _block(p + "0", ch, ch, False)
        spec += _block(p + "1", ch, ch, False)
        spec += [(p + "2.weight", (nout, ch, 1, 1), "conv"), (p + "2.bias", (nout,), "bias")]
    return spec


# FLOP accounting used by bench.py and DESIGN.md (2 * MAC).
FLOP_BACKBONE_PER_CROP = 969_228_288
FLOP_PER_HAND_FRAME_KNOWN = 1_957_607_712
FLOP_PER_HAND_FRAME_UNKNOWN = 1_956_022_560
