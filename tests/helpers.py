"""Shared helpers for the GPU parity tests."""
import os

import numpy as np
import torch

from oracle import umetrack_oracle as O
from umetrack_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def hm_dict():
    return synthetic.load_hand_model_json(os.path.join(GOLDEN, "generic_hand_model_nomesh.json"))


def hm_tensors(hm=None):
    hm = hm or hm_dict()
    return {k: torch.tensor(v) for k, v in hm.items() if isinstance(v, list)}


def crop_params_from_row(labels, row):
    """(src params, crop f, crop c, crop xf) of one tracker_golden crop_rows entry."""
    t, h, c = int(row[0]), int(row[1]), int(row[2])
    cam = labels["cameras"][c]
    src_xf = np.array(labels["camera_to_world_transforms"][t][c])
    coeffs = [cam[n] for n in ("k1", "k2", "k3", "k4", "p1", "p2", "k5", "k6")]
    return t, h, c, (cam["fx"], cam["fy"]), (cam["cx"], cam["cy"]), coeffs, src_xf, row[3:5], row[5:7], row[7:].reshape(4, 4)


def pack_resample_params(sf, sc, coeffs, sxf, cf, cc, cxf):
    return np.concatenate([np.asarray(sf, float), np.asarray(sc, float), np.asarray(coeffs, float),
                           np.asarray(sxf, float).reshape(-1), np.asarray(cf, float), np.asarray(cc, float),
                           np.asarray(cxf, float).reshape(-1)])


def aggregate_inputs():
    """The seeded eval results the aggregate goldens were generated from
    (tests/golden/make_aggregate_golden.py:aggregate_inputs)."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("make_aggregate_golden",
                                                  os.path.join(GOLDEN, "make_aggregate_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.aggregate_inputs()
