"""Batched recording driver (device crop cameras + batched resample/encode + time scan) against the
reference's per-frame loop."""
import numpy as np
import pytest
import torch

from umetrack_b200 import hand_model as hmod
from umetrack_b200 import synthetic
from umetrack_b200.batched import crop_cameras_device, track_recordings_batched
from umetrack_b200.evaluation import track_recording
from umetrack_b200.metrics import MetricSums
from umetrack_b200.model import UmeTrackModel
from tests.helpers import golden, hm_dict

pytestmark = pytest.mark.gpu


def test_device_crop_cameras_match_reference_golden():
    g = golden("tracker_golden.npz")
    hm = hm_dict()
    labels, _ = synthetic.make_recording(hm, seed=3, n_frames=12)
    dev = torch.device("cuda")
    cc = crop_cameras_device(hmod.load_hand_model_from_dict(hm), labels, dev)
    cnt = cc["sel_count"].cpu().numpy()
    sel = cc["sel_cam"].cpu().numpy()
    par = cc["resample_params"].cpu().numpy()
    rows = g["crop_rows"]
    assert int(cnt.sum()) == len(rows)          # same hands, same number of views (dropouts included)
    seen = {}
    for row in rows:
        t, h, c = int(row[0]), int(row[1]), int(row[2])
        k = seen.get((t, h), 0)
        seen[(t, h)] = k + 1
        assert sel[t, h, k] == c                # same cameras, in index order
        p = par[t, h, k]
        np.testing.assert_allclose(p[28:32], row[3:7], rtol=2e-6)                    # crop f, c
        np.testing.assert_allclose(p[32:48], row[7:], atol=2e-4, rtol=2e-6)          # crop camera_to_world (mm)


def test_batched_driver_matches_reference_and_per_frame_tracker():
    g = golden("tracker_golden.npz")
    hm = hm_dict()
    rec_a = synthetic.make_recording(hm, seed=3, n_frames=12)
    rec_b = synthetic.make_recording(hm, seed=4, n_frames=9)      # ragged lengths
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    sums = MetricSums("cuda")
    res = track_recordings_batched(model, [rec_a, rec_b], sums=sums)
    a = res[0]
    assert np.array_equal(a["valid_tracking"], g["valid"])
    d = np.abs(a["tracked_keypoints"] - g["tracked"])[g["valid"]]
    assert d.max() < 0.05, f"keypoints differ from the reference by {d.max()} mm"
    assert abs(a["per_frame_mean_error"].mean() - g["per_frame_error"].mean()) < 0.01
    # second recording against our own per-frame tracker (fresh model: independent memory)
    ref_b = track_recording(UmeTrackModel(synthetic.make_state_dict(0)).to("cuda"), *rec_b)
    assert np.array_equal(res[1]["valid_tracking"], ref_b["valid_tracking"])
    assert np.abs(res[1]["tracked_keypoints"] - ref_b["tracked_keypoints"]).max() < 0.05
    both = np.concatenate([a["per_frame_mean_error"], res[1]["per_frame_mean_error"]])
    assert abs(sums.result()["mpjpe"] - both.mean()) < 0.01


def test_frame_staging_variants_give_identical_results(tmp_path):
    """Recording ingest: frames as numpy (pinned ring), pinned tensor, pageable
    tensor or CUDA tensor, with a chunk size that does not divide T and forces slot reuse -- the
    tracked keypoints are bit-identical in every case; labels round-trip through the JSON file."""
    import json

    from umetrack_b200.ingest import FrameStager, load_recording_labels

    hm = hm_dict()
    labels, frames = synthetic.make_recording(hm, seed=5, n_frames=23)
    path = tmp_path / "rec.json"
    path.write_text(json.dumps({k: (np.asarray(v).tolist() if isinstance(v, np.ndarray) else v)
                                for k, v in labels.items()}))
    lab2 = load_recording_labels(str(path))
    with pytest.raises(KeyError):
        bad = tmp_path / "bad.json"
        bad.write_text(json.dumps({"cameras": []}))
        load_recording_labels(str(bad))
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    base = track_recordings_batched(model, [(labels, frames)], chunk_frames=64)[0]["tracked_keypoints"]
    variants = {"numpy": frames, "pinned": torch.from_numpy(frames).pin_memory(),
                "pageable": torch.from_numpy(frames), "cuda": torch.from_numpy(frames).cuda()}
    for name, fr in variants.items():
        out = track_recordings_batched(model, [(lab2, fr)], chunk_frames=5)[0]["tracked_keypoints"]
        assert np.array_equal(out, base), name
    # the stager itself: chunks cover [0, T) in order and carry the right bytes
    st = FrameStager("cuda", chunk_frames=4, n_slots=2)
    seen = 0
    for t0, t1, chunk, slot in st.chunks(frames):
        assert t0 == seen and chunk.shape[0] == t1 - t0
        assert torch.equal(chunk.cpu(), torch.from_numpy(frames[t0:t1]))
        st.release(slot)
        seen = t1
    assert seen == frames.shape[0] and st.bytes_uploaded == frames.size
    with pytest.raises(ValueError):
        list(st.chunks(frames.astype(np.float32)))


def test_batched_driver_edge_cases():
    """Recordings without any trackable hand, a single-frame recording and an empty list."""
    hm = hm_dict()
    model = UmeTrackModel(synthetic.make_state_dict(0)).to("cuda")
    lab, fr = synthetic.make_recording(hm, seed=9, n_frames=6)
    dead = dict(lab)
    dead["hand_confidences"] = (np.asarray(lab["hand_confidences"]) * 0.0).tolist()
    one_lab = {k: (v[:1] if k in ("joint_angles", "wrist_transforms", "hand_confidences", "camera_to_world_transforms") else v)
               for k, v in lab.items()}
    sums = MetricSums("cuda")
    res = track_recordings_batched(model, [(dead, fr), (lab, fr), (one_lab, fr[:1])], sums=sums)
    assert not res[0]["valid_tracking"].any() and res[0]["per_frame_mean_error"].shape == (0,)
    assert np.all(res[0]["tracked_keypoints"] == 0)
    ref = track_recordings_batched(model, [(lab, fr)])[0]
    assert np.array_equal(res[1]["tracked_keypoints"], ref["tracked_keypoints"])      # neighbours do not matter
    assert res[2]["valid_tracking"].shape == (2, 1)
    assert np.array_equal(res[2]["tracked_keypoints"][:, 0], ref["tracked_keypoints"][:, 0])   # frame 0: no memory yet
    r = sums.result()
    assert r["n_hand_frames"] == float(res[1]["valid_tracking"].sum() + res[2]["valid_tracking"].sum())
    assert r["n_slots"] == 2 * 6 + 2 * 6 + 2
