"""umetrack_b200: UmeTrack's per-frame multi-view hand-tracking inference path on B200 (sm_100a).

Drop-in Python surface (same names as the reference's lib/tracker and lib/models):
    load_pretrained_model, UmeTrackModel, InputFrameData, InputFrameDesc, InputSkeletonData,
    RegressorOutput, HandTracker, HandTrackerOpts, InputFrame, ViewData, SingleHandPose,
    TrackingResult, landmarks_from_hand_pose, skin_landmarks, HandModel, ...
The arithmetic lives in libumetrack_b200.so (include/umetrack_b200.h); importing the package
does not need a GPU, running it does.
"""
from .arch import ModelOpts  # noqa: F401

__all__ = ["ModelOpts"]


def __getattr__(name):
    # lazy re-exports so `import umetrack_b200` stays cheap and works without torch.cuda
    import importlib

    table = {
        "load_pretrained_model": "model", "UmeTrackModel": "model", "InputFrameData": "model",
        "InputFrameDesc": "model", "InputSkeletonData": "model", "RegressorOutput": "model",
        "HandTracker": "hand_tracker", "HandTrackerOpts": "hand_tracker", "InputFrame": "hand_tracker",
        "ViewData": "hand_tracker", "SingleHandPose": "hand_tracker", "TrackingResult": "hand_tracker",
        "landmarks_from_hand_pose": "hand_tracker", "gen_crop_cameras_from_pose": "hand_tracker",
        "skin_landmarks": "skinning", "HandModel": "hand_model", "scaled_hand_model": "hand_model",
        "mirrored_hand_model": "hand_model", "load_hand_model_from_dict": "hand_model",
        "read_camera_from_json": "cameras", "PinholePlaneCameraModel": "cameras",
        "Fisheye62CameraModel": "cameras", "MetricSums": "metrics",
        "aggregate_metrics": "aggregate", "save_eval_result": "aggregate", "PCK_curve": "aggregate",
        "normalized_AUC": "metrics", "PCK_THRESHOLDS": "metrics",
    }
    if name in table:
        return getattr(importlib.import_module(f".{table[name]}", __name__), name)
    raise AttributeError(name)
