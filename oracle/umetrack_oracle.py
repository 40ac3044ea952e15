"""CPU oracle for the UmeTrack per-frame multi-view inference path.

TEST INFRASTRUCTURE ONLY.  This file is a CPU restatement (numpy float64 / int32 and
torch-CPU float32) of the reference algorithm for the hot path.  It may be imported by
`tests/`, by `__graft_entry__.smoke()` and by `bench.py`'s cpu_baseline / `--impl reference`
legs, and nowhere else: the product (`umetrack_b200/`) never imports it and has no CPU fallback.

Parity status
-------------
* Pinned: every function below is checked against outputs of the reference itself
  (`/root/reference`, imported unchanged with a `pytorch3d.so3_exp_map` shim and an empty
  `av` stub) on seeded synthetic inputs; the outputs are committed under `tests/golden/`
  together with the generating script `tests/golden/make_golden.py`.
  `remap_u8_bilinear` is additionally checked against `cv2.remap` 4.13.0 (third-party,
  unpinned in the reference: README.md:10).
* Parity UNPINNED at one boundary: `so3_exp_map` (pytorch3d `@stable`, not vendored, not
  installable here; call site lib/common/hand_skinning.py:46).  It is restated from the
  published pytorch3d 0.7.x algorithm (clamped-angle Rodrigues formula); the golden FK vectors
  were produced with the same restatement injected into the reference.

All `file:line` citations are relative to `/root/reference/`.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

CROP = 96
FEAT_HW = 6
N_FEAT = 72
N_MEM = 18
N_SKEL = 4
BN_EPS = 1e-5

# ----------------------------------------------------------------------------------------
# Stage (a): fisheye -> pinhole crop resampling
# ----------------------------------------------------------------------------------------


def fisheye62_distort(p: np.ndarray, coeffs: Sequence[float]) -> np.ndarray:
    """lib/common/camera.py:120-141 (Fisheye62CameraModel.evaluate), float64."""
    k1, k2, k3, k4, p1, p2, k5, k6 = [float(c) for c in coeffs]
    r2 = (p * p).sum(axis=-1, keepdims=True)
    r2 = np.clip(r2, -np.pi**2, np.pi**2)
    r4 = r2 * r2
    r6 = r2 * r4
    r8 = r4 * r4
    r10 = r4 * r6
    r12 = r6 * r6
    radial = 1 + k1 * r2 + k2 * r4 + k3 * r6 + k4 * r8 + k5 * r10 + k6 * r12
    uv = p * radial
    x = uv[..., 0].copy()
    y = uv[..., 1].copy()
    x2, y2, xy = x * x, y * y, x * y
    rr = x2 + y2
    xo = x + (2 * p2 * xy + p1 * (rr + 2 * x2))
    yo = y + (2 * p1 * xy + p2 * (rr + 2 * y2))
    return np.stack((xo, yo), axis=-1)


def arctan_project(v: np.ndarray) -> np.ndarray:
    """lib/common/camera.py:79-85 (ArctanProjection.project)."""
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    r = np.sqrt(x * x + y * y)
    s = np.arctan2(r, z) / np.maximum(r, 2.0**-128)
    return np.stack((x * s, y * s), axis=-1)


def warp_map(
    src_f: Sequence[float],
    src_c: Sequence[float],
    src_coeffs: Sequence[float],
    src_cam_to_world: np.ndarray,
    crop_f: Sequence[float],
    crop_c: Sequence[float],
    crop_cam_to_world: np.ndarray,
    width: int = CROP,
    height: int = CROP,
) -> Tuple[np.ndarray, np.ndarray]:
    """Float32 remap tables of `_warp_image`, lib/tracker/tracker.py:54-78.

    window_to_eye (camera.py:275-281, 68-75; affine.py:22-24) -> eye_to_world (:263-267)
    -> world_to_eye (:257-261) -> eye_to_window (:269-273); points behind the source camera
    get map value -1 (tracker.py:71-73); cast to float32 (:75).
    """
    src_m = np.asarray(src_cam_to_world, dtype=np.float64)
    crop_m = np.asarray(crop_cam_to_world, dtype=np.float64)
    px, py = np.meshgrid(np.arange(width), np.arange(height))
    win = np.column_stack((px.flatten(), py.flatten())).astype(np.float64)
    q = (win - np.asarray(crop_c, dtype=np.float64)) / np.asarray(crop_f, dtype=np.float64)
    v = np.concatenate([q, np.ones((q.shape[0], 1))], axis=1)
    v = v / np.maximum(5.43e-20, (v * v).sum(axis=-1, keepdims=True) ** 0.5)
    world = v @ crop_m[:3, :3].T + crop_m[:3, 3]
    eye = (world - src_m[:3, 3]) @ src_m[:3, :3]
    pq = fisheye62_distort(arctan_project(eye), src_coeffs)
    w = pq * np.asarray(src_f, dtype=np.float64) + np.asarray(src_c, dtype=np.float64)
    w[eye[:, 2] < 0] = -1
    w = w.astype(np.float32)
    return w[:, 0].reshape(height, width), w[:, 1].reshape(height, width)


def remap_fixed_point_indices(map_x: np.ndarray, map_y: np.ndarray):
    """Fixed-point coordinates of cv2.remap(INTER_LINEAR) on 8-bit images (third-party
    opencv-python; call site lib/tracker/tracker.py:80).  Coordinates are quantised to 1/32
    pixel with round-half-to-even in float32: sx = cvRound(mx * 32)."""
    sx = np.rint(map_x.astype(np.float32) * np.float32(32.0)).astype(np.int64)
    sy = np.rint(map_y.astype(np.float32) * np.float32(32.0)).astype(np.int64)
    sx = np.clip(sx, -(2**31), 2**31 - 1)
    sy = np.clip(sy, -(2**31), 2**31 - 1)
    ix = np.clip(sx >> 5, -32768, 32767).astype(np.int32)
    iy = np.clip(sy >> 5, -32768, 32767).astype(np.int32)
    fx = (sx & 31).astype(np.int32)
    fy = (sy & 31).astype(np.int32)
    return ix, iy, fx, fy


def remap_u8_bilinear(src: np.ndarray, map_x: np.ndarray, map_y: np.ndarray) -> np.ndarray:
    """Bit-exact model of cv2.remap(src_u8, map_x, map_y, INTER_LINEAR), BORDER_CONSTANT(0).

    Four taps weighted by integer weights wx in [0,32], wy in [0,32] (OpenCV's int16 table is
    exactly 32*wx*wy for bilinear), per-tap zero border, out = (sum + 512) >> 10.
    """
    H, W = src.shape
    ix, iy, fx, fy = remap_fixed_point_indices(map_x, map_y)
    acc = np.zeros(ix.shape, dtype=np.int64)
    s = src.astype(np.int64)
    for dy in (0, 1):
        wy = (32 - fy) if dy == 0 else fy
        yy = iy.astype(np.int64) + dy
        for dx in (0, 1):
            wx = (32 - fx) if dx == 0 else fx
            xx = ix.astype(np.int64) + dx
            ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
            val = np.where(ok, s[np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)], 0)
            acc += val * wx * wy
    return ((acc + 512) >> 10).astype(np.uint8)


def warp_image(src_image, src_f, src_c, src_coeffs, src_cam_to_world, crop_f, crop_c,
               crop_cam_to_world, width=CROP, height=CROP) -> np.ndarray:
    """`_warp_image` end to end (lib/tracker/tracker.py:54-80): uint8 [H,W] -> uint8 [96,96]."""
    mx, my = warp_map(src_f, src_c, src_coeffs, src_cam_to_world, crop_f, crop_c,
                      crop_cam_to_world, width, height)
    return remap_u8_bilinear(np.ascontiguousarray(src_image), mx, my)


# ----------------------------------------------------------------------------------------
# Stage (b): encoder, fusion, temporal memory, regressor  (torch CPU float32)
# ----------------------------------------------------------------------------------------

BACKBONE_BLOCKS = (2, 3, 5, 2)       # "resnet_layers_2352" (lib/models/backbone_resnet.py:169-192)
BACKBONE_PLANES = (32, 64, 128, 256)  # "-f32" start planes (backbone_resnet.py:161-166)


def _bn(x, sd, prefix):
    return F.batch_norm(x, sd[prefix + ".running_mean"], sd[prefix + ".running_var"],
                        sd[prefix + ".weight"], sd[prefix + ".bias"], False, 0.0, BN_EPS)


def _basic_block(x, sd, prefix, stride):
    """lib/models/backbone_resnet.py:56-72."""
    out = F.conv2d(x, sd[prefix + ".conv1.weight"], None, stride, 1)
    out = F.relu(_bn(out, sd, prefix + ".bn1"))
    out = F.conv2d(out, sd[prefix + ".conv2.weight"], None, 1, 1)
    out = _bn(out, sd, prefix + ".bn2")
    if (prefix + ".downsample.0.weight") in sd:
        res = F.conv2d(x, sd[prefix + ".downsample.0.weight"], None, stride, 0)
        res = _bn(res, sd, prefix + ".downsample.1")
    else:
        res = x
    return F.relu(out + res)


def backbone_forward(sd: Dict[str, torch.Tensor], images: torch.Tensor,
                     taps: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
    """images [n,96,96] in [0,1] -> [n,72,6,6].  lib/models/model_utils.py:107-138,
    backbone_resnet.py:75-158.  `taps`, if given, receives intermediate activations."""
    p = "_feature_extractor._image_backbone."
    x = images.unsqueeze(1)
    x = F.conv2d(x, sd[p + "0._layers.0.0.weight"], sd[p + "0._layers.0.0.bias"], 1, 1)
    x = F.relu(_bn(x, sd, p + "0._layers.0.1"))
    x = F.max_pool2d(x, 2, 2)
    if taps is not None:
        taps["stem"] = x
    for li, nblk in enumerate(BACKBONE_BLOCKS):
        for b in range(nblk):
            stride = 2 if (b == 0 and li > 0) else 1
            x = _basic_block(x, sd, f"{p}0._layers.{li + 1}.{b}", stride)
            if taps is not None:
                taps[f"layer{li + 1}.{b}"] = x
    x = F.conv2d(x, sd[p + "1.weight"], sd[p + "1.bias"])
    return x


def apply_ftl(xfs: torch.Tensor, fmap: torch.Tensor) -> torch.Tensor:
    """Feature-transform layer with ratio 1.0: lib/models/model_utils.py:57-104."""
    n = fmap.shape[0]
    pts = fmap.reshape(n, 3, -1)
    pts = torch.matmul(xfs[:, :3, :3], pts) + xfs[:, :3, 3].unsqueeze(-1)
    return pts.reshape(fmap.shape)


def singlev_xfs(intrinsics: torch.Tensor) -> torch.Tensor:
    """S = diag(1,1,fx/200,1): lib/models/model_utils.py:166-192."""
    n = intrinsics.shape[0]
    s = torch.eye(4, dtype=intrinsics.dtype).unsqueeze(0).repeat(n, 1, 1)
    s[:, 2, 2] = intrinsics[:, 0, 0] / 200.0
    return s


def fusion_forward(sd, feats2: torch.Tensor, s2: torch.Tensor, e2: torch.Tensor) -> torch.Tensor:
    """feats2 [B,2,72,6,6], s2 [B,2,4,4], e2 [B,2,4,4] -> [B,72,6,6] in cam0 space.
    lib/models/feature_extractor.py:61-141 (useUnscaledAsCanonical=False)."""
    p = "_feature_extractor._multi_view_fusion."
    xf0 = e2[:, 0:1]
    to_world = torch.inverse(e2) @ s2
    s0_inv = torch.inverse(s2[:, 0:1])
    to_canon = s0_inv @ xf0 @ to_world
    B = feats2.shape[0]
    canon = apply_ftl(to_canon.reshape(-1, 4, 4), feats2.flatten(0, 1)).reshape(feats2.shape)
    x = canon.flatten(1, 2)
    x = F.relu(_bn(F.conv2d(x, sd[p + "0.weight"], sd[p + "0.bias"]), sd, p + "1"))
    x = F.relu(_bn(F.conv2d(x, sd[p + "3.weight"], sd[p + "3.bias"]), sd, p + "4"))
    x = F.conv2d(x, sd[p + "6.weight"], sd[p + "6.bias"])
    return apply_ftl(s2[:, 0], x)


def feature_extractor_forward(sd, images, intrinsics, extrinsics, sample_range) -> torch.Tensor:
    """lib/models/umetrack_model.py:123-168: per-sample 1- or 2-view features [B,72,6,6]."""
    feats = backbone_forward(sd, images)
    s = singlev_xfs(intrinsics)
    out = []
    for r0, r1 in sample_range.tolist():
        if r1 - r0 == 1:
            out.append(apply_ftl(s[r0:r1], feats[r0:r1]))
        else:
            out.append(fusion_forward(sd, feats[r0:r1].unsqueeze(0), s[r0:r1].unsqueeze(0),
                                      extrinsics[r0:r1].unsqueeze(0)))
    return torch.cat(out, dim=0)


class TemporalState:
    """Recurrent state of SimpleConvRNN (lib/models/temporal.py:41-42)."""

    def __init__(self):
        self.mem = torch.empty(0)
        self.prev_extr = torch.empty(0)


def temporal_forward(sd, state: TemporalState, img_features, cur_extr, memory_idx, use_memory):
    """lib/models/temporal.py:51-139."""
    p = "_temporal._temporal_module."
    need = int(memory_idx.max()) + 1
    mem, prev = state.mem, state.prev_extr
    if len(mem) < need:
        nm = torch.zeros(need, N_MEM, FEAT_HW, FEAT_HW)
        ne = torch.zeros(need, 4, 4)
        if len(mem):
            nm[: len(mem)] = mem
            ne[: len(prev)] = prev
        mem, prev = nm, ne
    use_idx = memory_idx[use_memory]
    if len(use_idx) != len(use_memory):
        z = memory_idx[~use_memory]
        mem[z] = 0
        prev[z] = 0
    if len(use_idx):
        xf = cur_extr[use_memory].bmm(torch.inverse(prev[use_idx]))
        mem[use_idx] = apply_ftl(xf, mem[use_idx])
    prev[memory_idx] = cur_extr
    x = torch.cat([mem[memory_idx], img_features], dim=1)
    x = F.relu(F.conv2d(x, sd[p + "0.weight"], sd[p + "0.bias"]))
    x = F.relu(F.conv2d(x, sd[p + "2.weight"], sd[p + "2.bias"]))
    x = F.conv2d(x, sd[p + "4.weight"], sd[p + "4.bias"])
    mem[memory_idx] = x[:, :N_MEM]
    state.mem, state.prev_extr = mem, prev
    return x[:, N_MEM:]


def skeleton_encode(sd, axes: torch.Tensor, rest: torch.Tensor) -> torch.Tensor:
    """lib/models/skeleton_encoder.py:28-53: [n,22,3]x2 -> [n,4,6,6]."""
    p = "_skeleton_enc._layers."
    n = int(np.prod(axes.shape[:-2]))
    x = torch.cat((axes, rest), dim=-1).reshape(n, -1)
    x = F.linear(x, sd[p + "0.weight"], sd[p + "0.bias"]).view(-1, N_SKEL, FEAT_HW, FEAT_HW)
    return F.relu(_bn(x, sd, p + "2"))


def rigid_sample_points() -> torch.Tensor:
    """lib/models/regressor.py:19-47."""
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, -1, 0], [-1, 0, -1],
                    [0, -1, -1]], dtype=np.float64)
    out = np.empty(pts.shape)
    for i, row in enumerate(pts):
        nrm = np.linalg.norm(row)
        out[i] = row if nrm == 0 else row / nrm * 0.1
    return torch.from_numpy(out).float()


def procrustes_align(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    """lib/models/model_utils.py:17-54."""
    B = src.shape[0]
    mu_s, mu_d = src.mean(dim=1), dst.mean(dim=1)
    cs, cd = src - mu_s.reshape(-1, 1, 3), dst - mu_d.reshape(-1, 1, 3)
    h = torch.matmul(cs.transpose(1, 2), cd)
    u, _, v = h.svd()
    vut = torch.matmul(v, u.transpose(1, 2))
    w = torch.eye(3).unsqueeze(0).repeat(B, 1, 1)
    w[:, 2, 2] = torch.det(vut)
    xf = torch.eye(4).unsqueeze(0).repeat(B, 1, 1)
    xf[:, :3, :3] = torch.matmul(torch.matmul(v, w), u.transpose(1, 2))
    xf[:, :3, 3] = mu_d - torch.matmul(xf[:, :3, :3], mu_s.unsqueeze(-1)).squeeze(-1)
    return xf


def regressor_forward(sd, prefix: str, x: torch.Tensor, with_scale: bool):
    """lib/models/regressor.py:163-186 + model_utils.py:195-208; returns dict of decoded outputs."""
    p = prefix + "._pose_regression_layers."
    x = _basic_block(x, sd, p + "0", 1)
    x = _basic_block(x, sd, p + "1", 1)
    x = F.conv2d(x, sd[p + "2.weight"], sd[p + "2.bias"])
    f = F.adaptive_avg_pool2d(x, (1, 1)).flatten(1)
    B = f.shape[0]
    out = {}
    out["joint_angles"] = torch.cat([f[:, 0:20], torch.zeros(B, 2)], dim=1)
    src = rigid_sample_points().unsqueeze(0).expand(B, 7, 3).clone()
    out["wrist_xfs"] = procrustes_align(src, f[:, 20:41].reshape(B, 7, 3))
    o = 41
    if with_scale:
        out["skel_scales"] = torch.exp(f[:, 41:42].reshape(-1))
        o = 42
    out["landmark_uncertainty_sigmas"] = torch.clamp(F.softplus(f[:, o:o + 21]), min=1e-5)
    return out


def recover_world(hand_idx, cam0_extr, wrist_cam0):
    """lib/models/umetrack_model.py:77-90."""
    w = torch.inverse(cam0_extr) @ wrist_cam0
    w = w.clone()
    w[hand_idx == 1, :, 0] *= -1
    return w


class OracleModel:
    """Functional mirror of UmeTrackModel (lib/models/umetrack_model.py:100-242) on a flat
    state_dict with the reference's key names (lib/models/model_loader.py:53-88)."""

    def __init__(self, state_dict: Dict[str, torch.Tensor]):
        self.sd = {k: v.detach().clone() for k, v in state_dict.items()}
        self.state = TemporalState()

    def _temporal_features(self, images, intrinsics, extrinsics, sample_range, memory_idx,
                           use_memory):
        f = feature_extractor_forward(self.sd, images, intrinsics, extrinsics, sample_range)
        cam0 = extrinsics[sample_range[:, 0]]
        return temporal_forward(self.sd, self.state, f, cam0, memory_idx, use_memory), cam0

    @torch.no_grad()
    def regress_pose_use_skeleton(self, images, intrinsics, extrinsics, sample_range,
                                  memory_idx, use_memory, hand_idx, skel_axes, skel_rest):
        t, cam0 = self._temporal_features(images, intrinsics, extrinsics, sample_range,
                                          memory_idx, use_memory)
        sk = skeleton_encode(self.sd, skel_axes, skel_rest)
        if sk.shape[0] == 1 and t.shape[0] > 1:
            sk = sk.expand(t.shape[0], *sk.shape[1:])
        out = regressor_forward(self.sd, "_regressor_k", torch.cat([t, sk], dim=1), False)
        out["wrist_xfs"] = recover_world(hand_idx, cam0, out["wrist_xfs"])
        return out

    @torch.no_grad()
    def regress_pose_pred_skel_scale(self, images, intrinsics, extrinsics, sample_range,
                                     memory_idx, use_memory, hand_idx):
        assert bool(((sample_range[:, 1] - sample_range[:, 0]) != 1).all())
        t, cam0 = self._temporal_features(images, intrinsics, extrinsics, sample_range,
                                          memory_idx, use_memory)
        out = regressor_forward(self.sd, "_regressor_u", t, True)
        out["wrist_xfs"] = recover_world(hand_idx, cam0, out["wrist_xfs"])
        return out


# ----------------------------------------------------------------------------------------
# Stage (c): forward kinematics, skinning, metrics
# ----------------------------------------------------------------------------------------


def so3_exp_map(log_rot: torch.Tensor, eps: float = 1e-4) -> torch.Tensor:
    """pytorch3d.transforms.so3_exp_map (third-party, `@stable`, unpinned; call site
    lib/common/hand_skinning.py:46).  PARITY UNPINNED: restated from pytorch3d 0.7.x."""
    x, y, z = log_rot.unbind(1)
    zero = torch.zeros_like(x)
    K = torch.stack([zero, -z, y, z, zero, -x, -y, x, zero], dim=1).reshape(-1, 3, 3)
    th = torch.clamp((log_rot * log_rot).sum(1), eps).sqrt()
    fac1 = (th.sin() / th)[:, None, None]
    fac2 = ((1 - th.cos()) / (th * th))[:, None, None]
    return fac1 * K + fac2 * torch.bmm(K, K) + torch.eye(3, dtype=log_rot.dtype)[None]


def skin_landmarks(axes, rest, lm_rest, lm_idx, lm_w, joint_angles, wrist) -> torch.Tensor:
    """lib/common/hand_skinning.py:17-209 for a batch: joint_angles [B,22], wrist [B,4,4],
    one hand model (axes/rest [22,3], lm_rest [21,3], lm_idx/lm_w [21,3]) -> [B,21,3]."""
    B = joint_angles.shape[0]
    axes, rest = axes.float(), rest.float()
    aa = (axes[None, :20] * joint_angles[:, :20, None].float()).reshape(-1, 3)
    R = so3_exp_map(aa)
    rp = rest[None, :20].expand(B, 20, 3).reshape(-1, 3)
    t = rp - torch.matmul(R, rp.unsqueeze(-1)).squeeze(-1)
    L = torch.eye(4).unsqueeze(0).repeat(B * 20, 1, 1)
    L[:, :3, :3] = R
    L[:, :3, 3] = t
    L = L.reshape(B, 20, 4, 4)
    wrist = wrist.float()
    frames = [wrist, wrist]
    for f in range(5):
        cur = wrist
        chain = []
        for j in range(4):
            cur = torch.matmul(cur, L[:, 4 * f + j])
            chain.append(cur)
        frames += chain[1:]
    T = torch.stack(frames, dim=1)  # [B,17,4,4]
    skin = torch.zeros(21, 17)
    for l in range(21):
        for k in range(3):
            if float(lm_w[l, k]) != 0:
                skin[l, int(lm_idx[l, k])] = float(lm_w[l, k])
    homo = torch.cat([lm_rest.float(), torch.ones(21, 1)], dim=1)            # [21,4]
    weighted = homo[None, :, None, :] * skin[None, :, :, None]                # [1,21,17,4]
    pts = torch.matmul(T.unsqueeze(1), weighted.unsqueeze(-1)).sum(dim=2).squeeze(-1)
    return pts[..., :3]


def landmarks_from_hand_pose(model: dict, joint_angles: np.ndarray, wrist: np.ndarray,
                             hand_idx: int) -> np.ndarray:
    """lib/tracker/perspective_crop.py:28-51 (casts to float32; right hand: column 0 negated)."""
    xf = np.array(wrist, copy=True)
    if hand_idx == 1:
        xf[:, 0] *= -1
    out = skin_landmarks(model["joint_rotation_axes"], model["joint_rest_positions"],
                         model["landmark_rest_positions"], model["landmark_rest_bone_indices"],
                         model["landmark_rest_bone_weights"],
                         torch.from_numpy(np.asarray(joint_angles)).float()[None],
                         torch.from_numpy(xf).float()[None])
    return out[0].numpy()


def mpjpe_per_frame(gt: np.ndarray, pred: np.ndarray, valid: np.ndarray) -> np.ndarray:
    """run_eval_known_skeleton.py:88-89; arrays [2,T,21,3], valid [2,T]."""
    d = (gt - pred)[valid]
    return np.linalg.norm(d, axis=-1).mean(axis=-1)


def accel_per_frame(pts: np.ndarray, valid: np.ndarray) -> np.ndarray:
    """load_eval.py:29-44: mean over 21 of |p[t]+p[t+2]-2p[t+1]|, 3-consecutive-valid mask."""
    acc = pts[:, 0:-2] + pts[:, 2:] - 2 * pts[:, 1:-1]
    a = np.linalg.norm(acc, axis=-1).mean(axis=-1)
    ok = valid[:, 0:-2] & valid[:, 1:-1] & valid[:, 2:]
    return a[ok]


# ----------------------------------------------------------------------------------------
# Stage (a), torch_data variant: pinhole -> pinhole homography resampling
# ----------------------------------------------------------------------------------------


def resample_matrix(k_orig: np.ndarray, orig_cam_to_world: np.ndarray, k_new: np.ndarray,
                    new_cam_to_world: np.ndarray) -> np.ndarray:
    """lib/batched_dataset/data_transform.py:57-75: K_orig * E_orig * E_new^-1 * K_new^-1 as
    float32 4x4, taking a pixel (u, v, 1) of the new image to a pixel of the original."""
    k_inv_new = np.eye(4)
    k_inv_new[:3, :3] = np.linalg.inv(k_new)
    k_o = np.eye(4)
    k_o[:3, :3] = k_orig
    m = k_o @ np.linalg.inv(orig_cam_to_world) @ new_cam_to_world @ k_inv_new
    return m.astype(np.float32)


def resample_homography(images_orig: np.ndarray, resample_xfs: np.ndarray, out_h: int = CROP,
                        out_w: int = CROP) -> np.ndarray:
    """lib/batched_dataset/data_transform.py:78-144 (`_resample_images_batched`): float32 images
    [n,h,w], float32 transforms [n,4,4] -> float32 [n,out_h,out_w]; float64 coordinates
    (float32 x int32 promotes), pixels outside 0 <= x < w-1, 0 <= y < h-1 stay 0, truncation to
    the tap origin, float64 bilinear blend stored as float32."""
    n, h, w = images_orig.shape
    r = resample_xfs[:, 0:3, 0:3].astype(np.float64)
    t = resample_xfs[:, 0:3, 3].astype(np.float64)
    v, u = np.mgrid[0:out_h, 0:out_w]
    grid = np.stack([u, v, np.ones_like(u)], axis=-1).astype(np.float64)      # [H,W,3]
    xyz = np.einsum("nij,hwj->nhwi", r, grid) + t[:, None, None, :]
    x = xyz[..., 0] / xyz[..., 2]
    y = xyz[..., 1] / xyz[..., 2]
    with np.errstate(invalid="ignore"):
        ok = (x >= 0) & (x < w - 1) & (y >= 0) & (y < h - 1)
    out = np.zeros((n, out_h, out_w), dtype=np.float32)
    idx = np.nonzero(ok)
    xs, ys = x[ok], y[ok]
    x0 = xs.astype(np.int32)
    y0 = ys.astype(np.int32)
    x1, y1 = x0 + 1, y0 + 1
    img = idx[0]
    f00 = images_orig[img, y0, x0]
    f01 = images_orig[img, y1, x0]
    f10 = images_orig[img, y0, x1]
    f11 = images_orig[img, y1, x1]
    val = (f00 * (x1 - xs) * (y1 - ys) + f10 * (xs - x0) * (y1 - ys) + f01 * (x1 - xs) * (ys - y0)
           + f11 * (xs - x0) * (ys - y0)) / ((x1 - x0) * (y1 - y0))
    out[idx] = val
    return out


# ----------------------------------------------------------------------------------------
# Aggregate metrics: PCK curve, normalised AUC, success rate
# ----------------------------------------------------------------------------------------

MAX_LANDMARK_ERROR_MM = 50                                   # lib/common/metric_utils.py:19
PCK_THRESHOLDS = np.linspace(0, MAX_LANDMARK_ERROR_MM, 101)  # lib/common/metric_utils.py:20


def pck_curve(errors: np.ndarray, thresholds: np.ndarray, mask: Optional[np.ndarray] = None,
              axis: Optional[int] = None) -> np.ndarray:
    """lib/common/metric_utils.py:36-88: fraction of (masked) errors <= each threshold; one curve
    per element along `axis` if given.  Rows with an empty mask give 0 (their _safe_div)."""
    if mask is None:
        mask = np.ones_like(errors)
    if axis is None:
        e, m = errors.reshape(-1), mask.reshape(-1)
    else:
        n = errors.shape[axis]
        e, m = np.moveaxis(errors, axis, 0).reshape(n, -1), np.moveaxis(mask, axis, 0).reshape(n, -1)
    rows = []
    for t in thresholds:
        num, den = ((e <= t) * m).sum(axis=-1), m.sum(axis=-1)
        with np.errstate(divide="ignore", invalid="ignore"):
            z = np.asarray(num / den, dtype=np.float64)
        rows.append(np.where(den < 1e-6, 0.0, z))
    return np.stack(rows).T


def normalized_auc(x: np.ndarray, y: np.ndarray, y_max: float = 1.0) -> np.ndarray:
    """lib/common/metric_utils.py:91-112: trapezoid area over (x[-1]-x[0]) * y_max."""
    shape = y.shape[:-1]
    y2 = y.reshape(-1, y.shape[-1])
    auc = ((x[1:] - x[:-1]).reshape(1, -1) * ((y2[..., 1:] + y2[..., :-1]) * 0.5)).sum(axis=-1)
    return (auc / ((x[-1] - x[0]) * y_max)).reshape(shape)


def aggregate_metrics(records) -> dict:
    """load_eval.py:47-89 on a list of {"tracked_keypoints","gt_keypoints","valid_tracking"} dicts:
    the five numbers it prints plus the PCK curve (in percent, as it passes it to the AUC)."""
    errs = np.concatenate([mpjpe_per_frame(r["gt_keypoints"], r["tracked_keypoints"], r["valid_tracking"]) for r in records])
    acc = np.concatenate([accel_per_frame(r["tracked_keypoints"], r["valid_tracking"]) for r in records])
    gacc = np.concatenate([accel_per_frame(r["gt_keypoints"], r["valid_tracking"]) for r in records])
    pck = pck_curve(errs, PCK_THRESHOLDS) * 100.0
    valid = np.concatenate([r["valid_tracking"] for r in records], axis=1)
    return {"n_valid": int(valid.sum()), "n_total": int(valid.size),
            "success_rate": float(valid.sum() / valid.size * 100), "mean_keypoint_error": float(errs.mean()),
            "auc": float(normalized_auc(PCK_THRESHOLDS, pck)), "pck_curve": pck,
            "mean_keypoint_acceleration": float(acc.mean()), "gt_mean_keypoint_acceleration": float(gacc.mean())}
