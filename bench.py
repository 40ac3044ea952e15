"""Throughput benchmark of the hot path: 96x96 crops -> encoder/fusion/temporal/regressor -> FK
keypoints (+ MPJPE accumulation), in hand-frames per second (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--mode fp32|bf16]
    python bench.py --impl reference ...      # the CPU path (oracle port) on the host cores

One "step" = one batch of B hand-frames (2 crops each; 4 crops per two-hand frame) of synthetic
crops of the model's native shape with random-init weights, known-skeleton head, recurrent memory
in steady state (BASELINE.json configs[3], "synthetic throughput sweep").  Under torchrun every
rank runs its own batch (weak scaling, no data-path collective) and the MPJPE sums are all-reduced
over NCCL once at the end.  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "hand_frames_per_sec"
UNIT = "hand-frames/s"
SETTLE_S = 1.0          # untimed steady-state run before each mode's timed legs (see measure_mode)
CPU_SAMPLE_BATCH = 160  # the reference script's batch size (run_inference_torch_data.py:165)


def conv_flops_per_crop():
    """Algorithmic FLOPs (2*MAC) of each of the 24 backbone convolutions for one crop, in
    execution order (downsample branch counted with the conv2 it is fused into)."""
    from umetrack_b200 import arch

    out = []
    cin = 32
    for g, (nblk, cout, hw) in enumerate(zip(arch.STAGE_BLOCKS, arch.STAGE_PLANES, arch.STAGE_HW)):
        for b in range(nblk):
            first = b == 0 and g > 0
            c1_in = cin if first else cout
            out.append(2 * hw * hw * cout * c1_in * 9)
            f2 = 2 * hw * hw * cout * cout * 9
            if first:
                f2 += 2 * hw * hw * cout * cin
            out.append(f2)
        cin = cout
    return out


def ncu_conv_traffic(mode):
    """dram__bytes_read + dram__bytes_write per conv_umma_kernel launch, averaged over the 24
    backbone launches of the newest committed `ncu --set full` capture of this numeric mode
    (profiles/r*_ncu_conv_full_summary_<mode>.csv, 4096 crops per launch); None if there is none.
    Bytes per launch, like `achieved`."""
    import csv
    import glob
    note = None
    if mode == "bf16" and not glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_conv_full_summary_bf16.csv")):
        mode, note = "fp16", "capture of the fp16 mode: same kernels, same 2-byte operands and activations as bf16"
    mode_id = {"fp32": 0, "bf16": 1, "fp16": 2}[mode]
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", f"r*_ncu_conv_full_summary_{mode}.csv")), reverse=True):
        tot, n, crops = 0.0, 0, 4096
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd = wr = None
        for r in csv.reader(open(path)):
            if r and r[0].startswith("# crops_per_launch"):
                crops = int(r[1])
            if r and r[0].startswith("Kernel Name"):
                for i, h in enumerate(r):
                    if h.startswith("dram__bytes_read.sum"):
                        rd = (i, scale[h.split("[")[1].rstrip("]")])
                    if h.startswith("dram__bytes_write.sum"):
                        wr = (i, scale[h.split("[")[1].rstrip("]")])
            if len(r) <= 6 or not ("conv_umma_kernel<" in r[0] or "conv_pair_kernel<" in r[0]) or rd is None or wr is None:
                continue
            targs = [int(v) for v in r[0].split("<")[1].split(">")[0].replace("true", "1").replace("false", "0").split(",")]
            if targs[0] in (32, 64, 128, 256) and targs[1] == mode_id and targs[2] != 2:     # backbone launches of this mode
                tot += float(r[rd[0]]) * rd[1] + float(r[wr[0]]) * wr[1]
                n += 1
        if n:
            out = {"bytes_per_launch": tot / n, "launches": n, "crops_per_launch": crops,
                   "source": os.path.relpath(path, ROOT)}
            if note:
                out["note"] = note
            return out
    return None


class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l.split(", ") for ts, l in self.lines if t0 - 0.05 <= ts <= t1 + 0.15]
        rows = [r for r in rows if len(r) >= 8]
        if not rows:
            rows = [l.split(", ") for ts, l in self.lines][-3:]
            rows = [r for r in rows if len(r) >= 8]
        if not rows:
            return None
        try:
            sm = [float(r[1]) for r in rows]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            reasons = [n for i, n in enumerate(names) if any(r[4 + i].strip().lower() == "active" for r in rows)]
            return {"sm_mhz": statistics.median(sm), "sm_max_mhz": float(rows[0][2]), "reasons": reasons,
                    "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}
        except Exception:
            return None


def measure_stages(model, hm, hm_dict, dev, peaks):
    """Device-timed throughput of the stages around the network (north_star: achieved HBM GB/s for
    resampling and FK), and of the batched recording pipeline that chains all of them."""
    import numpy as np
    import torch

    from umetrack_b200 import native, synthetic
    from umetrack_b200.batched import crop_cameras_device, track_recordings_batched
    from umetrack_b200.skinning import fk_landmarks_device

    hbm = peaks.get("hbm_gbs") or 6650.0
    out = {"hbm_peak_gbs": hbm, "hbm_peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback"}

    def timed_ms(fn, iters=10):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    # (a) fisheye -> pinhole resampling: crops of a 512-frame synthetic rig (frames exceed L2)
    T = 512
    labels, _ = synthetic.make_recording(hm_dict, seed=11, n_frames=T, dropouts=False, with_frames=False)
    cc = crop_cameras_device(hm, labels, dev)
    cnt = cc["sel_count"].cpu().numpy()
    sel = cc["sel_cam"].cpu().numpy()
    H, W = 480, 640
    item = [(t, h, k) for t in range(T) for h in range(2) for k in range(int(cnt[t, h]))]
    ti, hi, ki = (torch.tensor([i[j] for i in item], device=dev) for j in range(3))
    params = cc["resample_params"][ti, hi, ki].contiguous()
    offs = torch.tensor([t * H * 4 * W + int(sel[t, h, k]) * W for t, h, k in item], dtype=torch.int64, device=dev)
    frames = torch.randint(0, 256, (T, H, 4 * W), dtype=torch.uint8, device=dev)
    crops = torch.empty((len(item), 96, 96), dtype=torch.float32, device=dev)
    lib = native.lib()

    def run_resample():
        native.check(lib.umt_resample_crops(frames.data_ptr(), offs.data_ptr(), 4 * W, W, H, params.data_ptr(),
                                            len(item), 0, crops.data_ptr(), 0, native.current_stream_ptr()), "resample")

    ms = timed_ms(run_resample)
    p = params.cpu().numpy()
    footprint = (96.0 * p[:, 0] / p[:, 28]) ** 2            # source pixels under one crop ~ (96 f_src / f_crop)^2 bytes
    alg_bytes = float((96 * 96 * 4 + footprint).sum())
    out["resample"] = {"crops": len(item), "ms": ms, "crops_per_s": len(item) / (ms * 1e-3),
                       "algorithmic_bytes_per_crop": alg_bytes / len(item), "achieved_gbs": alg_bytes / (ms * 1e-3) / 1e9,
                       "frac_of_hbm_peak": alg_bytes / (ms * 1e-3) / 1e9 / hbm,
                       "note": "float64 camera chain per pixel: fp64-issue bound, not HBM bound"}
    # (c) forward kinematics
    n = 1 << 20
    g = torch.Generator(device="cpu").manual_seed(3)
    ja = torch.rand((n, 22), generator=g).to(dev)
    wr = torch.eye(4).repeat(n, 1, 1).to(dev)
    ms = timed_ms(lambda: fk_landmarks_device(hm, ja, wr))
    fk_bytes = n * (22 * 4 + 16 * 4 + 63 * 4)
    out["fk"] = {"poses": n, "ms": ms, "poses_per_s": n / (ms * 1e-3), "achieved_gbs": fk_bytes / (ms * 1e-3) / 1e9,
                 "frac_of_hbm_peak": fk_bytes / (ms * 1e-3) / 1e9 / hbm}
    # whole pipeline on recordings (BASELINE configs[1]/[4] shape): raw frames -> keypoints -> MPJPE
    recs = [synthetic.make_recording(hm_dict, seed=100 + r, n_frames=48) for r in range(8)]
    recs = [(lab, torch.from_numpy(fr).pin_memory()) for lab, fr in recs]     # decoded frames in pinned memory
    track_recordings_batched(model, recs[:2])
    torch.cuda.synchronize()
    dt = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        res = track_recordings_batched(model, recs)
        torch.cuda.synchronize()
        dt = min(dt, time.perf_counter() - t0)
    n_hf = int(sum(r["valid_tracking"].sum() for r in res))
    frame_bytes = int(sum(fr.numel() for _, fr in recs))
    out["recordings_pipeline"] = {"recordings": len(recs), "frames_each": 48, "hand_frames": n_hf, "wall_s": dt,
                                  "hand_frames_per_s": n_hf / dt, "h2d_frame_bytes": frame_bytes,
                                  "note": "raw frames in pinned host memory -> upload + crop cameras + resample + "
                                          "encode + per-frame time scan + FK + metrics, host loop included (best of 3)"}
    # the reference's own call pattern: one gen_crop_cameras + track_frame per frame (HandTracker)
    from umetrack_b200.evaluation import track_recording
    lab1, fr1 = synthetic.make_recording(hm_dict, seed=100, n_frames=48)
    track_recording(model, lab1, fr1)      # warm-up (captures the CUDA graphs of the frame shapes)
    torch.cuda.synchronize()
    dt1 = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        r1 = track_recording(model, lab1, fr1)
        torch.cuda.synchronize()
        dt1 = min(dt1, time.perf_counter() - t0)
    n1 = int(r1["valid_tracking"].sum())
    out["per_frame_tracker"] = {"frames": 48, "hand_frames": n1, "wall_s": dt1, "ms_per_frame": dt1 / 48 * 1e3,
                                "hand_frames_per_s": n1 / dt1,
                                "note": "HandTracker.gen_crop_cameras + track_frame per frame as run_eval_known_skeleton.py "
                                        "calls them, host objects in and out (best of 3): per frame two launches + one read "
                                        "for the crop cameras, then upload + resample + one CUDA-graph replay of the forward "
                                        "pass (29 kernels, ~0.40 ms on the device for 2-4 crops) + one read; landmark FK of "
                                        "all frames after the loop"}
    return out


def cpu_oracle_rate(steps, warmup, batch=CPU_SAMPLE_BATCH):
    """hand-frames/s of the CPU restatement of the reference path (oracle port: torch fp32 model +
    FK) using every host core.  This is the one place bench.py executes oracle/."""
    import torch

    from oracle import umetrack_oracle as O
    from umetrack_b200 import synthetic

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hm = synthetic.load_hand_model_json(os.path.join(ROOT, "tests", "golden", "generic_hand_model_nomesh.json"))
    t = {k: torch.tensor(v) for k, v in hm.items() if isinstance(v, list)}
    ax, rest = t["joint_rotation_axes"].float()[None], (t["joint_rest_positions"] * 1e-3).float()[None]
    model = O.OracleModel(synthetic.make_state_dict(0))
    b = synthetic.make_model_batch(batch, seed=1234)

    def step(use_mem):
        b["use_memory"][:] = use_mem
        out = model.regress_pose_use_skeleton(b["left_images"], b["intrinsics"], b["extrinsics_xf"],
                                              b["sample_range"], b["memory_idx"], b["use_memory"], b["hand_idx"], ax, rest)
        w = out["wrist_xfs"].clone()
        w[:, :3, 3] *= 1000.0
        return O.skin_landmarks(t["joint_rotation_axes"], t["joint_rest_positions"], t["landmark_rest_positions"],
                                t["landmark_rest_bone_indices"], t["landmark_rest_bone_weights"],
                                out["joint_angles"], w)

    for i in range(warmup):
        step(i > 0)
    t0 = time.perf_counter()
    for _ in range(steps):
        step(True)
    dt = time.perf_counter() - t0
    return batch * steps / dt, cores, dt / max(steps, 1) * 1e3


def run_reference(args, rank):
    if rank != 0:
        return
    rate, cores, ms = cpu_oracle_rate(args.steps, args.warmup)
    sample = (f"{CPU_SAMPLE_BATCH} hand-frames per step (the reference script's batch size), crops->keypoints "
              f"(model + FK), torch fp32 on {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[3] synthetic crops->keypoints, known-skeleton head, bounded sample",
                   "hand_frames_per_step": CPU_SAMPLE_BATCH, "crops_per_hand_frame": 2},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def recordings_sharded(model, hm_dict, dev, rank, world, n_rec, n_frames, group=64):
    """BASELINE configs[4]: `n_rec` synthetic recordings (seeds 0..n_rec-1, `n_frames` frames, 4 fisheye
    views, both hands, dropouts) sharded by recording over the ranks (longest-processing-time first),
    every rank running the batched driver on its shard -- crop cameras, resampling, encoder, time
    scan, FK, MPJPE / acceleration sums -- and ONE all-reduce of the metric sums at the end
    (run_eval_known_skeleton.py:113-119 is a Pool(8).map over recordings followed by a concatenate).
    Up to `group` recordings of a rank are tracked together (one resampling launch, one encoder batch, ONE time
    scan whose steps carry the hands of all of them).  Frames are seeded noise generated on the device ("decoded frames staged to the GPU"); labels are
    the reference's JSON schema.  Returns the stage record; collective: every rank must call it."""
    import torch
    import torch.distributed as dist

    from umetrack_b200 import synthetic
    from umetrack_b200.batched import label_arrays, track_recordings_batched
    from umetrack_b200.metrics import MetricSums, shard_indices

    H, W4 = 480, 2560

    def frames_of(r):
        g = torch.Generator(device=dev).manual_seed(7000 + r)
        return torch.randint(0, 256, (n_frames, H, W4), dtype=torch.uint8, device=dev, generator=g)

    labels_of = {}

    def recording(r):
        if r not in labels_of:
            labels_of[r] = synthetic.make_recording(hm_dict, seed=r, n_frames=n_frames, with_frames=False)[0]
            label_arrays(labels_of[r])        # JSON lists -> arrays once, like parsing the label file
        return labels_of[r], frames_of(r)

    def run(indices, sums, staged=None):
        n_valid = 0
        for i0 in range(0, len(indices), group):
            recs = [staged[r] if staged else recording(r) for r in indices[i0:i0 + group]]
            res = track_recordings_batched(model, recs, sums=sums)
            n_valid += int(sum(x["valid_tracking"].sum() for x in res))
        return n_valid

    mine = shard_indices(n_rec, rank, world, costs=[n_frames] * n_rec)
    run(mine[:1], MetricSums(dev))                       # warm-up (allocations, lazy module load)
    # the shard's inputs are staged before the clock starts: label dicts on the host, decoded frames in HBM
    # (0.31 GB per recording); the timed region is the driver itself
    staged = {r: recording(r) for r in mine}
    # two timed passes over the shard: the first one also pays the one-time growth of the allocator pools and
    # of the model's workspace for this batch shape (reported as wall_s_first), the second is the steady state
    times = []
    for _ in range(2):
        sums = MetricSums(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_valid = run(mine, sums, staged)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    dt = times[1]
    staged.clear()
    t = torch.tensor([dt, float(n_valid)], dtype=torch.float64, device=dev)
    t_first = torch.tensor([times[0]], dtype=torch.float64, device=dev)
    if world > 1:
        tm = t[:1].clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)          # time: max over ranks
        dist.all_reduce(t_first, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)           # hand-frames: sum over ranks
        t[0] = tm[0]
    sums.all_reduce()                                      # the path's only collective (NCCL)
    res = sums.result()
    out = {"recordings": n_rec, "frames_each": n_frames, "hand_frames": int(t[1].item()), "wall_s": float(t[0].item()),
           "wall_s_first": float(t_first[0].item()),
           "hand_frames_per_s": float(t[1].item() / t[0].item()), "ranks": world,
           "recordings_per_rank": [len(shard_indices(n_rec, r, world, costs=[n_frames] * n_rec)) for r in range(world)],
           "mpjpe_mm": res["mpjpe"], "mpjpa": res["mpjpa"], "gt_mpjpa": res["gt_mpjpa"],
           "note": "inputs staged before the clock (labels on the host, frames in HBM); host loop of the driver included; "
                   "second of two passes (wall_s_first = the pass that also grows the memory pools); "
                   "time = max over ranks, hand-frames = sum over ranks, metric sums all-reduced once"}
    if world > 1 and rank == 0:
        # the same recordings on ONE GPU (untimed): the all-reduced means must agree within 0.01
        one = MetricSums(dev)
        run(list(range(n_rec)), one)
        r1 = one.result()
        out["single_gpu_check"] = {"mpjpe_mm": r1["mpjpe"], "mpjpa": r1["mpjpa"],
                                   "abs_delta_mpjpe": abs(r1["mpjpe"] - res["mpjpe"]),
                                   "abs_delta_mpjpa": abs(r1["mpjpa"] - res["mpjpa"]),
                                   "within_0.01": bool(abs(r1["mpjpe"] - res["mpjpe"]) < 0.01
                                                       and abs(r1["mpjpa"] - res["mpjpa"]) < 0.01)}
    if world > 1:
        dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=2048, help="hand-frames per step per GPU")
    ap.add_argument("--mode", default="fp32", choices=["fp32", "bf16", "fp16"],
                    help="encoder numeric mode of the headline: fp32 (default; split-bf16 x3, fp32-equivalent = the "
                         "reference's precision), fp16 or bf16 (16-bit operands, fp32 accumulate; reported under `modes`)")
    ap.add_argument("--chunk", type=int, default=4096, help="crops per encoder pass (workspace size)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stages", action="store_true", help="skip the per-stage and recording measurements")
    ap.add_argument("--sharded-recordings", type=int, default=64)
    ap.add_argument("--sharded-frames", type=int, default=256)
    ap.add_argument("--settle", type=float, default=SETTLE_S,
                    help="seconds of untimed steps before each mode's timed legs (0 for profiler runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from umetrack_b200 import arch, synthetic
    from umetrack_b200 import hand_model as hmod
    from umetrack_b200.metrics import MetricSums
    from umetrack_b200.model import InputFrameData, InputFrameDesc, InputSkeletonData, UmeTrackModel
    from umetrack_b200.skinning import fk_landmarks_device

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; umetrack_b200 has no CPU path (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    B = args.batch
    W = max(args.warmup, 3)
    K = args.steps
    hm_dict = synthetic.load_hand_model_json(os.path.join(ROOT, "tests", "golden", "generic_hand_model_nomesh.json"))
    hm = hmod.load_hand_model_from_dict(hm_dict)
    model = UmeTrackModel(synthetic.make_state_dict(0), chunk_images=args.chunk).to(dev)
    model.set_mode(args.mode)
    host = synthetic.make_model_batch(B, seed=1234 + rank)
    # crops as cv2.remap leaves them: uint8.  The device-resident leg takes the float32 crops the
    # reference's model API takes ([0,1], tracker.py:211); the end-to-end leg uploads the uint8 crops.
    crops_u8 = (host["left_images"] * 255.0).round().to(torch.uint8)
    host["left_images"] = crops_u8.float() / 255.0
    d = {k: v.to(dev) for k, v in host.items()}
    pinned = {k: (crops_u8 if k == "left_images" else v).pin_memory() for k, v in host.items()}
    fd = InputFrameData(d["left_images"], d["intrinsics"], d["extrinsics_xf"])
    desc = InputFrameDesc(d["sample_range"], d["memory_idx"], d["use_memory"], d["hand_idx"], n_memory_slots=B)
    skel = InputSkeletonData(hm.joint_rotation_axes.float()[None].to(dev),
                             (hm.joint_rest_positions * 1e-3).float()[None].to(dev))
    g = np.random.default_rng(99)
    gt_kp = torch.from_numpy(g.normal(0, 80, size=(B, 21, 3)).astype(np.float32)).to(dev)
    mirror = (d["hand_idx"] == 1).to(torch.uint8)
    sums = MetricSums(dev)

    def device_step(frame_data, frame_desc):
        out = model.regress_pose_use_skeleton(frame_data, frame_desc, skel)
        wr = out.wrist_xfs.clone()
        wr[:, :3, 3] *= 1000.0                    # m -> mm (tracker.py:258)
        kp = fk_landmarks_device(hm, out.joint_angles, wr, mirror)
        sums.add_errors(kp, gt_kp)
        return kp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, collective=True):
        """K steps between CUDA events; with `collective`, bracketed by barriers and reduced to the
        max over ranks (every rank must call it); without, a purely local measurement."""
        if collective:
            barrier()
        else:
            torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if collective and world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        if collective:
            barrier()
        return ms

    # ---- end to end through the public API: pinned host inputs (uint8 crops + descriptors) in,
    # keypoints out, every step.  The host->device copy of step i+1 runs on a copy stream while step
    # i computes (two device input slots), the way a user feeding the model from a loader would write
    # it; every timed step still moves its own inputs (h2d_bytes_per_step) and results
    # (d2h_bytes_per_step). ----
    out_host = torch.empty((B, 21, 3), dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    slots = [{k: torch.empty_like(v, device=dev) for k, v in pinned.items()} for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    state = {"i": 0}

    def stage_inputs(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])         # previous user of this slot is done
            for k, v in pinned.items():
                slots[slot][k].copy_(v, non_blocking=True)
            ready[slot].record(copy_stream)

    def e2e_step():
        slot = state["i"] & 1
        state["i"] += 1
        stage_inputs(slot ^ 1)                             # prefetch the next step's inputs
        torch.cuda.current_stream().wait_event(ready[slot])
        dd = slots[slot]
        kp = device_step(InputFrameData(dd["left_images"], dd["intrinsics"], dd["extrinsics_xf"]),
                         InputFrameDesc(dd["sample_range"], dd["memory_idx"], dd["use_memory"], dd["hand_idx"],
                                        n_memory_slots=B))
        consumed[slot].record()
        out_host.copy_(kp, non_blocking=True)

    pinned["use_memory"][:] = True
    h2d = sum(v.numel() * v.element_size() for v in pinned.values())
    d2h = out_host.numel() * out_host.element_size()

    def measure_mode(mode, collective):
        """(device-resident hand-frames/s, end-to-end hand-frames/s, ms per step, e2e ms per step,
        launches per step) of one numeric mode; recurrent memory in steady state."""
        model.set_mode(mode)
        d["use_memory"][:] = False            # first step starts with empty memory, like a recording's first frame
        device_step(fd, desc)
        d["use_memory"][:] = True
        for _ in range(W - 1):
            device_step(fd, desc)
        # The boards run at their power cap: the first ~0.5 s of a burst is 2-4 % faster than the steady
        # state (tools/e2e_probe.py).  Both timed legs below therefore start from the steady state, so
        # that `e2e` against `value` shows the cost of the copies and not which leg ran on the cooler chip.
        torch.cuda.synchronize()
        t_settle, settle_steps = time.time(), 0
        while time.time() - t_settle < args.settle:
            device_step(fd, desc)
            torch.cuda.synchronize()
            settle_steps += 1
        state["settle_steps"] = settle_steps
        launches = model.last_launch_count + 4           # + FK, MPJPE sum, per-frame errors, PCK counts
        ms = timed(lambda: device_step(fd, desc), collective)
        state["i"] = 0
        for ev in consumed:
            ev.record()
        stage_inputs(0)
        e2e_step()
        ms_e2e = timed(e2e_step, collective)
        n = world if collective else 1
        return n * B * K / (ms * 1e-3), n * B * K / (ms_e2e * 1e-3), ms / K, ms_e2e / K, launches

    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    t_wall0 = time.time()
    value, e2e_value, ms_step, ms_e2e_step, launches = measure_mode(args.mode, True)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1)
    model.check_status()

    sums.all_reduce()   # the only collective on the path: NCCL all-reduce of the metric sums
    metrics = sums.result()

    dtype = {"fp32": "bf16x3 (split-bf16 operands, three tensor-core products per MAC, fp32 accumulate: "
                     "fp32-equivalent, the reference's precision)",
             "bf16": "bf16 (fp32 accumulate)", "fp16": "fp16 (fp32 accumulate)"}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": dtype[args.mode], "data": "synthetic",
        "config": {"workload": "configs[3] synthetic throughput: 96x96 crops -> keypoints, known-skeleton head, "
                               "memory steady state", "hand_frames_per_step_per_gpu": B, "crops_per_hand_frame": 2,
                   "crops_per_sec": 2 * value, "numeric_mode": args.mode,
                   "settle": "%d untimed steps (%.1f s) after the %d warm-up steps, before the timed legs" % (
                       state.get("settle_steps", 0), args.settle, W),
                   "l2": "inputs (151 MB of float32 crops at B=2048) and activations exceed the 126 MB L2; no flush",
                   "sharding": "independent batches per rank, NCCL all-reduce of metric sums only"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e_step,
                "inputs": "uint8 crops (as cv2.remap leaves them; /255 on the device) + float32 camera parameters + "
                          "int64 descriptors from pinned host memory; float32 keypoints back"},
        "gpu_launches": launches * K,
        "clocks": clocks,
        "mpjpe_vs_synthetic_gt_mm": metrics["mpjpe"],
    }

    # ---- BASELINE configs[4]: recordings sharded over the ranks (collective) ----
    sharded = None
    if not args.no_stages and args.sharded_recordings > 0:
        model.set_mode(args.mode)
        sharded = recordings_sharded(model, hm_dict, dev, rank, world, args.sharded_recordings, args.sharded_frames)
        model.reset_memory()

    if rank == 0:
        # ---- roofline of the dominant kernel family: the 24 tcgen05 convolutions ----
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        burst = peaks.get("bf16_tflops") or 1590.0
        sustained = peaks.get("bf16_tflops_sustained") or 1400.0
        src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md: 1.59 PFLOP/s burst, 1.4 sustained)"
        model.set_mode(args.mode)
        n_prof = min(d["left_images"].shape[0], args.chunk)      # the crops one encoder launch of the step processes
        reps = 3
        acc = np.zeros(26)
        model.profile_layers(d["left_images"][:n_prof])
        for _ in range(reps):
            acc += np.array(model.profile_layers(d["left_images"][:n_prof]))
        acc /= reps
        flops = np.array(conv_flops_per_crop(), dtype=np.float64) * n_prof
        conv_ms = acc[1:25]
        achieved = flops.sum() / (conv_ms.sum() * 1e-3) / 1e12
        products = 3 if args.mode == "fp32" else 1
        traffic = ncu_conv_traffic(args.mode)
        step_tflops = value / world * 1.957607712e9 / 1e12
        line["roofline"] = {
            "bound": "tensor", "achieved": achieved, "peak": burst, "unit": "TFLOP/s",
            "frac": achieved / burst,
            "peak_sustained": sustained, "frac_vs_sustained": achieved / sustained,
            "traffic": (traffic or {}).get("bytes_per_launch"),
            "traffic_source": traffic,
            "kernel": "conv_pair_kernel / conv_umma_kernel (the 24 tcgen05 convolutions of one encoder pass)",
            "peak_source": src + ": `peak` is the bf16 burst figure because the launches are timed one by one with "
                                 "CUDA events at burst clocks (see `clocks`); the sustained figure is given beside it",
            "note": "achieved = algorithmic conv FLOPs (962.6 MFLOP per crop x crops per launch) / sum of the 24 launch "
                    "durations (CUDA events between launches, umt_model_profile_layers, same crops per launch as the "
                    "timed step); the fp32-equivalent mode executes 3 bf16 tensor-core products per algorithmic MAC, "
                    "so its ceiling on this scale is 1/3",
            "tensor_products_per_mac": products,
            "executed_tensor_tflops": achieved * products, "executed_frac": achieved * products / burst,
            "executed_frac_vs_sustained": achieved * products / sustained,
            "whole_step_tflops": step_tflops, "whole_step_frac": step_tflops / burst,
            "whole_step_frac_vs_sustained": step_tflops / sustained,
            "conv_ms_per_pass": float(conv_ms.sum()), "stem_ms": float(acc[0]), "proj_ms": float(acc[25]),
            "crops_per_pass": n_prof,
            "per_layer_tflops": [float(f / (t * 1e-3) / 1e12) for f, t in zip(flops, conv_ms)],
        }
        # ---- every numeric mode as a first-class record (BASELINE.json configs[3]: 16-bit encoder vs fp32
        # tolerance check): device-resident and end-to-end rate on this GPU, keypoint deltas against the
        # fp32-equivalent mode from the same recurrent state ----
        model.set_mode("fp32")
        d["use_memory"][:] = False               # (the recordings stage reset the memory bank: refill it)
        device_step(fd, desc)
        d["use_memory"][:] = True
        device_step(fd, desc)
        mem_a, ext_a = model._mem_features.clone(), model._prev_extrinsics.clone()

        def restore():
            model._mem_features.copy_(mem_a)
            model._prev_extrinsics.copy_(ext_a)

        model.set_mode("fp32")
        restore()
        kp_ref = device_step(fd, desc).clone()
        line["modes"] = {}
        for mode in ("fp32", "fp16", "bf16"):
            if mode == args.mode and world == 1:
                v, e, ms1, ms2, _ = value, e2e_value, ms_step, ms_e2e_step, launches
            else:
                v, e, ms1, ms2, _ = measure_mode(mode, False)      # rank 0 only: no collectives here
            model.set_mode(mode)
            restore()
            kp_m = device_step(fd, desc)
            line["modes"][mode] = {"dtype": dtype[mode], "value": v, "e2e": e, "unit": UNIT, "n_gpus": 1,
                                   "ms_per_step": ms1, "e2e_ms_per_step": ms2,
                                   "keypoint_delta_vs_fp32_mm_max": float((kp_ref - kp_m).abs().max()),
                                   "keypoint_delta_vs_fp32_mm_mean": float((kp_ref - kp_m).abs().mean()),
                                   "keypoint_gate_mm": 0.05}
        model.set_mode(args.mode)
        restore()
        try:
            model.check_status()
            line["status_flags"] = "clear"
        except Exception as e:  # noqa: BLE001
            line["status_flags"] = str(e)
        if not args.no_stages:
            line["stages"] = measure_stages(model, hm, hm_dict, dev, peaks)
            if sharded is not None:
                line["stages"]["recordings_sharded"] = sharded
        if world == 1 and not args.no_cpu_baseline:
            rate, cores, _ = cpu_oracle_rate(steps=5, warmup=2)
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"5 steps of {CPU_SAMPLE_BATCH} hand-frames (crops->keypoints, torch fp32 model + FK) "
                          f"after 2 warm-up steps, {cores} threads"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
