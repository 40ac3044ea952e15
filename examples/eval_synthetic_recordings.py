"""Known-skeleton evaluation of a set of recordings, sharded over the GPUs of one box
(BASELINE configs[1] and [4]; the reference's run_eval_known_skeleton.py + load_eval.py).

    python examples/eval_synthetic_recordings.py --recordings 16 --frames 64 --out /tmp/umt_eval
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \\
        examples/eval_synthetic_recordings.py --recordings 64 --frames 256 --out /tmp/umt_eval

The dataset and the weights are absent from this environment, so the recordings are the seeded
synthetic ones of umetrack_b200.synthetic (same JSON schema as the reference's recordings; frames
as uint8 [T, H, 4*W] arrays instead of an mp4) and the weights are the seeded synthetic state_dict.
With real data: `load_pretrained_model(path)`, `ingest.load_recording_labels(json)` and the decoded
frames go in their place.  Each rank tracks its share of the recordings with the batched driver,
writes one result file per recording in the reference's pickle format, accumulates the metric sums
on its GPU and all-reduces them once (NCCL); rank 0 prints the reference's summary lines.
"""
import argparse
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from umetrack_b200 import synthetic  # noqa: E402
from umetrack_b200.aggregate import aggregate_metrics, save_eval_result  # noqa: E402
from umetrack_b200.batched import track_recordings_batched  # noqa: E402
from umetrack_b200.metrics import MetricSums, shard_indices  # noqa: E402
from umetrack_b200.model import UmeTrackModel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--recordings", type=int, default=16)
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--group", type=int, default=8, help="recordings tracked together by the batched driver")
    ap.add_argument("--mode", default="fp32", choices=["fp32", "bf16", "fp16"])
    ap.add_argument("--out", default="/tmp/umt_eval")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    hm = synthetic.load_hand_model_json(os.path.join(ROOT, "tests", "golden", "generic_hand_model_nomesh.json"))
    model = UmeTrackModel(synthetic.make_state_dict(0)).to(dev)
    model.set_mode(args.mode)
    mine = shard_indices(args.recordings, rank, world, costs=[args.frames] * args.recordings)
    sums = MetricSums(dev)
    n_hf, t_track = 0, 0.0
    for g0 in range(0, len(mine), args.group):
        ids = mine[g0:g0 + args.group]
        recs = [synthetic.make_recording(hm, seed=1000 + i, n_frames=args.frames) for i in ids]
        recs = [(lab, torch.from_numpy(fr).pin_memory()) for lab, fr in recs]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = track_recordings_batched(model, recs, sums=sums)
        torch.cuda.synchronize()
        t_track += time.perf_counter() - t0
        for i, r in zip(ids, res):
            save_eval_result(os.path.join(args.out, f"user{i % 4}", f"recording_{i:03d}.npy"), r)
            n_hf += int(r["valid_tracking"].sum())
    stats = torch.tensor([float(n_hf), t_track], dtype=torch.float64, device=dev)
    if world > 1:
        sums.all_reduce()
        mx = stats.clone()
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        stats[1] = mx[1]
        dist.barrier()
    if rank == 0:
        r = sums.result()
        print(f"{world} GPU(s): {int(stats[0])} hand-frames tracked in {float(stats[1]):.3f} s (slowest rank) = "
              f"{float(stats[0]) / float(stats[1]):.0f} hand-frames/s")
        print(f"all-reduced sums: MPJPE {r['mpjpe']:.6f} mm, MPJPA {r['mpjpa']:.6f}, AUC {r['auc']:.6f}, "
              f"success rate {r['success_rate']:.3f}%")
        print("from the result files (load_eval.py format):")
        agg = aggregate_metrics(args.out)
        assert agg is not None and abs(agg["mean_keypoint_error"] - r["mpjpe"]) < 1e-6
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
