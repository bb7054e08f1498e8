"""BASELINE.json configs[4]: RSN18+UDP 256x192 large-batch sweep, 1 Ki .. 1 Mi synthetic persons.

    python profiles/sweep.py [--config 5] [--reps 3]                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P profiles/sweep.py                                               # N GPUs

Strong scaling: a sweep point is N persons in total; rank r owns the contiguous block
[r*ceil(N/R), min(N, (r+1)*ceil(N/R))) (the reference's own rule).  Per rank the block
is processed as a stream of chunks of at most 16 Ki persons drawn from a resident pool (1 Mi persons
of network output would be 418 GB); one pass over a chunk = the five kernels of bench.py's step
(warp matrices, crop, joint map, 5-scale target encode, flip-fused decode) plus, for R > 1, the
all_gather of the chunk's keypoints.  A chunk's inputs (6.8 GB at 16 Ki) exceed L2, so repeating the
chunk does not warm anything.  Timed with CUDA events per rank, max over ranks, best of `reps`.
Rank 0 prints one JSON line per sweep point."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402

POINTS = [1 << 10, 1 << 12, 1 << 14, 1 << 16, 1 << 18, 1 << 20]
CHUNK = 1 << 14


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=5)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--graphs", action="store_true", help="replay each chunk pass from a CUDA graph (udp_pose_b200/graphs.py)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from udp_pose_b200 import _lib, device_api as D
    from udp_pose_b200.configs import CONFIGS
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.load()
    cfg = CONFIGS[args.config]
    K = cfg.num_joints
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    wl = bench.make_device_workload(cfg, CHUNK, rank, torch, D)
    tshape = (CHUNK, 5, K, h, w) if cfg.target_type == "rsn" else (CHUNK, cfg.channels, h, w)
    target = torch.empty(tshape, dtype=torch.float32, device="cuda")
    crop = torch.empty((CHUNK, H, W, 3), dtype=torch.uint8, device="cuda")
    j3 = wl["joints"].clone()

    def chunk_pass(m, gathered):
        """The codec pass over the first m persons of the resident chunk."""
        sl = slice(0, m)
        src = srcs[m]
        M = D.warp_matrices(wl["center"][sl], wl["scale"][sl], wl["rot"][sl], cfg.image_size)
        D.crop_batch(src, M, cfg.image_size, out=crop[sl])
        D.rotate_points_batch(wl["joints"][sl], wl["rot"][sl], wl["center"][sl], wl["scale"][sl], cfg.image_size, out=j3[sl])
        D.encode_batch(cfg, j3[sl], wl["vis"][sl], out=target[sl])
        r = D.decode_batch(cfg, wl["out"][sl], wl["flipped"][sl], wl["center"][sl], wl["scale"][sl], want_packed=True)
        if world > 1:
            dist.all_gather_into_tensor(gathered[: world * m], r["packed"])

    srcs, graphs = {}, {}
    gathered = torch.empty((world * CHUNK, K, 3), dtype=torch.float32, device="cuda") if world > 1 else None
    for N in POINTS:
        per = -(-N // world)                                   # ceil(N / R): every rank runs the padded block
        chunks = [CHUNK] * (per // CHUNK) + ([per % CHUNK] if per % CHUNK else [])
        for m in set(chunks):                                  # per-person image tables: host work, outside the timing
            if m not in srcs:
                srcs[m] = D.SourceImages(wl["images"], wl["image_index"][:m])
            if args.graphs and world == 1 and m not in graphs:   # NCCL collectives stay outside the capture
                from udp_pose_b200.graphs import CapturedPass
                graphs[m] = CapturedPass(lambda m=m: chunk_pass(m, None))
        best = None
        for rep in range(args.reps + 1):                       # first repetition is the warm-up
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for m in chunks:
                if m in graphs:
                    graphs[m].replay()
                else:
                    chunk_pass(m, gathered)
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rep:
                best = float(t[0]) if best is None else min(best, float(t[0]))
        if rank == 0:
            print(json.dumps({"cfg": args.config, "workload": cfg.name, "n_gpus": world, "persons": N, "persons_per_gpu": per,
                              "chunks_per_gpu": len(chunks), "ms": round(best, 4), "persons_per_s": round(N / (best * 1e-3)),
                              "scaling": "strong", "cuda_graphs": bool(graphs), "timing": "CUDA events, max over ranks, best of %d" % args.reps}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
