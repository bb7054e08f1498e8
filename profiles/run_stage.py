"""Run the codec stages of one config a few times on cuda:0 (no timing) -- the target program
for `ncu --set full -k regex:<kernel> -c 1 python profiles/run_stage.py ...`."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from udp_pose_b200 import device_api as D  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=2)
ap.add_argument("--batch", type=int, default=2048)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--stages", default="crop,encode,decode")
a = ap.parse_args()
cfg = CONFIGS[a.config]
bench.CFG_ID = a.config
wl = bench.make_device_workload(cfg, a.batch, 0, torch, D)
src = D.SourceImages(wl["images"], wl["image_index"])
st = a.stages.split(",")
for _ in range(a.iters):
    if "crop" in st:
        M = D.warp_matrices(wl["center"], wl["scale"], wl["rot"], cfg.image_size)
        D.crop_batch(src, M, cfg.image_size)
    if "fused" in st:  # crop + ToTensor + Normalize + mirrored input
        M = D.warp_matrices(wl["center"], wl["scale"], wl["rot"], cfg.image_size)
        D.crop_normalize_batch(src, M, cfg.image_size, want_flipped=True)
    if "encode" in st:
        j3 = D.rotate_points_batch(wl["joints"], wl["rot"], wl["center"], wl["scale"], cfg.image_size)
        D.encode_batch(cfg, j3, wl["vis"])
    if "decode" in st:
        D.decode_batch(cfg, wl["out"], wl["flipped"], wl["center"], wl["scale"], want_packed=True)
torch.cuda.synchronize()
print("ok")
