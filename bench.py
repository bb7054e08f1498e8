#!/usr/bin/env python
"""bench.py -- UDP heatmap codec throughput on B200 (person-instances/sec, encode+decode).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): COCO HRNet-W32 256x192 UDPv1 -- per person one pass of
  warp matrices -> 256x192 bilinear crop -> joint forward map -> classification/offset target
  (KPD 3.5, 51x64x48) -> flip-fused offset-vote decode of a 51x64x48 network output pair.
A step is that pass over one batch of synthetic persons per GPU (weak scaling); the persons
are sharded contiguously over ranks and the keypoints are gathered once per step (NCCL).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG_ID = 2
METRIC = "person_instances_per_sec_encode_decode"
UNIT = "persons/s"
N_IMAGES = 64


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=16384, help="persons per GPU per step")
    ap.add_argument("--e2e-batch", type=int, default=2048, help="persons per GPU per e2e step")
    ap.add_argument("--e2e-chunk", type=int, default=256)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline budget")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fused", action="store_true", help="skip the 8(f)-1 fused crop+normalize timing")
    ap.add_argument("--no-configs", action="store_true", help="skip the `configs` block (the other BASELINE.json configs)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the `sweep` block (cfg-5 strong-scaling sweep, 1 Ki .. 1 Mi persons)")
    ap.add_argument("--no-sweep-graphs", action="store_true", help="sweep: direct launches instead of CUDA-graph replay of each chunk pass")
    ap.add_argument("--config", type=int, default=CFG_ID)
    ap.add_argument("--images", type=int, default=N_IMAGES, help="source image pool (SURVEY 8d: 64; 256 exceeds L2)")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------
# algorithmic (compulsory) bytes per person, SURVEY.md 8(d)
# ----------------------------------------------------------------------------------------
def _clip_poly(poly, w, h):
    """Sutherland-Hodgman clip of a convex polygon to [0,w]x[0,h]; returns its area."""
    def clip(pts, inside, inter):
        out = []
        for i in range(len(pts)):
            a, b = pts[i], pts[(i + 1) % len(pts)]
            ia, ib = inside(a), inside(b)
            if ia:
                out.append(a)
            if ia != ib:
                out.append(inter(a, b))
        return out

    def ix(x):
        return lambda a, b: (x, a[1] + (b[1] - a[1]) * (x - a[0]) / (b[0] - a[0]))

    def iy(y):
        return lambda a, b: (a[0] + (b[0] - a[0]) * (y - a[1]) / (b[1] - a[1]), y)
    for inside, inter in ((lambda p: p[0] >= 0, ix(0.0)), (lambda p: p[0] <= w, ix(float(w))),
                          (lambda p: p[1] >= 0, iy(0.0)), (lambda p: p[1] <= h, iy(float(h)))):
        poly = clip(poly, inside, inter)
        if not poly:
            return 0.0
    a = 0.0
    for i in range(len(poly)):
        x0, y0 = poly[i]
        x1, y1 = poly[(i + 1) % len(poly)]
        a += x0 * y1 - x1 * y0
    return abs(a) * 0.5


def crop_src_area(center, scale, rot, src_w, src_h):
    """Per-person area (source pixels) of the rotated box footprint clipped to the image."""
    out = np.zeros(len(rot))
    for i in range(len(rot)):
        bw, bh = float(scale[i, 0]) * 200.0, float(scale[i, 1]) * 200.0
        th = np.deg2rad(rot[i])
        c, s = np.cos(th), np.sin(th)
        pts = []
        for dx, dy in ((-bw / 2, -bh / 2), (bw / 2, -bh / 2), (bw / 2, bh / 2), (-bw / 2, bh / 2)):
            pts.append((center[i, 0] + c * dx + s * dy, center[i, 1] - s * dx + c * dy))
        out[i] = _clip_poly(pts, src_w, src_h)
    return out


def algorithmic_bytes(cfg, src_area_mean):
    """Bytes per person of each stage: every input byte once, every output byte once."""
    K, C = cfg.num_joints, cfg.channels
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    hw4 = h * w * 4
    cs = 8 if cfg.cs_dtype == "float64" else 4
    tgt = (5 * K if cfg.target_type == "rsn" else C) * hw4
    b = {
        "warp_matrices": 2 * 2 * cs + 8 + 24,
        "rotate_points": K * 24 + 2 * 2 * cs + 8 + K * 16,
        "crop": 3.0 * src_area_mean + W * H * 3 + 24,
        "encode": K * 3 * 8 * 2 + tgt + K * 4,
        # SURVEY 8(d) figure: both network outputs in full + preds/maxvals/preds_in out
        "decode_survey": 2 * C * hw4 + K * 36,
    }
    if cfg.target_type == "offset":
        # what this decoder must read: the 2K classification maps in full, plus per joint two
        # 7x7 offset windows from each tensor (7 rows x one 32-byte sector x 2 channels x 2)
        b["decode"] = 2 * K * hw4 + K * 2 * 2 * 7 * 32 + K * 36
    else:
        b["decode"] = b["decode_survey"]
    return b


# ----------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not (t0 - 0.05 <= t <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except (ValueError, IndexError):
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------
# synthetic workload resident in HBM
# ----------------------------------------------------------------------------------------
def cfg_id_of(cfg):
    from udp_pose_b200.configs import CONFIGS
    return [k for k, v in CONFIGS.items() if v is cfg][0]


def make_device_workload(cfg, B, rank, torch, D):
    """Deterministic per-rank synthetic batch (SURVEY 8d distributions), generated outside
    the timed region.  Heat maps are produced on the device in slices."""
    from udp_pose_b200 import synth
    dev = torch.device("cuda")
    rng = np.random.default_rng(1000 * cfg_id_of(cfg) + rank)
    c, s, r = synth.boxes(cfg, B, rng, rotate=True)
    K = cfg.num_joints
    js = np.zeros((B, K, 3))
    js[..., 0] = c[:, None, 0] + rng.uniform(-0.6, 0.6, (B, K)) * s[:, None, 0] * 200.0
    js[..., 1] = c[:, None, 1] + rng.uniform(-0.6, 0.6, (B, K)) * s[:, None, 1] * 200.0
    v = (rng.uniform(size=(B, K)) < 0.8).astype(np.float64)
    vis = np.zeros((B, K, 3))
    vis[..., 0] = vis[..., 1] = v
    imgs = synth.source_images(N_IMAGES, rng)
    wl = dict(center=torch.from_numpy(c).to(dev), scale=torch.from_numpy(s).to(dev),
              rot=torch.from_numpy(r).to(dev), joints=torch.from_numpy(js).to(dev),
              vis=torch.from_numpy(vis).to(dev), images=torch.from_numpy(imgs).to(dev),
              image_index=np.arange(B) % N_IMAGES,
              src_area=crop_src_area(c, s, r, synth.SRC_W, synth.SRC_H).mean(),
              host=dict(center=c, scale=s, rot=r, joints=js, vis=vis, images=imgs))
    w, h = cfg.heatmap_size
    C = cfg.channels
    out = torch.empty((B, C, h, w), dtype=torch.float32, device=dev)
    fl = torch.empty_like(out)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    xs = torch.arange(w, device=dev).view(1, 1, 1, w)
    ys = torch.arange(h, device=dev).view(1, 1, h, 1)
    trip = cfg.target_type == "offset"
    for lo in range(0, B, 1024):
        m = min(1024, B - lo)
        mux = torch.rand((m, K, 1, 1), device=dev, generator=g) * (w + 3) - 2
        muy = torch.rand((m, K, 1, 1), device=dev, generator=g) * (h + 3) - 2
        amp = torch.rand((m, K, 1, 1), device=dev, generator=g) * 0.8 + 0.2
        d2 = (xs - mux) ** 2 + (ys - muy) ** 2
        if trip:
            keep = (d2 / cfg.kpd ** 2) <= 1
            clean = torch.stack([amp * torch.exp(-d2 / 8.0),
                                 torch.where(keep, (mux - xs) / cfg.kpd, torch.zeros_like(d2)),
                                 torch.where(keep, (muy - ys) / cfg.kpd, torch.zeros_like(d2))], 2)
            clean = clean.reshape(m, C, h, w)
            nscale = torch.tensor([0.02, 0.05, 0.05], device=dev).repeat(K).view(1, C, 1, 1)
        else:
            clean = amp * torch.exp(-d2 / (2.0 * float(cfg.sigma) ** 2))
            nscale = 0.02
        mul = 255.0 if cfg.pre_div else 1.0
        out[lo:lo + m] = (clean + (torch.rand(clean.shape, device=dev, generator=g) * 2 - 1) * nscale) * mul
        cf = D.flip_back(clean.float().contiguous(), cfg.flip_pairs, trip)  # what the mirrored input yields
        fl[lo:lo + m] = (cf + (torch.rand(clean.shape, device=dev, generator=g) * 2 - 1) * nscale) * mul
    wl["out"], wl["flipped"] = out, fl
    return wl


# ----------------------------------------------------------------------------------------
# CPU baseline: the numpy/cv2 port of the reference path on the host cores
# ----------------------------------------------------------------------------------------
CPU_MAP_POOL = 256  # distinct network-output persons per worker (cycled): 1024 persons of cfg-2 maps would be 1.3 GB per core


def _cpu_inputs(cfg_id, seed, n):
    """Seeded inputs of one worker, the distributions of make_device_workload (60 % rotated boxes,
    joint visibility 0.8, the N_IMAGES-image source pool); generated once, outside every timed region."""
    os.environ["OMP_NUM_THREADS"] = "1"
    import cv2
    cv2.setNumThreads(1)
    from udp_pose_b200 import synth
    from udp_pose_b200.configs import CONFIGS
    cfg = CONFIGS[cfg_id]
    rng = np.random.default_rng(seed)
    c, s, r = synth.boxes(cfg, n, rng, rotate=True)
    imgs = synth.source_images(N_IMAGES, rng)
    K = cfg.num_joints
    js = np.zeros((n, K, 3))
    js[..., 0] = c[:, None, 0] + rng.uniform(-0.6, 0.6, (n, K)) * s[:, None, 0] * 200.0
    js[..., 1] = c[:, None, 1] + rng.uniform(-0.6, 0.6, (n, K)) * s[:, None, 1] * 200.0
    v = (rng.uniform(size=(n, K)) < 0.8).astype(np.float64)
    vis = np.zeros((n, K, 3))
    vis[..., 0] = vis[..., 1] = v
    out, fl = synth.heatmaps(cfg, min(n, CPU_MAP_POOL), rng)
    return dict(cfg=cfg, n=n, c=c, s=s, r=r, imgs=imgs, js=js, vis=vis, out=out, fl=fl)


def _cpu_pass(st):
    """One worker = one core: the reference's per-person path over the worker's persons;
    (seconds, seconds per stage -- the four stages SURVEY 8d asks to be timed separately)."""
    from oracle import np_port as P
    cfg, n, c, s, r, imgs, js, vis = (st[k] for k in ("cfg", "n", "c", "s", "r", "imgs", "js", "vis"))
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    trip = cfg.target_type == "offset"
    pc = time.perf_counter
    st_s = {"warpAffine": 0.0, "generate_target": 0.0, "flip_back+average": 0.0, "get_final_preds": 0.0}
    t0 = pc()
    for i in range(n):  # encode: one person per __getitem__ call
        a = pc()
        P.crop_person(imgs[i % len(imgs)], float(r[i]), c[i], s[i], W, H)
        b = pc()
        j = js[i].copy()
        j[:, 0:2] = P.map_joints(j[:, 0:2], float(r[i]), c[i], s[i], W, H)
        if cfg.target_type == "gaussian":
            P.target_gaussian(j, vis[i], W, H, w, h, cfg.sigma)
        elif trip:
            P.target_offset(j, vis[i], W, H, w, h, cfg.kpd)
        else:
            P.labels_rsn(j, vis[i, :, 0], W, H, w, h)
        e = pc()
        st_s["warpAffine"] += b - a
        st_s["generate_target"] += e - b
    bs = 32  # decode: per validation batch, as validate() does
    with np.errstate(all="ignore"):
        pool = st["out"].shape[0]
        for lo in range(0, n, bs):
            a = pc()
            j0 = lo % pool                      # the worker's map pool, cycled (views: pool % bs == 0, no copy)
            m = min(bs, n - lo, pool - j0)
            avg = P.unflip_average(st["out"][j0:j0 + m], st["fl"][j0:j0 + m], cfg.flip_pairs, trip)
            b = pc()
            P.final_preds(avg, c[lo:lo + m], s[lo:lo + m], "offset" if trip else "gaussian",
                          cfg.kpd, cfg.post_process, cfg.pre_div)
            e = pc()
            st_s["flip_back+average"] += b - a
            st_s["get_final_preds"] += e - b
    return pc() - t0, st_s


def _cpu_worker_main(conn, cfg_id, seed, n):
    st = _cpu_inputs(cfg_id, seed, n)
    conn.send("ready")
    while conn.recv() is not None:
        conn.send(_cpu_pass(st))


class CpuArm:
    """The CPU port on all host cores: one single-threaded process per core (the reference's
    DataLoader-worker model, tools/train.py:145-151), each holding its own `per_worker` persons.
    step() = every worker runs the encode+decode pass over its persons once; returns the slowest
    worker's seconds."""

    def __init__(self, cfg_id, per_worker):
        import multiprocessing as mp
        self.cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.per_worker = per_worker
        ctx = mp.get_context("spawn")
        self.w = []
        for i in range(self.cores):
            a, b = ctx.Pipe()
            pr = ctx.Process(target=_cpu_worker_main, args=(b, cfg_id, 100 + i, per_worker), daemon=True)
            pr.start()
            self.w.append((pr, a))
        for _, a in self.w:
            a.recv()

    def step(self):
        for _, a in self.w:
            a.send(1)
        return [a.recv() for _, a in self.w]

    def close(self):
        for pr, a in self.w:
            a.send(None)
        for pr, _ in self.w:
            pr.join(5)


def _cpu_calibrate(cfg_id):
    """persons/s of one core, from 8 persons run in this process."""
    st = _cpu_inputs(cfg_id, 1, 8)
    return 8 / _cpu_pass(st)[0]


def _cpu_record(cfg_id, arm, times):
    """times: list (steps) of lists (workers) of seconds."""
    n = arm.per_worker * arm.cores
    busy = [max(x[0] for x in t) for t in times]
    rate = n * len(times) / sum(busy)
    core_rate = float(np.mean([arm.per_worker / x[0] for t in times for x in t]))
    stage_us = {k: round(1e6 * float(np.mean([x[1][k] for t in times for x in t])) / arm.per_worker, 1)
                for k in times[0][0][1]}
    return {"value": rate, "unit": UNIT, "cores": arm.cores, "kind": "port",
            "per_core": core_rate, "per_worker": arm.per_worker, "stage_us_per_person_per_core": stage_us,
            "sample": "%d step(s) of %d persons (%d per core x %d single-threaded processes) of the same cfg-%d "
                      "encode+decode pass with oracle/np_port.py (numpy+cv2 restatement of the "
                      "reference, checked against its golden vectors); value = persons / slowest worker's "
                      "timed pass (%.2f s per step), inputs generated once per worker outside the timing"
                      % (len(times), n, arm.per_worker, arm.cores, cfg_id, float(np.mean(busy)))}


def cpu_baseline(cfg_id, seconds):
    """persons/s of the CPU port on all host cores: one warm-up pass, one timed pass of about `seconds`."""
    per_worker = max(32, min(1024, (int(seconds * 0.6 * _cpu_calibrate(cfg_id)) // 32) * 32))
    arm = CpuArm(cfg_id, per_worker)
    try:
        arm.step()
        return _cpu_record(cfg_id, arm, [arm.step()])
    finally:
        arm.close()


def config_dict(cfg, B, world):
    """`config` of the JSON line: the workload definition, identical for both arms."""
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    return {"workload": cfg.name + ": " + workload_text(cfg), "cfg": cfg_id_of(cfg), "persons_per_gpu_per_step": B,
            "sharding": "contiguous persons per rank, one all_gather of [B,K,3] keypoints per step",
            "l2": "inputs larger than L2 (%.1f GB of network outputs per step per GPU; the %d source images, %.0f MB, are the SURVEY 8d pool)"
                  % (2 * cfg.channels * h * w * 4 * B / 1e9, N_IMAGES, N_IMAGES * 480 * 640 * 3 / 1e6)}


def run_reference(args):
    """--impl reference: the reference's CPU path (numpy/cv2 port; the Python tree itself cannot
    travel to the GPU box) on all host cores, on the SAME workload definition as our arm (`config` is
    identical: same config, same input distributions, same persons per step when that fits).  Rank 0 only.
    Exactly --warmup untimed and --steps timed steps.  A step is one pass of every core over its share of
    the step's persons; when --gpus * --batch persons per step would push the run past about 2.5 minutes,
    the step is a bounded sample of them (stated in cpu_baseline.sample)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from udp_pose_b200.configs import CONFIGS
    cfg = CONFIGS[args.config]
    steps, warm = max(1, args.steps), max(0, args.warmup)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    full = args.batch * max(1, args.gpus)                       # persons our arm processes per step
    per_core_rate = 0.7 * _cpu_calibrate(args.config)           # all cores busy: ~0.7 of the single-core rate
    budget = 150.0 / (steps + warm)                             # seconds per step
    per_worker = -(-full // cores)
    bounded = per_worker / per_core_rate > budget
    if bounded:
        per_worker = int(budget * per_core_rate)
    per_worker = max(32, (per_worker // 32) * 32)               # whole validation batches of 32 (function.py)
    arm = CpuArm(args.config, per_worker)
    try:
        for _ in range(warm):
            arm.step()
        t0 = time.perf_counter()
        times = [arm.step() for _ in range(steps)]
        wall = time.perf_counter() - t0
    finally:
        arm.close()
    rec = _cpu_record(args.config, arm, times)
    v = rec["value"]
    n_step = per_worker * arm.cores
    rec["sample"] += ("; %d of the %d persons of our arm's step" % (n_step, full)) if bounded or n_step != full else "; the full step"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": 1e3 * n_step / v, "wall_ms_per_step": 1e3 * wall / steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 maps / f64 joint prologue / u8 crop",
            "data": "synthetic",
            "config": config_dict(cfg, args.batch, max(1, args.gpus)),
            "sample_persons_per_step": n_step,
            "note": "CPU arm: numpy/cv2 port of the reference path on %d host cores; it does not scale with --gpus" % arm.cores,
            "cpu_baseline": rec,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------
def link_ceiling(torch, up_bytes, down_bytes, persons, sync=None):
    """What the host link alone allows: pinned H2D and D2H copies of one chunk's bytes, both
    directions at once on two streams, no kernels (same probe as profiles/pcie_probe.py)."""
    hu, hd = torch.empty(up_bytes, dtype=torch.uint8).pin_memory(), torch.empty(down_bytes, dtype=torch.uint8).pin_memory()
    du, dd = torch.empty(up_bytes, dtype=torch.uint8, device="cuda"), torch.empty(down_bytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()

    def both():
        with torch.cuda.stream(s1):
            du.copy_(hu, non_blocking=True)
        with torch.cuda.stream(s2):
            hd.copy_(dd, non_blocking=True)
    both()
    torch.cuda.synchronize()
    if sync:
        sync()  # all ranks start copying together (pinning the buffers takes a different time on each)
    reps = 40   # ~0.3 s: long against any residual skew between the ranks
    t0 = time.perf_counter()
    for _ in range(reps):
        both()
    torch.cuda.synchronize()
    t = (time.perf_counter() - t0) / reps
    return {"h2d_gbs": round(up_bytes / t / 1e9, 1), "d2h_gbs": round(down_bytes / t / 1e9, 1),
            "ceiling_persons_per_s": round(persons / t, 1),
            "note": "pinned copies of one chunk's bytes, both directions concurrently, no kernels: the e2e bound"}


STAGES = ["warp_matrices", "crop", "rotate_points", "encode", "decode"]


def workload_text(cfg):
    """One-line description of a step for the JSON line's config.workload."""
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    enc = {"offset": "UDPv1 offset target (KPD %g)" % cfg.kpd, "gaussian": "UDP Gaussian target (sigma %g)" % cfg.sigma,
           "rsn": "RSN 5-scale full-map target"}[cfg.target_type]
    dec = "flip-test offset-vote decode" if cfg.target_type == "offset" else "flip-test DARK-refined decode"
    return "crop %dx%d + %s + %s, %dx%dx%d maps" % (H, W, enc, dec, cfg.channels, h, w)


class StepBench:
    """One configuration resident on this rank's GPU: the synthetic workload, the output buffers and the
    step (warp matrices -> crop -> joint map -> target encode -> flip-fused decode [-> keypoint gather])."""

    def __init__(self, cfg, B, rank, world, torch, dist, D):
        self.cfg, self.B, self.rank, self.world, self.torch, self.dist, self.D = cfg, B, rank, world, torch, dist, D
        K = cfg.num_joints
        (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
        self.wl = wl = make_device_workload(cfg, B, rank, torch, D)
        self.src = D.SourceImages(wl["images"], wl["image_index"])
        self.crop = torch.empty((B, H, W, 3), dtype=torch.uint8, device="cuda")
        tshape = (B, 5, K, h, w) if cfg.target_type == "rsn" else (B, cfg.channels, h, w)
        self.target = torch.empty(tshape, dtype=torch.float32, device="cuda")
        self.gathered = torch.empty((world * B, K, 3), dtype=torch.float32, device="cuda") if world > 1 else None
        self.j3 = wl["joints"].clone()

    def step(self, e=None):
        D, wl, cfg = self.D, self.wl, self.cfg
        if e: e[0].record()
        M = D.warp_matrices(wl["center"], wl["scale"], wl["rot"], cfg.image_size)
        if e: e[1].record()
        D.crop_batch(self.src, M, cfg.image_size, out=self.crop)
        if e: e[2].record()
        D.rotate_points_batch(wl["joints"], wl["rot"], wl["center"], wl["scale"], cfg.image_size, out=self.j3)
        if e: e[3].record()
        D.encode_batch(cfg, self.j3, wl["vis"], out=self.target)
        if e: e[4].record()
        r = D.decode_batch(cfg, wl["out"], wl["flipped"], wl["center"], wl["scale"], want_packed=True)
        if e: e[5].record()
        if self.world > 1:  # the one collective of the path: gather the keypoints
            self.dist.all_gather_into_tensor(self.gathered, r["packed"])
        if e: e[6].record()
        return r

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def measure(self, steps, warmup, local):
        """`warmup` untimed steps, then exactly `steps` timed ones between barrier + synchronize on both sides;
        CUDA events per stage; device time = first event of step 0 to last event of the last step, max over
        ranks; nvidia-smi clocks sampled during the timed region (rank 0)."""
        torch = self.torch
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(STAGES) + 2)] for _ in range(steps)]
        for _ in range(warmup):
            self.step()
        self.barrier()
        sampler = ClockSampler(local)
        if self.rank == 0:
            sampler.start()
            time.sleep(0.3)
        self.barrier()
        t0 = time.time()
        for i in range(steps):
            self.step(ev[i])
        self.barrier()
        t1 = time.time()
        clocks = sampler.stop(t0, t1) if self.rank == 0 else None
        total_ms = ev[0][0].elapsed_time(ev[-1][-1])
        stage_ms = {nm: sum(e[i].elapsed_time(e[i + 1]) for e in ev) / steps for i, nm in enumerate(STAGES)}
        stage_ms["gather"] = sum(e[5].elapsed_time(e[6]) for e in ev) / steps
        tt = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return {"total_ms": float(tt[0]), "stage_ms": stage_ms, "clocks": clocks, "steps": steps}

    def tables(self, m, peak):
        """kernels / paths / value of a measure() result."""
        cfg, B = self.cfg, self.B
        ab = algorithmic_bytes(cfg, self.wl["src_area"])
        stage_ms = m["stage_ms"]
        kern = {}
        for nm in STAGES:
            gbs = ab[nm] * B / (stage_ms[nm] * 1e-3) / 1e9
            kern[nm] = {"ms": round(stage_ms[nm], 4), "bytes_per_person": round(ab[nm], 1),
                        "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4)}
        kern["decode"]["survey_bytes_per_person"] = ab["decode_survey"]
        kern["decode"]["frac_on_survey_bytes"] = round(ab["decode_survey"] * B / (stage_ms["decode"] * 1e-3) / 1e9 / peak, 4)
        # the north star states its target per PATH: encode = warp + target (four kernels), decode = one kernel
        enc_st = ["warp_matrices", "crop", "rotate_points", "encode"]
        enc_ms, enc_b = sum(stage_ms[n] for n in enc_st), sum(ab[n] for n in enc_st)
        paths = {"encode_path": {"stages": enc_st, "ms": round(enc_ms, 4), "bytes_per_person": round(enc_b, 1),
                                 "achieved_gbs": round(enc_b * B / (enc_ms * 1e-3) / 1e9, 1)},
                 "decode_path": {"stages": ["decode"], "ms": kern["decode"]["ms"], "bytes_per_person": kern["decode"]["bytes_per_person"],
                                 "achieved_gbs": kern["decode"]["achieved_gbs"]}}
        for v in paths.values():
            v["frac"] = round(v["achieved_gbs"] / peak, 4)
        value = self.world * B * m["steps"] / (m["total_ms"] * 1e-3)
        return kern, paths, value, ab

    def free(self):
        for k in ("wl", "src", "crop", "target", "gathered", "j3"):
            setattr(self, k, None)
        self.torch.cuda.empty_cache()


SWEEP_POINTS = [1 << 10, 1 << 12, 1 << 14, 1 << 16, 1 << 18, 1 << 20]
SWEEP_CHUNK = 1 << 14


def run_sweep(sb, local, use_graphs=True):
    """BASELINE.json configs[4]: the large-batch sweep, STRONG scaling -- a point is N persons in total, rank r
    owns the contiguous block of ceil(N/R) (RSN/lib/utils/dataloader.py:88-93), processed as a stream of chunks
    of at most 16 Ki persons from the resident pool of `sb` (1 Mi persons of network output would be 418 GB; a
    chunk's 6.8 GB exceed L2, so repeating it warms nothing), keypoints gathered per chunk.  Per point: one
    warm-up pass, then `reps` timed passes (CUDA events, max over ranks per pass), the MEDIAN reported, clocks
    sampled over the timed passes."""
    torch, dist, D, wl, cfg, world, rank = sb.torch, sb.dist, sb.D, sb.wl, sb.cfg, sb.world, sb.rank
    srcs, graph_note = {}, []

    graphs = {}

    def chunk_kernels(m):
        sl = slice(0, m)
        M = D.warp_matrices(wl["center"][sl], wl["scale"][sl], wl["rot"][sl], cfg.image_size)
        D.crop_batch(srcs[m], M, cfg.image_size, out=sb.crop[sl])
        D.rotate_points_batch(wl["joints"][sl], wl["rot"][sl], wl["center"][sl], wl["scale"][sl], cfg.image_size, out=sb.j3[sl])
        D.encode_batch(cfg, sb.j3[sl], wl["vis"][sl], out=sb.target[sl])
        return D.decode_batch(cfg, wl["out"][sl], wl["flipped"][sl], wl["center"][sl], wl["scale"][sl], want_packed=True)["packed"]

    def chunk_pass(m):
        # the five launches of a chunk replay from a CUDA graph (udp_pose_b200.graphs.CapturedPass: below a few thousand
        # persons a pass is launch-bound); the one collective stays outside the capture
        packed = graphs[m].replay() if m in graphs else chunk_kernels(m)
        if world > 1:
            dist.all_gather_into_tensor(sb.gathered[: world * m], packed)

    points = []
    for N in SWEEP_POINTS:
        per = -(-N // world)                                   # ceil(N / R): every rank runs the padded block
        chunks = [SWEEP_CHUNK] * (per // SWEEP_CHUNK) + ([per % SWEEP_CHUNK] if per % SWEEP_CHUNK else [])
        for m in set(chunks):                                  # per-person image tables: host work, outside the timing
            if m not in srcs:
                srcs[m] = D.SourceImages(wl["images"], wl["image_index"][:m])
                if use_graphs:
                    try:
                        from udp_pose_b200.graphs import CapturedPass
                        graphs[m] = CapturedPass(lambda m=m: chunk_kernels(m))
                    except Exception as exc:  # capture unavailable: direct launches (recorded in the block)
                        graph_note.append(str(exc)[:80])
        reps = 7 if per <= (1 << 16) else 3
        ts = []
        sampler = ClockSampler(local)
        t0 = t1 = time.time()
        for rep in range(reps + 1):                            # repetition 0 is the warm-up
            sb.barrier()
            if rep == 1:
                if rank == 0:
                    sampler.start()
                    time.sleep(0.25)
                sb.barrier()
                t0 = time.time()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for m in chunks:
                chunk_pass(m)
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rep:
                ts.append(float(t[0]))
        sb.barrier()
        t1 = time.time()
        clocks = sampler.stop(t0, t1) if rank == 0 else None
        ms = float(np.median(ts))
        points.append({"persons": N, "persons_per_gpu": per, "chunks_per_gpu": len(chunks), "ms": round(ms, 4),
                       "ms_min": round(min(ts), 4), "ms_max": round(max(ts), 4), "reps": reps,
                       "persons_per_s": round(N / (ms * 1e-3)), "clocks": clocks})
    return {"cfg": cfg_id_of(cfg), "workload": cfg.name, "scaling": "strong", "n_gpus": world,
            "timing": "CUDA events, max over ranks per pass, median of `reps` passes after one warm-up pass",
            "cuda_graphs": bool(graphs), "cuda_graph_errors": graph_note,
            "points": points}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from udp_pose_b200 import _lib, device_api as D
    from udp_pose_b200.configs import CONFIGS
    from udp_pose_b200.pipeline import HostBatch, HostPipeline

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.load()
    cfg = CONFIGS[args.config]
    B, K = args.batch, cfg.num_joints
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"

    sb = StepBench(cfg, B, rank, world, torch, dist, D)
    wl, src, barrier = sb.wl, sb.src, sb.barrier
    head = sb.measure(args.steps, max(3, args.warmup), local)
    total_ms, stage_ms, clocks = head["total_ms"], head["stage_ms"], head["clocks"]
    target = sb.target

    # ---- SURVEY 8(f)-1 (outside the timed step): crop + ToTensor + Normalize + mirrored input, one gather
    fused = None
    if world == 1 and not args.no_fused:
        M = D.warp_matrices(wl["center"], wl["scale"], wl["rot"], cfg.image_size)
        fo = torch.empty((B, 3, H, W), dtype=torch.float32, device="cuda")
        ff = torch.empty_like(fo)
        fused = {}
        for key, flip in (("crop_normalize", False), ("crop_normalize_flip", True)):
            for _ in range(3):
                D.crop_normalize_batch(src, M, cfg.image_size, want_flipped=flip, out=fo, out_flipped=ff if flip else None)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = max(5, min(args.steps, 20))
            a.record()
            for _ in range(reps):
                D.crop_normalize_batch(src, M, cfg.image_size, want_flipped=flip, out=fo, out_flipped=ff if flip else None)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / reps
            bpp = 3.0 * wl["src_area"] + 24 + 12.0 * W * H * (2 if flip else 1)
            fused[key] = {"ms": round(ms, 4), "bytes_per_person": round(bpp, 1)}
        del fo, ff
        torch.cuda.empty_cache()
        if cfg.target_type != "rsn":  # accuracy()/PCK of the network output against the target
            cs = 3 if cfg.target_type == "offset" else 1
            for _ in range(3):
                D.accuracy_pck(wl["out"], target, cs)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                D.accuracy_pck(wl["out"], target, cs)
            b.record()
            torch.cuda.synchronize()
            fused["accuracy_pck"] = {"ms": round(a.elapsed_time(b) / reps, 4), "bytes_per_person": 2.0 * K * h * w * 4 + 8 * K + 8 * K}

    if fused is not None:  # SURVEY 8(f)-4 (first half): the demo engines' quarter-offset decoder over the same network output
        qb = torch.cat([wl["center"].float(), wl["scale"].float()], 1).contiguous()
        for _ in range(3):
            D.decode_quarter(wl["out"], qb)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            D.decode_quarter(wl["out"], qb)
        b.record()
        torch.cuda.synchronize()
        fused["decode_quarter"] = {"ms": round(a.elapsed_time(b) / reps, 4), "bytes_per_person": cfg.channels * h * w * 4.0 + 16 + 8 * cfg.channels,
                                   "maps_per_person": cfg.channels}

    if fused is not None:  # SURVEY 8(f)-4 (second half): RSN's legacy get_results over the same network output
        qc, qs = wl["center"].float().contiguous(), wl["scale"].float().contiguous()
        ksz = 9 if cfg.num_joints == 16 else 5   # cfg.TEST.GAUSSIAN_KERNEL: RSN/dataset/attribute.py:49 (COCO), :92 (MPII)
        nb_l = min(B, 4096)
        for _ in range(2):
            D.rsn_get_results(wl["out"][:nb_l], qc[:nb_l], qs[:nb_l], (H, W), ksz)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            D.rsn_get_results(wl["out"][:nb_l], qc[:nb_l], qs[:nb_l], (H, W), ksz)
        b.record()
        torch.cuda.synchronize()
        lg_ms = a.elapsed_time(b) / 5
        legacy = {"ms": round(lg_ms, 4), "persons": nb_l, "maps_per_person": cfg.channels, "gaussian_kernel": ksz,
                  "persons_per_s": round(nb_l / (lg_ms * 1e-3)),
                  "achieved_gbs": round((cfg.channels * h * w * 4.0 + 16 + 24 * cfg.channels) * nb_l / (lg_ms * 1e-3) / 1e9, 1),
                  "note": "udp_rsn_get_results (RSN/exps/RSN18.coco/test.py:25-71): float64 blur of the zero-padded map + top-2 arg-max, "
                          "issue-bound, no HBM roofline claimed"}
    else:
        legacy = None

    # ---- SURVEY 8(f)-3 (outside the timed step): COCO rescoring + OKS-NMS of B pose results, one CTA per image
    nms = None
    if world == 1 and not args.no_fused:
        from udp_pose_b200 import synth as _synth
        n_img = max(1, B // 16)
        npreds, nboxes, nimg = _synth.nms_persons(n_img, 32, 17, np.random.default_rng(77))
        order = np.argsort(nimg, kind="stable")
        off = np.zeros(n_img + 1, np.int32)
        off[1:] = np.cumsum(np.bincount(nimg, minlength=n_img))
        dp = torch.from_numpy(npreds[order]).cuda()
        da, db = torch.from_numpy(nboxes[order, 4].copy()).cuda(), torch.from_numpy(nboxes[order, 5].copy()).cuda()
        doff = torch.from_numpy(off).cuda()
        nms = {"persons": int(len(npreds)), "images": int(n_img)}
        for key, soft in (("hard_ms", False), ("soft_ms", True)):
            for _ in range(3):
                r = D.oks_nms_batch(dp, da, D.coco_rescore(dp, db, 0.2), doff, 32, 0.9, soft=soft)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                r = D.oks_nms_batch(dp, da, D.coco_rescore(dp, db, 0.2), doff, 32, 0.9, soft=soft)
            b.record()
            torch.cuda.synchronize()
            nms[key] = round(a.elapsed_time(b) / 10, 4)
            nms["kept_" + key[:4]] = int(r["keep_cnt"].sum())
        nms["persons_per_s"] = round(nms["persons"] / (nms["hard_ms"] * 1e-3))
        nms["note"] = ("udp_coco_rescore + udp_oks_nms (coco.py:321-356, nms.py:75-180), in_vis_thre 0.2, oks_thre 0.9, up to 32 persons "
                       "per image; sequential greedy picks per image, float64 exp: latency/FP64-bound, 232 B per person, no HBM roofline")

    kern, paths, value, ab = sb.tables(head, peak)
    host = wl["host"]
    out_head, fl_head = wl["out"][:args.e2e_batch].clone(), wl["flipped"][:args.e2e_batch].clone()
    sb.free()
    del wl, src, target
    torch.cuda.empty_cache()

    # ---- end to end through host buffers (pinned), H2D and D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        nb = args.e2e_batch
        hb = HostBatch(cfg, nb, N_IMAGES)
        hb.images.copy_(torch.from_numpy(host["images"]))
        hb.image_index[:] = np.arange(nb) % N_IMAGES
        for dst, key in ((hb.center, "center"), (hb.scale, "scale"), (hb.rot, "rot"),
                         (hb.joints, "joints"), (hb.joints_vis, "vis")):
            dst.copy_(torch.from_numpy(host[key][:nb]))
        hb.out.copy_(out_head)
        hb.out_flipped.copy_(fl_head)
        pipe = HostPipeline(cfg, args.e2e_chunk, N_IMAGES)
        for _ in range(3):
            pipe.run(hb)
        barrier()
        e_steps = max(3, min(args.steps, 10))
        pipe.launches = 0
        ts = time.perf_counter()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(e_steps):
            pipe.run(hb)
        b.record()
        barrier()
        wall = time.perf_counter() - ts
        ems = max(a.elapsed_time(b), wall * 1e3)
        te = torch.tensor([ems], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        # the same steps enqueued back to back (a streaming caller): the uploads of step i+1 overlap the downloads of step i
        barrier()
        ts = time.perf_counter()
        for _ in range(e_steps):
            pipe.run(hb, sync=False)
        pipe.synchronize()
        stream_s = time.perf_counter() - ts
        barrier()
        tsr = torch.tensor([stream_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tsr, op=dist.ReduceOp.MAX)
        stream_s = float(tsr[0])
        e2e = {"value": world * nb * e_steps / (float(te[0]) * 1e-3), "unit": UNIT,
               "streaming_value": round(world * nb * e_steps / stream_s, 1),
               "h2d_bytes_per_step": pipe.h2d_bytes(hb), "d2h_bytes_per_step": hb.d2h_bytes(),
               "h2d_bytes_per_step_full_tensors": hb.h2d_bytes(),
               "persons_per_step_per_gpu": nb, "steps": e_steps, "chunk": args.e2e_chunk,
               "api": "udp_pose_b200.pipeline.HostPipeline.run (pinned host in -> H2D -> 5 kernels via the C ABI -> D2H: crop, target, target_weight, keypoints)"
                      + ("; offset configs upload the classification channels only and the decoder gathers its 7x7 offset windows in place from "
                         "the caller's pinned buffers (udp_decode_offset_split), counted in h2d_bytes_per_step" if pipe.split else ""),
               "launches": pipe.launches}
        # what the host links alone allow, all ranks copying at the same time (no kernels)
        barrier()
        link = link_ceiling(torch, pipe.h2d_bytes(hb) // (nb // args.e2e_chunk), hb.d2h_bytes() // (nb // args.e2e_chunk), args.e2e_chunk, barrier)
        mine = [link["h2d_gbs"], link["d2h_gbs"], link["ceiling_persons_per_s"]]
        lt = torch.tensor(mine, dtype=torch.float64, device="cuda")
        if world > 1:
            allr = [torch.empty_like(lt) for _ in range(world)]
            dist.all_gather(allr, lt)
            link["per_rank"] = [{"h2d_gbs": round(float(v[0]), 1), "d2h_gbs": round(float(v[1]), 1),
                                 "ceiling_persons_per_s": round(float(v[2]), 1)} for v in allr]
            dist.all_reduce(lt, op=dist.ReduceOp.SUM)
            link["note"] += "; summed over the %d ranks, which copy concurrently (per_rank: each rank's own figures)" % world
        link["h2d_gbs"], link["d2h_gbs"], link["ceiling_persons_per_s"] = (round(float(v), 1) for v in lt)
        e2e["link"] = link
        del hb, pipe
    del out_head, fl_head
    torch.cuda.empty_cache()

    # ---- the other BASELINE.json configs, same step, same batch, own clock record each
    cfg_block, sweep = None, None
    if not args.no_configs:
        cfg_block = []
        for cid in sorted(CONFIGS):
            if cid == args.config:
                continue
            o = StepBench(CONFIGS[cid], B, rank, world, torch, dist, D)
            m = o.measure(max(3, min(args.steps, 10)), 3, local)
            k2, p2, v2, _ = o.tables(m, peak)
            cfg_block.append({"cfg": cid, "workload": CONFIGS[cid].name + ": " + workload_text(CONFIGS[cid]),
                              "persons_per_gpu_per_step": B, "steps": m["steps"], "ms_per_step": round(m["total_ms"] / m["steps"], 4),
                              "value": round(v2, 1), "kernels": {n: {"ms": k2[n]["ms"], "frac": k2[n]["frac"], "achieved_gbs": k2[n]["achieved_gbs"]} for n in STAGES},
                              "encode_path_frac": p2["encode_path"]["frac"], "decode_path_frac": p2["decode_path"]["frac"],
                              "gather_ms": round(m["stage_ms"]["gather"], 4), "clocks": m["clocks"]})
            if cid == SWEEP_CFG and not args.no_sweep and B >= SWEEP_CHUNK:  # the sweep draws its chunks from this pool
                sweep = run_sweep(o, local, not args.no_sweep_graphs)
            o.free()
            del o
    if sweep is None and not args.no_sweep:
        o = StepBench(CONFIGS[SWEEP_CFG], SWEEP_CHUNK, rank, world, torch, dist, D)
        sweep = run_sweep(o, local, not args.no_sweep_graphs)
        o.free()
        del o

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    dom = max(STAGES, key=lambda n: stage_ms[n])
    # measured DRAM bytes of one launch (ncu dram__bytes_read.sum + dram__bytes_write.sum), committed in
    # profiles/traffic.json by profiles/make_traffic.py together with a hash of the kernel sources it was captured
    # from: used only if that hash is the tree's (a stale capture reads as null), scaled linearly to the batch size
    traffic, traffic_note = None, "no capture for this configuration"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("source_sha16") != kernel_source_sha16():
            traffic_note = "profiles/traffic.json was captured from other kernel sources (stale): not used"
        else:
            traffic = tj["cfg%d" % args.config][dom] * B / tj["persons"]
            traffic_note = tj.get("note", "")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_gbs"], "peak": peak,
                "unit": "GB/s", "frac": kern[dom]["frac"], "traffic": traffic, "traffic_note": traffic_note, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": ab[dom] * B,
                "note": "slowest stage of the step; per-stage figures under `kernels`, per-path under `paths` (DESIGN.md section 4 says what bounds each kernel)"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 maps / f64 joint prologue / u8 crop", "data": "synthetic",
            "config": config_dict(cfg, B, world),
            "roofline": roofline, "kernels": kern, "paths": paths,
            "gather_ms": round(stage_ms["gather"], 4),
            "gpu_launches": (len(STAGES) - 1 + D.crop_launches(cfg.image_size)) * args.steps,
            "clocks": clocks}
    if fused:
        for v in fused.values():
            v["achieved_gbs"] = round(v["bytes_per_person"] * B / (v["ms"] * 1e-3) / 1e9, 1)
            v["frac"] = round(v["achieved_gbs"] / peak, 4)
        line["next"] = dict(fused, note="SURVEY 8(f) ranks 1-4, timed outside the step: udp_warp_affine_normalize_f32 writes Normalize(ToTensor(crop)) "
                                        "as float32 CHW (and the x-mirrored flip-test input) straight from the gather; udp_accuracy_pck is evaluate.py:accuracy "
                                        "on device tensors; udp_decode_quarter is tools/infer_utils/decode.py:get_final_preds over every channel of the same output; "
                                        "oks_nms is coco.py:321-356; none is part of `value`")
    if nms:
        line.setdefault("next", {})["oks_nms"] = nms
    if legacy:
        line.setdefault("next", {})["rsn_get_results"] = legacy
    if e2e:
        line["e2e"] = e2e
    if cfg_block is not None:
        line["configs"] = cfg_block
    if sweep is not None:
        line["sweep"] = sweep
    if not args.no_cpu and world == 1:
        line["cpu_baseline"] = cpu_baseline(args.config, args.cpu_seconds)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


SWEEP_CFG = 5


def kernel_source_sha16():
    """sha256 (first 16 hex digits) over the CUDA sources and headers the library is built from."""
    import hashlib
    from udp_pose_b200 import build as B
    hs = hashlib.sha256()
    for f in sorted(B.SOURCES + B.HEADERS):
        with open(os.path.join(B.CSRC, f), "rb") as fh:
            hs.update(fh.read())
    return hs.hexdigest()[:16]


if __name__ == "__main__":
    a = parse()
    N_IMAGES = a.images
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
