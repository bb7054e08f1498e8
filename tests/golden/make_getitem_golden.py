"""Generate tests/golden/getitem.npz and tests/golden/fliplr.npz from the UNMODIFIED reference
(build container only: needs /root/reference):

    python tests/golden/make_getitem_golden.py

getitem.npz -- the reference's own `JointsDataset.__getitem__` (deep_hrnet/lib/dataset/JointsDataset.py:172-256)
in TRAIN mode under seeded `np.random` / `random`, on a record with more than NUM_JOINTS_HALF_BODY visible
joints, the image written to a temporary PNG (what cv2.imread + BGR2RGB hands the method is stored in
`images`, case n uses images[n % 2]), `flip=True`, `upper_body_ids` / `aspect_ratio` / `flip_pairs` set as lib/dataset/coco.py:66,90-95
does.  Cases cover: half-body taken (PROB_HALF_BODY = 1) and not (0), the flip branch taken and not,
rotation drawn and not, gaussian and offset targets.  Stored per case: seeds, switches, and the four
outputs (input, target, target_weight, meta joints / joints_vis / center / scale / rotation) plus the
next number of both generators (pins the number and order of draws).

fliplr.npz -- `fliplr_joints` (deep_hrnet/lib/utils/transforms.py:50-64) and RSN's `flip_joints`
(RSN/lib/utils/transforms.py:70-79) on seeded joints, COCO and MPII pair tables.
"""
import os
import random
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

from oracle import ref_import as R  # noqa: E402
from udp_pose_b200 import synth  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402

# (cfg id, PROB_HALF_BODY, seed).  Python's `random` decides rotation and flip (:218-220): seeds 0, 2, 5 take
# neither, 1 rotation only, 15, 22, 26, 37 the flip only, 4, 7, 16, 28 both.
CASES = [(1, 1.0, 0), (1, 1.0, 15), (1, 1.0, 4), (1, 1.0, 1), (1, 0.0, 22), (1, 0.0, 7), (1, 0.0, 2),
         (2, 1.0, 26), (2, 1.0, 16), (4, 1.0, 37), (4, 1.0, 28), (4, 1.0, 5)]


def record(cfg, rng):
    """One db record: a person standing in a 320x240 image, 14+ visible joints."""
    K = cfg.num_joints
    W, H = cfg.image_size
    ctr = np.array([rng.uniform(120, 200), rng.uniform(90, 150)])
    bh = rng.uniform(100, 200)
    bw = bh * W / H
    j = np.zeros((K, 3))
    j[:, 0] = ctr[0] + rng.uniform(-0.45, 0.45, K) * bw
    j[:, 1] = ctr[1] + np.linspace(-0.45, 0.45, K) * bh   # head to feet: upper ids up, lower ids down
    v = np.zeros((K, 3))
    vis = (rng.uniform(size=K) < 0.9).astype(np.float64)
    v[:, 0] = v[:, 1] = vis
    dt = np.dtype(cfg.cs_dtype)
    return dict(joints_3d=j, joints_3d_vis=v, center=ctr.astype(dt),
                scale=(np.array([bw, bh]) / 200.0 * 1.25).astype(dt), score=0.75)


def main():
    import cv2
    d = {"versions": np.array("numpy %s cv2 %s" % (np.__version__, cv2.__version__)),
         "cases": np.array(CASES, np.float64)}
    tmp = tempfile.mkdtemp()
    images = synth.source_images(2, np.random.default_rng(6999), h=240, w=320)   # RGB, what the method sees after BGR2RGB
    d["images"] = images
    for i in range(2):
        cv2.imwrite(os.path.join(tmp, "p%d.png" % i), cv2.cvtColor(images[i], cv2.COLOR_RGB2BGR))
    for n, (cid, phb, seed) in enumerate(CASES):
        cfg = CONFIGS[cid]
        rng = np.random.default_rng(7000 + seed)
        rec = record(cfg, rng)
        rec["image"] = os.path.join(tmp, "p%d.png" % (n % 2))
        tt = "offset" if cfg.target_type == "offset" else "gaussian"
        c = R.make_cfg(tt, cfg.image_size, cfg.heatmap_size, cfg.sigma, kpd=cfg.kpd)
        c.DATASET.PROB_HALF_BODY = phb
        c.DATASET.ROT_FACTOR = cfg.rot_factor
        ds = R.hrnet().joints_dataset.JointsDataset(c, "", "train", True)
        ds.num_joints = cfg.num_joints
        ds.flip_pairs = cfg.flip_pairs
        ds.upper_body_ids = (0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10) if cfg.num_joints == 17 else (7, 8, 9, 10, 11, 12, 13, 14, 15)
        ds.aspect_ratio = cfg.image_size[0] * 1.0 / cfg.image_size[1]
        ds.db = [rec]
        keep = {k: np.array(v, copy=True) for k, v in rec.items() if isinstance(v, np.ndarray)}
        np.random.seed(seed)
        random.seed(seed)
        inp, target, tw, meta = ds[0]
        nxt = np.array([np.random.rand(), random.random()])
        p = "c%d_" % n
        for k, v in keep.items():
            d[p + "rec_" + k] = v
        d[p + "input"] = inp
        d[p + "target"] = target.numpy()
        d[p + "target_weight"] = tw.numpy()
        for k in ("joints", "joints_vis", "center", "scale"):
            d[p + "meta_" + k] = np.asarray(meta[k])
        d[p + "meta_rotation"] = np.float64(meta["rotation"])
        d[p + "next"] = nxt
        d[p + "trans"] = R.hrnet().joints_dataset.get_warpmatrix(meta["rotation"], np.asarray(meta["center"]) * 2.0,
                                                                  ds.image_size - 1.0, meta["scale"])
        print(n, "cfg", cid, "phb", phb, "seed", seed, "rot %.2f" % meta["rotation"], "c", np.round(meta["center"], 1),
              "s", np.round(meta["scale"], 3), "tw", int(tw.sum()))
    np.savez_compressed(os.path.join(HERE, "getitem.npz"), **d)
    print("getitem.npz %.2f MB" % (os.path.getsize(os.path.join(HERE, "getitem.npz")) / 1e6))

    # ---- a4: fliplr_joints / flip_joints
    f = {}
    hr, rs = R.hrnet().transforms, R.rsn().transforms
    for name, cid, width in (("coco", 1, 320), ("mpii", 4, 517)):
        cfg = CONFIGS[cid]
        rng = np.random.default_rng(7100 + cid)
        K = cfg.num_joints
        j = np.zeros((6, K, 3))
        j[..., 0] = rng.uniform(-20, width + 20, (6, K))
        j[..., 1] = rng.uniform(0, 400, (6, K))
        j[0, :, 2] = rng.uniform(0, 1, K)            # a non-zero third column is multiplied by vis[:, 2] too
        v = np.zeros((6, K, 3))
        vv = rng.integers(0, 3, (6, K)).astype(np.float64)   # COCO visibility flags 0/1/2 multiply the coordinates
        v[..., 0] = v[..., 1] = vv
        v[0, :, 2] = 1.0
        f[name + "_joints"], f[name + "_vis"], f[name + "_width"] = j, v, np.int64(width)
        a = [hr.fliplr_joints(j[i].copy(), v[i].copy(), width, cfg.flip_pairs) for i in range(6)]
        b = [rs.flip_joints(j[i].copy(), v[i].copy(), width, cfg.flip_pairs) for i in range(6)]
        f[name + "_fliplr_joints"], f[name + "_fliplr_vis"] = np.stack([x[0] for x in a]), np.stack([x[1] for x in a])
        f[name + "_flip_joints"], f[name + "_flip_vis"] = np.stack([x[0] for x in b]), np.stack([x[1] for x in b])
    np.savez_compressed(os.path.join(HERE, "fliplr.npz"), **f)
    print("fliplr.npz %.3f MB" % (os.path.getsize(os.path.join(HERE, "fliplr.npz")) / 1e6))


if __name__ == "__main__":
    main()
