"""Generate tests/golden/cfg{1..5}.npz from the UNMODIFIED reference (run in the build
container, where /root/reference exists):

    python tests/golden/make_golden.py

Each fixture stores seeded synthetic inputs (udp_pose_b200/synth.py) and what the reference's
own functions return for them -- get_warpmatrix, cv2.warpAffine, rotate_points,
generate_target / RSN generate_heatmap, flip_back(_offset)+average -> get_final_preds.
Versions used are recorded in the fixture (`versions`).  The reference ships no golden
vectors of its own, so these ARE the pin for the oracle.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import refcalls as RC  # noqa: E402
from udp_pose_b200 import synth  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402

N_ENC, N_CROP, N_DEC = 6, 3, 2


def quant(x, q):
    return (np.round(x * q) / q).astype(np.float32)


def make(cid):
    import cv2
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(1000 * cid)
    d = {"versions": np.array("numpy %s cv2 %s" % (np.__version__, cv2.__version__))}
    # --- crop side: boxes with rotation, small source images partly left by the boxes
    c, s, r = synth.boxes(cfg, N_ENC, rng, rotate=True)
    r[0] = 0.0
    c = (c * 0.5).astype(c.dtype)
    s = (s * 0.5).astype(s.dtype)  # boxes scaled to the 320x240 fixture images
    d.update(center=c, scale=s, rot=r)
    d["warp_matrix"] = RC.warp_matrices(cfg, c, s, r)
    imgs = synth.source_images(2, rng, h=240, w=320)
    ii = np.arange(N_CROP) % 2
    d.update(src_images=imgs, src_index=ii.astype(np.int32))
    d["crop"] = RC.warp_affine(cfg, imgs, ii, d["warp_matrix"][:N_CROP])
    K = cfg.num_joints
    js = np.zeros((N_ENC, K, 3))
    js[..., 0] = rng.uniform(0, 320, (N_ENC, K))
    js[..., 1] = rng.uniform(0, 240, (N_ENC, K))
    d["joints_src"] = js
    d["joints_rot"] = RC.rotate_points(cfg, js, c, s, r)
    # --- target side
    j, v = synth.joints(cfg, N_ENC, rng, rsn=cfg.target_type == "rsn")
    # edge cases: window clipped at each border / fully outside, negative mu truncation
    W, H = cfg.image_size
    j[0, 0, :2] = (-0.3 * (W - 1) / 47, 5.0)
    j[0, 1, :2] = (W - 1 + 20.0, H - 1 + 20.0)
    j[0, 2, :2] = (-40.0, -40.0)
    j[0, 3, :2] = (0.0, 0.0)
    j[0, 4, :2] = (W - 1.0, H - 1.0)
    v[0, :5, :2] = 1.0
    d.update(joints=j, joints_vis=v)
    if cfg.target_type == "rsn":
        d["labels"] = RC.rsn_labels(cfg, j, v)
        j2, v2, Mi = RC.rsn_transform_joints(cfg, js, v, c, s, r)
        d.update(rsn_joints=j2, rsn_vis=v2, rsn_minv=Mi)
    else:
        t, tw = RC.generate_target(cfg, j, v)
        d.update(target=t, target_weight=tw)
        jw = np.linspace(1.0, 1.5, K).astype(np.float32)
        _, tw2 = RC.generate_target(cfg, j, v, joints_weight=jw)
        d.update(joints_weight=jw, target_weight_jw=tw2)
    # --- decode side
    out, fl = synth.heatmaps(cfg, N_DEC, rng)
    q = 16.0 if cfg.pre_div else 4096.0
    out, fl = quant(out, q), quant(fl, q)
    cd, sd, _ = synth.boxes(cfg, N_DEC, rng)
    d.update(dec_out=out, dec_flipped=fl, dec_center=cd, dec_scale=sd)
    avg = RC.flip_average(cfg, out, fl)
    p, mv, pin = RC.final_preds(cfg, avg, cd, sd)
    d.update(preds=p, maxvals=mv, preds_in=pin)
    p, mv, pin = RC.final_preds(cfg, out, cd, sd)
    d.update(preds_noflip=p, maxvals_noflip=mv, preds_in_noflip=pin)
    if cfg.target_type != "offset":
        p, mv, pin = RC.final_preds(cfg, avg, cd, sd, post_process=False)
        d.update(preds_nopost=p, maxvals_nopost=mv, preds_in_nopost=pin)
    if cid == 1:
        adv = quant(synth.adversarial_heatmaps(cfg, rng), 65536.0)
        ca, sa, _ = synth.boxes(cfg, adv.shape[0], rng)
        p, mv, pin = RC.final_preds(cfg, adv, ca, sa)
        d.update(adv_hm=adv, adv_center=ca, adv_scale=sa, adv_preds=p, adv_maxvals=mv,
                 adv_preds_in=pin)
    path = os.path.join(HERE, "cfg%d.npz" % cid)
    np.savez_compressed(path, **d)
    print(path, "%.2f MB" % (os.path.getsize(path) / 1e6))


if __name__ == "__main__":
    for cid in CONFIGS:
        make(cid)
