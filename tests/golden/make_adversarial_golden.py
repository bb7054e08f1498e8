"""Generate tests/golden/adv_cfg{2,3,4}.npz and tests/golden/dec8.npz from the UNMODIFIED reference (build container
only: needs /root/reference):

    python tests/golden/make_adversarial_golden.py

adv_cfg2 -- adversarial maps for the offset-voting decoder (synth.adversarial_offset_maps), with and without a
            flipped partner tensor: ties after the 15x15 blur, all-zero / all-negative classification maps, peaks on
            every border and corner (reflecting 15x15 and 7x7 windows), single pixel, plateaus.
adv_cfg3, adv_cfg4 -- the cfg-1 adversarial set (synth.adversarial_heatmaps) at the 96x72 shape and at the MPII
            64x64 shape with float64 centre / scale.
dec8     -- eight realistic persons per config through flip_back(_offset) + average + get_final_preds (the cfg*.npz
            fixtures hold two).
Inputs are quantised so that they survive compression; outputs are what the reference's own functions return."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import refcalls as RC  # noqa: E402
from udp_pose_b200 import synth  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402


def quant(x, q):
    return (np.round(x * q) / q).astype(np.float32)


def main():
    import cv2
    ver = np.array("numpy %s cv2 %s" % (np.__version__, cv2.__version__))
    # ---- offset decoder
    cfg = CONFIGS[2]
    rng = np.random.default_rng(8200)
    # classification channels on a fine grid, offset noise on a coarse one (fixture size)
    raw = synth.adversarial_offset_maps(cfg, rng)
    adv = quant(raw, 1024.0)
    adv[:, 0::3] = quant(raw[:, 0::3], 65536.0)
    rawf = synth._flip_of(adv.astype(np.float64), cfg) + rng.standard_normal(adv.shape) * 0.01
    fl = quant(rawf, 1024.0)
    fl[:, 0::3] = quant(rawf[:, 0::3], 65536.0)
    c, s, _ = synth.boxes(cfg, adv.shape[0], rng)
    d = {"versions": ver, "adv_out": adv, "adv_flipped": fl, "adv_center": c, "adv_scale": s}
    p, mv, pin = RC.final_preds(cfg, adv, c, s)
    d.update(preds_noflip=p, maxvals_noflip=mv, preds_in_noflip=pin)
    p, mv, pin = RC.final_preds(cfg, RC.flip_average(cfg, adv, fl), c, s)
    d.update(preds=p, maxvals=mv, preds_in=pin)
    np.savez_compressed(os.path.join(HERE, "adv_cfg2.npz"), **d)
    # ---- gaussian decoders at the other shapes
    for cid in (3, 4):
        cfg = CONFIGS[cid]
        rng = np.random.default_rng(8000 + 100 * cid)
        adv = quant(synth.adversarial_heatmaps(cfg, rng), 65536.0)
        c, s, _ = synth.boxes(cfg, adv.shape[0], rng)
        p, mv, pin = RC.final_preds(cfg, adv, c, s)
        np.savez_compressed(os.path.join(HERE, "adv_cfg%d.npz" % cid), versions=ver, adv_hm=adv, adv_center=c, adv_scale=s,
                            adv_preds=p, adv_maxvals=mv, adv_preds_in=pin)
    # ---- eight realistic persons per config
    d = {"versions": ver}
    for cid in sorted(CONFIGS):
        cfg = CONFIGS[cid]
        rng = np.random.default_rng(8500 + cid)
        out, fl = synth.heatmaps(cfg, 8, rng)
        q = 16.0 if cfg.pre_div else 1024.0
        out, fl = quant(out, q), quant(fl, q)
        c, s, _ = synth.boxes(cfg, 8, rng)
        p, mv, pin = RC.final_preds(cfg, RC.flip_average(cfg, out, fl), c, s)
        k = "cfg%d_" % cid
        d.update({k + "out": out, k + "flipped": fl, k + "center": c, k + "scale": s, k + "preds": p, k + "maxvals": mv,
                  k + "preds_in": pin})
    np.savez_compressed(os.path.join(HERE, "dec8.npz"), **d)
    for f in ("adv_cfg2.npz", "adv_cfg3.npz", "adv_cfg4.npz", "dec8.npz"):
        print(f, "%.2f MB" % (os.path.getsize(os.path.join(HERE, f)) / 1e6))


if __name__ == "__main__":
    main()
