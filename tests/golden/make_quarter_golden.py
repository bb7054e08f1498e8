"""Golden vectors for the demo engines' quarter-offset decoder (first half), produced by
the UNMODIFIED reference function deep_hrnet/tools/infer_utils/decode.py:get_final_preds (:19-40, with
its own get_max_preds :43-52 and the biased transform_preds :55-62) in this container:

    python tests/golden/make_quarter_golden.py          # writes tests/golden/quarter.npz

Inputs are regenerated from seeds by the tests (`inputs(case)` is imported by them); the fixture holds
only the reference's outputs (int64 [B,C,2])."""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# case -> (B, C, H, W, seed, boxes dtype)
CASES = {"coco": (24, 17, 64, 48, 1, "float32"), "mpii": (9, 16, 64, 64, 2, "float32"),
         "w48": (5, 17, 96, 72, 3, "float32"), "edge": (8, 17, 64, 48, 4, "float32"), "f64boxes": (6, 17, 64, 48, 5, "float64")}


def inputs(case):
    """heatmaps f32 [B,C,H,W]: noisy Gaussian blobs; boxes [B,4] = (cx, cy, sx, sy) with scale in the
    reference's /200 units.  'edge': peaks on rows/columns 0, 1, H-2, H-1 (no quarter shift there, :30),
    all-negative maps (coordinates -1, :51), flat neighbourhoods (sign 0) and exact ties."""
    B, C, H, W, seed, dt = CASES[case]
    rng = np.random.default_rng(seed)
    mu = np.stack([rng.uniform(0, W - 1, (B, C)), rng.uniform(0, H - 1, (B, C))], -1)
    xs, ys = np.arange(W)[None, None, None, :], np.arange(H)[None, None, :, None]
    hm = np.exp(-((xs - mu[..., 0, None, None]) ** 2 + (ys - mu[..., 1, None, None]) ** 2) / 8.0)
    hm = hm * rng.uniform(0.2, 1.0, (B, C, 1, 1)) + rng.uniform(-0.02, 0.02, hm.shape)
    if case == "edge":
        hm[0] = -np.abs(hm[0]) - 0.01                      # nothing positive
        for k, (y, x) in enumerate([(0, 5), (1, 5), (H - 2, 5), (H - 1, 5), (5, 0), (5, 1), (5, W - 2), (5, W - 1)]):
            hm[1, k] = 0.0
            hm[1, k, y, x] = 1.0
        hm[2] = 0.0
        hm[2, :, 10, 10] = 1.0                             # single pixel: both differences are 0
        hm[3] = 0.0
        hm[3, :, 20, 7] = hm[3, :, 30, 40] = 0.5           # tie: first index wins
        hm[3, :, 20, 8] = 0.25                             # and a one-sided neighbour
    c = rng.uniform(100, 500, (B, 2))
    s = rng.uniform(0.4, 3.0, (B, 1)) * np.array([[0.75, 1.0]])
    return hm.astype(np.float32), np.concatenate([c, s], 1).astype(dt)


def reference_fn():
    root = os.environ.get("UDP_REFERENCE_ROOT", "/root/reference")
    spec = importlib.util.spec_from_file_location("_ref_decode", os.path.join(root, "deep_hrnet", "tools", "infer_utils", "decode.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.get_final_preds


if __name__ == "__main__":
    import torch
    fn = reference_fn()
    out = {}
    for case in CASES:
        hm, boxes = inputs(case)
        out[case] = fn(torch.from_numpy(hm), torch.from_numpy(boxes))
        print(case, out[case].dtype, out[case].shape, out[case][0, :3].tolist())
    np.savez_compressed(os.path.join(HERE, "quarter.npz"), **out)
