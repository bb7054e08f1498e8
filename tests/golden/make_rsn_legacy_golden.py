"""Golden vectors for RSN's legacy decoder `get_results` (second half), produced by the
UNMODIFIED reference function RSN/exps/RSN18.coco/test.py:25-71 in this container:

    python tests/golden/make_rsn_legacy_golden.py          # writes tests/golden/rsn_legacy.npz

test.py cannot be imported as a module (its top level builds the network and reads the experiment
config), so the script compiles the `get_results` FunctionDef node straight from the reference file
(no source is copied) and runs it with `np`, `cv2` and a stand-in `cfg` holding the three fields it
reads (DATASET.KEYPOINT.NUM, OUTPUT_SHAPE, INPUT_SHAPE; RSN/dataset/attribute.py:12,19-20,57-63).

Inputs are regenerated from seeds by the tests (`inputs(case)` is imported by them); the fixture holds
only the reference's outputs (preds f64 [N,K,2], maxvals f64 [N,K,1])."""
import ast
import os
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# case -> (N, K, (h, w), (H, W), gaussian kernel, seed).  attribute.py:49 sets TEST.GAUSSIAN_KERNEL = 5 for
# COCO and :92 sets 9 for MPII; 11 is the function's own default.
CASES = {"coco_k5": (12, 17, (64, 48), (256, 192), 5, 1), "coco_k11": (12, 17, (64, 48), (256, 192), 11, 2),
         "mpii_k9": (6, 16, (64, 64), (256, 256), 9, 3), "edge_k5": (6, 17, (64, 48), (256, 192), 5, 4),
         "edge_k11": (6, 17, (64, 48), (256, 192), 11, 5)}


def inputs(case):
    """outputs f32 [N,K,h,w] in the RSN network's units (the labels are 255 * Gaussian blobs,
    RSN/dataset/JointsDataset.py:276-291) plus noise, centers/scales f32 [N,2] (scales in pixels/200 as the loader
    yields them).  'edge': all-negative maps (the arg-max falls in the zero padding), a single pixel
    (four-way tie for the second maximum), peaks on the border and in the corner, flat map."""
    N, K, (h, w), _, _, seed = CASES[case]
    rng = np.random.default_rng(seed)
    mu = np.stack([rng.uniform(-1, w, (N, K)), rng.uniform(-1, h, (N, K))], -1)
    xs, ys = np.arange(w)[None, None, None, :], np.arange(h)[None, None, :, None]
    hm = np.exp(-((xs - mu[..., 0, None, None]) ** 2 + (ys - mu[..., 1, None, None]) ** 2) / 8.0)
    out = 255.0 * hm * rng.uniform(0.2, 1.0, (N, K, 1, 1)) + rng.uniform(-3, 3, hm.shape)
    if case.startswith("edge"):
        out[0] = -np.abs(out[0]) - 1.0
        out[1] = 0.0
        out[1, :, 20, 30] = 100.0
        for k, (y, x) in enumerate([(0, 0), (0, w - 1), (h - 1, 0), (h - 1, w - 1), (0, 7), (9, 0)]):
            out[2, k] = -50.0
            out[2, k, y, x] = 200.0
        out[3] = 7.0
    c = rng.uniform(100, 500, (N, 2))
    s = rng.uniform(0.4, 3.0, (N, 1)) * np.array([[0.75, 1.0]])
    return out.astype(np.float32), c.astype(np.float32), s.astype(np.float32)


def reference_fn(K, out_shape, in_shape):
    import cv2
    root = os.environ.get("UDP_REFERENCE_ROOT", "/root/reference")
    path = os.path.join(root, "RSN", "exps", "RSN18.coco", "test.py")
    tree = ast.parse(open(path).read(), path)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "get_results"]
    S = types.SimpleNamespace
    ns = {"np": np, "cv2": cv2,
          "cfg": S(DATASET=S(KEYPOINT=S(NUM=K)), OUTPUT_SHAPE=tuple(out_shape), INPUT_SHAPE=tuple(in_shape))}
    exec(compile(ast.Module(body=fn, type_ignores=[]), path, "exec"), ns)
    return ns["get_results"]


if __name__ == "__main__":
    res = {}
    for case, (N, K, osh, ish, ksz, _) in CASES.items():
        out, c, s = inputs(case)
        preds, maxvals = reference_fn(K, osh, ish)(out.copy(), c.copy(), s.copy(), ksz, [0.25])
        res[case + "_preds"], res[case + "_maxvals"] = preds, maxvals
        print(case, preds.dtype, preds[0, :2].tolist(), maxvals[0, :2, 0].tolist())
    np.savez_compressed(os.path.join(HERE, "rsn_legacy.npz"), **res)
