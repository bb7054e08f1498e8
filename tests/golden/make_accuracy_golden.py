"""Golden vectors for accuracy()/PCK, produced by the UNMODIFIED reference function
deep_hrnet/lib/core/evaluate.py:accuracy in this container:

    python tests/golden/make_accuracy_golden.py          # writes tests/golden/accuracy.npz

Inputs are regenerated from seeds by the tests (`inputs(case)` below is imported by them), so the
fixture holds only the reference's outputs."""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {"cfg1": (40, 17, 64, 48, 1), "mpii": (24, 16, 64, 64, 2), "small": (7, 5, 16, 12, 3), "degenerate": (6, 17, 64, 48, 4)}


def inputs(case):
    """(output, target) float32 [N,K,h,w]: target = Gaussian blobs (about 15 % of the joints empty,
    as target_weight = 0 leaves them), output = blobs displaced by a few pixels plus noise."""
    N, K, h, w, seed = CASES[case]
    rng = np.random.default_rng(seed)
    mu = np.stack([rng.uniform(0, w - 1, (N, K)), rng.uniform(0, h - 1, (N, K))], -1)
    xs, ys = np.arange(w)[None, None, None, :], np.arange(h)[None, None, :, None]
    blob = lambda m: np.exp(-((xs - m[..., 0, None, None]) ** 2 + (ys - m[..., 1, None, None]) ** 2) / 8.0)
    target = blob(np.round(mu)) * (rng.uniform(size=(N, K)) > 0.15)[..., None, None]
    output = blob(mu + rng.normal(0, 2.5, mu.shape)) * rng.uniform(0.3, 1.0, (N, K))[..., None, None]
    output = output + rng.uniform(-0.05, 0.05, output.shape)
    if case == "degenerate":
        target[0] = 0.0                      # a person with no visible joint
        target[:, 3] = 0.0                   # a joint nobody has: acc[4] = -1
        output[1] = -np.abs(output[1])       # all-negative prediction: coords (0, 0)
        target[2, :, 1, 1] = 5.0             # peak at (1, 1): not > 1, ignored
    return output.astype(np.float32), target.astype(np.float32)


def reference_accuracy():
    root = os.environ.get("UDP_REFERENCE_ROOT", "/root/reference")
    if not hasattr(np, "float"):
        np.float = float
    sys.path.insert(0, os.path.join(root, "deep_hrnet"))
    sys.path.insert(0, os.path.join(root, "deep_hrnet", "lib"))
    spec = importlib.util.spec_from_file_location("_ref_evaluate", os.path.join(root, "deep_hrnet", "lib", "core", "evaluate.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.accuracy


if __name__ == "__main__":
    acc_fn = reference_accuracy()
    out = {}
    for case in CASES:
        o, t = inputs(case)
        acc, avg, cnt, pred = acc_fn(o, t)
        out[case + "_acc"], out[case + "_avg"], out[case + "_cnt"], out[case + "_pred"] = acc, np.float64(avg), np.int64(cnt), pred
        print(case, avg, cnt, acc[:5])
    np.savez_compressed(os.path.join(HERE, "accuracy.npz"), **out)
