"""Golden vectors for COCO OKS rescoring + oks_nms / soft_oks_nms, produced by the
UNMODIFIED reference in this container:

    python tests/golden/make_nms_golden.py          # writes tests/golden/nms.npz

The rescoring loop is inline in `COCODataset.evaluate` (deep_hrnet/lib/dataset/coco.py:289-356), so
the script calls that method itself on a stand-in `self` that only captures what it hands to
`_write_coco_keypoint_results`; `oks_nms` / `soft_oks_nms` (lib/nms/nms.py:97-180) run inside it.
Harness-side stubs, nothing in the reference is edited: `pycocotools`, `json_tricks` (not installed)
and the two Cython box-NMS extensions `nms.cpu_nms` / `nms.gpu_nms` (never built; imported at
nms.py:13-14 but not called on this path).

Inputs are regenerated from seeds by the tests (`inputs(case)` is imported by them); the fixture
holds only the reference's outputs: the rescored score of every person, and per image the kept
persons in the reference's order."""
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# case -> (images, max persons per image, K, seed, soft)
CASES = {
    "coco_hard": (40, 24, 17, 1, False),
    "coco_soft": (40, 48, 17, 2, True),
    "crowd_hard": (3, 300, 17, 3, False),
    "single": (5, 1, 17, 4, False),
    "lowvis_soft": (12, 30, 17, 5, True),
}
IN_VIS_THRE, OKS_THRE = 0.2, 0.9


def inputs(case):
    """preds f32 [N,K,3] (x, y, maxval), all_boxes f64 [N,6] (cx, cy, sx, sy, area, box score),
    image id per person int64 [N] (persons of one image are NOT contiguous: the rows are shuffled,
    as a detector's box list is), all in the layout validate() builds (function.py:208-214).
    Every image holds a few base poses, each with near-duplicates a few pixels away (what a box
    detector with overlapping boxes yields), so both suppression and survival happen."""
    I, nmax, K, seed, _ = CASES[case]
    rng = np.random.default_rng(seed)
    preds, boxes, img = [], [], []
    for i in range(I):
        n = int(rng.integers(1, nmax + 1)) if nmax > 1 else 1
        nbase = max(1, n // 4)
        base = rng.uniform(40, 600, (nbase, 1, 2)) + rng.normal(0, 40, (nbase, K, 2))
        size = rng.uniform(60, 260, nbase)
        for p in range(n):
            b = int(rng.integers(nbase))
            jitter = rng.choice([0.5, 2.0, 6.0, 25.0]) * size[b] / 100.0
            xy = base[b] + rng.normal(0, jitter, (K, 2))
            mv = rng.uniform(0.05, 1.0, (K, 1)) * (rng.uniform(size=(K, 1)) > (0.5 if case.startswith("lowvis") else 0.1))
            preds.append(np.concatenate([xy, mv], 1))
            s = np.array([size[b] * 0.75, size[b]]) / 200.0 * rng.uniform(0.9, 1.1)
            boxes.append([xy[:, 0].mean(), xy[:, 1].mean(), s[0], s[1], np.prod(s * 200.0), rng.uniform(0.05, 1.0)])
            img.append(1000 + 7 * i)
    perm = rng.permutation(len(img))
    preds = np.asarray(preds, np.float32)[perm]
    boxes = np.asarray(boxes, np.float64)[perm]
    img = np.asarray(img, np.int64)[perm]
    if case == "lowvis_soft":
        # one person per image with no joint above in_vis_thre: score 0 (one only -- equal scores have
        # no defined order under numpy's default argsort, so ties are kept out of the fixtures)
        preds[np.unique(img, return_index=True)[1], :, 2] = 0.1
    return preds, boxes, img


def reference_evaluate(preds, all_boxes, img, soft, oks_thre=OKS_THRE):
    """Runs coco.py:evaluate of the reference; returns (score f64 [N], [kept person rows per image]).
    With oks_thre > 1 and soft=False nothing is suppressed, which exposes every rescored score."""
    root = os.environ.get("UDP_REFERENCE_ROOT", "/root/reference")
    lib = os.path.join(root, "deep_hrnet", "lib")
    if not hasattr(np, "float"):
        np.float = float
    for name in ("pycocotools", "pycocotools.coco", "pycocotools.cocoeval", "json_tricks", "nms.cpu_nms", "nms.gpu_nms"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.COCO = m.COCOeval = m.cpu_nms = m.gpu_nms = None
            sys.modules[name] = m
    if lib not in sys.path:
        sys.path.insert(0, lib)
    from dataset.coco import COCODataset  # noqa: E402  (the unmodified class)
    captured = {}
    me = types.SimpleNamespace(num_joints=preds.shape[1], in_vis_thre=IN_VIS_THRE, oks_thre=oks_thre,
                               soft_nms=soft, image_set="test",
                               _write_coco_keypoint_results=lambda kpts, f: captured.setdefault("k", kpts))
    cfg = types.SimpleNamespace(RANK=0)
    # person row is recoverable from the record: the centre x is replaced by the row number
    boxes = all_boxes.copy()
    boxes[:, 0] = np.arange(len(boxes))
    paths = ["x/%012d.jpg" % i for i in img]
    with tempfile.TemporaryDirectory() as d:
        COCODataset.evaluate(me, cfg, preds.copy(), d, boxes, paths)
    score = np.full(len(preds), np.nan)
    kept = []
    for per_image in captured["k"]:
        rows = [int(r["center"][0]) for r in per_image]
        for r in per_image:
            score[int(r["center"][0])] = r["score"]
        kept.append(rows)
    return score, kept


DIRECT = {"direct_hard_vis": (False, 0.7, 0.3, 6), "direct_soft_vis": (True, 0.7, 0.3, 7), "direct_hard_sig": (False, 0.5, None, 8)}


def direct_inputs(case):
    """One image's kpts_db for a direct oks_nms / soft_oks_nms call with the optional arguments the
    COCO evaluator never passes (sigmas, in_vis_thre): (preds, area, score, sigmas or None)."""
    soft, thresh, vis, seed = DIRECT[case]
    rng = np.random.default_rng(seed)
    n, K = 60, 17
    base = rng.uniform(100, 400, (6, 1, 2)) + rng.normal(0, 30, (6, K, 2))
    b = rng.integers(6, size=n)
    xy = base[b] + rng.normal(0, 1, (n, K, 2)) * rng.choice([0.3, 2.0, 10.0], (n, 1, 1))
    mv = rng.uniform(0.0, 1.0, (n, K, 1))
    mv[3] = 0.0                                   # nobody home: e[ind] is empty -> iou 0
    preds = np.concatenate([xy, mv], 2).astype(np.float32)
    area = rng.uniform(3000, 20000, n)
    score = rng.uniform(0.01, 1.0, n)
    sigmas = rng.uniform(0.02, 0.11, K) if case.endswith("sig") else None
    return preds, area, score, sigmas


def reference_direct(case):
    reference_evaluate(*inputs("single"), False)      # imports the reference with the stubs in place
    from nms.nms import oks_nms, soft_oks_nms
    soft, thresh, vis, _ = DIRECT[case]
    preds, area, score, sigmas = direct_inputs(case)
    db = [{"keypoints": preds[i], "area": area[i], "score": score[i]} for i in range(len(preds))]
    return np.asarray((soft_oks_nms if soft else oks_nms)(db, thresh, sigmas, vis), np.int64)


if __name__ == "__main__":
    out = {}
    for case in DIRECT:
        out[case + "_keep"] = reference_direct(case)
        print(case, "kept", len(out[case + "_keep"]), "of 60")
    for case, (_, _, _, _, soft) in CASES.items():
        p, b, img = inputs(case)
        score, kept = reference_evaluate(p, b, img, soft)
        out[case + "_kept"] = np.asarray([r for k in kept for r in k], np.int64)
        out[case + "_kept_cnt"] = np.asarray([len(k) for k in kept], np.int64)
        score, all_rows = reference_evaluate(p, b, img, False, oks_thre=2.0)   # nothing suppressed: every score
        assert not np.isnan(score).any() and sum(len(k) for k in all_rows) == len(p)
        out[case + "_score"] = score
        print(case, "persons", len(p), "images", len(kept), "kept", int(out[case + "_kept_cnt"].sum()))
    np.savez_compressed(os.path.join(HERE, "nms.npz"), **out)
