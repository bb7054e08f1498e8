"""Golden vectors for the AID augmentations, produced by the UNMODIFIED reference
classes deep_hrnet/lib/utils/transforms.py:Cutout / HideAndSeek in this container:

    python tests/golden/make_aid_golden.py          # writes tests/golden/aid.npz

For each case numpy's global generator is seeded, the reference object is called on a seeded image,
and the fixture keeps the erased-pixel mask (packed bits) plus the next rand() of the generator --
which pins how many random numbers the reference consumed."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
# case -> (kind, constructor args, (H, W), seed)
CASES = {
    "cutout_default": ("cutout", (), (256, 192), 1),
    "cutout_3": ("cutout", (1.0, 0.35, 3), (256, 192), 2),
    "cutout_gate": ("cutout", (0.3, 0.2, 1), (256, 192), 3),
    "hide_default": ("hide", (), (256, 192), 4),
    "hide_44": ("hide", (1.0, 0.5, (44, 44, 0)), (256, 192), 5),
    "hide_square": ("hide", (1.0, 0.3, (32, 56, 0)), (256, 256), 6),
    "hide_wide": ("hide", (1.0, 0.5, (16, 16, 0)), (128, 192), 7),
}
REPEATS = 6   # images per case (consecutive draws from one seeded stream)


def image(H, W, seed):
    return np.random.default_rng(seed).integers(1, 256, (H, W, 3)).astype(np.uint8)   # no zero pixels: erased <=> 0


def reference_objects():
    from oracle import ref_import
    tr = ref_import.hrnet().transforms
    return tr.Cutout, tr.HideAndSeek


if __name__ == "__main__":
    RC, RH = reference_objects()
    out = {}
    for case, (kind, args, (H, W), seed) in CASES.items():
        obj = (RC if kind == "cutout" else RH)(*args)
        np.random.seed(seed)
        masks = []
        for k in range(REPEATS):
            res = obj(image(H, W, 100 * seed + k).copy())
            masks.append(np.packbits((res == 0).all(2)))
        out[case + "_mask"] = np.stack(masks)
        out[case + "_next"] = np.float64(np.random.rand())
        print(case, [int(np.unpackbits(m).sum()) for m in masks])
    np.savez_compressed(os.path.join(HERE, "aid.npz"), **out)
