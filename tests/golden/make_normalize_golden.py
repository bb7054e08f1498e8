"""Golden vectors for the fused crop + ToTensor + Normalize, generated HERE with
torchvision's own transforms (the reference's tools/train.py:127-143 composes exactly these):

    python tests/golden/make_normalize_golden.py        # writes tests/golden/normalize.npz

lut[c, v] = Normalize(ToTensor(pixel value v in channel c)) for all 256 values; `crop`/`out` = a
random uint8 HWC crop and its transform; `mean2/std2/out2` = the same with non-ImageNet constants."""
import os

import numpy as np
import torch
from torchvision import transforms

MEAN, STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]


def run(img, mean, std):
    t = transforms.Compose([transforms.ToTensor(), transforms.Normalize(mean=mean, std=std)])
    return t(img).numpy()


if __name__ == "__main__":
    ramp = np.arange(256, dtype=np.uint8).reshape(256, 1, 1).repeat(3, 2)
    lut = run(ramp, MEAN, STD)[:, :, 0]
    rng = np.random.default_rng(5)
    crop = rng.integers(0, 256, (64, 96, 3), dtype=np.uint8)
    mean2, std2 = [0.5, 0.25, 0.4321], [0.5, 0.3, 0.07]
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "normalize.npz"),
                        mean=np.float32(MEAN), std=np.float32(STD), lut=lut, crop=crop, out=run(crop, MEAN, STD),
                        mean2=np.float32(mean2), std2=np.float32(std2), out2=run(crop, mean2, std2),
                        lut2=run(ramp, mean2, std2)[:, :, 0],
                        versions=np.array([torch.__version__, __import__("torchvision").__version__]))
