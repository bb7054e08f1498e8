"""the AID augmentations Cutout / HideAndSeek (deep_hrnet/lib/utils/transforms.py:144-223).
tests/golden/aid.npz holds what the reference's own classes erase under a seeded numpy generator, and the
generator's next number afterwards (tests/golden/make_aid_golden.py).  Our host classes must draw the same
parameters from the same stream (CPU test, with the oracle's erase) and the device erase must produce the
same pixels (GPU test).  uint8: exact."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_aid_golden as G  # noqa: E402

from oracle import oracle as O  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "aid.npz"))


def _run(case, apply):
    from udp_pose_b200 import reference_api as R
    kind, args, (H, W), seed = G.CASES[case]
    obj = (R.Cutout if kind == "cutout" else R.HideAndSeek)(*args)
    np.random.seed(seed)
    for k in range(G.REPEATS):
        img = G.image(H, W, 100 * seed + k)
        res = apply(kind, obj, img)
        assert np.array_equal(np.packbits((res == 0).all(2)), GOLD[case + "_mask"][k]), (case, k)
        assert np.array_equal(res[res != 0], img[res != 0])
    assert np.random.rand() == float(GOLD[case + "_next"])    # same number of draws as the reference


@pytest.mark.parametrize("case", sorted(G.CASES))
def test_host_draws_and_oracle_erase_match_reference(case):
    def apply(kind, obj, img):
        H, W, _ = img.shape
        if kind == "cutout":
            return O.aid_erase(img, ellipses=obj.draw(H, W))
        g, bits = obj.draw(H, W)
        return O.aid_erase(img, grid=g, hidden=bits)
    _run(case, apply)


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.CASES))
def test_device_erase_matches_reference(case):
    _run(case, lambda kind, obj, img: obj(img))


@pytest.mark.gpu
def test_device_erase_batch_both():
    """A batch with both augmentations, per-person parameters, against the oracle."""
    import torch
    from udp_pose_b200 import device_api as D, reference_api as R
    np.random.seed(11)
    cut, hide = R.Cutout(0.7, 0.25, 2), R.HideAndSeek(0.8, 0.4, (0, 16, 32, 44, 56))
    imgs = np.stack([G.image(256, 192, 500 + k) for k in range(24)])
    cuts = [cut.draw(256, 192) for _ in range(24)]
    hides = [hide.draw(256, 192) for _ in range(24)]
    t = torch.from_numpy(imgs.copy()).cuda()
    D.aid_erase(t, *R._aid_pack(cuts, hides, 24))
    got = t.cpu().numpy()
    for k in range(24):
        want = O.aid_erase(O.aid_erase(imgs[k], ellipses=cuts[k]), grid=hides[k][0], hidden=hides[k][1])
        assert np.array_equal(got[k], want), k


@pytest.mark.gpu
def test_device_hide_and_seek_on_the_384x288_crop_with_the_default_grids():
    """grid 16 on a 384x288 crop has 18 * 24 = 432 cells, more than the old 256-bit mask
    (one draw in four raised ValueError where the reference works).  The mask now holds 512 cells."""
    import torch
    from udp_pose_b200 import device_api as D, reference_api as R
    np.random.seed(21)
    hide = R.HideAndSeek(1.0, 0.5, (16, 16, 0))   # always grid 16 (the last size is never drawn)
    imgs = np.stack([G.image(384, 288, 700 + k) for k in range(6)])
    hides = [hide.draw(384, 288) for _ in range(6)]
    assert all(g == 16 and len(bits) == 432 for g, bits in hides)
    t = torch.from_numpy(imgs.copy()).cuda()
    D.aid_erase(t, *R._aid_pack(None, hides, 6))
    got = t.cpu().numpy()
    for k in range(6):
        assert np.array_equal(got[k], O.aid_erase(imgs[k], grid=16, hidden=hides[k][1])), k


def test_hide_and_seek_grid_that_cannot_fit_fails_at_construction():
    """... and a combination that cannot fit raises when the dataset is built, not at a random draw."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("reference_api needs the built library's package import only; constructed on the GPU box")
    from udp_pose_b200 import reference_api as R
    from udp_pose_b200.configs import CONFIGS
    c = CONFIGS[3].as_cfg()
    c.DATASET.CUTOUT, c.DATASET.HIDE_AND_SEEK = None, [1.0, 0.5, (8, 16, 0)]   # grid 8 on 384x288: 1728 cells
    with pytest.raises(ValueError):
        R.JointsDataset(c, "", "train", True)
    c.DATASET.HIDE_AND_SEEK = [1.0, 0.5, (0, 16, 32, 44, 56)]                  # the reference's defaults fit
    R.JointsDataset(c, "", "train", True)
