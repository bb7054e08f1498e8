"""a8 / f5 -- the TRAIN-time branches of JointsDataset.__getitem__ (deep_hrnet/lib/dataset/JointsDataset.py:
204-235: half-body box, scale / rotation draws, flip + fliplr_joints + centre mirror) and
half_body_transform (:124-167), pinned by outputs of the reference's own __getitem__ run in train mode
under seeded generators (tests/golden/getitem.npz, made by tests/golden/make_getitem_golden.py).

udp_pose_b200.reference_api.JointsDataset.__getitem__ is run under the same seeds on the same record
and must return: the same number of generator draws, bit-identical scale / centre / rotation and
joints_vis, the u8 crop bit-exact (where the device warp matrix equals the reference's -- always at
rotation 0; a rotated matrix may differ by one float32 ulp in device sincos, in which case at most a few
pixels may differ by one grey level), the target within 1e-5, target_weight exact, meta joints within 1e-4."""
import os
import random

import numpy as np
import pytest

from parity import GOLDEN, same_bits, TOL_HM, TOL_PX
from udp_pose_b200.configs import CONFIGS

G = np.load(os.path.join(GOLDEN, "getitem.npz"))
CASES = [(int(c), float(p), int(s)) for c, p, s in G["cases"]]


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available()
    from udp_pose_b200 import _lib, reference_api
    _lib.load()
    return reference_api


def _dataset(R, cfg, phb):
    c = cfg.as_cfg()
    c.DATASET.PROB_HALF_BODY = phb
    ds = R.JointsDataset(c, "", "train", True)
    ds.num_joints, ds.flip_pairs = cfg.num_joints, cfg.flip_pairs
    # what the subclasses set (lib/dataset/coco.py:65-66,94, mpii.py:33)
    ds.upper_body_ids = (0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10) if cfg.num_joints == 17 else (7, 8, 9, 10, 11, 12, 13, 14, 15)
    ds.aspect_ratio = cfg.image_size[0] * 1.0 / cfg.image_size[1]
    return ds


@pytest.mark.gpu
@pytest.mark.parametrize("n", range(len(CASES)))
def test_getitem_train_matches_reference(R, n):
    cid, phb, seed = CASES[n]
    cfg = CONFIGS[cid]
    p = "c%d_" % n
    ds = _dataset(R, cfg, phb)
    ds.db = [dict(image="mem://%d" % n, joints_3d=G[p + "rec_joints_3d"].copy(), joints_3d_vis=G[p + "rec_joints_3d_vis"].copy(),
                  center=G[p + "rec_center"].copy(), scale=G[p + "rec_scale"].copy(), score=0.75)]
    np.random.seed(seed)
    random.seed(seed)
    x, t, tw, meta = ds.__getitem__(0, image=G["images"][n % 2])
    nxt = np.array([np.random.rand(), random.random()])
    assert np.array_equal(nxt, G[p + "next"])                       # same number of draws from both generators
    assert meta["rotation"] == float(G[p + "meta_rotation"])
    assert same_bits(np.asarray(meta["scale"]), G[p + "meta_scale"])  # half-body box and the scale draw, dtype included
    assert same_bits(np.asarray(meta["center"]), G[p + "meta_center"])  # incl. the mirrored centre of the flip branch
    assert np.array_equal(meta["joints_vis"], G[p + "meta_joints_vis"])  # pair swap of the flip branch
    assert np.abs(meta["joints"] - G[p + "meta_joints"]).max() <= TOL_PX
    assert meta["score"] == 0.75
    # target: weights exact, values within the heat-map tolerance
    assert same_bits(tw.numpy(), G[p + "target_weight"])
    assert t.shape == G[p + "target"].shape and np.abs(t.numpy() - G[p + "target"]).max() <= TOL_HM
    # the crop
    M = R.get_warpmatrix(meta["rotation"], np.asarray(meta["center"]) * 2.0, ds.image_size - 1.0, meta["scale"])
    want = G[p + "input"]
    assert x.dtype == np.uint8 and x.shape == want.shape
    if same_bits(M, G[p + "trans"]):
        assert np.array_equal(x, want)
    else:  # one-ulp matrix (device sincos vs glibc): a handful of pixels may move by one grey level
        np.testing.assert_allclose(M, G[p + "trans"], rtol=1.2e-7, atol=1e-30)
        d = np.abs(x.astype(np.int16) - want)
        assert d.max() <= 1 and (d != 0).mean() < 1e-3


def test_cases_cover_every_branch():
    """The fixture must exercise: half-body taken / not, flip taken / not, rotation drawn / not."""
    half = flip = rot = nohalf = noflip = norot = 0
    for n, (cid, phb, seed) in enumerate(CASES):
        p = "c%d_" % n
        # the half-body box has its own centre: without it the row coordinate of the centre is the record's
        took_half = float(G[p + "rec_center"][1]) != float(G[p + "meta_center"][1])
        assert took_half == (phb == 1.0)
        random.seed(seed)
        r = random.random() <= 0.6
        f = random.random() <= 0.5
        half += took_half; nohalf += not took_half
        flip += f; noflip += not f
        rot += r; norot += not r
        assert (float(G[p + "meta_rotation"]) != 0) == r
    assert min(half, nohalf, flip, noflip, rot, norot) >= 2
