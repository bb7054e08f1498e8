"""a4 -- fliplr_joints (deep_hrnet/lib/utils/transforms.py:50-64) and RSN's flip_joints
(RSN/lib/utils/transforms.py:70-79), pinned by outputs of the reference's own two functions
(tests/golden/fliplr.npz, made by tests/golden/make_getitem_golden.py).

CPU: the oracle restatement against the golden.  GPU: udp_fliplr_joints through the C ABI, both
flavours (mul_vis = 1 / 0), batched and through the reference-named wrappers, bit-exact."""
import os

import numpy as np
import pytest

from parity import GOLDEN, same_bits
from udp_pose_b200.configs import CONFIGS

G = np.load(os.path.join(GOLDEN, "fliplr.npz"))
SETS = [("coco", 1), ("mpii", 4)]


@pytest.mark.parametrize("name,cid", SETS)
def test_oracle_fliplr_joints(name, cid):
    from oracle import oracle as O
    pairs, width = CONFIGS[cid].flip_pairs, int(G[name + "_width"])
    j, v = G[name + "_joints"], G[name + "_vis"]
    a, av = O.fliplr_joints(j, v, width, pairs, mul_vis=True)
    assert same_bits(a, G[name + "_fliplr_joints"]) and same_bits(av, G[name + "_fliplr_vis"])
    b, bv = O.fliplr_joints(j, v, width, pairs, mul_vis=False)
    assert same_bits(b, G[name + "_flip_joints"]) and same_bits(bv, G[name + "_flip_vis"])
    # the two flavours differ exactly by the visibility product (transforms.py:64)
    assert same_bits(a, b * bv)


def test_oracle_fliplr_is_an_involution_without_product():
    from oracle import oracle as O
    rng = np.random.default_rng(3)
    j = rng.integers(0, 300, (5, 17, 3)).astype(np.float64)  # integer coordinates: the two subtractions are exact
    v = rng.integers(0, 3, (5, 17, 3)).astype(np.float64)
    a, av = O.fliplr_joints(j, v, 301, CONFIGS[1].flip_pairs, mul_vis=False)
    b, bv = O.fliplr_joints(a, av, 301, CONFIGS[1].flip_pairs, mul_vis=False)
    assert np.array_equal(b, j) and np.array_equal(bv, v)


@pytest.mark.gpu
@pytest.mark.parametrize("name,cid", SETS)
def test_gpu_fliplr_joints_batched(name, cid):
    import torch
    from udp_pose_b200 import _lib, device_api as D
    _lib.load()
    pairs, width = CONFIGS[cid].flip_pairs, int(G[name + "_width"])
    j, v = G[name + "_joints"], G[name + "_vis"]
    a, av = D.fliplr_joints_batch(j, v, width, pairs, mul_vis=True)
    assert same_bits(a.cpu().numpy(), G[name + "_fliplr_joints"]) and same_bits(av.cpu().numpy(), G[name + "_fliplr_vis"])
    b, bv = D.fliplr_joints_batch(j, v, width, pairs, mul_vis=False)   # RSN flavour: no visibility product
    assert same_bits(b.cpu().numpy(), G[name + "_flip_joints"]) and same_bits(bv.cpu().numpy(), G[name + "_flip_vis"])
    # per-person widths (images of different sizes in one batch)
    wd = torch.tensor([width] * j.shape[0], dtype=torch.int32)
    c, _ = D.fliplr_joints_batch(j, v, wd, pairs, mul_vis=True)
    assert same_bits(c.cpu().numpy(), G[name + "_fliplr_joints"])


@pytest.mark.gpu
@pytest.mark.parametrize("name,cid", SETS)
def test_gpu_reference_named_wrappers(name, cid):
    from udp_pose_b200 import _lib, reference_api as R
    _lib.load()
    pairs, width = CONFIGS[cid].flip_pairs, int(G[name + "_width"])
    j, v = G[name + "_joints"], G[name + "_vis"]
    for i in range(j.shape[0]):
        a, av = R.fliplr_joints(j[i].copy(), v[i].copy(), width, pairs)
        assert same_bits(a, G[name + "_fliplr_joints"][i]) and same_bits(av, G[name + "_fliplr_vis"][i])
        b, bv = R.flip_joints(j[i].copy(), v[i].copy(), width, pairs)
        assert same_bits(b, G[name + "_flip_joints"][i]) and same_bits(bv, G[name + "_flip_vis"][i])


def test_large_oracle_vs_formula():
    """N = 1000: the batched oracle against the reference's per-person formula written out."""
    from oracle import oracle as O
    rng = np.random.default_rng(11)
    N, K = 1000, 17
    j, v = rng.uniform(-50, 700, (N, K, 3)), (rng.uniform(size=(N, K, 3)) < 0.7).astype(np.float64)
    wd = rng.integers(100, 700, N)
    a, av = O.fliplr_joints(j, v, wd, CONFIGS[1].flip_pairs, mul_vis=True)
    perm = np.arange(K)
    for p, q in CONFIGS[1].flip_pairs:
        perm[p], perm[q] = q, p
    jf = j.copy()
    jf[..., 0] = wd[:, None] - jf[..., 0] - 1
    assert np.array_equal(av, v[:, perm]) and np.array_equal(a, jf[:, perm] * v[:, perm])
