"""accuracy()/PCK (deep_hrnet/lib/core/evaluate.py:16-71).

tests/golden/accuracy.npz holds the outputs of the reference's own accuracy() on seeded inputs
(tests/golden/make_accuracy_golden.py; the inputs are regenerated here from the same seeds).
CPU: the oracle restatement against them.  GPU: the device kernels against them and against
the oracle at a larger batch, counts and coordinates exact."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_accuracy_golden as G  # noqa: E402

from oracle import oracle as O  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "accuracy.npz"))


def _check(case, acc, avg, cnt, pred):
    assert np.array_equal(np.asarray(acc, np.float64), GOLD[case + "_acc"]), case
    assert float(avg) == float(GOLD[case + "_avg"]) and int(cnt) == int(GOLD[case + "_cnt"])
    assert np.array_equal(np.asarray(pred, np.float32), GOLD[case + "_pred"])


@pytest.mark.parametrize("case", sorted(G.CASES))
def test_oracle_accuracy_matches_reference(case):
    o, t = G.inputs(case)
    _check(case, *O.accuracy_pck(o, t))


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.CASES))
def test_device_accuracy_matches_reference(case):
    from udp_pose_b200 import reference_api as R
    o, t = G.inputs(case)
    _check(case, *R.accuracy(o, t))


@pytest.mark.gpu
def test_device_accuracy_large_and_offset_views():
    """N = 2048 persons against the oracle (counts exact), CUDA tensors in, and the offset mode's
    [:, ::3] views (function.py:190-191) through ch_stride = 3 without materialising them."""
    import torch
    from udp_pose_b200 import device_api as D, reference_api as R, synth
    from udp_pose_b200.configs import CONFIGS
    rng = np.random.default_rng(9)
    cfg = CONFIGS[1]
    out, _ = synth.heatmaps(cfg, 2048, rng, flip=False)
    tgt, _ = synth.heatmaps(cfg, 2048, np.random.default_rng(10), flip=False)
    tgt[::7] = 0.0
    a = O.accuracy_pck(out, tgt)
    acc, avg, cnt, pred = R.accuracy(torch.from_numpy(out).cuda(), torch.from_numpy(tgt).cuda())
    assert np.array_equal(acc, a[0]) and avg == a[1] and cnt == a[2] and np.array_equal(pred, a[3])
    cfg2 = CONFIGS[2]
    o3, _ = synth.heatmaps(cfg2, 64, rng, flip=False)
    t3, _ = synth.heatmaps(cfg2, 64, np.random.default_rng(11), flip=False)
    b = O.accuracy_pck(o3[:, ::3], t3[:, ::3])
    r = D.accuracy_pck(torch.from_numpy(o3).cuda(), torch.from_numpy(t3).cuda(), ch_stride=3)
    assert np.array_equal(r["acc"].cpu().numpy(), b[0]) and int(r["cnt"][0]) == b[2]
    assert np.array_equal(r["pred"].cpu().numpy(), b[3])


@pytest.mark.gpu
def test_device_accuracy_nan_and_ties():
    """np.argmax rules: a NaN wins and masks the coordinates to 0; equal maxima -> lowest index."""
    import torch
    from udp_pose_b200 import reference_api as R
    o, t = G.inputs("small")
    o[0, 0, 3, 4] = np.nan
    o[1, 1] = 0.25
    t[2, 2] = 0.0
    t[2, 2, 5, 6] = t[2, 2, 9, 2] = 1.0
    with np.errstate(all="ignore"):
        a = O.accuracy_pck(o, t)
    acc, avg, cnt, pred = R.accuracy(o, t)
    assert np.array_equal(acc, a[0]) and cnt == a[2] and np.array_equal(pred, a[3])
    assert pred[0, 0].tolist() == [0.0, 0.0] and pred[1, 1].tolist() == [0.0, 0.0]
