"""SURVEY 8(f)-4 (first half): the quarter-offset decoder of the demo engines
(deep_hrnet/tools/infer_utils/decode.py:19-62).  tests/golden/quarter.npz holds the outputs of the
reference's own get_final_preds (tests/golden/make_quarter_golden.py; inputs regenerated from the same
seeds).  Integer pixel coordinates: exact."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_quarter_golden as G  # noqa: E402

from oracle import oracle as O  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "quarter.npz"))


@pytest.mark.parametrize("case", sorted(G.CASES))
def test_oracle_quarter_matches_reference(case):
    hm, boxes = G.inputs(case)
    assert np.array_equal(O.decode_quarter(hm, boxes), GOLD[case])


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.CASES))
def test_device_quarter_matches_reference(case):
    from udp_pose_b200 import reference_api as R
    hm, boxes = G.inputs(case)
    got = R.demo_get_final_preds(hm, boxes)
    assert got.dtype == np.int64 and np.array_equal(got, GOLD[case])


@pytest.mark.gpu
def test_device_quarter_large_and_nan():
    """2048 persons of cfg-1 maps against the oracle, CUDA tensors in; a NaN wins the arg-max
    (np.argmax) and `maxval > 0` is then false: coordinates -1."""
    import torch
    from udp_pose_b200 import reference_api as R, synth
    from udp_pose_b200.configs import CONFIGS
    rng = np.random.default_rng(5)
    hm, _ = synth.heatmaps(CONFIGS[1], 2048, rng, flip=False)
    c, s, _ = synth.boxes(CONFIGS[1], 2048, rng)
    boxes = np.concatenate([c, s], 1).astype(np.float32)
    hm[3, 2, 10, 11] = np.nan
    with np.errstate(all="ignore"):
        want = O.decode_quarter(hm, boxes)
    got = R.demo_get_final_preds(torch.from_numpy(hm).cuda(), torch.from_numpy(boxes))
    assert np.array_equal(got, want)
