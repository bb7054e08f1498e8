"""COCO OKS rescoring + oks_nms / soft_oks_nms (deep_hrnet/lib/dataset/coco.py:321-356,
lib/nms/nms.py:75-180).

tests/golden/nms.npz holds the outputs of the reference's own COCODataset.evaluate / oks_nms /
soft_oks_nms on seeded inputs (tests/golden/make_nms_golden.py; inputs regenerated here from the
same seeds).  CPU: the oracle restatement against them.  GPU: the device kernels against them
(kept rows and their order exact, scores bit-exact) and against the oracle at a larger size."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_nms_golden as G  # noqa: E402

from oracle import oracle as O  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "nms.npz"))


def _check(case, score, kept):
    assert np.array_equal(np.asarray(score, np.float64), GOLD[case + "_score"]), case
    assert [len(k) for k in kept] == GOLD[case + "_kept_cnt"].tolist(), case
    assert [r for k in kept for r in k] == GOLD[case + "_kept"].tolist(), case


@pytest.mark.parametrize("case", sorted(G.CASES))
def test_oracle_rescore_nms_matches_reference(case):
    p, b, img = G.inputs(case)
    _check(case, *O.coco_rescore_nms(p, b, img, G.IN_VIS_THRE, G.OKS_THRE, soft=G.CASES[case][4]))


@pytest.mark.parametrize("case", sorted(G.DIRECT))
def test_oracle_direct_nms_matches_reference(case):
    soft, thresh, vis, _ = G.DIRECT[case]
    p, area, score, sig = G.direct_inputs(case)
    keep = (O.soft_oks_nms if soft else O.oks_nms)(p, area, score, thresh, sig, vis)
    assert keep == GOLD[case + "_keep"].tolist()
