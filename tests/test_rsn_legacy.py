"""SURVEY 8(f)-4 (second half): RSN's legacy decoder get_results (RSN/exps/RSN18.coco/test.py:25-71).
tests/golden/rsn_legacy.npz holds the outputs of the reference's own function (compiled from the
reference file by tests/golden/make_rsn_legacy_golden.py; inputs regenerated from the same seeds).
Arg-max positions are integers, so coordinates and scores are compared bit for bit."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_rsn_legacy_golden as G  # noqa: E402

from oracle import oracle as O  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "rsn_legacy.npz"))


def _check(case, preds, maxvals):
    assert np.array_equal(np.asarray(preds, np.float64), GOLD[case + "_preds"]), case
    assert np.array_equal(np.asarray(maxvals, np.float64), GOLD[case + "_maxvals"]), case


@pytest.mark.parametrize("n", [5, 9, 11])
def test_blur_f64_restatement_matches_installed_cv2(n):
    """The float64 filter path (fused row chain, unfused symmetric column sum) against cv2 itself on this
    host, and the kernel tables against cv2.getGaussianKernel."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(n)
    a = np.zeros((84, 68))
    a[10:-10, 10:-10] = rng.uniform(-20, 255, (64, 48)).astype(np.float32)
    assert np.array_equal(O.gauss_kernel_f64(n), cv2.getGaussianKernel(n, 0, cv2.CV_64F).ravel())
    assert np.array_equal(O.gaussian_blur_f64(a, n), cv2.GaussianBlur(a, (n, n), 0))


@pytest.mark.parametrize("case", sorted(G.CASES))
def test_oracle_get_results_matches_reference(case):
    N, K, osh, ish, ksz, _ = G.CASES[case]
    out, c, s = G.inputs(case)
    _check(case, *O.rsn_get_results(out, c, s, ish, ksz))
