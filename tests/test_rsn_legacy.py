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


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.CASES))
def test_device_get_results_matches_reference(case):
    from udp_pose_b200 import reference_api as R
    N, K, osh, ish, ksz, _ = G.CASES[case]
    out, c, s = G.inputs(case)
    s0 = s.copy()
    _check(case, *R.rsn_legacy_get_results(out, c, s, ksz, [0.25], input_shape=ish))
    assert np.array_equal(s, s0)


@pytest.mark.gpu
def test_device_get_results_flip_fused_and_large():
    """1024 persons with the flip test fused, against the oracle run on the host-side average
    (flip_back of RSN/lib/utils/transforms.py:59-67, then (a + b) * 0.5)."""
    import torch
    from udp_pose_b200 import device_api as D, reference_api as R, synth
    from udp_pose_b200.configs import CONFIGS
    cfg = CONFIGS[5]
    rng = np.random.default_rng(8)
    out, fl = synth.heatmaps(cfg, 1024, rng)
    c, s, _ = synth.boxes(cfg, 1024, rng)
    c, s = c.astype(np.float32), s.astype(np.float32)
    avg = O.flip_average(out, fl, cfg.flip_pairs)
    want = O.rsn_get_results(avg[:96], c[:96], s[:96], (256, 192), 5)
    got = R.rsn_legacy_get_results(torch.from_numpy(out).cuda(), c, s, 5, [0.25], outputs_flipped=torch.from_numpy(fl).cuda(),
                                   flip_pairs=cfg.flip_pairs)
    assert np.array_equal(got[0][:96], want[0]) and np.array_equal(got[1][:96], want[1])
    # the fused flip equals decoding the device-side average
    d_avg = D.flip_back(torch.from_numpy(fl).cuda(), cfg.flip_pairs, False)
    both = (torch.from_numpy(out).cuda() + d_avg) * 0.5
    again = R.rsn_legacy_get_results(both, c, s, 5, [0.25])
    assert np.array_equal(got[0], again[0]) and np.array_equal(got[1], again[1])


def test_product_tap_tables_match_cv2():
    cv2 = pytest.importorskip("cv2")
    from udp_pose_b200 import device_api as D
    for n in (5, 7, 9, 11, 13, 15):
        assert np.array_equal(D.cv_gauss_taps_f64(n), cv2.getGaussianKernel(n, 0, cv2.CV_64F).ravel())


@pytest.mark.gpu
def test_device_get_results_unaligned_input_takes_scalar_loads():
    """A network output that does not start on a 16-byte boundary goes through the scalar staging loop
    (the float4 loop needs aligned maps); same results."""
    import torch
    from udp_pose_b200 import device_api as D
    N, K, osh, ish, ksz, _ = G.CASES["coco_k5"]
    out, c, s = G.inputs("coco_k5")
    buf = torch.empty(out.size + 1, dtype=torch.float32, device="cuda")
    view = buf[1:].view(*out.shape)
    view.copy_(torch.from_numpy(out))
    assert view.data_ptr() % 16 != 0
    preds, maxvals = D.rsn_get_results(view, c, s, ish, ksz)
    _check("coco_k5", preds.cpu().numpy(), maxvals.cpu().numpy())
