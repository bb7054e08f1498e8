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


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.CASES))
def test_device_rescore_nms_matches_reference(case):
    from udp_pose_b200 import reference_api as R
    p, b, img = G.inputs(case)
    _check(case, *R.coco_rescore_nms(p, b, img, G.IN_VIS_THRE, G.OKS_THRE, soft_nms=G.CASES[case][4]))


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G.DIRECT))
def test_device_direct_nms_matches_reference(case):
    """The reference's own signatures: oks_nms / soft_oks_nms(kpts_db, thresh, sigmas, in_vis_thre)."""
    from udp_pose_b200 import reference_api as R
    soft, thresh, vis, _ = G.DIRECT[case]
    p, area, score, sig = G.direct_inputs(case)
    db = [{"keypoints": p[i], "area": area[i], "score": score[i]} for i in range(len(p))]
    keep = (R.soft_oks_nms if soft else R.oks_nms)(db, thresh, sig, vis)
    assert list(keep) == GOLD[case + "_keep"].tolist()
    assert R.oks_nms([], 0.9) == [] and list(R.soft_oks_nms([], 0.9)) == []


@pytest.mark.gpu
def test_device_oks_values_and_large_batch():
    """512 images x up to 64 persons (16 489 in all) against the oracle: kept rows exact; the OKS of every person
    against its image's top pick within 4 ulp of the numpy value (CUDA exp vs numpy exp)."""
    import torch
    from udp_pose_b200 import device_api as D, reference_api as R, synth
    preds, boxes, img = synth.nms_persons(512, 64, 17, np.random.default_rng(3))
    score, kept = R.coco_rescore_nms(preds, boxes, img, 0.2, 0.9)
    o_score, o_kept = O.coco_rescore_nms(preds, boxes, img, 0.2, 0.9)
    assert np.array_equal(score, o_score) and kept == o_kept
    s_score, s_kept = R.coco_rescore_nms(preds, boxes, img, 0.2, 0.9, soft_nms=True)
    assert s_kept == O.coco_rescore_nms(preds, boxes, img, 0.2, 0.9, soft=True)[1]
    # one image, OKS values
    rows = np.nonzero(img == img[0])[0]
    r = D.oks_nms_batch(preds[rows], boxes[rows, 4], o_score[rows], np.array([0, len(rows)], np.int32), len(rows), 0.9,
                        want_oks=True)
    top = int(r["keep"][0])
    rest = [j for j in range(len(rows)) if j != top]
    want = O.oks_row(preds[rows], boxes[rows, 4], top, rest)
    got = r["oks_top"].cpu().numpy()[rest]
    assert np.all(np.abs(got - want) <= 4 * np.spacing(want))
    assert np.isnan(r["oks_top"].cpu().numpy()[top])


@pytest.mark.gpu
def test_device_nms_rejects_oversized_image():
    from udp_pose_b200 import device_api as D, synth
    preds, boxes, img = synth.nms_persons(1, 8, 17, np.random.default_rng(4), exact=True)
    r = D.oks_nms_batch(preds, boxes[:, 4], boxes[:, 5], np.array([0, 8], np.int32), 4, 0.9)
    assert int(r["keep_cnt"][0]) == -1


def test_install_rebinds_nms_names():
    """install(nms=...) substitutes oks_nms / soft_oks_nms in the modules that hold the names
    (lib/nms/nms.py and lib/dataset/coco.py, which imports them at :20); no GPU needed."""
    import types
    from udp_pose_b200 import reference_api as R
    a, b = types.ModuleType("nms.nms"), types.ModuleType("dataset.coco")
    a.oks_nms = a.soft_oks_nms = b.oks_nms = b.soft_oks_nms = None
    saved = R.install(nms=[a, b])
    assert a.oks_nms is R.oks_nms and b.soft_oks_nms is R.soft_oks_nms
    assert ("nms.nms", "oks_nms") in saved and ("dataset.coco", "soft_oks_nms") in saved
