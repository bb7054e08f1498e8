"""Adversarial and larger decode fixtures produced by the UNMODIFIED reference (tests/golden/make_adversarial_golden.py):
adv_cfg2 (offset voting: ties after the 15x15 blur, all-zero / all-negative maps, border and corner peaks, plateaus),
adv_cfg3 / adv_cfg4 (the cfg-1 adversarial set at 96x72 and at the MPII float64-centre shape), dec8 (eight persons per
config).  CPU: the oracle.  GPU: the CUDA path through the C ABI.  Offset mode is float32 end to end: bit-exact."""
import os

import numpy as np
import pytest

import parity
from parity import same_bits, TOL_PX
from udp_pose_b200.configs import CONFIGS


def _g(name):
    return np.load(os.path.join(parity.GOLDEN, name))


def check_adv_offset(impl):
    cfg, g = CONFIGS[2], _g("adv_cfg2.npz")
    for fl, sfx in ((None, "_noflip"), (g["adv_flipped"], "")):
        r = impl.decode(cfg, g["adv_out"], fl, g["adv_center"], g["adv_scale"])
        assert same_bits(r["maxvals"], g["maxvals" + sfx])                      # maxima of the blurred maps
        assert same_bits(r["preds"].astype(np.float32), g["preds" + sfx])      # arg-max + vote + inverse affine
        assert same_bits(r["preds_in"].astype(np.float32), g["preds_in" + sfx])
    # row 3: all-zero classification maps -> maxval 0, mask false, coordinates = the vote at pixel (0, 0) only
    assert (g["maxvals_noflip"][3] == 0).all()
    # row 4: two bit-identical maxima after the blur -> the first one (lower flat index) wins
    r = impl.decode(cfg, g["adv_out"], None, g["adv_center"], g["adv_scale"])
    if "idx" in r:
        w, h = cfg.heatmap_size
        cls = g["adv_out"][4, 0::3]
        for k in range(cfg.num_joints):
            ys, xs = np.nonzero(cls[k])
            assert int(r["idx"][4, k]) == int(ys[0]) * w + int(xs[0])


def check_adv_gaussian(impl, cid):
    cfg, g = CONFIGS[cid], _g("adv_cfg%d.npz" % cid)
    with np.errstate(all="ignore"):
        r = impl.decode(cfg, g["adv_hm"], None, g["adv_center"], g["adv_scale"])
    assert same_bits(r["maxvals"], g["adv_maxvals"])
    assert np.array_equal(np.isnan(r["preds"]), np.isnan(g["adv_preds"]))
    for key in ("preds", "preds_in"):
        err = np.nanmax(np.abs(r[key] - g["adv_" + key]), axis=(1, 2))
        # rows: 0 U(0,1) noise, 1 N(0,1) noise (ill-conditioned Taylor steps amplify the reference's own float32 log
        # noise), 2 all-negative, 3 ties, 4 border peaks, 5 single pixel
        assert err[2] == 0.0 and err[3] <= 1e-3 and err[4] <= TOL_PX and err[5] <= TOL_PX, err
        assert err[0] <= 5e-2 and err[1] <= 5e-1, err


def check_dec8(impl, cid):
    cfg, g, k = CONFIGS[cid], _g("dec8.npz"), "cfg%d_" % cid
    r = impl.decode(cfg, g[k + "out"], g[k + "flipped"], g[k + "center"], g[k + "scale"])
    assert same_bits(r["maxvals"], g[k + "maxvals"])
    assert np.abs(r["preds"] - g[k + "preds"]).max() <= TOL_PX and np.abs(r["preds_in"] - g[k + "preds_in"]).max() <= TOL_PX
    if cfg.target_type == "offset":
        assert same_bits(r["preds"].astype(np.float32), g[k + "preds"])


@pytest.fixture(scope="module")
def orc():
    return parity.OracleImpl()


@pytest.fixture(scope="module")
def cuda():
    from test_gpu_parity import CudaImpl
    return CudaImpl()


def test_oracle_adv_offset(orc):
    check_adv_offset(orc)


@pytest.mark.parametrize("cid", [3, 4])
def test_oracle_adv_gaussian(orc, cid):
    check_adv_gaussian(orc, cid)


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_oracle_dec8(orc, cid):
    check_dec8(orc, cid)


@pytest.mark.gpu
def test_gpu_adv_offset(cuda):
    check_adv_offset(cuda)


@pytest.mark.gpu
def test_gpu_adv_offset_generic_and_prune_paths(cuda, monkeypatch):
    """The same fixture through the shape-generic kernel and through the opt-in branch-and-bound kernel."""
    monkeypatch.setenv("UDP_FORCE_GENERIC", "1")
    check_adv_offset(cuda)
    monkeypatch.delenv("UDP_FORCE_GENERIC")
    monkeypatch.setenv("UDP_OFFSET_PRUNE", "1")
    check_adv_offset(cuda)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", [3, 4])
def test_gpu_adv_gaussian(cuda, cid):
    check_adv_gaussian(cuda, cid)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_gpu_dec8(cuda, cid):
    check_dec8(cuda, cid)


@pytest.mark.gpu
@pytest.mark.parametrize("cid", [1, 3, 4, 5])
def test_gpu_nonfinite_maps_take_the_generic_path_in_every_shape(cuda, orc, cid):
    """Maps holding NaN / inf are re-done by the shape-generic, NaN-exact code from INSIDE the streaming kernels --
    whose CTA size differs per shape (96, 144, 128 threads; 144 has a partial last warp, which the generic code's
    butterfly reductions must cope with).  np.argmax: a NaN wins; masks, maxvals and NaN-ness as the oracle."""
    from udp_pose_b200 import synth
    cfg = CONFIGS[cid]
    w, h = cfg.heatmap_size
    rng = np.random.default_rng(600 + cid)
    N, K = 12, cfg.num_joints
    out, fl = synth.heatmaps(cfg, N, rng)
    out[1, 2, h // 2, w // 3] = np.nan
    out[4, 0, 0, 0] = np.inf
    out[7, K - 1, h - 1, w - 1] = -np.inf
    fl[9, 3, 5, 5] = np.nan
    c, s, _ = synth.boxes(cfg, N, rng)
    with np.errstate(all="ignore"):
        a = orc.decode(cfg, out, fl, c, s)
    b = cuda.decode(cfg, out, fl, c, s)
    assert np.array_equal(a["idx"], b["idx"])
    assert np.array_equal(np.isnan(a["maxvals"]), np.isnan(b["maxvals"]))
    ok = ~np.isnan(a["maxvals"])
    assert np.array_equal(a["maxvals"][ok], b["maxvals"][ok])
    assert np.array_equal(np.isnan(a["preds"]), np.isnan(b["preds"]))
    fin = np.isfinite(a["preds"])
    assert np.array_equal(fin, np.isfinite(b["preds"]))
    assert np.abs(a["preds"][fin] - b["preds"][fin]).max() <= TOL_PX
