"""Shared parity checks: any implementation (the C oracle, or the CUDA codec through its
C-ABI) is run on the committed golden inputs and compared with what the unmodified
reference returned (tests/golden/cfg*.npz, made by tests/golden/make_golden.py).

Tolerances (BASELINE.json north_star): crops, argmax indices, masks, target_weight and
matrices bit-exact; heatmaps/targets <= 1e-5 absolute; coordinates <= 1e-4 px.
"""
import os

import numpy as np

from udp_pose_b200.configs import CONFIGS

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_HM = 1e-5
TOL_PX = 1e-4


def load(cid):
    return np.load(os.path.join(GOLDEN, "cfg%d.npz" % cid))


def same_bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a.view(np.uint8), b.view(np.uint8))


class OracleImpl:
    """Adapter giving oracle/oracle.py the interface the checks below expect."""

    def __init__(self):
        from oracle import oracle as O
        self.O = O

    def warp_matrices(self, rot, c, s, W, H):
        return self.O.warp_matrices(rot, c, s, W, H)

    def rotate_points(self, jxy, rot, c, s, W, H):
        return self.O.rotate_points(jxy, rot, c, s, W, H)

    def warp_affine(self, images, index, M, W, H):
        return np.stack([self.O.warp_affine_u8(images[index[i]], M[i], W, H) for i in range(len(M))])

    def encode_gaussian(self, j, v, W, H, w, h, sigma, jw=None):
        return self.O.encode_gaussian(j, v, W, H, w, h, sigma, jw)

    def encode_offset(self, j, v, W, H, w, h, kpd, jw=None):
        return self.O.encode_offset(j, v, W, H, w, h, kpd, jw)

    def encode_rsn(self, j, valid, W, H, w, h):
        return self.O.encode_rsn(j, valid, W, H, w, h)

    def rsn_transform_joints(self, j, v, rot, c, s, W, H):
        return self.O.rsn_transform_joints(j, v, rot, c, s, W, H)

    def decode(self, cfg, out, flipped, c, s, post_process=True):
        O = self.O
        trip = cfg.target_type == "offset"
        hm = out if flipped is None else O.flip_average(out, flipped, cfg.flip_pairs, trip)
        if trip:
            return O.decode_offset(hm, c, s, cfg.kpd)
        return O.decode_gaussian(hm, c, s, post_process, cfg.pre_div)


def check_crop_side(impl, cid):
    cfg, g = CONFIGS[cid], load(cid)
    W, H = cfg.image_size
    M = impl.warp_matrices(g["rot"], g["center"], g["scale"], W, H)
    # r == 0 rows are exact by construction; rotated rows depend on sin/cos of the platform
    # (glibc on the host, CUDA libdevice on the GPU): bit-exact expected, <= 1 float32 ulp allowed.
    assert same_bits(M[0], g["warp_matrix"][0])
    np.testing.assert_allclose(M, g["warp_matrix"], rtol=1.2e-7, atol=1e-30)
    n = len(g["src_index"])
    crop = impl.warp_affine(g["src_images"], g["src_index"], g["warp_matrix"][:n], W, H)
    assert crop.dtype == np.uint8 and np.array_equal(crop, g["crop"])
    rp = impl.rotate_points(g["joints_src"][..., :2], g["rot"], g["center"], g["scale"], W, H)
    assert np.abs(rp - g["joints_rot"]).max() <= TOL_PX
    return dict(matrix_bit_exact=bool(same_bits(M, g["warp_matrix"])))


def check_encode(impl, cid):
    cfg, g = CONFIGS[cid], load(cid)
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    j, v = g["joints"], g["joints_vis"]
    if cfg.target_type == "rsn":
        lab = impl.encode_rsn(j, v[..., 0], W, H, w, h)
        assert lab.shape == g["labels"].shape and lab.dtype == np.float32
        # values reach 255: one float32 ulp there is 1.5e-5, so the bound is 1 ulp-or-1e-5
        err = np.abs(lab.astype(np.float64) - g["labels"])
        assert (err <= np.maximum(TOL_HM, np.spacing(np.abs(g["labels"])))).all(), err.max()
        j2, v2, Mi = impl.rsn_transform_joints(g["joints_src"], v, g["rot"], g["center"], g["scale"], W, H)
        assert np.array_equal(v2, g["rsn_vis"])
        assert np.abs(j2 - g["rsn_joints"]).max() <= TOL_PX
        np.testing.assert_allclose(Mi, g["rsn_minv"], rtol=1.2e-7, atol=1e-30)
        return
    if cfg.target_type == "gaussian":
        t, tw = impl.encode_gaussian(j, v, W, H, w, h, cfg.sigma)
        _, tw2 = impl.encode_gaussian(j, v, W, H, w, h, cfg.sigma, g["joints_weight"])
    else:
        t, tw = impl.encode_offset(j, v, W, H, w, h, cfg.kpd)
        _, tw2 = impl.encode_offset(j, v, W, H, w, h, cfg.kpd, g["joints_weight"])
        assert np.array_equal(t[:, 0::3], g["target"][:, 0::3])  # classification channel bit-exact
    assert t.shape == g["target"].shape and t.dtype == np.float32
    assert same_bits(tw, g["target_weight"])
    assert same_bits(tw2, g["target_weight_jw"])
    assert np.array_equal(t != 0, g["target"] != 0)
    assert np.abs(t - g["target"]).max() <= TOL_HM


def _cmp_decode(r, preds, maxvals, pin, tol=TOL_PX, exact_maxvals=True):
    assert r["maxvals"].shape == maxvals.shape
    if exact_maxvals:
        assert same_bits(r["maxvals"], maxvals)
    assert np.array_equal(np.isnan(r["preds"]), np.isnan(preds))
    e1 = np.nanmax(np.abs(r["preds"] - preds)) if preds.size else 0.0
    e2 = np.nanmax(np.abs(r["preds_in"] - pin)) if pin.size else 0.0
    assert e1 <= tol and e2 <= tol, (e1, e2)
    return max(e1, e2)


def check_decode(impl, cid):
    cfg, g = CONFIGS[cid], load(cid)
    c, s = g["dec_center"], g["dec_scale"]
    errs = {}
    r = impl.decode(cfg, g["dec_out"], g["dec_flipped"], c, s)
    errs["flip"] = _cmp_decode(r, g["preds"], g["maxvals"], g["preds_in"])
    r = impl.decode(cfg, g["dec_out"], None, c, s)
    errs["noflip"] = _cmp_decode(r, g["preds_noflip"], g["maxvals_noflip"], g["preds_in_noflip"])
    if cfg.target_type != "offset":
        r = impl.decode(cfg, g["dec_out"], g["dec_flipped"], c, s, post_process=False)
        # no refinement: float32 arithmetic end to end, expected bit-exact
        assert same_bits(r["preds"].astype(np.float32), g["preds_nopost"])
        assert same_bits(r["preds_in"].astype(np.float32), g["preds_in_nopost"])
        assert same_bits(r["maxvals"], g["maxvals_nopost"])
    else:
        assert same_bits(r["preds"].astype(np.float32), g["preds_noflip"])
    return errs


def check_adversarial(impl):
    """cfg 1 adversarial maps: masks/maxvals exact; coordinates compared where the Taylor
    step is well conditioned (the reference's own float32 log noise is amplified by
    cond(H) elsewhere -- reported, not asserted)."""
    cfg, g = CONFIGS[1], load(1)
    r = impl.decode(cfg, g["adv_hm"], None, g["adv_center"], g["adv_scale"])
    assert same_bits(r["maxvals"], g["adv_maxvals"])
    err = np.abs(r["preds"] - g["adv_preds"]).max(axis=(1, 2))
    # rows: 0 U(0,1) noise, 1 N(0,1) noise, 2 all-negative, 3 ties, 4 border peaks, 5 single pixel
    assert err[2] == 0.0            # mask = 0: coords zeroed, zero shift (flat clipped log map)
    assert err[3] <= 1e-3 and err[4] <= TOL_PX and err[5] <= TOL_PX
    assert err[0] <= 5e-2 and err[1] <= 5e-1
    return err
