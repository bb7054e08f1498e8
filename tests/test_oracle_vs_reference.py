"""CPU, container only: the C oracle against the UNMODIFIED reference imported live from
/root/reference on fresh seeds (skipped on the GPU box, where the tree does not exist)."""
import numpy as np
import pytest

from oracle import ref_import

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="/root/reference not present")

from oracle import oracle as O  # noqa: E402
from parity import same_bits  # noqa: E402
from udp_pose_b200 import synth  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_live(cid):
    import refcalls as RC
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(1000 * cid + 77)
    N = 10
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    c, s, r = synth.boxes(cfg, N, rng, rotate=True)
    M = RC.warp_matrices(cfg, c, s, r)
    assert same_bits(M, O.warp_matrices(r, c, s, W, H))
    imgs = synth.source_images(2, rng)
    for i in range(4):
        import cv2
        ref = cv2.warpAffine(imgs[i % 2], M[i], (W, H), flags=cv2.WARP_INVERSE_MAP | cv2.INTER_LINEAR)
        assert np.array_equal(ref, O.warp_affine_u8(imgs[i % 2], M[i], W, H))
    j, v = synth.joints(cfg, N, rng, rsn=cfg.target_type == "rsn")
    js = j.copy()
    js[..., 0] = rng.uniform(0, 640, js.shape[:2])
    js[..., 1] = rng.uniform(0, 480, js.shape[:2])
    assert same_bits(RC.rotate_points(cfg, js, c, s, r), O.rotate_points(js[..., :2], r, c, s, W, H))
    if cfg.target_type == "rsn":
        assert same_bits(RC.rsn_labels(cfg, j, v), O.encode_rsn(j, v[..., 0], W, H, w, h))
        j2, v2, Mi = RC.rsn_transform_joints(cfg, js, v, c, s, r)
        j2o, v2o, Mio = O.rsn_transform_joints(js, v, r, c, s, W, H)
        assert same_bits(Mi, Mio) and np.array_equal(v2, v2o) and np.abs(j2 - j2o).max() < 1e-9
    else:
        t, tw = RC.generate_target(cfg, j, v)
        if cfg.target_type == "gaussian":
            to, two = O.encode_gaussian(j, v, W, H, w, h, cfg.sigma)
        else:
            to, two = O.encode_offset(j, v, W, H, w, h, cfg.kpd)
        assert same_bits(tw, two) and np.abs(t - to).max() <= 1e-7 and np.array_equal(t != 0, to != 0)
    out, fl = synth.heatmaps(cfg, N, rng)
    trip = cfg.target_type == "offset"
    avg = RC.flip_average(cfg, out, fl)
    assert same_bits(avg, O.flip_average(out, fl, cfg.flip_pairs, trip))
    p, mv, pin = RC.final_preds(cfg, avg, c, s)
    d = O.decode_offset(avg, c, s, cfg.kpd) if trip else O.decode_gaussian(avg, c, s, True, cfg.pre_div)
    assert same_bits(mv, d["maxvals"])
    assert np.abs(p - d["preds"]).max() <= 1e-4 and np.abs(pin - d["preds_in"]).max() <= 1e-4
    if trip:
        assert same_bits(p, d["preds"].astype(np.float32))
    # the blurred, renormalised map post() writes back into its input (inference.py:80)
    if not trip:
        inf = ref_import.hrnet().inference
        x = (avg / cfg.pre_div if cfg.pre_div else avg).copy()
        coords, _ = inf.get_max_preds(x)
        with np.errstate(all="ignore"):
            inf.post(coords, x)
        d = O.decode_gaussian(avg, c, s, True, cfg.pre_div, want_blurred=True)
        assert same_bits(x, d["blurred"])
