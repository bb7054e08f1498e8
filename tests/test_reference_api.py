"""GPU: the reference's own call surface (udp_pose_b200.reference_api) against the golden
vectors of the unmodified reference -- one person per call, numpy in / numpy out, written the
way a test of deep_hrnet/lib would call these names."""
import numpy as np
import pytest

import parity
from parity import same_bits, TOL_HM, TOL_PX
from udp_pose_b200.configs import CONFIGS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available()
    from udp_pose_b200 import _lib, reference_api
    _lib.load()
    return reference_api


def _dataset(R, cfg):
    ds = R.JointsDataset(cfg.as_cfg(), "", "val", False)
    ds.num_joints = cfg.num_joints
    ds.flip_pairs = cfg.flip_pairs
    return ds


@pytest.mark.parametrize("cid", [1, 2, 3, 4])
def test_get_warpmatrix_rotate_points_warp_affine(R, cid):
    cfg, g = CONFIGS[cid], parity.load(cid)
    img_size = np.array(cfg.image_size)
    for i in range(len(g["rot"])):
        c, s, r = g["center"][i], g["scale"][i], float(g["rot"][i])
        M = R.get_warpmatrix(r, c * 2.0, img_size - 1.0, s)
        assert M.dtype == np.float32 and M.shape == (2, 3)
        np.testing.assert_allclose(M, g["warp_matrix"][i], rtol=1.2e-7, atol=1e-30)
        p = R.rotate_points(g["joints_src"][i, :, 0:2], r, c, img_size, s, False)
        assert p.dtype == np.float64 and np.abs(p - g["joints_rot"][i]).max() <= TOL_PX
    for i, si in enumerate(g["src_index"]):
        crop = R.warp_affine(g["src_images"][si], g["warp_matrix"][i], (int(img_size[0]), int(img_size[1])))
        assert np.array_equal(crop, g["crop"][i])


@pytest.mark.parametrize("cid", [1, 2, 3, 4])
def test_generate_target(R, cid):
    cfg, g = CONFIGS[cid], parity.load(cid)
    ds = _dataset(R, cfg)
    for i in range(g["joints"].shape[0]):
        t, tw = ds.generate_target(g["joints"][i], g["joints_vis"][i])
        assert t.dtype == np.float32 and tw.dtype == np.float32 and tw.shape == (cfg.num_joints, 1)
        assert same_bits(tw, g["target_weight"][i])
        assert np.abs(t - g["target"][i]).max() <= TOL_HM
    ds.use_different_joints_weight = True
    ds.joints_weight = g["joints_weight"].reshape(-1, 1)
    _, tw = ds.generate_target(g["joints"][0], g["joints_vis"][0])
    assert same_bits(tw, g["target_weight_jw"][0])


@pytest.mark.parametrize("cid", [1, 2, 3, 4])
def test_flip_test_and_get_final_preds(R, cid):
    """validate(): flip_back(_offset) -> average -> get_final_preds (function.py:154-171,205)."""
    cfg, g = CONFIGS[cid], parity.load(cid)
    c = cfg.as_cfg()
    out, fl = g["dec_out"], g["dec_flipped"]
    fb = (R.flip_back_offset if cfg.target_type == "offset" else R.flip_back)(fl.copy(), cfg.flip_pairs)
    avg = (out + fb) * np.float32(0.5)
    preds, maxvals, pin = R.get_final_preds(c, avg.copy(), g["dec_center"], g["dec_scale"])
    assert maxvals.dtype == np.float32 and same_bits(maxvals, g["maxvals"])
    want_dt = np.float64 if cfg.target_type == "gaussian" else np.float32
    assert preds.dtype == want_dt and pin.dtype == want_dt
    assert np.abs(preds - g["preds"]).max() <= TOL_PX and np.abs(pin - g["preds_in"]).max() <= TOL_PX
    # the fused form gives the same answer in one launch
    p2, m2, i2 = R.get_final_preds(c, out, g["dec_center"], g["dec_scale"], output_flipped=fl,
                                   flip_pairs=cfg.flip_pairs)
    assert same_bits(m2, maxvals) and same_bits(p2, preds) and same_bits(i2, pin)
    # no flip
    p3, m3, _ = R.get_final_preds(c, out.copy(), g["dec_center"], g["dec_scale"])
    assert same_bits(m3, g["maxvals_noflip"]) and np.abs(p3 - g["preds_noflip"]).max() <= TOL_PX


def test_get_max_preds_post_transform_preds(R):
    cfg, g = CONFIGS[1], parity.load(1)
    hm = g["dec_out"].copy()
    coords, maxvals = R.get_max_preds(hm)
    assert coords.dtype == np.float32 and coords.shape == (2, 17, 2) and maxvals.shape == (2, 17, 1)
    flat = hm.reshape(2, 17, -1)
    assert np.array_equal(coords[..., 0], (flat.argmax(2) % 48).astype(np.float32) * (flat.max(2) > 0))
    assert np.array_equal(maxvals[..., 0], flat.max(2))
    before = hm.copy()
    res = R.post(coords, hm)
    assert res.dtype == np.float64
    assert not np.array_equal(before, hm)  # post() writes the blurred maps back (inference.py:80)
    for i in range(2):
        t = R.transform_preds(res[i], g["dec_center"][i], g["dec_scale"][i], [48, 64])
        assert t.dtype == np.float64 and np.abs(t - g["preds_noflip"][i]).max() <= TOL_PX
    with pytest.raises(ValueError):
        R.post(coords + 1, g["dec_out"].copy())
    with pytest.raises(AssertionError):
        R.get_max_preds(hm[0])          # ndim != 4 (inference.py:35-37)
    with pytest.raises(AssertionError):
        R.get_max_preds(hm.tolist())    # not an ndarray


def test_fliplr_joints(R):
    cfg, g = CONFIGS[1], parity.load(1)
    j, v = g["joints_src"][0].copy(), g["joints_vis"][0].copy()
    jo, vo = R.fliplr_joints(j.copy(), v.copy(), 320, cfg.flip_pairs)
    perm = np.arange(17)
    for a, b in cfg.flip_pairs:
        perm[a], perm[b] = b, a
    jf = j.copy()
    jf[:, 0] = 320 - jf[:, 0] - 1
    assert np.array_equal(vo, v[perm]) and np.array_equal(jo, jf[perm] * v[perm])


def test_getitem_eval_and_train_rng_order(R):
    """__getitem__ (JointsDataset.py:172-256): eval record reproduces crop/target of the golden
    set; the train path draws its random numbers in the reference's order."""
    import random
    cfg, g = CONFIGS[1], parity.load(1)
    ds = _dataset(R, cfg)
    n = len(g["src_index"])
    ds.db = [dict(image="mem://%d" % i, joints_3d=g["joints_src"][i].copy(), joints_3d_vis=g["joints_vis"][i].copy(),
                  center=g["center"][i].copy(), scale=g["scale"][i].copy(), score=0.5) for i in range(n)]
    for i in range(n):
        if g["rot"][i] != 0:
            continue
        x, t, tw, meta = ds.__getitem__(i, image=g["src_images"][g["src_index"][i]])
        assert np.array_equal(x, g["crop"][i])
        assert np.abs(meta["joints"][:, :2] - g["joints_rot"][i]).max() <= TOL_PX
        assert meta["rotation"] == 0 and meta["score"] == 0.5 and t.shape == (17, 64, 48)
    ds.is_train = True
    np.random.seed(3); random.seed(3)
    _, _, _, meta = ds.__getitem__(0, image=g["src_images"][0])
    np.random.seed(3); random.seed(3)
    sf, rf = ds.scale_factor, ds.rotation_factor
    if np.sum(g["joints_vis"][0][:, 0]) > ds.num_joints_half_body:
        np.random.rand()  # the half-body draw (short-circuit `and`, JointsDataset.py:207-208)
    s = g["scale"][0] * np.clip(np.random.randn() * sf + 1, 1 - sf, 1 + sf)
    r = np.clip(np.random.randn() * rf, -rf * 2, rf * 2) if random.random() <= 0.6 else 0
    assert np.array_equal(meta["scale"], s) and meta["rotation"] == r


def test_getitem_train_with_aid(R):
    """JointsDataset.py:96-106, 230-235: with DATASET.CUTOUT / HIDE_AND_SEEK set, the train path erases
    the crop with parameters drawn after the warp draws, Cutout before HideAndSeek, and leaves numpy's
    generator where the reference would."""
    import random
    from oracle import oracle as O
    cfg, g = CONFIGS[1], parity.load(1)
    ycfg = cfg.as_cfg()
    ycfg.DATASET.CUTOUT, ycfg.DATASET.HIDE_AND_SEEK = [1.0, 0.3, 2], [1.0, 0.5, (16, 32, 0)]
    ds = R.JointsDataset(ycfg, "", "train", True)
    plain = R.JointsDataset(cfg.as_cfg(), "", "train", True)
    for d in (ds, plain):
        d.num_joints, d.flip_pairs = cfg.num_joints, cfg.flip_pairs
        d.db = [dict(image="mem://0", joints_3d=g["joints_src"][0].copy(), joints_3d_vis=g["joints_vis"][0].copy(),
                     center=g["center"][0].copy(), scale=g["scale"][0].copy(), score=1)]
    img = g["src_images"][g["src_index"][0]]
    np.random.seed(5); random.seed(5)
    x_aid, _, _, _ = ds.__getitem__(0, image=img)
    after = np.random.rand()
    np.random.seed(5); random.seed(5)
    x_plain, _, _, _ = plain.__getitem__(0, image=img)
    ell = R.Cutout(1.0, 0.3, 2).draw(256, 192)
    grid, bits = R.HideAndSeek(1.0, 0.5, (16, 32, 0)).draw(256, 192)
    want = O.aid_erase(O.aid_erase(x_plain, ellipses=ell), grid=grid, hidden=bits)
    assert np.array_equal(x_aid, want) and not np.array_equal(x_aid, x_plain)
    assert np.random.rand() == after
    with pytest.raises(NotImplementedError):
        ds.get_batch([0], [img], normalize=((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)))


def test_get_batch_with_fused_normalize(R):
    """get_batch(normalize=(mean, std)) returns what the reference's DataLoader hands the model:
    Normalize(ToTensor(crop)) float32 CHW (tools/train.py:127-143), plus np.flip(input, 3)
    (core/function.py:154) -- bit-exact against the transform of the golden crops."""
    import torch
    from oracle import oracle as O
    from udp_pose_b200 import device_api as D
    cfg, g = CONFIGS[1], parity.load(1)
    ds = _dataset(R, cfg)
    n = len(g["src_index"])
    ds.db = [dict(image="mem://%d" % i, joints_3d=g["joints_src"][i].copy(), joints_3d_vis=g["joints_vis"][i].copy(),
                  center=g["center"][i].copy(), scale=g["scale"][i].copy(), score=0.5) for i in range(n)]
    idx = [i for i in range(n) if g["rot"][i] == 0]
    imgs = [g["src_images"][g["src_index"][i]] for i in idx]
    (x, xf), t, tw, metas = ds.get_batch(idx, imgs, normalize=(D.IMAGENET_MEAN, D.IMAGENET_STD), want_flipped=True)
    assert x.dtype == torch.float32 and tuple(x.shape[1:]) == (3, 256, 192)
    assert torch.equal(xf, x.flip(3))
    xh = x.cpu().numpy()
    for k, i in enumerate(idx):
        want = O.to_tensor_normalize(g["crop"][i], D.IMAGENET_MEAN, D.IMAGENET_STD)
        assert np.array_equal(xh[k].view(np.uint32), want.view(np.uint32))
