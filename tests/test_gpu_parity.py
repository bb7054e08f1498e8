"""GPU: the CUDA path, called through the C ABI, against (1) the committed golden vectors of
the unmodified reference and (2) the C oracle on fresh seeded inputs, plus size-independent
properties at larger N."""
import numpy as np
import pytest

import parity
from parity import same_bits
from udp_pose_b200 import synth
from udp_pose_b200.configs import CONFIGS, RSN_SIGMAS

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


class CudaImpl:
    """Adapter over udp_pose_b200.device_api with the interface parity.py expects."""

    def __init__(self):
        import torch
        from udp_pose_b200 import _lib, device_api
        assert torch.cuda.is_available()
        _lib.load()  # raises if libudp_codec.so is missing: no fallback
        self.D = device_api
        self.torch = torch

    def warp_matrices(self, rot, c, s, W, H):
        return _np(self.D.warp_matrices(c, s, np.asarray(rot, np.float64), (W, H)))

    def rotate_points(self, jxy, rot, c, s, W, H):
        return _np(self.D.rotate_points_batch(np.ascontiguousarray(jxy), np.asarray(rot, np.float64), c, s, (W, H)))

    def warp_affine(self, images, index, M, W, H):
        src = self.D.SourceImages([self.torch.from_numpy(np.ascontiguousarray(im)) for im in images], index)
        return _np(self.D.crop_batch(src, M, (W, H)))

    def encode_gaussian(self, j, v, W, H, w, h, sigma, jw=None):
        t, tw = self.D.encode_gaussian(j, v, (W, H), (w, h), sigma, jw)
        return _np(t), _np(tw)

    def encode_offset(self, j, v, W, H, w, h, kpd, jw=None):
        t, tw = self.D.encode_offset(j, v, (W, H), (w, h), kpd, jw)
        return _np(t), _np(tw)

    def encode_rsn(self, j, valid, W, H, w, h):
        return _np(self.D.encode_rsn(j, np.ascontiguousarray(valid), (W, H), (w, h), RSN_SIGMAS))

    def rsn_transform_joints(self, j, v, rot, c, s, W, H):
        a, b, m = self.D.rsn_transform_joints_batch(j, v, np.asarray(rot, np.float64), c, s, (W, H))
        return _np(a), _np(b), _np(m)

    def decode(self, cfg, out, flipped, c, s, post_process=True, **kw):
        r = self.D.decode_batch(cfg, out, flipped, c, s, post_process=post_process, want_idx=True, **kw)
        return {k: _np(v).astype(np.float64) if k in ("preds", "preds_in") else _np(v) for k, v in r.items()}


@pytest.fixture(scope="module")
def impl():
    return CudaImpl()


@pytest.fixture(scope="module")
def orc():
    return parity.OracleImpl()


# ------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_golden_crop_side(impl, cid):
    parity.check_crop_side(impl, cid)


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_golden_encode(impl, cid):
    parity.check_encode(impl, cid)


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_golden_decode(impl, cid):
    parity.check_decode(impl, cid)


def test_golden_adversarial(impl):
    parity.check_adversarial(impl)


# ------------------------------------------------------------------ oracle, fresh seeds
@pytest.mark.parametrize("cid", sorted(CONFIGS))
@pytest.mark.parametrize("N", [1, 7, 32])
def test_crop_vs_oracle(impl, orc, cid, N):
    cfg = CONFIGS[cid]
    W, H = cfg.image_size
    rng = np.random.default_rng(1000 * cid + N)
    c, s, r = synth.boxes(cfg, N, rng, rotate=True)
    M = orc.warp_matrices(r, c, s, W, H)
    Mg = impl.warp_matrices(r, c, s, W, H)
    assert same_bits(M[r == 0], Mg[r == 0])
    np.testing.assert_allclose(Mg, M, rtol=1.2e-7, atol=1e-30)
    imgs = synth.source_images(3, rng)
    idx = np.arange(N) % 3
    want = orc.warp_affine(imgs, idx, M, W, H)
    got = impl.warp_affine(imgs, idx, M, W, H)
    assert np.array_equal(want, got)  # u8 crop bit-exact
    K = cfg.num_joints
    js = np.stack([rng.uniform(0, 640, (N, K)), rng.uniform(0, 480, (N, K))], -1)
    assert np.abs(orc.rotate_points(js, r, c, s, W, H) - impl.rotate_points(js, r, c, s, W, H)).max() <= 1e-9


@pytest.mark.parametrize("cid", sorted(CONFIGS))
@pytest.mark.parametrize("N", [1, 7, 64])
def test_encode_vs_oracle(impl, orc, cid, N):
    cfg = CONFIGS[cid]
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    rng = np.random.default_rng(2000 * cid + N)
    j, v = synth.joints(cfg, N, rng, rsn=cfg.target_type == "rsn")
    if cfg.target_type == "rsn":
        a, b = orc.encode_rsn(j, v[..., 0], W, H, w, h), impl.encode_rsn(j, v[..., 0], W, H, w, h)
        err = np.abs(a.astype(np.float64) - b)
        assert (err <= np.maximum(1e-5, np.spacing(np.abs(a)))).all()
        # The GPU forms the product of the separable factors as a float-float sum on the packed FP32
        # pipe (2^-46 relative before rounding): bits differ from the oracle only at near-ties, and
        # in the band where the value itself is a float32 denormal-range number (< 2^-80, i.e. an
        # absolute difference below 1e-38).
        mism, big = a != b, np.abs(a) >= 2.0 ** -80
        assert (mism & big).mean() < 1e-5
        if (mism & ~big).any():
            assert np.abs(a[mism & ~big].astype(np.float64) - b[mism & ~big]).max() < 1e-30
        return
    if cfg.target_type == "gaussian":
        (a, aw), (b, bw) = orc.encode_gaussian(j, v, W, H, w, h, cfg.sigma), impl.encode_gaussian(j, v, W, H, w, h, cfg.sigma)
    else:
        (a, aw), (b, bw) = orc.encode_offset(j, v, W, H, w, h, cfg.kpd), impl.encode_offset(j, v, W, H, w, h, cfg.kpd)
        assert same_bits(a, b)  # offset target: every channel bit-exact
    assert same_bits(aw, bw)
    assert np.array_equal(a != 0, b != 0)
    assert np.abs(a - b).max() <= 1e-7


@pytest.mark.parametrize("cid", sorted(CONFIGS))
@pytest.mark.parametrize("flip", [True, False])
def test_decode_vs_oracle(impl, orc, cid, flip):
    cfg = CONFIGS[cid]
    N = 33
    rng = np.random.default_rng(3000 * cid + flip)
    out, fl = synth.heatmaps(cfg, N, rng, flip=flip)
    c, s, _ = synth.boxes(cfg, N, rng)
    a = orc.decode(cfg, out, fl, c, s)
    b = impl.decode(cfg, out, fl, c, s)
    assert np.array_equal(a["idx"], b["idx"])            # argmax indices bit-exact
    assert same_bits(a["maxvals"], b["maxvals"])         # incl. the blurred maxima (offset mode)
    assert np.abs(a["preds"] - b["preds"]).max() <= 1e-4
    assert np.abs(a["preds_in"] - b["preds_in"]).max() <= 1e-4
    if cfg.target_type == "offset":
        assert same_bits(a["preds"].astype(np.float32), b["preds"].astype(np.float32))


@pytest.mark.parametrize("cid", [1, 3, 4, 5])
def test_blurred_writeback(impl, orc, cid):
    """post() overwrites its input with the blurred, renormalised maps (inference.py:80)."""
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(4000 + cid)
    out, _ = synth.heatmaps(cfg, 5, rng, flip=False)
    c, s, _ = synth.boxes(cfg, 5, rng)
    want = orc.O.decode_gaussian(out, c, s, True, cfg.pre_div, want_blurred=True)["blurred"]
    got = impl.decode(cfg, out, None, c, s, want_blurred=True)["blurred"]
    assert same_bits(want, got)


def test_edge_maps(impl, orc):
    """Constant map -> NaN coords (0/0 at inference.py:79); NaN-ness must match the oracle."""
    cfg = CONFIGS[1]
    w, h = cfg.heatmap_size
    hm = np.zeros((2, cfg.num_joints, h, w), np.float32)
    hm[0] = 0.25
    hm[1, :, 10, 10] = 1.0
    c, s, _ = synth.boxes(cfg, 2, np.random.default_rng(5))
    a, b = orc.decode(cfg, hm, None, c, s), impl.decode(cfg, hm, None, c, s)
    assert np.array_equal(np.isnan(a["preds"]), np.isnan(b["preds"]))
    assert np.isnan(b["preds"][0]).all() and not np.isnan(b["preds"][1]).any()
    assert np.abs(a["preds"][1] - b["preds"][1]).max() <= 1e-4


def test_empty_batch(impl):
    cfg = CONFIGS[1]
    w, h = cfg.heatmap_size
    r = impl.decode(cfg, np.zeros((0, 17, h, w), np.float32), None, np.zeros((0, 2), np.float32),
                    np.zeros((0, 2), np.float32))
    assert r["preds"].shape == (0, 17, 2)
    t, tw = impl.encode_gaussian(np.zeros((0, 17, 3)), np.zeros((0, 17, 3)), 192, 256, w, h, 2)
    assert t.shape == (0, 17, h, w)


# ------------------------------------------------------------------ properties at size
@pytest.mark.parametrize("cid", [1, 2])
def test_roundtrip_large(impl, cid):
    """encode -> decode round trip at N = 4096: decoding the encoder's own target recovers
    the joint positions (size-independent property; no oracle involved)."""
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[cid]
    N = 4096
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    rng = np.random.default_rng(77 + cid)
    j, v = synth.joints(cfg, N, rng)
    j[..., 0] = rng.uniform(0.1, 0.9, j.shape[:2]) * (W - 1)
    j[..., 1] = rng.uniform(0.1, 0.9, j.shape[:2]) * (H - 1)
    v[:] = 1.0
    tgt, tw = D.encode_batch(cfg, j, v)
    assert float(tw.min()) == 1.0
    r = D.decode_batch(cfg, tgt, None, None, None)
    hm_xy = r["preds"].double().cpu().numpy()
    fs = np.array([(W - 1.0) / (w - 1.0), (H - 1.0) / (h - 1.0)])
    err = np.abs(hm_xy * fs - j[..., :2])
    tol = 0.02 if cid == 1 else 1.0  # Taylor refinement of an exact Gaussian / blurred offset vote
    assert err.max() <= tol, err.max()
    torch.cuda.synchronize()


def test_flip_idempotence_large(impl):
    """flip_back(flip_back(x)) == x and decode(x, flipped=flip(x)) == decode(x) at N = 2048."""
    import torch
    from udp_pose_b200 import device_api as D
    for cid in (1, 2):
        cfg = CONFIGS[cid]
        trip = cfg.target_type == "offset"
        w, h = cfg.heatmap_size
        g = torch.Generator(device="cuda").manual_seed(cid)
        x = torch.rand((2048, cfg.channels, h, w), device="cuda", generator=g)
        f = D.flip_back(x, cfg.flip_pairs, trip)
        assert torch.equal(D.flip_back(f, cfg.flip_pairs, trip), x)
        a = D.decode_batch(cfg, x, None, None, None, want_idx=True)
        b = D.decode_batch(cfg, x, f, None, None, want_idx=True)  # (x + x) * 0.5 == x exactly
        assert torch.equal(a["idx"], b["idx"]) and torch.equal(a["preds"], b["preds"])


@pytest.mark.parametrize("case", ["odd_width", "misaligned_base", "tiny_image", "far_outside", "huge_downscale", "upscale"])
def test_crop_alignment_and_edges(impl, orc, case):
    """The tiled crop copies 16-byte chunks that keep the global phase: exercise every image
    alignment (odd widths, misaligned base pointers), images smaller than the crop, boxes far
    outside the image, and extreme scales (per-pixel fallback) -- all bit-exact vs the oracle."""
    import torch
    from udp_pose_b200 import device_api as D
    rng = np.random.default_rng(hash(case) % 1000)
    cfg = CONFIGS[1]
    W, H = cfg.image_size
    sizes = {"odd_width": [(375, 427), (333, 501), (240, 333)], "misaligned_base": [(480, 640), (300, 401)],
             "tiny_image": [(17, 23), (40, 31), (5, 4)], "far_outside": [(200, 300)],
             "huge_downscale": [(600, 800)], "upscale": [(480, 640)]}[case]
    N = 12
    imgs, dev_imgs = [], []
    for k, (hs, ws) in enumerate(sizes):
        im = rng.integers(0, 256, (hs, ws, 3), dtype=np.uint8)
        imgs.append(im)
        if case == "misaligned_base":
            buf = torch.zeros(im.size + 64, dtype=torch.uint8, device="cuda")
            off = 1 + 2 * k + 4 * (k % 2)
            buf[off:off + im.size] = torch.from_numpy(im.reshape(-1)).cuda()
            dev_imgs.append(buf[off:off + im.size].view(hs, ws, 3))
            assert dev_imgs[-1].data_ptr() % 4 != 0
        else:
            dev_imgs.append(torch.from_numpy(im).cuda())
    idx = np.arange(N) % len(sizes)
    c = np.zeros((N, 2), np.float32)
    s = np.zeros((N, 2), np.float32)
    r = np.where(np.arange(N) % 3 == 0, 0.0, np.clip(rng.standard_normal(N) * 45, -90, 90))
    for i in range(N):
        hs, ws = sizes[idx[i]]
        bh = rng.uniform(0.3, 1.2) * max(hs, 30)
        if case == "huge_downscale":
            bh = rng.uniform(8, 40) * hs
        if case == "upscale":
            bh = rng.uniform(6, 30)
        s[i] = np.array([bh * W / H, bh]) / 200.0 * 1.25
        c[i] = [rng.uniform(-0.1, 1.1) * ws, rng.uniform(-0.1, 1.1) * hs]
        if case == "far_outside":
            c[i] += np.array([ws, hs]) * rng.choice([-3.0, 3.0], 2)
    M = orc.warp_matrices(r, c, s, W, H)
    want = np.stack([orc.O.warp_affine_u8(imgs[idx[i]], M[i], W, H) for i in range(N)])
    got = D.crop_batch(D.SourceImages(dev_imgs, idx), M, (W, H)).cpu().numpy()
    assert np.array_equal(want, got)
    if case == "far_outside":
        assert got.max() == 0


@pytest.mark.parametrize("cid", [2, 3])
def test_crop_tma_and_cpasync_paths_agree(impl, orc, cid):
    """Images with a 16-byte stride go through the TMA kernel (device-patched tensor map, hardware
    zero fill), the rest through the cp.async kernel; UDP_CROP_NO_TMA=1 forces the latter for all.
    Same bytes from both, and from the oracle on a subset.  Covers: contiguous 640-wide images,
    a padded pitch with garbage in the padding, a misaligned base with an aligned pitch (the
    bytes before the first pixel are not zero-filled by TMA), a stride that is not a multiple of 16,
    every tile height (scales up to 3.3), rotated and unrotated boxes crossing all four edges."""
    import os
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[cid]
    W, H = cfg.image_size
    rng = np.random.default_rng(100 + cid)
    N = 384
    hw = [(480, 640), (300, 401), (360, 638), (375, 500)]
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in hw]
    dev = [torch.from_numpy(imgs[0]).cuda()]
    pad = torch.full((300, 1216), 255, dtype=torch.uint8, device="cuda")          # pitch 1216 = 16 * 76 > 3 * 401
    pad[:, :1203] = torch.from_numpy(imgs[1].reshape(300, 1203)).cuda()
    dev.append(pad.as_strided((300, 401, 3), (1216, 3, 1)))
    buf = torch.full((360 * 1920 + 64,), 255, dtype=torch.uint8, device="cuda")  # base + 3, pitch 1920
    inner = buf[3:3 + 360 * 1920].view(360, 1920)
    inner[:, :1914] = torch.from_numpy(imgs[2].reshape(360, 1914)).cuda()
    dev.append(inner.as_strided((360, 638, 3), (1920, 3, 1)))
    assert dev[2].data_ptr() % 16 == 3
    dev.append(torch.from_numpy(imgs[3]).cuda())                                  # stride 1500: cp.async path
    idx = np.arange(N) % 4
    c = np.zeros((N, 2), np.float32)
    s = np.zeros((N, 2), np.float32)
    r = np.where(rng.uniform(size=N) < 0.4, 0.0, np.clip(rng.standard_normal(N) * 40, -80, 80))
    for i in range(N):
        h, w = hw[idx[i]]
        bh = rng.uniform(0.2, 1.4) * h * (H / 256.0) ** 0 if i % 7 else rng.uniform(1.4, 1.8) * h
        s[i] = np.array([bh * W / H, bh]) / 200.0 * 1.25
        c[i] = [rng.uniform(-0.15, 1.15) * w, rng.uniform(-0.15, 1.15) * h]
    M = orc.warp_matrices(r, c, s, W, H)
    src = D.SourceImages(dev, idx)
    assert D.crop_launches((W, H)) == 2
    a = D.crop_batch(src, M, (W, H))
    os.environ["UDP_CROP_NO_TMA"] = "1"
    try:
        assert D.crop_launches((W, H)) == 1
        b = D.crop_batch(src, M, (W, H))
    finally:
        del os.environ["UDP_CROP_NO_TMA"]
    assert torch.equal(a, b)
    got = a.cpu().numpy()
    for i in range(0, N, 5):
        assert np.array_equal(orc.O.warp_affine_u8(imgs[idx[i]], M[i], W, H), got[i]), i


def test_crop_generic_path_matches_tiled(impl):
    """UDP_FORCE_GENERIC=1 routes the crop through the first-generation kernel: same bytes."""
    import os
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[3]
    rng = np.random.default_rng(11)
    c, s, r = synth.boxes(cfg, 9, rng, rotate=True)
    imgs = torch.from_numpy(synth.source_images(2, rng)).cuda()
    src = D.SourceImages(imgs, np.arange(9) % 2)
    M = D.warp_matrices(c, s, r, cfg.image_size)
    a = D.crop_batch(src, M, cfg.image_size)
    os.environ["UDP_FORCE_GENERIC"] = "1"
    try:
        b = D.crop_batch(src, M, cfg.image_size)
    finally:
        del os.environ["UDP_FORCE_GENERIC"]
    assert torch.equal(a, b)


@pytest.mark.parametrize("cid", [1, 2])
def test_decode_generic_and_misaligned_match_streaming(impl, cid):
    """UDP_FORCE_GENERIC=1 routes the decoders through the shape-generic kernels: identical bits.
    Map tensors must be 16-byte aligned (vector loads): a misaligned one is rejected, not misread."""
    import os
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(60 + cid)
    N = 9
    out, fl = synth.heatmaps(cfg, N, rng)
    c, s, _ = synth.boxes(cfg, N, rng)
    a = D.decode_batch(cfg, out, fl, c, s, want_idx=True)
    os.environ["UDP_FORCE_GENERIC"] = "1"
    try:
        b = D.decode_batch(cfg, out, fl, c, s, want_idx=True)
    finally:
        del os.environ["UDP_FORCE_GENERIC"]
    def shifted(x):  # same values at an address that is only 4-byte aligned
        buf = torch.empty(x.size + 1, dtype=torch.float32, device="cuda")
        buf[1:] = torch.from_numpy(x.reshape(-1)).cuda()
        v = buf[1:].view(*x.shape)
        assert v.data_ptr() % 16 != 0
        return v
    from udp_pose_b200._lib import UdpCodecError
    with pytest.raises(UdpCodecError):
        D.decode_batch(cfg, shifted(out), shifted(fl), c, s, want_idx=True)
    for k in ("preds", "maxvals", "preds_in", "idx"):
        assert torch.equal(a[k], b[k]), k


def test_decode_nonfinite_maps_inside_a_stream(impl, orc):
    """A NaN / inf map in the middle of a batch is re-done by the NaN-exact generic path inside the
    persistent kernel; its neighbours (prefetched while it ran) must be unaffected."""
    cfg = CONFIGS[1]
    rng = np.random.default_rng(71)
    N = 40
    out, fl = synth.heatmaps(cfg, N, rng)
    out[3, 5, 10, 7] = np.nan
    out[3, 6, 0, 0] = np.inf
    out[17, 0, 63, 47] = -np.inf
    fl[29, 16, 1, 1] = np.nan
    c, s, _ = synth.boxes(cfg, N, rng)
    with np.errstate(all="ignore"):
        a = orc.decode(cfg, out, fl, c, s)
    b = impl.decode(cfg, out, fl, c, s)
    assert np.array_equal(a["idx"], b["idx"])
    nan = np.isnan(a["maxvals"])                         # NaN payload/sign bits are not part of the contract
    assert np.array_equal(nan, np.isnan(b["maxvals"])) and nan.sum() == 2
    assert same_bits(a["maxvals"][~nan], b["maxvals"][~nan])
    assert np.array_equal(np.isnan(a["preds"]), np.isnan(b["preds"]))
    ok = ~np.isnan(a["preds"])
    assert np.abs(a["preds"][ok] - b["preds"][ok]).max() <= 1e-4


def test_full_size_batch_independence(impl):
    """At the bench's full size (16 384 persons, cfg 2: 20 GB of network outputs) every stage is
    person-independent: one launch over the whole batch must return exactly what eight launches over
    its 2 048-person chunks return (the small-N chunks are what the oracle tests pin), and the
    decoded keypoints of a map must not depend on which CTA of the persistent grid met it."""
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[2]
    N, C = 16384, 2048
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    rng = np.random.default_rng(2024)
    c, s, r = synth.boxes(cfg, N, rng, rotate=True)
    imgs = torch.from_numpy(synth.source_images(8, rng)).cuda()
    idx = np.arange(N) % 8
    j, v = synth.joints(cfg, N, rng)
    g = torch.Generator(device="cuda").manual_seed(5)
    out = torch.rand((N, cfg.channels, h, w), device="cuda", generator=g)
    fl = torch.rand((N, cfg.channels, h, w), device="cuda", generator=g)
    cd, sd, rd = (torch.from_numpy(a).cuda() for a in (c, s, r))
    M = D.warp_matrices(cd, sd, rd, (W, H))
    crop = D.crop_batch(D.SourceImages(imgs, idx), M, (W, H))
    jd, vd = torch.from_numpy(j).cuda(), torch.from_numpy(v).cuda()
    tgt, tw = D.encode_batch(cfg, jd, vd)
    dec = D.decode_batch(cfg, out, fl, cd, sd, want_idx=True)
    for lo in range(0, N, C):
        sl = slice(lo, lo + C)
        Mc = D.warp_matrices(cd[sl], sd[sl], rd[sl], (W, H))
        assert torch.equal(Mc, M[sl])
        assert torch.equal(D.crop_batch(D.SourceImages(imgs, idx[sl]), Mc, (W, H)), crop[sl])
        t2, w2 = D.encode_batch(cfg, jd[sl], vd[sl])
        assert torch.equal(t2, tgt[sl]) and torch.equal(w2, tw[sl])
        d2 = D.decode_batch(cfg, out[sl], fl[sl], cd[sl], sd[sl], want_idx=True)
        for k in ("preds", "maxvals", "preds_in", "idx"):
            assert torch.equal(d2[k], dec[k][sl]), k
