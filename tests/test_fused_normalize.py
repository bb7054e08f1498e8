"""crop + ToTensor + Normalize (+ mirrored flip-test input) in one kernel.

CPU: the numpy restatement against torchvision's own output (tests/golden/normalize.npz, made by
tests/golden/make_normalize_golden.py).  GPU: the fused kernel against that restatement applied
to the oracle's bit-exact crop -- float32 outputs compared bit for bit."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from udp_pose_b200.configs import CONFIGS

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "normalize.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_oracle_normalize_matches_torchvision():
    ramp = np.arange(256, dtype=np.uint8).reshape(256, 1, 1).repeat(3, 2)
    assert np.array_equal(bits(O.to_tensor_normalize(ramp, GOLD["mean"], GOLD["std"])[:, :, 0]), bits(GOLD["lut"]))
    assert np.array_equal(bits(O.to_tensor_normalize(ramp, GOLD["mean2"], GOLD["std2"])[:, :, 0]), bits(GOLD["lut2"]))
    assert np.array_equal(bits(O.to_tensor_normalize(GOLD["crop"], GOLD["mean"], GOLD["std"])), bits(GOLD["out"]))
    assert np.array_equal(bits(O.to_tensor_normalize(GOLD["crop"], GOLD["mean2"], GOLD["std2"])), bits(GOLD["out2"]))
    x = O.to_tensor_normalize(GOLD["crop"], GOLD["mean"], GOLD["std"])[None]
    assert np.array_equal(O.flip_input(x)[0, :, :, 0], x[0, :, :, -1])


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["imagenet", "other", "awkward"])
def test_every_byte_value_matches_torchvision(which):
    """Identity warp of an image holding all 256 values in every channel: the fused output must be
    torchvision's value for that byte, bit for bit (the reciprocal + FMA correction is exact for
    the ImageNet constants; other constants are checked on the host per call and fall back to a
    true division -- 'awkward' are constants chosen without regard to that)."""
    import torch
    from udp_pose_b200 import device_api as D
    y, x, c = np.mgrid[0:32, 0:32, 0:3]
    img = ((y * 32 + x + 37 * c) % 256).astype(np.uint8)
    if which == "imagenet":
        mean, std, lut = GOLD["mean"], GOLD["std"], GOLD["lut"]
    elif which == "other":
        mean, std, lut = GOLD["mean2"], GOLD["std2"], GOLD["lut2"]
    else:
        mean, std = np.float32([0.1234567, 0.7654321, 0.3333333]), np.float32([0.0123457, 3.1415927, 0.9999999])
        ramp = np.arange(256, dtype=np.uint8).reshape(256, 1, 1).repeat(3, 2)
        lut = O.to_tensor_normalize(ramp, mean, std)[:, :, 0]
    M = torch.tensor([[[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]]], dtype=torch.float32, device="cuda")
    src = D.SourceImages([torch.from_numpy(img).cuda()], [0])
    out, fl = D.crop_normalize_batch(src, M, (32, 32), mean, std, want_flipped=True)
    got = out[0].cpu().numpy()
    want = np.stack([lut[ch][img[:, :, ch]] for ch in range(3)])
    assert np.array_equal(bits(got), bits(want))
    assert torch.equal(fl, out.flip(3))


@pytest.mark.gpu
@pytest.mark.parametrize("cid", [2, 3, 4])
def test_fused_crop_normalize_vs_oracle(cid):
    """Rotated / border-crossing boxes, images on the TMA path (stride % 16 == 0) and on the
    cp.async path (stride 1500): fused float32 CHW output == Normalize(ToTensor(oracle crop))."""
    import torch
    from udp_pose_b200 import device_api as D, synth
    cfg = CONFIGS[cid]
    W, H = cfg.image_size
    rng = np.random.default_rng(40 + cid)
    N = 24
    c, s, r = synth.boxes(cfg, N, rng, rotate=True)
    imgs = [synth.source_images(1, rng)[0], rng.integers(0, 256, (375, 500, 3), dtype=np.uint8)]
    c[1::2] *= np.array([500 / 640, 375 / 480], c.dtype)
    idx = np.arange(N) % 2
    src = D.SourceImages([torch.from_numpy(i).cuda() for i in imgs], idx)
    M = D.warp_matrices(c, s, r, cfg.image_size)
    out, fl = D.crop_normalize_batch(src, M, (W, H), want_flipped=True)
    only = D.crop_normalize_batch(src, M, (W, H))
    assert torch.equal(only, out) and torch.equal(fl, out.flip(3))
    u8 = D.crop_batch(src, M, (W, H)).cpu().numpy()
    Mh = M.cpu().numpy()
    got = out.cpu().numpy()
    for i in range(N):
        crop = O.warp_affine_u8(imgs[idx[i]], Mh[i], W, H)
        assert np.array_equal(crop, u8[i])
        assert np.array_equal(bits(got[i]), bits(O.to_tensor_normalize(crop, D.IMAGENET_MEAN, D.IMAGENET_STD))), i


@pytest.mark.gpu
def test_fused_crop_normalize_rejects_bad_shapes():
    import torch
    from udp_pose_b200 import device_api as D
    from udp_pose_b200._lib import UdpCodecError
    img = torch.zeros((64, 64, 3), dtype=torch.uint8, device="cuda")
    M = torch.tensor([[[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]]], dtype=torch.float32, device="cuda")
    src = D.SourceImages([img], [0])
    with pytest.raises(UdpCodecError):
        D.crop_normalize_batch(src, M, (48, 64))          # W not a multiple of 32
    with pytest.raises(UdpCodecError):
        D.crop_normalize_batch(src, M, (32, 32), std=(0.2, 0.0, 0.2))
