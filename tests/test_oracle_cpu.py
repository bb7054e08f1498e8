"""CPU: the oracle's OpenCV restatements against the installed cv2 (bit-exact), so the
third-party arithmetic the reference delegates to is pinned on whatever host runs the tests
(re-verify on the GPU box's host)."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from oracle import oracle as O  # noqa: E402


def test_gauss_kernel_tables():
    k7 = cv2.getGaussianKernel(7, 0, cv2.CV_32F).ravel()
    k15 = cv2.getGaussianKernel(15, 0, cv2.CV_32F).ravel()
    import ctypes as C
    for n, ref in ((7, k7), (15, k15)):
        k = np.zeros(15, np.float32)
        O.lib().udp_o_gauss_kernel(n, k.ctypes.data_as(C.c_void_p))
        assert np.array_equal(k[:n].view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("n", [7, 15])
@pytest.mark.parametrize("shape", [(64, 48), (96, 72), (64, 64), (12, 16)])
def test_blur_bit_exact(n, shape):
    # NOTE: the restated summation order is what cv2's AVX2 path does for 8-wide column
    # groups; it holds for every width that is a multiple of 8 (all reference heatmap widths:
    # 48, 64, 72).  For other widths cv2's scalar tail columns use another order (<= 2.4e-7).
    if min(shape) < n // 2 + 1:
        pytest.skip("map smaller than the kernel radius")
    rng = np.random.default_rng(n * 1000 + shape[0])
    for t in range(6):
        a = (rng.uniform(0, 1, shape) if t % 2 == 0 else rng.standard_normal(shape) * 3).astype(np.float32)
        ref = cv2.GaussianBlur(a, (n, n), 0)
        got = O.gaussian_blur(a, n)
        assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))


def test_warp_affine_bit_exact():
    rng = np.random.default_rng(7)
    total = 0
    for t in range(12):
        hs, ws = int(rng.integers(100, 400)), int(rng.integers(100, 500))
        img = rng.integers(0, 256, (hs, ws, 3), dtype=np.uint8)
        W, H = [(192, 256), (288, 384), (256, 256)][t % 3]
        c = np.array([rng.uniform(-20, ws + 20), rng.uniform(-20, hs + 20)], np.float32)
        bh = rng.uniform(30, 1.3 * hs)
        s = np.array([bh * W / H, bh], np.float32) / 200
        r = float(np.clip(rng.standard_normal() * 45, -90, 90)) if t % 2 else 0.0
        M = O.warp_matrices([r], c[None], s[None], W, H)[0]
        ref = cv2.warpAffine(img, M, (W, H), flags=cv2.WARP_INVERSE_MAP | cv2.INTER_LINEAR)
        got = O.warp_affine_u8(img, M, W, H)
        assert np.array_equal(ref, got)
        total += ref.size
    assert total > 1e6
