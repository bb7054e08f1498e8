"""The offset-vote decoder has one kernel per map shape (decode_offset.cu: teams of band-owning warps for 64-row maps
-- 64x48 with five 3-warp teams, 64x64 with four 4-warp teams --, the barrier-per-map pipe kernel for 96x72, the
shape-generic kernel for anything else) and, in the team kernel, its own non-finite rule (a band whose row-pass results
are not all finite runs its column pass with np.argmax's NaN rule).  Every one of them against the oracle, whose decode
is pinned to the reference (tests/test_oracle_vs_reference.py), on the same seeded maps: arg-max index and blurred
maximum bit-exact, coordinates bit-exact in float32 (offset mode is float32 end to end).  Sizes are chosen so that a CTA
walks many maps (the ring of raw slots wraps, teams drift apart, the vote mailbox is flushed several times)."""
import dataclasses

import numpy as np
import pytest

import parity
from udp_pose_b200 import synth
from udp_pose_b200.configs import CONFIGS

SHAPES = [(48, 64), (64, 64), (72, 96), (40, 56)]  # (w, h): team x5, team x4, pipe, generic


def _cfg(w, h):
    return dataclasses.replace(CONFIGS[2], name="offset_%dx%d" % (w, h), heatmap_size=(w, h), image_size=(4 * w, 4 * h))


@pytest.fixture(scope="module")
def orc():
    return parity.OracleImpl()


@pytest.fixture(scope="module")
def cuda():
    from test_gpu_parity import CudaImpl
    return CudaImpl()


def _compare(a, b):
    assert np.array_equal(a["idx"], b["idx"])
    assert np.array_equal(np.isnan(a["maxvals"]), np.isnan(b["maxvals"]))
    ok = ~np.isnan(a["maxvals"])
    assert parity.same_bits(a["maxvals"][ok], b["maxvals"][ok])
    fin = np.isfinite(a["preds"])
    assert np.array_equal(fin, np.isfinite(b["preds"]))
    assert parity.same_bits(a["preds"][fin].astype(np.float32), b["preds"][fin].astype(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("flip", [True, False])
def test_gpu_offset_decode_every_shape(cuda, orc, shape, flip):
    cfg = _cfg(*shape)
    N = 700 if shape[0] * shape[1] <= 4096 else 300  # 148 CTAs x (4 or 5) teams: ~16 maps per team at 64x48
    rng = np.random.default_rng(9100 + shape[0] + (1 if flip else 0))
    out, fl = synth.heatmaps(cfg, N, rng)
    c, s, _ = synth.boxes(cfg, N, rng)
    a = orc.decode(cfg, out, fl if flip else None, c, s)
    b = cuda.decode(cfg, out, fl if flip else None, c, s)
    _compare(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", SHAPES)
def test_gpu_offset_decode_nonfinite_maps(cuda, orc, shape):
    """NaN / +inf / -inf in the classification maps, in the map itself or in its mirrored partner, at band borders, in
    the first and in the last row: np.argmax on the blurred map (the first NaN wins), maxvals NaN where the oracle's are."""
    cfg = _cfg(*shape)
    w, h = cfg.heatmap_size
    K, N = cfg.num_joints, 40
    rng = np.random.default_rng(9200 + w)
    out, fl = synth.heatmaps(cfg, N, rng)
    cls = lambda t, n, k: t[n, 3 * k]  # classification channel of joint k
    cls(out, 1, 2)[h // 2, w // 3] = np.nan
    cls(out, 4, 0)[0, 0] = np.inf
    cls(out, 7, K - 1)[h - 1, w - 1] = -np.inf
    cls(fl, 9, 3)[5, 5] = np.nan
    cls(out, 11, 5)[h // 3, 15] = np.nan          # last column of the first 16-column band
    cls(out, 12, 6)[h - 1, 16] = np.inf           # first column of the second band, last row
    cls(out, 13, 7)[0, w - 1] = np.nan
    cls(out, 14, 8)[10, 20] = np.inf
    cls(out, 14, 8)[40, 30] = -np.inf             # +inf and -inf under one blur window further down: NaN appears
    cls(out, 15, 9)[:, :] = -np.inf
    cls(fl, 15, 9)[:, :] = -np.inf
    cls(out, 16, 10)[:, :] = np.nan
    c, s, _ = synth.boxes(cfg, N, rng)
    with np.errstate(all="ignore"):
        a = orc.decode(cfg, out, fl, c, s)
    b = cuda.decode(cfg, out, fl, c, s)
    _compare(a, b)
    with np.errstate(all="ignore"):
        a = orc.decode(cfg, out, None, c, s)
    b = cuda.decode(cfg, out, None, c, s)
    _compare(a, b)


@pytest.mark.gpu
def test_gpu_offset_team_and_pipe_kernels_agree(cuda, monkeypatch):
    """UDP_OFFSET_PIPE=1 selects the previous (barrier-per-map) kernel: same bits."""
    cfg = CONFIGS[2]
    rng = np.random.default_rng(9300)
    out, fl = synth.heatmaps(cfg, 500, rng)
    c, s, _ = synth.boxes(cfg, 500, rng)
    a = cuda.decode(cfg, out, fl, c, s)
    monkeypatch.setenv("UDP_OFFSET_PIPE", "1")
    b = cuda.decode(cfg, out, fl, c, s)
    for k in ("idx", "maxvals", "preds", "preds_in"):
        assert parity.same_bits(a[k], b[k]), k


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(48, 64), (64, 64)])
def test_gpu_offset_team_kernel_sizes_and_repeats(shape):
    """Batch sizes around the team kernel's corners (fewer maps than teams, a partial last group, one map more or less
    than CTAs x teams, several ring revolutions), each decoded three times: the hand-off between teams is dynamic, the
    result must not be.  Reference: the barrier-per-map kernel (UDP_OFFSET_PIPE=1); NaN payloads aside, same bits."""
    import os
    import torch
    from udp_pose_b200 import device_api as D
    w, h = shape
    cfg = _cfg(w, h)
    old = os.environ.get("UDP_OFFSET_PIPE")
    try:
        for N in (1, 3, 17, 44, 45, 175, 871):
            rng = np.random.default_rng(N * 7 + w)
            out, fl = synth.heatmaps(cfg, N, rng)
            c, s, _ = synth.boxes(cfg, N, rng)
            if N >= 17:
                out[N // 2, 3 * 2, h // 2, w // 3] = np.nan
                out[N // 3, 0, 0, 0] = np.inf
                fl[N - 1, 3 * 5, h - 1, w - 1] = -np.inf
            o, f = torch.from_numpy(out).cuda(), torch.from_numpy(fl).cuda()
            ct, st = torch.from_numpy(c).cuda(), torch.from_numpy(s).cuda()
            for flip in (True, False):
                os.environ["UDP_OFFSET_PIPE"] = "1"
                ref = {k: v.clone() for k, v in D.decode_batch(cfg, o, f if flip else None, ct, st, want_packed=True).items() if v is not None}
                os.environ["UDP_OFFSET_PIPE"] = "0"
                for _ in range(3):
                    r = D.decode_batch(cfg, o, f if flip else None, ct, st, want_packed=True)
                    for k, a in ref.items():
                        b = r[k]
                        if a.dtype == torch.float32:
                            na, nb = torch.isnan(a), torch.isnan(b)
                            assert torch.equal(na, nb), (N, flip, k)
                            assert torch.equal(a[~na].view(torch.int32), b[~nb].view(torch.int32)), (N, flip, k)
                        else:
                            assert torch.equal(a, b), (N, flip, k)
    finally:
        if old is None:
            os.environ.pop("UDP_OFFSET_PIPE", None)
        else:
            os.environ["UDP_OFFSET_PIPE"] = old
