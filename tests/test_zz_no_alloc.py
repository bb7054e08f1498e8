"""udp_accuracy_pck_noalloc -- the allocation-free variant of udp_accuracy_pck (include/udp_codec.h; no
counterpart in the reference): its [K,N] float64 `dists`
workspace is a static device array inside the library, so the call itself allocates nothing.

GPU: called through the C ABI (ctypes, raw device pointers) on two calls in a row -- the second with other
inputs and a smaller batch, so the shared workspace still holds the first call's values -- and compared with
the existing entry point on the same inputs: acc / avg_acc / cnt / pred bit-exact (the tolerance of
tests/test_accuracy.py).  Requests beyond UDP_PCK_NOALLOC_MAX_DISTS are rejected with UDP_ESCRATCH.

This file sorts last on purpose (a failure here must not hide another test);
it passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

from udp_pose_b200.configs import CONFIGS

UDP_ESCRATCH = -2
MAX_DISTS = 512 * 1024  # UDP_PCK_NOALLOC_MAX_DISTS


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _both(L, torch, out, tgt, ch_stride, thr=0.5):
    """(existing entry point, no-alloc entry point) results on the same device tensors."""
    N, Cc, h, w = out.shape
    K = Cc // ch_stride
    dev = out.device
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    res = []
    for noalloc in (False, True):
        pred = torch.full((N, K, 2), -7.0, dtype=torch.float32, device=dev)
        acc = torch.full((K + 1,), -7.0, dtype=torch.float64, device=dev)
        avg = torch.full((1,), -7.0, dtype=torch.float64, device=dev)
        cnt = torch.full((1,), -7, dtype=torch.int32, device=dev)
        if noalloc:
            rc = L.udp_accuracy_pck_noalloc(_vp(out), _vp(tgt), ch_stride, K, w, h, N, thr, _vp(pred), _vp(acc), _vp(avg),
                                            _vp(cnt), stream)
        else:
            dists = torch.empty((K, N), dtype=torch.float64, device=dev)
            rc = L.udp_accuracy_pck(_vp(out), _vp(tgt), ch_stride, K, w, h, N, thr, _vp(pred), _vp(dists), _vp(acc), _vp(avg),
                                    _vp(cnt), stream)
        assert rc == 0, rc
        torch.cuda.synchronize()
        res.append((acc.cpu().numpy(), avg.cpu().numpy(), cnt.cpu().numpy(), pred.cpu().numpy()))
    return res


@pytest.mark.gpu
def test_noalloc_matches_the_existing_entry_point_on_two_calls_in_a_row():
    import torch
    from udp_pose_b200 import _lib, synth
    assert torch.cuda.is_available()
    L = _lib.load()
    # call 1: cfg 1 (17 joints, 64x48), N = 96; call 2: the offset mode's [:, ::3] views (ch_stride 3), N = 40 --
    # fewer (joint, person) pairs than call 1 left in the workspace, other values, other row length
    cases = [(CONFIGS[1], 96, 1, 11), (CONFIGS[2], 40, 3, 12), (CONFIGS[4], 64, 1, 13)]
    for cfg, N, cs, seed in cases:
        out, _ = synth.heatmaps(cfg, N, np.random.default_rng(seed), flip=False)
        tgt, _ = synth.heatmaps(cfg, N, np.random.default_rng(seed + 100), flip=False)
        tgt[::5] = 0.0                      # target peaks that are not > 1: distance -1, joint ignored for that person
        o, t = torch.from_numpy(out).cuda(), torch.from_numpy(tgt).cuda()
        (a0, v0, c0, p0), (a1, v1, c1, p1) = _both(L, torch, o, t, cs)
        assert np.array_equal(a0, a1) and np.array_equal(v0, v1) and np.array_equal(c0, c1) and np.array_equal(p0, p1)
        assert int(c1[0]) > 0 and (a1[1:] >= 0).any()   # a non-trivial answer, not the -7 fill
    # the same inputs twice in a row give the same answer (the workspace's previous contents do not leak)
    (_, _, _, _), (a2, v2, c2, p2) = _both(L, torch, o, t, cs)
    assert np.array_equal(a1, a2) and np.array_equal(v1, v2) and np.array_equal(c1, c2) and np.array_equal(p1, p2)


@pytest.mark.gpu
def test_noalloc_rejects_what_does_not_fit_and_takes_the_largest_that_does():
    import torch
    from udp_pose_b200 import _lib
    L = _lib.load()
    K, h, w = 32, 2, 2
    N_max = MAX_DISTS // K
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(3)
    out = torch.rand((N_max, K, h, w), device="cuda", generator=g) * 4
    tgt = torch.rand((N_max, K, h, w), device="cuda", generator=g) * 4
    pred = torch.empty((N_max, K, 2), dtype=torch.float32, device="cuda")
    acc = torch.full((K + 1,), -7.0, dtype=torch.float64, device="cuda")
    avg = torch.empty((1,), dtype=torch.float64, device="cuda")
    cnt = torch.empty((1,), dtype=torch.int32, device="cuda")
    # one person too many: a clear error, nothing launched (acc keeps its fill)
    rc = L.udp_accuracy_pck_noalloc(_vp(out), _vp(tgt), 1, K, w, h, N_max + 1, 0.5, _vp(pred), _vp(acc), _vp(avg), _vp(cnt), stream)
    torch.cuda.synchronize()
    assert rc == UDP_ESCRATCH and float(acc[0]) == -7.0
    # exactly the maximum: accepted, and equal to the existing entry point
    (a0, v0, c0, p0), (a1, v1, c1, p1) = _both(L, torch, out, tgt, 1)
    assert np.array_equal(a0, a1) and np.array_equal(c0, c1) and np.array_equal(p0, p1)


def test_noalloc_size_check_precedes_every_cuda_call():
    """No GPU needed: the limit is checked before the library touches CUDA (dummy, never dereferenced pointers)."""
    from udp_pose_b200 import _lib
    L = _lib.load()
    fake = C.c_void_p(0x1000)
    assert L.udp_accuracy_pck_noalloc(fake, fake, 1, 17, 48, 64, MAX_DISTS // 17 + 1, 0.5, fake, fake, fake, fake, None) == UDP_ESCRATCH
    assert L.udp_accuracy_pck_noalloc(fake, fake, 1, 33, 48, 64, 1, 0.5, fake, fake, fake, fake, None) == -1  # K > 32: UDP_EINVAL
