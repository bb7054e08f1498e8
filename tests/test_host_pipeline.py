"""HostPipeline (what bench.py times as `e2e`): pinned host buffers in and out, chunks flowing
through upload / compute / download streams over a ring of device slots.  The chunked,
overlapped pass must return exactly what one direct pass through the device API returns --
also when two passes are enqueued back to back without a synchronisation in between."""
import numpy as np
import pytest

from udp_pose_b200 import synth
from udp_pose_b200.configs import CONFIGS

pytestmark = pytest.mark.gpu


def _fill(hb, cfg, n, rng, n_images):
    import torch
    c, s, r = synth.boxes(cfg, n, rng, rotate=True)
    K = cfg.num_joints
    js = np.zeros((n, K, 3))
    js[..., 0] = c[:, None, 0] + rng.uniform(-0.6, 0.6, (n, K)) * s[:, None, 0] * 200.0
    js[..., 1] = c[:, None, 1] + rng.uniform(-0.6, 0.6, (n, K)) * s[:, None, 1] * 200.0
    vis = np.zeros((n, K, 3))
    vis[..., 0] = vis[..., 1] = (rng.uniform(size=(n, K)) < 0.8)
    out, fl = synth.heatmaps(cfg, n, rng)
    hb.images.copy_(torch.from_numpy(synth.source_images(n_images, rng)))
    hb.image_index[:] = rng.integers(0, n_images, n)
    for dst, src in ((hb.center, c), (hb.scale, s), (hb.rot, r), (hb.joints, js), (hb.joints_vis, vis),
                     (hb.out, out), (hb.out_flipped, fl)):
        dst.copy_(torch.from_numpy(np.ascontiguousarray(src)))


def _direct(hb, cfg):
    import torch
    from udp_pose_b200 import device_api as D
    imgs = hb.images.cuda()
    c, s, r = hb.center.cuda(), hb.scale.cuda(), hb.rot.cuda()
    M = D.warp_matrices(c, s, r, cfg.image_size)
    crop = D.crop_batch(D.SourceImages(imgs, hb.image_index), M, cfg.image_size)
    j3 = D.rotate_points_batch(hb.joints.cuda(), r, c, s, cfg.image_size)
    if cfg.target_type == "rsn":
        tgt, _ = D.encode_batch(cfg, j3, hb.joints_vis.cuda())
        tgt = tgt.reshape(hb.n, -1, *tgt.shape[-2:])
        tw = hb.joints_vis[:, :, 0:1].float()
    else:
        tgt, tw = D.encode_batch(cfg, j3, hb.joints_vis.cuda())
    dec = D.decode_batch(cfg, hb.out.cuda(), hb.out_flipped.cuda(), c, s, want_packed=True)
    torch.cuda.synchronize()
    return crop.cpu(), tgt.cpu(), tw.cpu(), dec["packed"].cpu()


@pytest.mark.parametrize("cid", [1, 2, 5])
def test_pipeline_matches_direct_calls(cid):
    import torch
    from udp_pose_b200.pipeline import HostBatch, HostPipeline
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(300 + cid)
    n, chunk, n_images = 333, 64, 5          # 6 chunks over 3 slots, ragged tail of 13
    hb = HostBatch(cfg, n, n_images)
    _fill(hb, cfg, n, rng, n_images)
    want = _direct(hb, cfg)
    pipe = HostPipeline(cfg, chunk, n_images)
    for t in (hb.crop, hb.target, hb.target_weight, hb.keypoints):
        t.fill_(0)
    pipe.run(hb)
    got = (hb.crop, hb.target, hb.target_weight, hb.keypoints)
    for a, b in zip(want, got):
        assert torch.equal(a, b)
    assert pipe.launches == len(pipe._schedule(n)) * (4 + 2)
    assert sum(m for _, m in pipe._schedule(2048)) == 2048 and pipe._schedule(2048)[0][1] == chunk // 4
    # two passes without a synchronisation in between (streaming callers), then one wait
    for t in got:
        t.fill_(0)
    pipe.run(hb, sync=False)
    pipe.run(hb, sync=False)
    pipe.synchronize()
    for a, b in zip(want, got):
        assert torch.equal(a, b)


def test_split_offsets_reads_the_windows_from_the_host_buffers_only():
    """cfg 2, split mode (the default): only the classification channels are uploaded; the decoder gathers its
    7x7 offset windows from the caller's pinned buffers.  The device copies' offset channels are poisoned with NaN
    before the run -- if the decoder read them, the coordinates would be NaN.  Both modes equal the direct pass, and
    the split mode moves about a third of the bytes towards the device."""
    import torch
    from udp_pose_b200.pipeline import HostBatch, HostPipeline
    cfg = CONFIGS[2]
    rng = np.random.default_rng(77)
    n, chunk, n_images = 200, 64, 3
    hb = HostBatch(cfg, n, n_images)
    _fill(hb, cfg, n, rng, n_images)
    want = _direct(hb, cfg)
    got = (hb.crop, hb.target, hb.target_weight, hb.keypoints)
    for split in (True, False):
        pipe = HostPipeline(cfg, chunk, n_images, split_offsets=split)
        assert pipe.split == split
        for sl in pipe.slots:
            sl["out"].fill_(float("nan"))
            sl["fl"].fill_(float("nan"))
        torch.cuda.synchronize()
        for t in got:
            t.fill_(0)
        pipe.run(hb)
        for a, b in zip(want, got):
            assert torch.equal(a, b)
        assert not torch.isnan(hb.keypoints).any()
    full, less = hb.h2d_bytes(), HostPipeline(cfg, chunk, n_images).h2d_bytes(hb)
    img = hb.images.numel()
    assert (less - img) < 0.36 * (full - img)


def test_decode_offset_split_entry_point_with_nan_maps():
    """udp_decode_offset_split on device tensors: classification channels in one pair of tensors, offsets in another
    (whose classification channels are poisoned); a map holding NaN takes the generic path, which must use the split
    sources too.  Bit-identical to udp_decode_offset on the complete tensors."""
    import torch
    from udp_pose_b200 import device_api as D
    cfg = CONFIGS[2]
    rng = np.random.default_rng(78)
    N, K = 9, cfg.num_joints
    out, fl = synth.heatmaps(cfg, N, rng)
    out[3, 3 * 5, 10, 10] = np.nan          # person 3, joint 5: NaN in the classification map
    c, s, _ = synth.boxes(cfg, N, rng)
    o, f = torch.from_numpy(out).cuda(), torch.from_numpy(fl).cuda()
    want = D.decode_offset(o, c, s, cfg.kpd, f, cfg.flip_pairs, want_idx=True)
    cls_o, cls_f, off_o, off_f = o.clone(), f.clone(), o.clone(), f.clone()
    for t in (cls_o, cls_f):
        t.view(N, K, 3, *t.shape[2:])[:, :, 1:] = float("nan")     # offsets must not be read from here
    for t in (off_o, off_f):
        t.view(N, K, 3, *t.shape[2:])[:, :, 0] = float("nan")      # nor classification maps from here
    got = D.decode_offset(cls_o, c, s, cfg.kpd, cls_f, cfg.flip_pairs, want_idx=True, off_ptrs=(off_o.data_ptr(), off_f.data_ptr()))
    for k in ("idx", "maxvals", "preds", "preds_in"):
        a, b = want[k].cpu().numpy(), got[k].cpu().numpy()
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), k
