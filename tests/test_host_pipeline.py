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
