"""CUDA-graph replay of a codec pass (udp_pose_b200/graphs.py): the replayed pass must produce exactly
what direct launches produce, also after the static inputs were overwritten with another batch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_captured_pass_matches_direct_calls():
    import torch
    from udp_pose_b200 import device_api as D, synth
    from udp_pose_b200.configs import CONFIGS
    from udp_pose_b200.graphs import CapturedPass
    cfg = CONFIGS[2]
    n = 96
    W, H = cfg.image_size

    def batch(seed):
        rng = np.random.default_rng(seed)
        c, s, r = synth.boxes(cfg, n, rng, rotate=True)
        j, v = synth.joints(cfg, n, rng)
        out, fl = synth.heatmaps(cfg, n, rng)
        return dict(c=c, s=s, r=r, j=j, v=v, out=out, fl=fl)

    imgs = torch.from_numpy(synth.source_images(4, np.random.default_rng(0))).cuda()
    src = D.SourceImages(imgs, np.arange(n) % 4)
    b0 = batch(1)
    st = {k: torch.from_numpy(np.ascontiguousarray(x)).cuda() for k, x in b0.items()}

    def run():
        M = D.warp_matrices(st["c"], st["s"], st["r"], cfg.image_size)
        crop = D.crop_batch(src, M, cfg.image_size)
        jc = D.rotate_points_batch(st["j"], st["r"], st["c"], st["s"], cfg.image_size)
        target, tw = D.encode_batch(cfg, jc, st["v"])
        r = D.decode_batch(cfg, st["out"], st["fl"], st["c"], st["s"], want_packed=True)
        return crop, target, tw, r["packed"]

    cap = CapturedPass(run)
    for seed in (1, 2):
        b = batch(seed)
        for k, x in b.items():
            st[k].copy_(torch.from_numpy(np.ascontiguousarray(x)))
        want = [t.clone() for t in run()]
        torch.cuda.synchronize()
        got = cap.replay()
        torch.cuda.synchronize()
        for a, g in zip(want, got):
            assert torch.equal(a, g)
