"""GPU: the Python surface that INTEGRATION.md advertises and that had no direct test --
reference_api.RSNJointsEncoder (RSN/dataset/JointsDataset.py:193-220,276-291), reference_api.rsn_get_results
(RSN/exps/RSN18.coco.e1.se.36x8x132000_prm/test.py:183-196), install() / uninstall() on a stand-in module
triple, batched_loader(), the per-sample names inside spawned DataLoader workers, sharding.decode_sharded."""
import os
import random
import types

import numpy as np
import pytest

import parity
from parity import same_bits, TOL_HM, TOL_PX
from udp_pose_b200.configs import CONFIGS, RSN_SIGMAS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available()
    from udp_pose_b200 import _lib, reference_api
    _lib.load()
    return reference_api


# ------------------------------------------------------------------------------------------- RSN encoder
def test_rsn_joints_encoder_against_reference_goldens(R):
    cfg, g = CONFIGS[5], parity.load(5)
    (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
    enc = R.RSNJointsEncoder(cfg.num_joints, (H, W), (h, w))
    j, v = g["joints"], g["joints_vis"]
    for i in range(j.shape[0]):
        lab = enc.labels(j[i], v[i, :, 0])                        # the five scales at once (:214-220)
        assert lab.shape == (5, cfg.num_joints, h, w) and lab.dtype == np.float32
        err = np.abs(lab.astype(np.float64) - g["labels"][i])
        assert (err <= np.maximum(TOL_HM, np.spacing(np.abs(g["labels"][i])))).all()
        one = enc.generate_heatmap(j[i], v[i, :, 0:1], kernel=(9, 9))   # one scale, the reference's own signature (:276)
        assert same_bits(one, lab[2])
        # joint map + visibility cull (:203-211)
        j2, v2, Mi = enc.transform_joints(g["joints_src"][i], v[i], g["rot"][i], g["center"][i], g["scale"][i])
        assert np.array_equal(v2, g["rsn_vis"][i])
        assert np.abs(j2 - g["rsn_joints"][i]).max() <= TOL_PX
        np.testing.assert_allclose(Mi, g["rsn_minv"][i], rtol=1.2e-7, atol=1e-30)
    assert [R.RSN_KERNEL_SIGMA[k[0]] for k in enc.gaussian_kernels] == list(RSN_SIGMAS)


def test_rsn_get_results_against_reference_goldens(R):
    """prm decode glue: outputs/255 -> get_max_preds -> post -> transform_preds, with and without the flip pair."""
    cfg, g = CONFIGS[5], parity.load(5)
    preds, maxvals = R.rsn_get_results(g["dec_out"], g["dec_center"], g["dec_scale"], g["dec_flipped"], cfg.flip_pairs)
    assert preds.dtype == np.float64 and maxvals.dtype == np.float32
    assert same_bits(maxvals, g["maxvals"]) and np.abs(preds - g["preds"]).max() <= TOL_PX
    preds, maxvals = R.rsn_get_results(g["dec_out"], g["dec_center"], g["dec_scale"])
    assert same_bits(maxvals, g["maxvals_noflip"]) and np.abs(preds - g["preds_noflip"]).max() <= TOL_PX
    # the same through get_final_preds on pre-divided maps (what inference.py sees in the reference's prm test)
    p2, m2, _ = R.get_final_preds(cfg.as_cfg(), g["dec_out"] / np.float32(255.0), g["dec_center"], g["dec_scale"])
    assert same_bits(m2, maxvals) and np.abs(p2 - preds).max() <= TOL_PX


# ------------------------------------------------------------------------------------------- install()
def _stand_in_modules():
    """Module objects shaped like lib.core.inference, lib.utils.transforms, lib.dataset.JointsDataset,
    lib.core.evaluate, lib.nms.nms: every name install() touches, bound to sentinels."""
    def sentinel(name):
        def f(*a, **k):
            raise AssertionError("original %s called" % name)
        f.__name__ = name
        return f
    inf = types.ModuleType("core.inference")
    for n in ("get_max_preds", "post", "get_final_preds", "transform_preds"):
        setattr(inf, n, sentinel(n))
    tr = types.ModuleType("utils.transforms")
    for n in ("flip_back", "flip_back_offset", "fliplr_joints", "transform_preds"):
        setattr(tr, n, sentinel(n))
    jd = types.ModuleType("dataset.JointsDataset")
    for n in ("get_warpmatrix", "rotate_points", "fliplr_joints"):
        setattr(jd, n, sentinel(n))

    class JointsDataset:  # the attributes of JointsDataset.py:76-117, set the way a subclass does
        def __init__(self, cfg, is_train):
            c = cfg.as_cfg()
            self.is_train, self.num_joints, self.flip_pairs = is_train, cfg.num_joints, cfg.flip_pairs
            self.scale_factor, self.rotation_factor, self.flip = c.DATASET.SCALE_FACTOR, c.DATASET.ROT_FACTOR, c.DATASET.FLIP
            self.num_joints_half_body, self.prob_half_body = c.DATASET.NUM_JOINTS_HALF_BODY, c.DATASET.PROB_HALF_BODY
            self.color_rgb, self.data_format = True, "jpg"
            self.target_type = c.MODEL.TARGET_TYPE
            self.image_size, self.heatmap_size = np.array(c.MODEL.IMAGE_SIZE), np.array(c.MODEL.HEATMAP_SIZE)
            self.sigma, self.kpd = c.MODEL.SIGMA, c.LOSS.KPD
            self.use_different_joints_weight, self.joints_weight = False, 1
            self.cutout = self.hide_and_seek = None
            self.transform = None
            self.db = []

        def __len__(self):
            return len(self.db)

        generate_target = sentinel("generate_target")
    jd.JointsDataset = JointsDataset
    ev = types.ModuleType("core.evaluate")
    ev.accuracy = sentinel("accuracy")
    nm = types.ModuleType("nms.nms")
    nm.oks_nms, nm.soft_oks_nms = sentinel("oks_nms"), sentinel("soft_oks_nms")
    return inf, tr, jd, ev, nm


def test_install_rebinds_and_uninstall_restores(R):
    cfg, g = CONFIGS[1], parity.load(1)
    inf, tr, jd, ev, nm = _stand_in_modules()
    before = {m.__name__: dict(vars(m)) for m in (inf, tr, jd, ev, nm)}
    saved = R.install(inf, tr, jd, evaluate=ev, nms=nm)
    # decode side: the module-level names now run the CUDA codec
    avg = ((g["dec_out"] + tr.flip_back(g["dec_flipped"].copy(), cfg.flip_pairs)) * np.float32(0.5)).astype(np.float32)
    preds, maxvals, pin = inf.get_final_preds(cfg.as_cfg(), avg, g["dec_center"], g["dec_scale"])
    assert same_bits(maxvals, g["maxvals"]) and np.abs(preds - g["preds"]).max() <= TOL_PX
    coords, mv = inf.get_max_preds(avg)
    assert same_bits(mv, g["maxvals"])
    assert ev.accuracy is R.accuracy and nm.oks_nms is R.oks_nms and nm.soft_oks_nms is R.soft_oks_nms
    # encode side: per-sample names are NOT rebound by default (they would run in forked DataLoader workers) ...
    assert jd.get_warpmatrix is before[jd.__name__]["get_warpmatrix"] and tr.fliplr_joints is before[tr.__name__]["fliplr_joints"]
    assert jd.JointsDataset.generate_target is before[jd.__name__]["JointsDataset"].__dict__["generate_target"]
    # ... the batched methods are added instead, and work on an instance of the (stand-in) reference class
    ds = jd.JointsDataset(cfg, False)
    n = len(g["src_index"])
    ds.db = [dict(image="mem://%d" % i, joints_3d=g["joints_src"][i].copy(), joints_3d_vis=g["joints_vis"][i].copy(),
                  center=g["center"][i].copy(), scale=g["scale"][i].copy(), score=0.5) for i in range(n)]
    idx = [i for i in range(n) if g["rot"][i] == 0]
    inp, target, tw, metas = ds.get_batch(idx, [g["src_images"][g["src_index"][i]] for i in idx])
    for k, i in enumerate(idx):
        assert np.array_equal(inp[k].cpu().numpy(), g["crop"][i])
        assert np.abs(metas[k]["joints"][:, :2] - g["joints_rot"][i]).max() <= TOL_PX
    R.uninstall(saved, [inf, tr, jd, ev, nm])
    for m in (inf, tr, jd, ev, nm):
        for k, v in before[m.__name__].items():
            assert getattr(m, k) is v, (m.__name__, k)
    assert not hasattr(jd.JointsDataset, "get_batch")
    # per_sample_encode=True rebinds the per-sample names as well (num_workers=0 or spawn, see the docstring)
    saved = R.install(joints_dataset=jd, transforms=tr, per_sample_encode=True)
    M = jd.get_warpmatrix(0.0, g["center"][0] * 2.0, np.array(cfg.image_size) - 1.0, g["scale"][0])
    assert same_bits(M, g["warp_matrix"][0])
    t, w8 = jd.JointsDataset(cfg, False).generate_target(g["joints"][0], g["joints_vis"][0])
    assert same_bits(w8, g["target_weight"][0]) and np.abs(t - g["target"][0]).max() <= TOL_HM
    R.uninstall(saved, [tr, jd])
    assert jd.get_warpmatrix is before[jd.__name__]["get_warpmatrix"]


def _file_dataset(R, cfg, tmp_path, train):
    import cv2
    g = parity.load(cfg_id_of(cfg))
    ds = R.JointsDataset(cfg.as_cfg(), "", "train" if train else "val", train)
    ds.num_joints, ds.flip_pairs = cfg.num_joints, cfg.flip_pairs
    ds.upper_body_ids, ds.aspect_ratio = tuple(range(11)), cfg.image_size[0] / cfg.image_size[1]
    paths = []
    for i in range(2):
        p = os.path.join(str(tmp_path), "img%d.png" % i)
        cv2.imwrite(p, cv2.cvtColor(g["src_images"][i], cv2.COLOR_RGB2BGR))
        paths.append(p)
    n = len(g["center"])
    ds.db = [dict(image=paths[i % 2], joints_3d=g["joints_src"][i].copy(), joints_3d_vis=g["joints_vis"][i].copy(),
                  center=g["center"][i].copy(), scale=g["scale"][i].copy(), score=1) for i in range(n)]
    return ds, g


def cfg_id_of(cfg):
    return [k for k, v in CONFIGS.items() if v is cfg][0]


def test_batched_loader_equals_per_sample_getitem(R, tmp_path):
    """batched_loader (main process, one launch per stage and batch) against the per-sample __getitem__ under the
    same seeds, train mode: identical crops, targets and metas; the default-collated meta has the DataLoader's shape."""
    import torch
    cfg = CONFIGS[1]
    ds, g = _file_dataset(R, cfg, tmp_path, train=True)
    n = len(ds)
    np.random.seed(9); random.seed(9)
    single = [ds[i] for i in range(n)]
    np.random.seed(9); random.seed(9)
    batches = list(R.batched_loader(ds, 4))
    assert [b[0].shape[0] for b in batches] == [4, n - 4]
    k = 0
    for inp, target, tw, meta in batches:
        assert inp.is_cuda and target.is_cuda and isinstance(meta["joints"], torch.Tensor) and len(meta["image"]) == inp.shape[0]
        for b in range(inp.shape[0]):
            x, t, w8, m = single[k]
            assert np.array_equal(inp[b].cpu().numpy(), x)
            assert torch.equal(target[b].cpu(), t) and torch.equal(tw[b].cpu(), w8)
            assert np.array_equal(meta["joints"][b].numpy(), m["joints"]) and float(meta["rotation"][b]) == m["rotation"]
            k += 1
    assert k == n


def test_per_sample_names_in_spawned_dataloader_workers(R, tmp_path):
    """The reference's own DataLoader pattern (tools/test.py:122-128) with workers > 0: the per-sample
    __getitem__ needs the GPU in every worker, which works with multiprocessing_context='spawn' (a forked
    child of a CUDA parent cannot initialise CUDA -- the reason install() does not rebind these by default)."""
    import torch
    cfg = CONFIGS[1]
    ds, g = _file_dataset(R, cfg, tmp_path, train=False)
    n = len(g["src_index"])
    ds.db = ds.db[:n]
    want = [ds[i] for i in range(n)]
    dl = torch.utils.data.DataLoader(ds, batch_size=2, shuffle=False, num_workers=2, multiprocessing_context="spawn")
    got = list(dl)
    k = 0
    for inp, target, tw, meta in got:
        for b in range(inp.shape[0]):
            assert np.array_equal(inp[b].numpy(), want[k][0]) and torch.equal(target[b], want[k][1])
            k += 1
    assert k == n
    for i in range(n):
        if g["rot"][i] == 0:  # (the eval path has rotation 0: the golden crops of the unrotated boxes apply)
            assert np.array_equal(want[i][0], g["crop"][i])


def test_decode_sharded_single_process_is_decode_batch(R):
    import torch
    from udp_pose_b200 import device_api as D, sharding as S
    cfg, g = CONFIGS[2], parity.load(2)
    want = D.decode_batch(cfg, g["dec_out"], g["dec_flipped"], g["dec_center"], g["dec_scale"], want_packed=True)["packed"]
    got = S.decode_sharded(cfg, torch.from_numpy(g["dec_out"]).cuda(), torch.from_numpy(g["dec_flipped"]).cuda(),
                           g["dec_center"], g["dec_scale"])
    assert torch.equal(got, want) and got.shape == (g["dec_out"].shape[0], cfg.num_joints, 3)
    assert same_bits(got[..., 2:].cpu().numpy(), g["maxvals"])


def test_accuracy_ignores_thr_like_the_reference(R):
    """evaluate.py:63 never forwards thr to dist_acc: any thr gives the thr = 0.5 numbers (the golden ones)."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_accuracy_golden as MG
    gold = np.load(os.path.join(parity.GOLDEN, "accuracy.npz"))
    out, tgt = MG.inputs("cfg1")
    for thr in (0.5, 0.1, 0.9):
        acc, avg, cnt, pred = R.accuracy(out, tgt, thr=thr)
        assert np.array_equal(acc, gold["cfg1_acc"]) and float(avg) == float(gold["cfg1_avg"]) and cnt == int(gold["cfg1_cnt"])


@pytest.mark.parametrize("cid", [1, 4])
def test_pack_all_boxes_equals_the_reference_lines(R, cid):
    """lib/core/function.py:211-214, executed literally in numpy on float32 (COCO) and float64 (MPII) records."""
    from udp_pose_b200 import device_api as D, synth
    cfg = CONFIGS[cid]
    rng = np.random.default_rng(40 + cid)
    c, s, _ = synth.boxes(cfg, 257, rng)
    score = rng.uniform(0, 1, 257)
    all_boxes = np.zeros((257, 6))
    all_boxes[:, 0:2] = c[:, 0:2]
    all_boxes[:, 2:4] = s[:, 0:2]
    all_boxes[:, 4] = np.prod(s * 200, 1)
    all_boxes[:, 5] = score
    got = D.pack_all_boxes(c, s, score).cpu().numpy()
    assert got.dtype == np.float64 and same_bits(got, all_boxes)
    assert same_bits(D.pack_all_boxes(c, s).cpu().numpy()[:, :5], all_boxes[:, :5])
