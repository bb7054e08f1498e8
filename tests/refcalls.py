"""Batched calls into the UNMODIFIED reference (container only; needs /root/reference).

Used by tests/golden/make_golden.py to produce the committed fixtures and by
tests/test_oracle_vs_reference.py.  Each helper loops the reference's per-person function
exactly the way its callers do (JointsDataset.__getitem__, validate()).
"""
import numpy as np

from oracle import ref_import as R


def _ds(cfg):
    tt = "offset" if cfg.target_type == "offset" else "gaussian"
    c = R.make_cfg(tt, cfg.image_size, cfg.heatmap_size, cfg.sigma, kpd=cfg.kpd,
                   post_process=cfg.post_process)
    return R.make_dataset(c, cfg.num_joints), c


def warp_matrices(cfg, center, scale, rot):
    jd = R.hrnet().joints_dataset
    img = np.array(cfg.image_size)
    return np.stack([jd.get_warpmatrix(float(rot[i]), center[i] * 2.0, img - 1.0, scale[i])
                     for i in range(len(rot))])


def rotate_points(cfg, joints, center, scale, rot, do_clip=False):
    jd = R.hrnet().joints_dataset
    img = np.array(cfg.image_size)
    return np.stack([jd.rotate_points(joints[i, :, 0:2].copy(), float(rot[i]), center[i], img,
                                      scale[i], do_clip) for i in range(len(rot))])


def warp_affine(cfg, images, img_index, M):
    import cv2
    W, H = cfg.image_size
    return np.stack([cv2.warpAffine(images[img_index[i]], M[i], (int(W), int(H)),
                                    flags=cv2.WARP_INVERSE_MAP | cv2.INTER_LINEAR)
                     for i in range(len(M))])


def generate_target(cfg, joints, vis, joints_weight=None):
    ds, _ = _ds(cfg)
    if joints_weight is not None:
        ds.use_different_joints_weight = True
        ds.joints_weight = np.asarray(joints_weight, np.float32).reshape(-1, 1)
    t, tw = [], []
    for i in range(joints.shape[0]):
        a, b = ds.generate_target(joints[i].copy(), vis[i].copy())
        t.append(a)
        tw.append(b)
    return np.stack(t), np.stack(tw)


def rsn_labels(cfg, joints, vis):
    import torch
    W, H = cfg.image_size
    w, h = cfg.heatmap_size
    ds = R.make_rsn_dataset(cfg.num_joints, (H, W), (h, w))
    out = []
    for i in range(joints.shape[0]):
        valid = torch.from_numpy(vis[i]).float()
        labels = np.zeros((len(ds.gaussian_kernels), cfg.num_joints, h, w))
        for s, kern in enumerate(ds.gaussian_kernels):
            labels[s] = ds.generate_heatmap(joints[i], valid[:, 0:1], kernel=kern)
        out.append(torch.from_numpy(labels).float().numpy())
    return np.stack(out)


def rsn_transform_joints(cfg, joints, vis, center, scale, rot):
    """RSN/dataset/JointsDataset.py:195,203-211."""
    rs = R.rsn()
    jd, tr = rs.joints_dataset, rs.transforms
    W, H = cfg.image_size
    joints = joints.copy()
    vis = vis.copy()
    Ms = []
    for i in range(joints.shape[0]):
        Minv = jd.get_warpmatrix_inverse(float(rot[i]), center[i] * 2.0,
                                         np.array([W, H], dtype=float) - 1.0, scale[i])
        Ms.append(Minv)
        for k in range(cfg.num_joints):
            if vis[i, k, 0] > 0:
                joints[i, k, 0:2] = tr.affine_transform(joints[i, k, 0:2], Minv)
                if joints[i, k, 0] < 0 or joints[i, k, 0] > W - 1 or joints[i, k, 1] < 0 \
                        or joints[i, k, 1] > H - 1:
                    vis[i, k, 0] = 0
    return joints, vis, np.stack(Ms)


def flip_average(cfg, out, flipped):
    """function.py:163-171: flip_back(_offset) on a copy, then (a+b)*0.5 in float32."""
    tr = R.hrnet().transforms
    f = flipped.copy()
    if cfg.target_type == "offset":
        fb = tr.flip_back_offset(f, cfg.flip_pairs)
    else:
        fb = tr.flip_back(f, cfg.flip_pairs)
    return ((out + fb.copy()) * np.float32(0.5)).astype(np.float32)


def final_preds(cfg, hm, center, scale, post_process=None):
    """get_final_preds on a copy (post() mutates its input). RSN: /255.0 first (prm test.py:185)."""
    inf = R.hrnet().inference
    _, c = _ds(cfg)
    if post_process is not None:
        c.TEST.POST_PROCESS = post_process
    x = hm.copy()
    if cfg.pre_div:
        x = x / cfg.pre_div
    with np.errstate(all="ignore"):
        preds, maxvals, pin = inf.get_final_preds(c, x, center, scale)
    return preds, maxvals, pin
