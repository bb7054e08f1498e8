"""Deterministic synthetic workloads for the codec.

Host-side numpy generators shared by tests/, tests/golden/make_golden.py and bench.py so
that the oracle, the reference and the CUDA path all see identical inputs.  Nothing here is
on the product path.  seed convention: 1000 * cfg_id + trial.
"""
import numpy as np

SRC_W, SRC_H = 640, 480


def boxes(cfg, N, rng, rotate=False):
    """Person boxes in a 640x480 source image.  Returns center [N,2], scale [N,2] (cfg.cs_dtype,
    scale = box/200*1.25 as lib/dataset/coco.py:223-227) and rot [N] float64 degrees."""
    W, H = cfg.image_size
    bh = rng.uniform(0.25, 1.1, N) * SRC_H
    bw = bh * W / H
    c = rng.uniform(0.1, 0.9, (N, 2)) * np.array([SRC_W, SRC_H])
    s = np.stack([bw, bh], 1) / 200.0 * 1.25
    dt = np.dtype(cfg.cs_dtype)
    if rotate:
        rf = cfg.rot_factor
        r = np.clip(rng.standard_normal(N) * rf, -2 * rf, 2 * rf)  # JointsDataset.py:217-218
        r = np.where(rng.uniform(size=N) <= 0.6, r, 0.0)
    else:
        r = np.zeros(N)
    return c.astype(dt), s.astype(dt), r.astype(np.float64)


def source_images(P, rng, h=SRC_H, w=SRC_W):
    """uint8 [P,h,w,3]: iid noise blended 50/50 with a smooth gradient."""
    noise = rng.integers(0, 256, (P, h, w, 3)).astype(np.float32)
    gy, gx = np.mgrid[0:h, 0:w].astype(np.float32)
    grad = np.stack([gx / (w - 1) * 255, gy / (h - 1) * 255, (gx + gy) / (w + h - 2) * 255], -1)
    return np.clip(0.5 * noise + 0.5 * grad[None], 0, 255).astype(np.uint8)


def joints(cfg, N, rng, rsn=False):
    """joints [N,K,3] float64 in crop coordinates (about 17% outside), vis [N,K,3] float64."""
    W, H = cfg.image_size
    K = cfg.num_joints
    j = np.zeros((N, K, 3))
    j[..., 0] = rng.uniform(-0.1, 1.1, (N, K)) * (W - 1)
    j[..., 1] = rng.uniform(-0.1, 1.1, (N, K)) * (H - 1)
    if rsn:
        v = rng.integers(0, 3, (N, K)).astype(np.float64)
    else:
        v = (rng.uniform(size=(N, K)) < 0.8).astype(np.float64)
    vis = np.zeros((N, K, 3))
    vis[..., 0] = v
    vis[..., 1] = v
    return j, vis


def _flip_of(maps, cfg):
    """What a model would emit for the mirrored input: x-mirrored, pair-swapped
    (and x-offset negated in offset mode) -- the inverse of flip_back*."""
    K = cfg.num_joints
    perm = np.arange(K)
    for a, b in cfg.flip_pairs:
        perm[a], perm[b] = b, a
    if cfg.target_type == "offset":
        N, C3, h, w = maps.shape
        m = maps.reshape(N, K, 3, h, w)[:, perm][..., ::-1].copy()
        m[:, :, 1] = -m[:, :, 1]
        return m.reshape(N, C3, h, w)
    return maps[:, perm][..., ::-1].copy()


def heatmaps(cfg, N, rng, flip=True):
    """Realistic decode inputs.  Returns (out, flipped or None), float32 [N,C,h,w]."""
    w, h = cfg.heatmap_size
    K = cfg.num_joints
    mu = np.stack([rng.uniform(-2, w + 1, (N, K)), rng.uniform(-2, h + 1, (N, K))], -1)
    amp = rng.uniform(0.2, 1.0, (N, K))
    xs = np.arange(w)[None, None, None, :]
    ys = np.arange(h)[None, None, :, None]
    d2 = (xs - mu[..., 0, None, None]) ** 2 + (ys - mu[..., 1, None, None]) ** 2
    if cfg.target_type == "offset":
        cls = amp[..., None, None] * np.exp(-d2 / (2 * 2.0 ** 2))
        keep = (d2 / cfg.kpd ** 2) <= 1
        ox = np.where(keep, (mu[..., 0, None, None] - xs) / cfg.kpd, 0.0)
        oy = np.where(keep, (mu[..., 1, None, None] - ys) / cfg.kpd, 0.0)
        clean = np.stack([cls, ox, oy], 2).reshape(N, 3 * K, h, w)

        def noisy():
            nz = np.empty((N, K, 3, h, w))
            nz[:, :, 0] = rng.uniform(-0.02, 0.02, (N, K, h, w))
            nz[:, :, 1:] = rng.standard_normal((N, K, 2, h, w)) * 0.05
            return nz.reshape(N, 3 * K, h, w)
    else:
        sig = float(cfg.sigma)
        clean = amp[..., None, None] * np.exp(-d2 / (2 * sig ** 2))

        def noisy():
            return rng.uniform(-0.02, 0.02, (N, K, h, w))
    mul = 255.0 if cfg.pre_div else 1.0
    out = ((clean + noisy()) * mul).astype(np.float32)
    if not flip:
        return out, None
    fl = (_flip_of(clean, cfg) + noisy()) * mul
    return out, fl.astype(np.float32)


def adversarial_heatmaps(cfg, rng):
    """Small adversarial parity set for the gaussian decoder: float32 [M,K,h,w].
    Rows: U(0,1) noise, N(0,1) noise, all-negative, exact ties, border/corner peaks,
    single non-zero pixel.  (Constant maps -> NaN in the reference, tested separately.)"""
    w, h = cfg.heatmap_size
    K = cfg.num_joints
    rows = []
    rows.append(rng.uniform(0, 1, (K, h, w)))
    rows.append(rng.standard_normal((K, h, w)))
    rows.append(-rng.uniform(0.1, 1, (K, h, w)))
    t = rng.uniform(0, 0.5, (K, h, w))
    for k in range(K):
        p = rng.integers(0, h * w, 2)
        t[k].reshape(-1)[p] = 0.75
    rows.append(t)
    b = rng.uniform(0, 0.05, (K, h, w))
    spots = [(0, 0), (0, w - 1), (h - 1, 0), (h - 1, w - 1), (0, w // 2), (h - 1, w // 2),
             (h // 2, 0), (h // 2, w - 1)]
    for k in range(K):
        y, x = spots[k % len(spots)]
        b[k, y, x] = 1.0
    rows.append(b)
    z = np.zeros((K, h, w))
    for k in range(K):
        z[k, rng.integers(0, h), rng.integers(0, w)] = 0.9
    rows.append(z)
    return np.stack(rows).astype(np.float32)


def nms_persons(images, max_per_image, K, rng, exact=False):
    """Pose results for the COCO rescoring + OKS-NMS step, in the layout validate()
    builds (lib/core/function.py:208-214): preds f32 [N,K,3] (x, y, maxval), all_boxes f64 [N,6]
    (cx, cy, sx, sy, area, box score), image id int64 [N], rows shuffled across images.  Every image
    has n ~ U{1..max_per_image} persons (exactly max_per_image with exact=True) drawn around n/4 base
    poses with jitters from 0.5 % to 25 % of the body size, so near-duplicates (suppressed) and
    distinct poses (kept) both occur."""
    n_img = np.full(images, max_per_image) if exact else rng.integers(1, max_per_image + 1, images)
    img = np.repeat(np.arange(images, dtype=np.int64), n_img)
    N = len(img)
    start = np.repeat(np.cumsum(n_img) - n_img, n_img)
    nbase = np.maximum(1, np.repeat(n_img, n_img) // 4)
    b = start + rng.integers(0, 1 << 30, N) % nbase                 # row that acts as this person's base pose
    centre = rng.uniform(40, 600, (N, 1, 2))
    shape = rng.normal(0, 40, (N, K, 2))
    size = rng.uniform(60, 260, N)
    jitter = rng.choice([0.5, 2.0, 6.0, 25.0], N) * size[b] / 100.0
    xy = centre[b] + shape[b] + rng.normal(0, 1, (N, K, 2)) * jitter[:, None, None]
    mv = rng.uniform(0.05, 1.0, (N, K, 1)) * (rng.uniform(size=(N, K, 1)) > 0.1)
    preds = np.concatenate([xy, mv], 2).astype(np.float32)
    s = np.stack([size[b] * 0.75, size[b]], 1) / 200.0 * rng.uniform(0.9, 1.1, (N, 1))
    boxes = np.concatenate([xy.mean(1), s, np.prod(s * 200.0, 1, keepdims=True), rng.uniform(0.05, 1.0, (N, 1))], 1)
    perm = rng.permutation(N)
    return preds[perm], boxes[perm].astype(np.float64), img[perm]


def adversarial_offset_maps(cfg, rng):
    """Adversarial parity set for the offset-voting decoder (inference.py:156-174): float32 [M,3K,h,w].
    The arg-max is taken on the 15x15-BLURRED classification map, so the rows stress exactly that:
    0 U(0,1) noise, 1 N(0,1) noise, 2 all-negative, 3 all-zero classification maps (maxval 0: mask false),
    4 two equal isolated pixels (the blurred map has two bit-identical maxima: first index must win),
    5 peaks on every border / corner (the 15x15 window and the 7x7 offset windows reflect),
    6 a single pixel, 7 a plateau of equal values (ties between neighbours after the blur).
    Offset channels: N(0, 0.3) noise (so a wrong window or a wrong reflection changes the vote)."""
    w, h = cfg.heatmap_size
    K = cfg.num_joints
    cls = []
    cls.append(rng.uniform(0, 1, (K, h, w)))
    cls.append(rng.standard_normal((K, h, w)))
    cls.append(-rng.uniform(0.1, 1, (K, h, w)))
    cls.append(np.zeros((K, h, w)))
    t = np.zeros((K, h, w))
    for k in range(K):
        y0, x0 = rng.integers(8, h // 2 - 8), rng.integers(8, w - 8)
        y1, x1 = rng.integers(h // 2 + 8, h - 8), rng.integers(8, w - 8)
        t[k, y0, x0] = t[k, y1, x1] = 0.75
    cls.append(t)
    b = rng.uniform(0, 0.01, (K, h, w))
    spots = [(0, 0), (0, w - 1), (h - 1, 0), (h - 1, w - 1), (0, w // 2), (h - 1, w // 2), (h // 2, 0), (h // 2, w - 1),
             (1, 1), (h - 2, w - 2), (3, w - 1), (h - 1, 3)]
    for k in range(K):
        y, x = spots[k % len(spots)]
        b[k, y, x] = 1.0
    cls.append(b)
    z = np.zeros((K, h, w))
    for k in range(K):
        z[k, rng.integers(0, h), rng.integers(0, w)] = 0.9
    cls.append(z)
    p = np.zeros((K, h, w))
    for k in range(K):
        y, x = rng.integers(0, h - 12), rng.integers(0, w - 12)
        p[k, y:y + 4 + k % 8, x:x + 4 + (k * 3) % 8] = 0.5
    cls.append(p)
    cls = np.stack(cls)
    M = cls.shape[0]
    out = np.zeros((M, K, 3, h, w))
    out[:, :, 0] = cls
    out[:, :, 1:] = rng.standard_normal((M, K, 2, h, w)) * 0.3
    return out.reshape(M, 3 * K, h, w).astype(np.float32)
