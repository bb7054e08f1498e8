"""numpy / OpenCV port of the reference codec path, kept at the reference's COST STRUCTURE:
one person at a time, one cv2 call per map, per-joint Python loops, float64 temporaries.

TEST INFRASTRUCTURE ONLY.  This is what bench.py times as the CPU baseline
(`cpu_baseline.kind = "port"` and the `--impl reference` arm): the reference itself is a
Python tree that cannot travel to the GPU box, so its per-person algorithm is restated here
with the same library calls it makes (cv2.warpAffine, cv2.GaussianBlur, np.linalg.inv).
Parity status: pinned -- tests/test_np_port.py checks every function against the golden
vectors generated from the unmodified reference (tests/golden/*.npz).

Reference lines restated (paths relative to the reference root):
  crop_person        deep_hrnet/lib/dataset/JointsDataset.py:29-49, :226-227
  map_joints         deep_hrnet/lib/dataset/JointsDataset.py:51-73, :228
  target_gaussian    deep_hrnet/lib/dataset/JointsDataset.py:301-348
  target_offset      deep_hrnet/lib/dataset/JointsDataset.py:349-381
  labels_rsn         RSN/dataset/JointsDataset.py:214-220, :276-291
  unflip_average     deep_hrnet/lib/utils/transforms.py:15-47, lib/core/function.py:171
  peaks              deep_hrnet/lib/core/inference.py:30-58
  refine             deep_hrnet/lib/core/inference.py:60-145
  final_preds        deep_hrnet/lib/core/inference.py:149-186, :20-27
"""
import math

import cv2
import numpy as np

RSN_SIGMAS = (2.6, 2.0, 1.7, 1.4105, 1.1105)


def unbiased_matrix(rot, c, s, W, H):
    box = s * 200.0
    th = rot / 180.0 * math.pi
    cs, sn = math.cos(th), math.sin(th)
    m = np.zeros((2, 3), np.float32)
    kx, ky = box[0] / (W - 1.0), box[1] / (H - 1.0)
    m[0, 0], m[0, 1] = cs * kx, sn * ky
    m[1, 0], m[1, 1] = -sn * kx, cs * ky
    c2 = c * 2.0
    m[0, 2] = -0.5 * box[0] * cs - 0.5 * box[1] * sn + 0.5 * c2[0]
    m[1, 2] = 0.5 * box[0] * sn - 0.5 * box[1] * cs + 0.5 * c2[1]
    return m


def crop_person(img, rot, c, s, W, H):
    m = unbiased_matrix(rot, c, s, W, H)
    return cv2.warpAffine(img, m, (int(W), int(H)), flags=cv2.WARP_INVERSE_MAP | cv2.INTER_LINEAR)


def map_joints(pts, rot, c, s, W, H):
    box = s * 200.0
    th = rot / 180.0 * math.pi
    ns, cs = -math.sin(th), math.cos(th)
    out = np.zeros(pts.shape, pts.dtype)
    dx, dy = pts[:, 0] - c[0], pts[:, 1] - c[1]
    out[:, 0] = cs * dx + ns * dy
    out[:, 1] = -ns * dx + cs * dy
    out[:, 0] += box[0] * 0.5
    out[:, 1] += box[1] * 0.5
    out[:, 0] *= (W - 1.0) / box[0]
    out[:, 1] *= (H - 1.0) / box[1]
    return out


def target_gaussian(joints, vis, W, H, w, h, sigma):
    K = joints.shape[0]
    weight = np.ones((K, 1), np.float32)
    weight[:, 0] = vis[:, 0]
    tgt = np.zeros((K, h, w), np.float32)
    rad = sigma * 3
    stride = (np.array([W, H]) - 1.0) / (np.array([w, h]) - 1.0)
    for k in range(K):
        ax, ay = joints[k][0] / stride[0], joints[k][1] / stride[1]
        mx, my = int(ax + 0.5), int(ay + 0.5)
        lo = [int(mx - rad), int(my - rad)]
        hi = [int(mx + rad + 1), int(my + rad + 1)]
        if lo[0] >= w or lo[1] >= h or hi[0] < 0 or hi[1] < 0:
            weight[k] = 0
            continue
        n = 2 * rad + 1
        xs = np.arange(0, n, 1, np.float32)
        ys = xs[:, None]
        x0 = n // 2 + (ax - mx)
        y0 = n // 2 + (ay - my)
        g = np.exp(-((xs - x0) ** 2 + (ys - y0) ** 2) / (2 * sigma ** 2))
        gx = max(0, -lo[0]), min(hi[0], w) - lo[0]
        gy = max(0, -lo[1]), min(hi[1], h) - lo[1]
        ix = max(0, lo[0]), min(hi[0], w)
        iy = max(0, lo[1]), min(hi[1], h)
        if weight[k] > 0.5:
            tgt[k][iy[0]:iy[1], ix[0]:ix[1]] = g[gy[0]:gy[1], gx[0]:gx[1]]
    return tgt, weight


def target_offset(joints, vis, W, H, w, h, kpd):
    K = joints.shape[0]
    weight = np.ones((K, 1), np.float32)
    weight[:, 0] = vis[:, 0]
    tgt = np.zeros((K, 3, h * w), np.float32)
    gx, gy = np.meshgrid(np.arange(0, w), np.arange(0, h))
    gx, gy = gx.reshape(-1), gy.reshape(-1)
    stride = (np.array([W, H]) - 1.0) / (np.array([w, h]) - 1.0)
    for k in range(K):
        ox = (joints[k][0] / stride[0] - gx) / kpd
        oy = (joints[k][1] / stride[1] - gy) / kpd
        d = ox ** 2 + oy ** 2
        keep = np.where((d <= 1) & (d >= 0))[0]
        if weight[k] > 0.5:
            tgt[k, 0, keep] = 1
            tgt[k, 1, keep] = ox[keep]
            tgt[k, 2, keep] = oy[keep]
    return tgt.reshape(K * 3, h, w), weight


def labels_rsn(joints, valid, W, H, w, h, sigmas=RSN_SIGMAS):
    K = joints.shape[0]
    gx, gy = np.meshgrid(np.arange(0, w), np.arange(0, h))
    gx, gy = gx.reshape(-1), gy.reshape(-1)
    out = np.zeros((len(sigmas), K, h, w))
    for si, sg in enumerate(sigmas):
        maps = np.zeros((K, h * w), dtype="float32")
        for k in range(K):
            if valid[k] < 1:
                continue
            ty = joints[k, 1] * (h - 1.0) / (H - 1.0)
            tx = joints[k, 0] * (w - 1.0) / (W - 1.0)
            d = (tx - gx) ** 2 + (ty - gy) ** 2
            maps[k] = 255.0 * np.exp(d / (-2.0 * sg ** 2))
        out[si] = maps.reshape(K, h, w)
    return out.astype(np.float32)


def unflip_average(out, flipped, pairs, triplets):
    """Batch-level, as validate() does it: mirror, (negate x offsets,) swap pairs, average."""
    f = flipped.copy()[:, :, :, ::-1]
    if triplets:
        shp = f.shape
        f[:, 1::3] = -f[:, 1::3]
        f = f.reshape(shp[0], -1, 3, shp[2], shp[3])
    for a, b in pairs:
        t = f[:, a].copy()
        f[:, a] = f[:, b]
        f[:, b] = t
    if triplets:
        f = f.reshape(shp)
    return (out + f.copy()) * np.float32(0.5)


def peaks(hm):
    B, K, h, w = hm.shape
    flat = hm.reshape(B, K, -1)
    idx = np.argmax(flat, 2).reshape(B, K, 1)
    val = np.amax(flat, 2).reshape(B, K, 1)
    xy = np.tile(idx, (1, 1, 2)).astype(np.float32)
    xy[:, :, 0] = xy[:, :, 0] % w
    xy[:, :, 1] = np.floor(xy[:, :, 1] / w)
    xy *= np.tile(np.greater(val, 0.0), (1, 1, 2)).astype(np.float32)
    return xy, val


def refine(coords, hm):
    B, K, h, w = hm.shape
    for i in range(B):
        for j in range(K):
            m = hm[i, j]
            top = np.max(m)
            m = cv2.GaussianBlur(m, (7, 7), 0)
            lo, hi = np.min(m), np.max(m)
            hm[i, j] = (m - lo) / (hi - lo) * top
    lg = np.log(np.clip(hm, 0.001, 50))
    pad = np.zeros((B, K, h + 2, w + 2), dtype=float)
    pad[:, :, 1:-1, 1:-1] = lg
    pad[:, :, 1:-1, -1], pad[:, :, 1:-1, 0] = lg[:, :, :, -1], lg[:, :, :, 0]
    pad[:, :, -1, 1:-1], pad[:, :, 0, 1:-1] = lg[:, :, -1, :], lg[:, :, 0, :]
    pad[:, :, -1, -1], pad[:, :, 0, 0] = lg[:, :, -1, -1], lg[:, :, 0, 0]
    pad[:, :, 0, -1], pad[:, :, -1, 0] = lg[:, :, 0, -1], lg[:, :, -1, 0]
    ci = coords.astype(np.int32)
    res = np.zeros(coords.shape)
    for i in range(B):
        for j in range(K):
            x, y = ci[i, j, 0] + 1, ci[i, j, 1] + 1
            p = pad[i, j]
            I, xp, xm, yp, ym = p[y, x], p[y, x + 1], p[y, x - 1], p[y + 1, x], p[y - 1, x]
            pp, mm = p[y + 1, x + 1], p[y - 1, x - 1]
            D = np.array([0.5 * (xp - xm), 0.5 * (yp - ym)])
            dxy = 0.5 * (pp - xp - yp + I + I - xm - ym + mm)
            Hm = np.array([[xp - 2 * I + xm, dxy], [dxy, yp - 2 * I + ym]])
            try:
                inv = np.linalg.inv(Hm)
            except np.linalg.LinAlgError:
                inv = np.zeros((2, 2))
            res[i, j] = ci[i, j].astype(float) - np.matmul(inv, D[:, None]).reshape(-1)
    return res


def to_image(coords, c, s, w, h):
    box = s * 200.0
    kx, ky = box[0] / (w - 1.0), box[1] / (h - 1.0)
    out = np.zeros(coords.shape)
    out[:, 0] = coords[:, 0] * kx + c[0] - box[0] * 0.5
    out[:, 1] = coords[:, 1] * ky + c[1] - box[1] * 0.5
    return out


def final_preds(hm, center, scale, mode, kpd=4.0, post_process=True, pre_div=0.0):
    """mode 'gaussian' | 'offset'.  Returns (preds, maxvals, preds_in_input_space)."""
    if pre_div:
        hm = hm / pre_div
    h, w = hm.shape[2], hm.shape[3]
    if mode == "gaussian":
        coords, val = peaks(hm)
        if post_process:
            coords = refine(coords, hm)
    else:
        net = hm.copy()
        cls = net[:, ::3, :]
        ox = net[:, 1::3, :] * kpd
        oy = net[:, 2::3, :] * kpd
        for i in range(cls.shape[0]):
            for j in range(cls.shape[1]):
                cls[i, j] = cv2.GaussianBlur(cls[i, j], (15, 15), 0)
                ox[i, j] = cv2.GaussianBlur(ox[i, j], (7, 7), 0)
                oy[i, j] = cv2.GaussianBlur(oy[i, j], (7, 7), 0)
        coords, val = peaks(cls)
        for i in range(coords.shape[0]):
            for j in range(coords.shape[1]):
                px, py = int(coords[i][j][0]), int(coords[i][j][1])
                coords[i][j][0] += ox[i, j, py, px]
                coords[i][j][1] += oy[i, j, py, px]
    preds = coords.copy()
    pin = preds.copy()
    pin[:, :, 0] = pin[:, :, 0] / (w - 1.0) * (4 * w - 1.0)
    pin[:, :, 1] = pin[:, :, 1] / (h - 1.0) * (4 * h - 1.0)
    for i in range(coords.shape[0]):
        preds[i] = to_image(coords[i], center[i], scale[i], w, h)
    return preds, val, pin
