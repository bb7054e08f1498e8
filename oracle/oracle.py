"""ctypes front end of the C oracle (oracle/libudp_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the cpu_baseline
leg of bench.py.  The product package never imports this module.

All functions are batched over persons with plain Python loops around the per-person C
restatement; arrays are numpy, layouts follow the reference.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_build.build())
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _cs(center, scale):
    """center/scale [N,2] of one dtype (float32 or float64) -> contiguous arrays + flag."""
    center = np.ascontiguousarray(center)
    scale = np.ascontiguousarray(scale)
    f64 = center.dtype == np.float64
    dt = np.float64 if f64 else np.float32
    return center.astype(dt, copy=False), scale.astype(dt, copy=False), int(f64)


def warp_matrices(rot_deg, center, scale, W, H):
    """a1: [N,2,3] float32.  deep_hrnet/lib/dataset/JointsDataset.py:29-49 called as :226."""
    center, scale, f64 = _cs(center, scale)
    N = center.shape[0]
    rot = np.broadcast_to(np.asarray(rot_deg, np.float64), (N,))
    M = np.zeros((N, 2, 3), np.float32)
    L = lib()
    for i in range(N):
        L.udp_o_warp_matrix(C.c_double(float(rot[i])), _p(center[i]), _p(scale[i]), f64,
                            int(W), int(H), _p(M[i]))
    return M


def rotate_points(joints_xy, rot_deg, center, scale, W, H, do_clip=False):
    """a3: joints_xy [N,K,2] float64 -> [N,K,2] float64.  JointsDataset.py:51-73."""
    center, scale, f64 = _cs(center, scale)
    pts = np.ascontiguousarray(joints_xy, np.float64)
    N, K = pts.shape[:2]
    rot = np.broadcast_to(np.asarray(rot_deg, np.float64), (N,))
    out = np.zeros_like(pts)
    L = lib()
    for i in range(N):
        L.udp_o_rotate_points(_p(pts[i]), K, 2, C.c_double(float(rot[i])), _p(center[i]),
                              _p(scale[i]), f64, int(W), int(H), int(do_clip), _p(out[i]), 2)
    return out


def fliplr_joints(joints, joints_vis, width, pairs, mul_vis=True):
    """a4: deep_hrnet/lib/utils/transforms.py:50-64 (mul_vis=True: returns joints*joints_vis, every column
    multiplied, :64) and RSN/lib/utils/transforms.py:70-79 (`flip_joints`, mul_vis=False: no product).
    x = width - x - 1 (two float64 subtractions, :55), then the rows of every pair swapped in both arrays
    (:58-62).  joints, joints_vis float64 [N,K,3] -> new arrays (the reference works in place); width
    scalar or [N].  Pinned by tests/golden/fliplr.npz (outputs of the reference's two functions)."""
    j = np.array(joints, np.float64, copy=True)
    v = np.array(joints_vis, np.float64, copy=True)
    N = j.shape[0]
    wd = np.broadcast_to(np.asarray(width, np.float64), (N,))
    j[..., 0] = wd[:, None] - j[..., 0] - 1
    for a, b in pairs:
        j[:, [a, b]] = j[:, [b, a]]
        v[:, [a, b]] = v[:, [b, a]]
    return (j * v if mul_vis else j), v


def accuracy_pck(output, target, thr=0.5):
    """accuracy(output, target, 'gaussian', thr) of deep_hrnet/lib/core/evaluate.py:41-71
    (calc_dists :16-28, dist_acc :31-38), restated with array operations in the same dtypes:
    get_max_preds on both tensors (core/inference.py:30-58), float32 coords / float64 norm
    (h, w)/10, float64 distances, -1 where the target peak is not > 1 in x and y.
    Returns (acc [K+1] f64, avg_acc, cnt, pred f32 [N,K,2]).  Pinned by tests/golden/accuracy.npz
    (outputs of the reference's own accuracy())."""
    def max_preds(hm):
        N, K, h, w = hm.shape
        flat = hm.reshape(N, K, -1)
        idx = np.argmax(flat, 2)
        mv = np.take_along_axis(flat, idx[..., None], 2)[..., 0]
        p = np.stack([idx % w, idx // w], -1).astype(np.float32)
        return p * (mv > 0.0)[..., None].astype(np.float32)
    output, target = np.asarray(output, np.float32), np.asarray(target, np.float32)
    N, K, h, w = output.shape
    pred, tgt = max_preds(output), max_preds(target)
    norm = np.ones((N, 2)) * np.array([h, w]) / 10
    d = pred / norm[:, None, :] - tgt / norm[:, None, :]
    dist = np.sqrt(d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1])
    dist = np.where((tgt[..., 0] > 1) & (tgt[..., 1] > 1), dist, -1.0).T  # [K, N]
    acc = np.zeros(K + 1)
    s, cnt = 0, 0
    for j in range(K):
        cal = dist[j] != -1
        a = (dist[j][cal] < thr).sum() * 1.0 / cal.sum() if cal.sum() > 0 else -1
        acc[j + 1] = a
        if a >= 0:
            s, cnt = s + a, cnt + 1
    avg = s / cnt if cnt != 0 else 0
    if cnt != 0:
        acc[0] = avg
    return acc, avg, cnt, pred


def decode_quarter(heatmaps, boxes):
    """SURVEY 8(f)-4 (first half): the demo engines' quarter-offset decoder,
    deep_hrnet/tools/infer_utils/decode.py:19-40 with its get_max_preds (:43-52: coordinates -1 where
    maxval <= 0) and the biased transform_preds (:55-62: scale/output_size, not size-1).  Restated
    with array operations in the dtypes the reference's scalars produce: float32 coordinates, +-0.25
    where both neighbours exist (1 < px < W-1 and 1 < py < H-1), float32 affine for float32 boxes
    (float64 products rounded into the float32 result for float64 boxes), truncation to int64.
    heatmaps f32 [B,C,H,W], boxes [B,4] = (cx, cy, sx, sy) -> int64 [B,C,2]."""
    hm = np.asarray(heatmaps, np.float32)
    boxes = np.asarray(boxes)
    B, C, H, W = hm.shape
    flat = hm.reshape(B, C, -1)
    idx = np.argmax(flat, 2)
    mv = np.take_along_axis(flat, idx[..., None], 2)[..., 0]
    coords = np.stack([idx % W, idx // W], -1).astype(np.float32)
    coords = np.where((mv > 0.0)[..., None], coords, np.float32(-1))
    px = np.floor(coords[..., 0] + np.float32(0.5)).astype(np.int64)
    py = np.floor(coords[..., 1] + np.float32(0.5)).astype(np.int64)
    inner = (1 < px) & (px < W - 1) & (1 < py) & (py < H - 1)
    cx, cy = np.clip(px, 1, W - 2), np.clip(py, 1, H - 2)
    b, c = np.meshgrid(np.arange(B), np.arange(C), indexing="ij")
    dx = hm[b, c, cy, cx + 1] - hm[b, c, cy, cx - 1]
    dy = hm[b, c, cy + 1, cx] - hm[b, c, cy - 1, cx]
    shift = np.stack([np.sign(dx), np.sign(dy)], -1) * np.float32(0.25)
    coords = np.where(inner[..., None], coords + shift, coords).astype(np.float32)
    center, scale = boxes[:, None, :2], boxes[:, None, 2:] * 200          # boxes' dtype
    per_px = scale / np.array([W, H])                                     # python ints: dtype kept
    out = np.ones_like(coords)
    out[...] = coords * per_px + center - scale * 0.5                     # float32 store (rounds a float64 result)
    return out.astype(np.int64)


# cv2.getGaussianKernel(n, 0, cv2.CV_64F)[: n // 2 + 1] as produced by OpenCV 4.13 (hex floats; the 5/7/9
# tables are its fixed small kernels, 11/13/15 come from its bit-exact exp)
GAUSS_F64 = {
    5: ['0x1.0000000000000p-4', '0x1.0000000000000p-2', '0x1.8000000000000p-2'],
    7: ['0x1.0000000000000p-5', '0x1.c000000000000p-4', '0x1.c000000000000p-3', '0x1.2000000000000p-2'],
    9: ['0x1.0000000000000p-6', '0x1.a000000000000p-5', '0x1.e000000000000p-4', '0x1.9800000000000p-3', '0x1.e000000000000p-3'],
    11: ['0x1.20c2564ee6772p-7', '0x1.bcb86a082c301p-6', '0x1.0ab50979aaf94p-4', '0x1.f2464c62edaf4p-4', '0x1.6a7e1d504a91dp-3',
         '0x1.9ac20a36ea596p-3'],
    13: ['0x1.7c090614dc85bp-8', '0x1.0cb82167c249dp-6', '0x1.3a8fe850a8f47p-5', '0x1.30cd237f01688p-4', '0x1.e8f1e9d4d4f8bp-4',
         '0x1.449ed5bf3520ap-3', '0x1.64ccba4a2c88fp-3'],
    15: ['0x1.0d2c46a814066p-8', '0x1.600a4b3638a7bp-7', '0x1.8d1b96696e28fp-6', '0x1.825954ed343f3p-5', '0x1.4431ed85d4175p-4',
         '0x1.d543dae40f820p-4', '0x1.24ec9c4a6087bp-3', '0x1.3b69611f1d587p-3'],
}


def gauss_kernel_f64(n):
    half = [float.fromhex(v) for v in GAUSS_F64[n]]
    return np.array(half + half[-2::-1])


def gaussian_blur_f64(a, n):
    """cv2.GaussianBlur(a float64 [h,w], (n,n), 0), restated (udp_o_gaussian_blur_f64)."""
    a = np.ascontiguousarray(a, np.float64)
    out = np.empty_like(a)
    k = gauss_kernel_f64(n)
    lib().udp_o_gaussian_blur_f64(_p(a), _p(out), C.c_int(a.shape[0]), C.c_int(a.shape[1]), C.c_int(n), _p(k))
    return out


def rsn_get_results(outputs, centers, scales, in_shape, kernel=11, shift=0.25, border=10):
    """SURVEY 8(f)-4 (second half): get_results of RSN/exps/RSN18.coco/test.py:25-71 (same function in the
    Res18 / 4XRSN18 / RSN18.mpii experiments).  Per map: zero-padded float64 copy (border 10), Gaussian
    blur, first arg-max, that element zeroed, second arg-max, a `shift`-pixel step from the first towards
    the second, clamp, `x*4+2`, score = (out/255 + 0.5) at the rounded position (float32), then the
    biased affine `kps / input * scale*200 + center - scale*200*0.5`.  Arrays restated per map with the
    scalar types the reference's expressions produce; does not mutate `scales` (the reference does,
    :26).  outputs f32 [N,K,h,w], centers/scales f32 [N,2], in_shape (H, W).
    Returns (preds f64 [N,K,2], maxvals f64 [N,K,1])."""
    outputs = np.asarray(outputs, np.float32)
    N, K, h, w = outputs.shape
    sc = np.asarray(scales, np.float32) * 200            # float32, as `scales *= 200` on the loader's array
    ce = np.asarray(centers, np.float32)
    preds, maxvals = np.zeros((N, K, 2)), np.zeros((N, K, 1))
    for i in range(N):
        score_map = outputs[i] / 255 + 0.5               # float32
        for j in range(K):
            dr = np.zeros((h + 2 * border, w + 2 * border))
            dr[border:-border, border:-border] = outputs[i, j]
            dr = gaussian_blur_f64(dr, kernel)
            y, x = np.unravel_index(dr.argmax(), dr.shape)
            dr[y, x] = 0
            x, y = x - border, y - border
            py, px = np.unravel_index(dr.argmax(), dr.shape)
            px, py = px - (border + x), py - (border + y)
            ln = (px ** 2 + py ** 2) ** 0.5
            if ln > 1e-3:
                x, y = x + shift * px / ln, y + shift * py / ln
            x, y = max(0, min(x, w - 1)), max(0, min(y, h - 1))
            preds[i, j] = x * 4 + 2, y * 4 + 2
            maxvals[i, j, 0] = score_map[j, int(round(y) + 1e-9), int(round(x) + 1e-9)]
        preds[i, :, 0] = preds[i, :, 0] / in_shape[1] * sc[i][0] + ce[i][0] - sc[i][0] * 0.5
        preds[i, :, 1] = preds[i, :, 1] / in_shape[0] * sc[i][1] + ce[i][1] - sc[i][1] * 0.5
    return preds, maxvals


def aid_erase(img, ellipses=(), grid=0, hidden=()):
    """Cutout._cutout (deep_hrnet/lib/utils/transforms.py:199-216) and
    HideAndSeek._hide_and_seek (:163-177) for DRAWN parameters (the random draws are the caller's):
    ellipses = [(cx, cy, rx, ry), ...]; grid and `hidden` = one flag per cell in draw order (outer loop
    over range(0, width, grid), inner over range(0, height, grid)); the reference's transposed slice
    img[x:x_end, y:y_end] is kept.  img uint8 [H,W,3]; returns a new array."""
    out = np.array(img, copy=True)
    H, W, _ = out.shape
    xs, ys = np.meshgrid(np.arange(W), np.arange(H))
    for cx, cy, rx, ry in ellipses:
        out[((cx - xs) / rx) ** 2 + ((cy - ys) / ry) ** 2 <= 1] = 0
    if grid:
        flags = iter(hidden)
        for x in range(0, W, grid):
            for y in range(0, H, grid):
                if next(flags):
                    out[x:min(W, x + grid), y:min(H, y + grid), :] = 0
    return out


COCO_SIGMAS = np.array([.26, .25, .25, .35, .35, .79, .79, .72, .72, .62, .62, 1.07, 1.07, .87, .87, .89, .89]) / 10.0


def coco_rescore(preds, box_score, in_vis_thre):
    """SURVEY 8(f)-3, rescoring of deep_hrnet/lib/dataset/coco.py:326-339: mean of the joint scores above
    `in_vis_thre` (accumulated in float32 in joint order, as the reference's scalar loop over a float32
    record does) times the box score (float64).  preds f32 [N,K,3], box_score f64 [N] -> f64 [N]."""
    preds = np.asarray(preds, np.float32)
    out = np.zeros(len(preds))
    for n in range(len(preds)):
        acc, cnt = np.float32(0), 0
        for t in preds[n, :, 2]:
            if t > in_vis_thre:
                acc, cnt = np.float32(acc + t), cnt + 1
        mean = np.float32(acc / np.float32(cnt)) if cnt else 0
        out[n] = mean * np.float64(box_score[n])
    return out


def oks_row(preds, area, i, rest, sigmas=None, in_vis_thre=None):
    """oks_iou of lib/nms/nms.py:75-94 between person i and persons `rest` of one image: float32
    coordinate differences and squares, float64 from the division by vars on, np.exp, and np.sum's
    pairwise order over the K (or the masked) joints.  The optional mask keeps joints whose score
    in the *candidate* is above in_vis_thre (`list(a) and list(b)` at :89 evaluates to list(b))."""
    sig = COCO_SIGMAS if sigmas is None else np.asarray(sigmas, np.float64)
    var = (sig * 2) ** 2
    g = preds[i]
    out = np.zeros(len(rest))
    for k, r in enumerate(rest):
        d = preds[r]
        dx, dy = d[:, 0] - g[:, 0], d[:, 1] - g[:, 1]
        e = (dx * dx + dy * dy) / var / ((area[i] + area[r]) / 2 + np.spacing(1)) / 2
        if in_vis_thre is not None:
            e = e[d[:, 2] > in_vis_thre]
        out[k] = np.sum(np.exp(-e)) / e.shape[0] if e.shape[0] else 0.0
    return out


def _order_desc(scores):
    """`scores.argsort()[::-1]` (nms.py:113): descending; among equal scores the higher row first,
    which is what reversing a stable ascending sort gives (numpy's default sort leaves tie order
    unspecified; the fixtures have no ties)."""
    return np.argsort(scores, kind="stable")[::-1]


def oks_nms(preds, area, score, thresh, sigmas=None, in_vis_thre=None):
    """oks_nms of lib/nms/nms.py:97-124 for ONE image; returns the kept rows in keep order."""
    preds = np.asarray(preds, np.float32)
    order = _order_desc(np.asarray(score, np.float64))
    keep = []
    while order.size:
        i = order[0]
        keep.append(int(i))
        ovr = oks_row(preds, area, i, order[1:], sigmas, in_vis_thre)
        order = order[1:][ovr <= thresh]
    return keep


def soft_oks_nms(preds, area, score, thresh, sigmas=None, in_vis_thre=None, max_dets=20):
    """soft_oks_nms of lib/nms/nms.py:139-180 (gaussian rescore :127-136) for ONE image.  The full
    re-sort of every round only matters through its head, so each round takes the arg-max of the
    decayed scores."""
    preds = np.asarray(preds, np.float32)
    score = np.asarray(score, np.float64)
    order = _order_desc(score)
    s = score[order]
    keep = []
    while order.size and len(keep) < max_dets:
        i = order[0]
        ovr = oks_row(preds, area, i, order[1:], sigmas, in_vis_thre)
        s = s[1:] * np.exp(-ovr ** 2 / thresh)
        order = order[1:]
        t = np.argsort(s, kind="stable")[::-1]
        order, s = order[t], s[t]
        keep.append(int(i))
    return keep


def coco_rescore_nms(preds, all_boxes, image_ids, in_vis_thre, oks_thre, soft=False):
    """coco.py:308-356: group the persons by image (first-appearance order, rows in input order),
    rescore, then oks_nms / soft_oks_nms per image.  Returns (score f64 [N], [kept global rows per image])."""
    preds = np.asarray(preds, np.float32)
    score = coco_rescore(preds, all_boxes[:, 5], in_vis_thre)
    groups = {}
    for r, im in enumerate(image_ids):
        groups.setdefault(int(im), []).append(r)
    kept = []
    for rows in groups.values():
        rows = np.asarray(rows)
        fn = soft_oks_nms if soft else oks_nms
        k = fn(preds[rows], all_boxes[rows, 4], score[rows], oks_thre)
        kept.append([int(rows[j]) for j in k])
    return score, kept


def to_tensor_normalize(crop, mean, std):
    """`transforms.Compose([ToTensor(), Normalize(mean, std)])` on the uint8 HWC crop
    (deep_hrnet/tools/train.py:127-143, applied at JointsDataset.py:237-238).  The transforms live
    in torchvision (pinned 0.9.1 by deep_hrnet/requirements.txt, not vendored in the reference):
    ToTensor = HWC uint8 -> CHW float32 `.div(255)`; Normalize = `.sub_(mean).div_(std)` with
    mean/std as float32 tensors.  Every step is one IEEE float32 operation, restated here in numpy
    float32; pinned by tests/golden/normalize.npz (generated with torchvision itself)."""
    x = np.ascontiguousarray(crop, np.uint8).astype(np.float32) / np.float32(255)
    x = (x - np.asarray(mean, np.float32)) / np.asarray(std, np.float32)
    return np.ascontiguousarray(x.transpose(2, 0, 1))


def flip_input(x):
    """The mirrored network input of the flip test: np.flip(input, 3) (core/function.py:154)."""
    return np.ascontiguousarray(x[..., ::-1])


def warp_affine_u8(src, M, W, H):
    """a2: src uint8 [Hs,Ws,3], M float32 [2,3] -> uint8 [H,W,3] (cv2.warpAffine,
    WARP_INVERSE_MAP|INTER_LINEAR, constant-0 border; JointsDataset.py:227)."""
    src = np.ascontiguousarray(src, np.uint8)
    M = np.ascontiguousarray(M, np.float32)
    dst = np.empty((H, W, 3), np.uint8)
    lib().udp_o_warp_affine_u8(_p(src), src.shape[0], src.shape[1], src.strides[0], _p(M),
                               _p(dst), int(W), int(H))
    return dst


def _jw(joints_weight, K):
    if joints_weight is None:
        return None, None
    jw = np.ascontiguousarray(np.asarray(joints_weight, np.float32).reshape(K))
    return jw, _p(jw)


def encode_gaussian(joints, joints_vis, W, H, w, h, sigma, joints_weight=None):
    """a5: joints, joints_vis [N,K,3] float64 -> target [N,K,h,w] f32, target_weight [N,K,1] f32."""
    joints = np.ascontiguousarray(joints, np.float64)
    vis = np.ascontiguousarray(joints_vis, np.float64)
    N, K = joints.shape[:2]
    tgt = np.empty((N, K, h, w), np.float32)
    tw = np.empty((N, K, 1), np.float32)
    jw, jwp = _jw(joints_weight, K)
    L = lib()
    for i in range(N):
        L.udp_o_encode_gaussian(_p(joints[i]), _p(vis[i]), K, int(W), int(H), int(w), int(h),
                                C.c_double(sigma), jwp, _p(tgt[i]), _p(tw[i]))
    return tgt, tw


def encode_offset(joints, joints_vis, W, H, w, h, kpd, joints_weight=None):
    """a6: -> target [N,3K,h,w] f32, target_weight [N,K,1] f32."""
    joints = np.ascontiguousarray(joints, np.float64)
    vis = np.ascontiguousarray(joints_vis, np.float64)
    N, K = joints.shape[:2]
    tgt = np.empty((N, 3 * K, h, w), np.float32)
    tw = np.empty((N, K, 1), np.float32)
    jw, jwp = _jw(joints_weight, K)
    L = lib()
    for i in range(N):
        L.udp_o_encode_offset(_p(joints[i]), _p(vis[i]), K, int(W), int(H), int(w), int(h),
                              C.c_double(kpd), jwp, _p(tgt[i]), _p(tw[i]))
    return tgt, tw


RSN_SIGMAS = (2.6, 2.0, 1.7, 1.4105, 1.1105)  # RSN/dataset/JointsDataset.py:19-25, kernels 15,11,9,7,5


def encode_rsn(joints, valid, W, H, w, h, sigmas=RSN_SIGMAS):
    """a7: joints [N,K,>=2] float64 (crop coords), valid [N,K] -> labels [N,S,K,h,w] f32."""
    joints = np.ascontiguousarray(joints, np.float64)
    valid = np.ascontiguousarray(valid, np.float32).reshape(joints.shape[0], -1)
    N, K, js = joints.shape
    sig = np.ascontiguousarray(sigmas, np.float64)
    S = sig.shape[0]
    out = np.empty((N, S, K, h, w), np.float32)
    L = lib()
    for i in range(N):
        L.udp_o_encode_rsn(_p(joints[i]), js, _p(valid[i]), K, int(W), int(H), int(w), int(h),
                           _p(sig), S, _p(out[i]))
    return out


def rsn_transform_joints(joints, vis, rot_deg, center, scale, W, H):
    """a3': returns (joints [N,K,3] with x,y mapped where vis>0, vis [N,K,3] culled, Minv [N,2,3])."""
    joints = np.array(joints, np.float64, copy=True)
    vis = np.array(vis, np.float64, copy=True)
    center = np.ascontiguousarray(center, np.float32)
    scale = np.ascontiguousarray(scale, np.float32)
    N, K = joints.shape[:2]
    rot = np.broadcast_to(np.asarray(rot_deg, np.float64), (N,))
    Minv = np.zeros((N, 2, 3), np.float32)
    L = lib()
    for i in range(N):
        L.udp_o_rsn_warp_matrix_inverse(C.c_double(float(rot[i])), _p(center[i]), _p(scale[i]),
                                        int(W), int(H), _p(Minv[i]))
        L.udp_o_rsn_transform_joints(_p(joints[i]), joints.shape[2], _p(vis[i]), vis.shape[2], K,
                                     _p(Minv[i]), int(W), int(H))
    return joints, vis, Minv


def gaussian_blur(a, n):
    """cv2.GaussianBlur(a, (n,n), 0) for float32 [h,w], n in {7,15}."""
    a = np.ascontiguousarray(a, np.float32)
    out = np.empty_like(a)
    lib().udp_o_gaussian_blur(_p(a), _p(out), a.shape[0], a.shape[1], int(n))
    return out


def flip_perm(K, pairs):
    perm = np.arange(K, dtype=np.int32)
    for a, b in pairs:
        perm[a], perm[b] = b, a
    return perm


def flip_average(out, flipped, pairs, triplets=False):
    """a9-a11: (out + flip_back(flipped)) * 0.5, [N,C,h,w] float32."""
    out = np.ascontiguousarray(out, np.float32)
    flipped = np.ascontiguousarray(flipped, np.float32)
    N, Cc, h, w = out.shape
    K = Cc // 3 if triplets else Cc
    perm = flip_perm(K, pairs)
    avg = np.empty_like(out)
    L = lib()
    for i in range(N):
        L.udp_o_flip_average(_p(out[i]), _p(flipped[i]), _p(perm), K, h, w, int(triplets), _p(avg[i]))
    return avg


def decode_gaussian(hm, center, scale, post_process=True, pre_div=0.0, want_blurred=False):
    """a12-a14,a16: hm [N,K,h,w] f32 -> dict(preds f64 [N,K,2], maxvals f32 [N,K,1],
    preds_in f64 [N,K,2], idx int32 [N,K][, blurred])."""
    hm = np.ascontiguousarray(hm, np.float32)
    center, scale, f64 = _cs(center, scale)
    N, K, h, w = hm.shape
    preds = np.empty((N, K, 2), np.float64)
    pin = np.empty((N, K, 2), np.float64)
    mv = np.empty((N, K, 1), np.float32)
    idx = np.empty((N, K), np.int32)
    bl = np.empty_like(hm) if want_blurred else None
    L = lib()
    for i in range(N):
        L.udp_o_decode_gaussian(_p(hm[i]), K, w, h, _p(center[i]), _p(scale[i]), f64,
                                C.c_float(pre_div), int(post_process), _p(preds[i]), _p(mv[i]),
                                _p(pin[i]), _p(idx[i]), _p(bl[i]) if want_blurred else None)
    r = dict(preds=preds, maxvals=mv, preds_in=pin, idx=idx)
    if want_blurred:
        r["blurred"] = bl
    return r


def decode_offset(out, center, scale, kpd):
    """a15,a16: out [N,3K,h,w] f32 -> dict as decode_gaussian (values are float32-exact)."""
    out = np.ascontiguousarray(out, np.float32)
    center, scale, f64 = _cs(center, scale)
    N, C3, h, w = out.shape
    K = C3 // 3
    preds = np.empty((N, K, 2), np.float64)
    pin = np.empty((N, K, 2), np.float64)
    mv = np.empty((N, K, 1), np.float32)
    idx = np.empty((N, K), np.int32)
    L = lib()
    for i in range(N):
        L.udp_o_decode_offset(_p(out[i]), K, w, h, C.c_double(kpd), _p(center[i]), _p(scale[i]),
                              f64, _p(preds[i]), _p(mv[i]), _p(pin[i]), _p(idx[i]))
    return dict(preds=preds, maxvals=mv, preds_in=pin, idx=idx)
