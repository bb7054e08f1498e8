"""Batched device API: CUDA tensors in, CUDA tensors out, one kernel launch per call.

This is the fast path that bench.py times.  Every function maps 1:1 onto a C-ABI entry point
of include/udp_codec.h; torch is used only for device memory and streams.  All launches go
to torch's current CUDA stream.  Nothing here touches the CPU oracle; a missing
libudp_codec.so raises (udp_pose_b200._lib.load).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

_F64 = torch.float64
_F32 = torch.float32


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _dev(t, dtype=None):
    """Contiguous CUDA tensor of the wanted dtype (numpy arrays are uploaded)."""
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(t))
    if not t.is_cuda:
        t = t.cuda(non_blocking=True)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _cs(center, scale):
    """center/scale keep their dtype (float32 or float64): it selects the arithmetic."""
    center = _dev(center)
    scale = _dev(scale)
    if center.dtype not in (_F32, _F64):
        center = center.to(_F64)
    scale = scale.to(center.dtype)
    return center, scale, int(center.dtype == _F64)


_perm_cache = {}


def flip_perm(K, pairs, device):
    """int32 [K] partner-joint table from a flip_pairs list (cached per device)."""
    key = (K, tuple(map(tuple, pairs)), str(device))
    if key not in _perm_cache:
        perm = list(range(K))
        for a, b in pairs:
            perm[a], perm[b] = b, a
        _perm_cache[key] = torch.tensor(perm, dtype=torch.int32, device=device)
    return _perm_cache[key]


# ---------------------------------------------------------------------------- crop side
def warp_matrices(center, scale, rot_deg, image_size):
    """a1 get_warpmatrix(r, c*2.0, image_size-1.0, s), batched -> float32 [N,2,3]."""
    center, scale, f64 = _cs(center, scale)
    N = center.shape[0]
    rot = _dev(rot_deg, _F64) if rot_deg is not None else None
    M = torch.empty((N, 2, 3), dtype=_F32, device=center.device)
    _lib.call("udp_warp_matrices", _ptr(center), _ptr(scale), f64, _ptr(rot), int(image_size[0]),
              int(image_size[1]), _ptr(M), N, _stream())
    return M


class SourceImages:
    """Device-resident source images for the crop: per-person pointer/shape tables.

    images: list of uint8 CUDA tensors [Hs,Ws,3] (HWC, last two dims dense), or one uint8
    tensor [P,Hs,Ws,3]; index: which image each person uses (len N)."""

    def __init__(self, images, index):
        if isinstance(images, torch.Tensor) and images.dim() == 4:
            images = [images[i] for i in range(images.shape[0])]
        self.images = [_dev(im, torch.uint8) for im in images]
        for im in self.images:
            assert im.dim() == 3 and im.shape[2] == 3 and im.stride(2) == 1 and im.stride(1) == 3
        index = np.asarray(index.cpu() if isinstance(index, torch.Tensor) else index, dtype=np.int64)
        dev = self.images[0].device
        ptrs = np.array([self.images[i].data_ptr() for i in index], dtype=np.int64)
        hs = np.array([self.images[i].shape[0] for i in index], dtype=np.int32)
        ws = np.array([self.images[i].shape[1] for i in index], dtype=np.int32)
        st = np.array([self.images[i].stride(0) for i in index], dtype=np.int32)
        self.ptrs = torch.from_numpy(ptrs).to(dev)
        self.h = torch.from_numpy(hs).to(dev)
        self.w = torch.from_numpy(ws).to(dev)
        self.stride = torch.from_numpy(st).to(dev)
        self.n = len(index)

    @classmethod
    def from_tables(cls, images, ptrs, h, w, stride):
        """Wrap per-person tables that are already on the device (ptrs int64 [N], h/w/stride int32
        [N]); `images` is kept only to hold the storage alive.  Used by the host pipeline, which
        uploads the pointer table with each chunk instead of building it per call."""
        o = cls.__new__(cls)
        o.images, o.ptrs, o.h, o.w, o.stride, o.n = images, ptrs, h, w, stride, int(ptrs.numel())
        return o


def crop_batch(src, M, image_size, out=None):
    """a2 cv2.warpAffine(img, M, (W,H), WARP_INVERSE_MAP|INTER_LINEAR) for N persons.
    src: SourceImages; M float32 [N,2,3] -> uint8 [N,H,W,3]."""
    W, H = int(image_size[0]), int(image_size[1])
    M = _dev(M, _F32)
    N = M.shape[0]
    assert src.n == N
    if out is None:
        out = torch.empty((N, H, W, 3), dtype=torch.uint8, device=M.device)
    _lib.call("udp_warp_affine_u8", _ptr(src.ptrs), _ptr(src.h), _ptr(src.w), _ptr(src.stride),
              _ptr(M), _ptr(out), W, H, N, _stream())
    return out


def copy2d_async(dst_ptr, dpitch, src_ptr, spitch, width, height, to_device=True):
    """cudaMemcpy2DAsync on torch's current stream (udp_copy2d_async): `height` rows of `width` bytes."""
    _lib.call("udp_copy2d_async", C.c_void_p(dst_ptr), int(dpitch), C.c_void_p(src_ptr), int(spitch), int(width),
              int(height), 1 if to_device else 2, _stream())


def host_device_pointer(t):
    """Device-side address of a pinned (mapped) host tensor; raises if the tensor is not mapped."""
    assert not t.is_cuda and t.is_pinned()
    out = C.c_void_p()
    _lib.call("udp_host_device_pointer", C.c_void_p(t.data_ptr()), C.byref(out))
    return int(out.value)


IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)  # tools/train.py:127-129


def crop_normalize_batch(src, M, image_size, mean=IMAGENET_MEAN, std=IMAGENET_STD, want_flipped=False,
                         out=None, out_flipped=None):
    """a2 with ToTensor + Normalize fused (tools/train.py:127-143,
    JointsDataset.py:237-238) -> float32 [N,3,H,W]; with want_flipped also np.flip(input, 3)
    (core/function.py:154) -- the two network inputs of the flip test from one gather."""
    W, H = int(image_size[0]), int(image_size[1])
    M = _dev(M, _F32)
    N = M.shape[0]
    assert src.n == N
    if out is None:
        out = torch.empty((N, 3, H, W), dtype=_F32, device=M.device)
    if want_flipped and out_flipped is None:
        out_flipped = torch.empty((N, 3, H, W), dtype=_F32, device=M.device)
    mean_c = (C.c_float * 3)(*[float(v) for v in mean])
    std_c = (C.c_float * 3)(*[float(v) for v in std])
    _lib.call("udp_warp_affine_normalize_f32", _ptr(src.ptrs), _ptr(src.h), _ptr(src.w), _ptr(src.stride),
              _ptr(M), C.cast(mean_c, C.c_void_p), C.cast(std_c, C.c_void_p), _ptr(out),
              _ptr(out_flipped) if want_flipped else None, W, H, N, _stream())
    return (out, out_flipped) if want_flipped else out


def crop_launches(image_size):
    """Kernel launches one crop_batch() makes for this crop size (TMA kernel + cp.async kernel, or one)."""
    return int(_lib.load().udp_warp_affine_launches(int(image_size[0]), int(image_size[1])))


def _joints3(joints, vis):
    """generate_target reads joints[:, 0:2] of a [K,3] array (JointsDataset.py:313-314): the C ABI takes
    rows of 3 doubles.  [N,K,2] coordinates (rotate_points' own output) are widened here."""
    joints, vis = _dev(joints, _F64), _dev(vis, _F64)
    if joints.dim() != 3 or joints.shape[2] not in (2, 3):
        raise ValueError("joints must be [N,K,2] or [N,K,3], got %s" % (tuple(joints.shape),))
    if joints.shape[2] == 2:
        joints = torch.cat([joints, torch.zeros_like(joints[..., :1])], 2).contiguous()
    if tuple(vis.shape) != (joints.shape[0], joints.shape[1], 3):
        raise ValueError("joints_vis must be [N,K,3], got %s" % (tuple(vis.shape),))
    return joints, vis


def rotate_points_batch(joints, rot_deg, center, scale, image_size, do_clip=False, out=None):
    """a3 rotate_points on joints [N,K,>=2] float64 -> float64 [N,K,2] (or into `out`
    [N,K,>=2], whose columns 0,1 are written)."""
    joints = _dev(joints, _F64)
    center, scale, f64 = _cs(center, scale)
    N, K, js = joints.shape
    rot = _dev(rot_deg, _F64) if rot_deg is not None else None
    if out is None:
        out = torch.empty((N, K, 2), dtype=_F64, device=joints.device)
    assert out.dtype == _F64 and out.is_contiguous() and out.shape[:2] == (N, K)
    _lib.call("udp_rotate_points", _ptr(joints), js, _ptr(rot), _ptr(center), _ptr(scale), f64,
              int(image_size[0]), int(image_size[1]), int(do_clip), _ptr(out), out.shape[2], K, N,
              _stream())
    return out


def rsn_transform_joints_batch(joints, vis, rot_deg, center, scale, image_size):
    """a3' RSN joint map + visibility cull.  Returns new (joints [N,K,3], vis [N,K,3], Minv [N,2,3])."""
    joints = _dev(joints, _F64).clone()
    vis = _dev(vis, _F64).clone()
    center = _dev(center, _F32)
    scale = _dev(scale, _F32)
    N, K, js = joints.shape
    rot = _dev(rot_deg, _F64) if rot_deg is not None else None
    Minv = torch.empty((N, 2, 3), dtype=_F32, device=joints.device)
    _lib.call("udp_rsn_transform_joints", _ptr(joints), js, _ptr(vis), vis.shape[2], _ptr(rot),
              _ptr(center), _ptr(scale), int(image_size[0]), int(image_size[1]), _ptr(Minv), K, N,
              _stream())
    return joints, vis, Minv


def fliplr_joints_batch(joints, vis, width, pairs, mul_vis=True):
    """a4 fliplr_joints for N persons; width int or [N]."""
    joints = _dev(joints, _F64)
    vis = _dev(vis, _F64)
    N, K, _ = joints.shape
    if np.isscalar(width):
        width = torch.full((N,), int(width), dtype=torch.int32, device=joints.device)
    width = _dev(width, torch.int32)
    perm = flip_perm(K, pairs, joints.device)
    jo, vo = torch.empty_like(joints), torch.empty_like(vis)
    _lib.call("udp_fliplr_joints", _ptr(joints), _ptr(vis), _ptr(perm), _ptr(width), K, N,
              int(mul_vis), _ptr(jo), _ptr(vo), _stream())
    return jo, vo


# -------------------------------------------------------------------------- encode side
def encode_gaussian(joints, vis, image_size, heatmap_size, sigma, joints_weight=None, out=None):
    """a5 -> (target f32 [N,K,h,w], target_weight f32 [N,K,1])."""
    joints, vis = _joints3(joints, vis)
    N, K, _ = joints.shape
    (W, H), (w, h) = image_size, heatmap_size
    tgt = out if out is not None else torch.empty((N, K, h, w), dtype=_F32, device=joints.device)
    tw = torch.empty((N, K, 1), dtype=_F32, device=joints.device)
    jw = _dev(joints_weight, _F32).reshape(-1) if joints_weight is not None else None
    _lib.call("udp_encode_gaussian", _ptr(joints), _ptr(vis), _ptr(jw), K, int(W), int(H), int(w),
              int(h), float(sigma), _ptr(tgt), _ptr(tw), N, _stream())
    return tgt, tw


def encode_offset(joints, vis, image_size, heatmap_size, kpd, joints_weight=None, out=None):
    """a6 -> (target f32 [N,3K,h,w], target_weight f32 [N,K,1])."""
    joints, vis = _joints3(joints, vis)
    N, K, _ = joints.shape
    (W, H), (w, h) = image_size, heatmap_size
    tgt = out if out is not None else torch.empty((N, 3 * K, h, w), dtype=_F32, device=joints.device)
    tw = torch.empty((N, K, 1), dtype=_F32, device=joints.device)
    jw = _dev(joints_weight, _F32).reshape(-1) if joints_weight is not None else None
    _lib.call("udp_encode_offset", _ptr(joints), _ptr(vis), _ptr(jw), K, int(W), int(H), int(w),
              int(h), float(kpd), _ptr(tgt), _ptr(tw), N, _stream())
    return tgt, tw


def encode_rsn(joints, valid, image_size, heatmap_size, sigmas, out=None):
    """a7 -> labels f32 [N,S,K,h,w].  joints [N,K,>=2] f64 crop coords, valid [N,K]."""
    joints = _dev(joints, _F64)
    N, K, js = joints.shape
    valid = _dev(valid, _F32).reshape(N, K)
    (W, H), (w, h) = image_size, heatmap_size
    S = len(sigmas)
    sig = (C.c_double * S)(*[float(s) for s in sigmas])
    lab = out if out is not None else torch.empty((N, S, K, h, w), dtype=_F32, device=joints.device)
    _lib.call("udp_encode_rsn", _ptr(joints), js, _ptr(valid), K, int(W), int(H), int(w), int(h),
              C.cast(sig, C.c_void_p), S, _ptr(lab), N, _stream())
    return lab


def encode_batch(cfg, joints, vis, joints_weight=None, out=None):
    """Target encode for a CodecConfig (udp_pose_b200.configs)."""
    if cfg.target_type == "gaussian":
        return encode_gaussian(joints, vis, cfg.image_size, cfg.heatmap_size, cfg.sigma, joints_weight, out)
    if cfg.target_type == "offset":
        return encode_offset(joints, vis, cfg.image_size, cfg.heatmap_size, cfg.kpd, joints_weight, out)
    from .configs import RSN_SIGMAS
    v = vis[..., 0] if vis.dim() == 3 else vis
    return encode_rsn(joints, v, cfg.image_size, cfg.heatmap_size, RSN_SIGMAS, out), None


# -------------------------------------------------------------------------- decode side
def flip_back(flipped, pairs, triplets=False, out=None):
    """a9/a10 (out=None) or a9-a11 fused: (out + flip_back(flipped)) * 0.5."""
    flipped = _dev(flipped, _F32)
    N, Cc, h, w = flipped.shape
    K = Cc // 3 if triplets else Cc
    perm = flip_perm(K, pairs, flipped.device)
    o = _dev(out, _F32) if out is not None else None
    dst = torch.empty_like(flipped)
    _lib.call("udp_flip_back", _ptr(o), _ptr(flipped), _ptr(perm), K, w, h, int(triplets), _ptr(dst),
              N, _stream())
    return dst


def decode_gaussian(hm, center, scale, flipped=None, pairs=None, post_process=True, pre_div=0.0,
                    want_idx=False, want_packed=False, want_blurred=False):
    """a9,a11-a14,a16(,a17): fused flip-average + get_final_preds('gaussian').
    center/scale None -> heatmap-space coordinates (get_max_preds / post).
    Returns dict(preds, maxvals, preds_in[, idx, packed, blurred])."""
    hm = _dev(hm, _F32)
    N, K, h, w = hm.shape
    dev = hm.device
    fl = perm = None
    if flipped is not None:
        fl = _dev(flipped, _F32)
        assert fl.shape == hm.shape
        perm = flip_perm(K, pairs or [], dev)
    c = s = None
    f64 = 0
    if center is not None:
        c, s, f64 = _cs(center, scale)
    dt = _F64 if post_process else _F32
    preds = torch.empty((N, K, 2), dtype=dt, device=dev)
    pin = torch.empty((N, K, 2), dtype=dt, device=dev)
    mv = torch.empty((N, K, 1), dtype=_F32, device=dev)
    idx = torch.empty((N, K), dtype=torch.int32, device=dev) if want_idx else None
    packed = torch.empty((N, K, 3), dtype=_F32, device=dev) if want_packed else None
    bl = torch.empty_like(hm) if want_blurred else None
    _lib.call("udp_decode_gaussian", _ptr(hm), _ptr(fl), _ptr(perm), _ptr(c), _ptr(s), f64,
              float(pre_div), int(bool(post_process)), K, w, h, _ptr(preds), _ptr(mv), _ptr(pin),
              _ptr(idx), _ptr(packed), _ptr(bl), N, _stream())
    r = dict(preds=preds, maxvals=mv, preds_in=pin)
    if want_idx:
        r["idx"] = idx
    if want_packed:
        r["packed"] = packed
    if want_blurred:
        r["blurred"] = bl
    return r


def decode_offset(out, center, scale, kpd, flipped=None, pairs=None, want_idx=False,
                  want_packed=False, off_ptrs=None):
    """a10,a11,a15,a16: fused flip-average + get_final_preds('offset').
    off_ptrs = (address of the tensor the offset windows of `out` are gathered from, same for `flipped`):
    udp_decode_offset_split -- `out` / `flipped` then need to hold the classification channels only."""
    out = _dev(out, _F32)
    N, C3, h, w = out.shape
    K = C3 // 3
    dev = out.device
    fl = perm = None
    if flipped is not None:
        fl = _dev(flipped, _F32)
        assert fl.shape == out.shape
        perm = flip_perm(K, pairs or [], dev)
    c = s = None
    f64 = 0
    if center is not None:
        c, s, f64 = _cs(center, scale)
    preds = torch.empty((N, K, 2), dtype=_F32, device=dev)
    pin = torch.empty((N, K, 2), dtype=_F32, device=dev)
    mv = torch.empty((N, K, 1), dtype=_F32, device=dev)
    idx = torch.empty((N, K), dtype=torch.int32, device=dev) if want_idx else None
    packed = torch.empty((N, K, 3), dtype=_F32, device=dev) if want_packed else None
    if off_ptrs is not None:
        _lib.call("udp_decode_offset_split", _ptr(out), _ptr(fl), C.c_void_p(off_ptrs[0]),
                  C.c_void_p(off_ptrs[1]) if fl is not None else None, _ptr(perm), float(kpd), _ptr(c), _ptr(s),
                  f64, K, w, h, _ptr(preds), _ptr(mv), _ptr(pin), _ptr(idx), _ptr(packed), N, _stream())
    else:
        _lib.call("udp_decode_offset", _ptr(out), _ptr(fl), _ptr(perm), float(kpd), _ptr(c), _ptr(s),
                  f64, K, w, h, _ptr(preds), _ptr(mv), _ptr(pin), _ptr(idx), _ptr(packed), N, _stream())
    r = dict(preds=preds, maxvals=mv, preds_in=pin)
    if want_idx:
        r["idx"] = idx
    if want_packed:
        r["packed"] = packed
    return r


def decode_batch(cfg, out, flipped, center, scale, post_process=None, **kw):
    """Decode for a CodecConfig; `flipped` may be None (no flip test)."""
    if cfg.target_type == "offset":
        return decode_offset(out, center, scale, cfg.kpd, flipped, cfg.flip_pairs, **kw)
    pp = cfg.post_process if post_process is None else post_process
    return decode_gaussian(out, center, scale, flipped, cfg.flip_pairs, pp, cfg.pre_div, **kw)


def transform_preds_batch(coords, center, scale, heatmap_size):
    """a16 transform_preds for coords [N,K,2] (float32 or float64) -> float64 [N,K,2]."""
    coords = _dev(coords)
    if coords.dtype not in (_F32, _F64):
        coords = coords.to(_F64)
    center, scale, f64 = _cs(center, scale)
    N, K, _ = coords.shape
    out = torch.empty((N, K, 2), dtype=_F64, device=coords.device)
    _lib.call("udp_transform_preds", _ptr(coords), int(coords.dtype == _F64), _ptr(center),
              _ptr(scale), f64, int(heatmap_size[0]), int(heatmap_size[1]), _ptr(out), K, N, _stream())
    return out


def accuracy_pck(output, target, ch_stride=1, thr=0.5):
    """accuracy() of lib/core/evaluate.py:41-71 on device tensors.
    output, target float32 CUDA [N, K*ch_stride, h, w] (offset mode: ch_stride=3, the [:, ::3]
    views of function.py:190-191).  Returns dict(acc f64 [K+1], avg_acc f64 [1], cnt i32 [1],
    pred f32 [N,K,2], dists f64 [K,N]) -- CUDA tensors, nothing synchronised."""
    output, target = _dev(output, _F32), _dev(target, _F32)
    assert output.shape == target.shape and output.dim() == 4 and output.shape[1] % ch_stride == 0
    N, C, h, w = output.shape
    K = C // ch_stride
    dev = output.device
    pred = torch.empty((N, K, 2), dtype=_F32, device=dev)
    dists = torch.empty((K, N), dtype=_F64, device=dev)
    acc = torch.empty((K + 1,), dtype=_F64, device=dev)
    avg = torch.empty((1,), dtype=_F64, device=dev)
    cnt = torch.empty((1,), dtype=torch.int32, device=dev)
    _lib.call("udp_accuracy_pck", _ptr(output), _ptr(target), int(ch_stride), K, int(w), int(h), N, float(thr),
              _ptr(pred), _ptr(dists), _ptr(acc), _ptr(avg), _ptr(cnt), _stream())
    return dict(acc=acc, avg_acc=avg, cnt=cnt, pred=pred, dists=dists)


def pack_all_boxes(center, scale, score=None):
    """lib/core/function.py:211-214: all_boxes rows (cx, cy, sx, sy, prod(s * 200), score) -> float64 CUDA [N,6]."""
    center, scale, f64 = _cs(center, scale)
    N = center.shape[0]
    sc = _dev(score, _F64) if score is not None else None
    out = torch.empty((N, 6), dtype=_F64, device=center.device)
    _lib.call("udp_pack_all_boxes", _ptr(center), _ptr(scale), f64, _ptr(sc), _ptr(out), N, _stream())
    return out


def coco_rescore(preds, box_score, in_vis_thre):
    """the rescoring of lib/dataset/coco.py:326-339.  preds f32 CUDA [N,K,3]
    (x, y, maxval), box_score f64 [N] -> score f64 CUDA [N]."""
    preds, box_score = _dev(preds, _F32), _dev(box_score, _F64)
    N, K, _ = preds.shape
    score = torch.empty((N,), dtype=_F64, device=preds.device)
    _lib.call("udp_coco_rescore", _ptr(preds), _ptr(box_score), int(K), int(N), float(in_vis_thre), _ptr(score), _stream())
    return score


def oks_nms_batch(preds, area, score, img_off, max_per_image, thresh, soft=False, sigmas=None, in_vis_thre=None,
                  max_dets=20, want_oks=False):
    """oks_nms / soft_oks_nms of lib/nms/nms.py:97-180 for every image at once.
    preds f32 [N,K,3], area/score f64 [N], img_off int32 [I+1] (persons of image i are rows
    img_off[i]:img_off[i+1]); max_per_image is a host int (no device sync here).  sigmas: host
    sequence of K floats (already /10) or None = COCO's.  Returns dict(keep i32 [N] -- image i's
    kept global rows, in keep order, start at keep[img_off[i]] -- keep_cnt i32 [I], oks_top f64 [N]
    or None); CUDA tensors, nothing synchronised."""
    preds, area, score = _dev(preds, _F32), _dev(area, _F64), _dev(score, _F64)
    img_off = _dev(img_off, torch.int32)
    N, K, _ = preds.shape
    I = img_off.numel() - 1
    dev = preds.device
    keep = torch.full((N,), -1, dtype=torch.int32, device=dev)
    keep_cnt = torch.zeros((max(I, 0),), dtype=torch.int32, device=dev)
    oks = torch.full((N,), float("nan"), dtype=_F64, device=dev) if want_oks else None
    sig = None
    if sigmas is not None:
        sig = np.ascontiguousarray(sigmas, np.float64)
        assert sig.shape == (K,)
    _lib.call("udp_oks_nms", _ptr(preds), _ptr(area), _ptr(score), _ptr(img_off), int(I), int(max_per_image), int(K),
              float(thresh), sig.ctypes.data_as(C.c_void_p) if sig is not None else None,
              int(in_vis_thre is not None), float(in_vis_thre or 0.0), int(bool(soft)), int(max_dets),
              _ptr(keep), _ptr(keep_cnt), _ptr(oks), _stream())
    return dict(keep=keep, keep_cnt=keep_cnt, oks_top=oks)


def decode_quarter(heatmaps, boxes, want_float=False):
    """SURVEY 8(f)-4 (first half): get_final_preds of tools/infer_utils/decode.py:19-40 (arg-max,
    +-0.25 towards the higher neighbour, scale/output_size affine, truncation).  heatmaps f32 CUDA
    [N,K,h,w]; boxes [N,4] (cx, cy, sx, sy) float32 or float64 (dtype kept: it selects the arithmetic).
    Returns int32 CUDA [N,K,2] (and the float32 values before truncation, and maxvals [N,K])."""
    heatmaps = _dev(heatmaps, _F32)
    boxes = _dev(boxes)
    if boxes.dtype not in (_F32, _F64):
        boxes = boxes.to(_F32)
    N, K, h, w = heatmaps.shape
    assert boxes.shape == (N, 4)
    dev = heatmaps.device
    out = torch.empty((N, K, 2), dtype=torch.int32, device=dev)
    cf = torch.empty((N, K, 2), dtype=_F32, device=dev) if want_float else None
    mv = torch.empty((N, K), dtype=_F32, device=dev) if want_float else None
    _lib.call("udp_decode_quarter", _ptr(heatmaps), _ptr(boxes), int(boxes.dtype == _F64), int(K), int(w), int(h), int(N),
              _ptr(out), _ptr(cf), _ptr(mv), _stream())
    return (out, cf, mv) if want_float else out


# cv2.getGaussianKernel(n, 0, cv2.CV_64F)[: n // 2 + 1] (OpenCV 4.x: fixed tables up to 9, bit-exact exp beyond;
# tests/test_rsn_legacy.py checks them against the installed cv2)
_CV_GAUSS_F64 = {
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


def cv_gauss_taps_f64(n):
    if n not in _CV_GAUSS_F64:
        raise ValueError("GAUSSIAN_KERNEL %r: supported sizes are %s" % (n, sorted(_CV_GAUSS_F64)))
    half = [float.fromhex(v) for v in _CV_GAUSS_F64[n]]
    return np.array(half + half[-2::-1], np.float64)


def rsn_get_results(outputs, centers, scales, input_shape, kernel=11, shift=0.25, outputs_flipped=None, flip_pairs=None):
    """SURVEY 8(f)-4 (second half): get_results of RSN/exps/RSN18.coco/test.py:25-71 for a batch, the flip
    average of :88-96 fused when outputs_flipped is given.  outputs f32 CUDA [N,K,h,w]; centers/scales f32
    [N,2] (scales as the loader yields them, not yet x200); input_shape = (H, W) = cfg.INPUT_SHAPE.
    Returns (preds f64 [N,K,2], maxvals f64 [N,K,1]) CUDA tensors."""
    outputs = _dev(outputs, _F32)
    centers, scales = _dev(centers, _F32), _dev(scales, _F32)
    N, K, h, w = outputs.shape
    fl = perm = None
    if outputs_flipped is not None:
        fl = _dev(outputs_flipped, _F32)
        assert fl.shape == outputs.shape
        perm = flip_perm(K, flip_pairs or [], outputs.device)
    taps = cv_gauss_taps_f64(int(kernel))
    preds = torch.empty((N, K, 2), dtype=_F64, device=outputs.device)
    maxvals = torch.empty((N, K, 1), dtype=_F64, device=outputs.device)
    _lib.call("udp_rsn_get_results", _ptr(outputs), _ptr(fl), _ptr(perm), _ptr(centers), _ptr(scales), int(K), int(w), int(h),
              int(input_shape[1]), int(input_shape[0]), int(kernel), taps.ctypes.data_as(C.c_void_p), float(shift),
              _ptr(preds), _ptr(maxvals), int(N), _stream())
    return preds, maxvals


AID_MASK_WORDS = 16  # UDP_AID_MASK_WORDS of include/udp_codec.h: 512 cells


def aid_erase(crops, ellipses=None, n_ellipses=None, grid_size=None, hide_mask=None, min_grid=0):
    """apply drawn Cutout / HideAndSeek parameters to uint8 CUDA crops [N,H,W,3] IN PLACE
    (lib/utils/transforms.py:144-223 at JointsDataset.py:230-235).  ellipses f64 [N,P,4] (cx, cy, rx, ry),
    n_ellipses i32 [N]; grid_size i32 [N], hide_mask u32/i32 [N,AID_MASK_WORDS] (bit c = cell c in draw order)."""
    assert crops.is_cuda and crops.dtype == torch.uint8 and crops.dim() == 4 and crops.shape[3] == 3 and crops.is_contiguous()
    N, H, W, _ = crops.shape
    P = 0
    if ellipses is not None:
        ellipses, n_ellipses = _dev(ellipses, _F64), _dev(n_ellipses, torch.int32)
        P = ellipses.shape[1]
        assert ellipses.shape == (N, P, 4) and n_ellipses.shape == (N,)
    if grid_size is not None:
        grid_size = _dev(grid_size, torch.int32)
        hide_mask = _dev(hide_mask, torch.int32)
        assert grid_size.shape == (N,) and hide_mask.shape == (N, AID_MASK_WORDS)
    _lib.call("udp_aid_erase", _ptr(crops), int(N), int(H), int(W), _ptr(ellipses), _ptr(n_ellipses), int(P),
              _ptr(grid_size), _ptr(hide_mask), int(min_grid), _stream())
    return crops
