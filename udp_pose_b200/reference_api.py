"""The reference's own call surface for the codec path, served by the CUDA library.

Every name below has the signature, argument meaning, return dtypes and error behaviour of
the reference function it replaces; numpy arrays in, numpy arrays out.  Each
call uploads its arguments, launches the kernels through the C ABI (udp_pose_b200.device_api
-> libudp_codec.so) and downloads the result.  There is no CPU implementation behind these
names: without a GPU or without the built library they raise (udp_pose_b200._lib.load).

`install()` substitutes them into an imported deep_hrnet/lib tree (lib.core.inference,
lib.utils.transforms, lib.dataset.JointsDataset); INTEGRATION.md shows the maintainer-side
patch.  Differences from the reference that a caller can observe are listed in each
docstring; the only deliberate ones are (1) inputs are never mutated except where the
reference's callers rely on it (`post` still writes the blurred maps back), and (2)
`get_final_preds` accepts CUDA tensors too and then skips the host round trip.

For throughput use the batched device API directly: one launch per batch instead of one
Python call per person (the reference's DataLoader-worker model).
"""
import copy
import random

import numpy as np
import torch

from . import device_api as D

RSN_KERNEL_SIGMA = {15: 2.6, 11: 2.0, 9: 1.7, 7: 1.4105, 5: 1.1105}  # RSN/dataset/JointsDataset.py:19-25


def _np(t):
    return t.detach().cpu().numpy()


def _cs1(center, scale):
    """One person's center/scale -> [1,2] arrays of one float dtype (float32 kept, else float64)."""
    c, s = np.asarray(center), np.asarray(scale)
    dt = np.float32 if (c.dtype == np.float32 and s.dtype == np.float32) else np.float64
    return c.astype(dt).reshape(1, 2), s.astype(dt).reshape(1, 2)


def _csN(center, scale):
    c, s = np.asarray(center), np.asarray(scale)
    dt = np.float32 if (c.dtype == np.float32 and s.dtype == np.float32) else np.float64
    return np.ascontiguousarray(c, dt).reshape(-1, 2), np.ascontiguousarray(s, dt).reshape(-1, 2)


# ------------------------------------------------------------------------------- crop side
def get_warpmatrix(theta, size_input, size_dst, size_target):
    """deep_hrnet/lib/dataset/JointsDataset.py:29-49.  theta degrees, size_input = 2*center,
    size_dst = image_size - 1, size_target = scale (box/200).  -> float32 [2,3] (dst->src)."""
    center = np.asarray(size_input) * 0.5  # exact inverse of the caller's c*2.0 (:226)
    c, s = _cs1(center.astype(np.asarray(size_input).dtype, copy=False), size_target)
    dst = np.asarray(size_dst, np.float64)
    W, H = int(round(float(dst[0]) + 1.0)), int(round(float(dst[1]) + 1.0))
    M = D.warp_matrices(c, s, np.array([float(theta)]), (W, H))
    return _np(M)[0]


def rotate_points(src_points, angle, c, dst_img_shape, size_target, do_clip=True):
    """JointsDataset.py:51-73.  src_points [K,2] -> [K,2] of the same dtype."""
    pts = np.asarray(src_points)
    cc, ss = _cs1(c, size_target)
    out = D.rotate_points_batch(np.ascontiguousarray(pts[None, :, :2], np.float64),
                                np.array([float(angle)]), cc, ss,
                                (int(dst_img_shape[0]), int(dst_img_shape[1])), do_clip=bool(do_clip))
    return _np(out)[0].astype(pts.dtype, copy=False)


def warp_affine(data_numpy, trans, dsize):
    """cv2.warpAffine(data_numpy, trans, dsize, flags=WARP_INVERSE_MAP|INTER_LINEAR) for a
    uint8 HWC3 image (call sites JointsDataset.py:227, RSN/dataset/JointsDataset.py:197)."""
    img = np.asarray(data_numpy)
    if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
        raise ValueError("warp_affine: expected a uint8 [H,W,3] image")
    src = D.SourceImages([torch.from_numpy(np.ascontiguousarray(img))], [0])
    M = np.ascontiguousarray(np.asarray(trans, np.float32).reshape(1, 2, 3))
    return _np(D.crop_batch(src, M, (int(dsize[0]), int(dsize[1]))))[0]


def fliplr_joints(joints, joints_vis, width, matched_parts):
    """deep_hrnet/lib/utils/transforms.py:50-64 -> (joints*joints_vis, joints_vis).
    The reference also flips its arguments in place; this does not."""
    j, v = D.fliplr_joints_batch(np.asarray(joints, np.float64)[None], np.asarray(joints_vis, np.float64)[None],
                                 int(width), matched_parts, mul_vis=True)
    return _np(j)[0], _np(v)[0]


def flip_joints(joints, joints_vis, width, pairs):
    """RSN/lib/utils/transforms.py:70-79 (no visibility product)."""
    j, v = D.fliplr_joints_batch(np.asarray(joints, np.float64)[None], np.asarray(joints_vis, np.float64)[None],
                                 int(width), pairs, mul_vis=False)
    return _np(j)[0], _np(v)[0]


# ----------------------------------------------------------------------------- encode side
class Cutout:
    """lib/utils/transforms.py:186-223, same constructor.  `draw(height, width)` consumes numpy's global
    generator exactly as the reference's __call__ does (one rand() for the gate, then per patch two for
    the centre and rand(2) for the radii) and returns the drawn ellipses [(cx, cy, rx, ry), ...] (empty
    when the gate fails); the erase itself runs on the device (udp_aid_erase)."""

    def __init__(self, prob=1.0, radius_factor=0.2, num_patch=1):
        self.prob, self.radius_factor, self.num_patch = prob, radius_factor, num_patch
        if num_patch > 8:
            raise ValueError("Cutout: at most 8 patches")

    def draw(self, height, width):
        out = []
        if np.random.rand() < self.prob:
            for _ in range(self.num_patch):
                center = [np.random.rand() * width, np.random.rand() * height]
                radius = self.radius_factor * (1 + np.random.rand(2)) * width
                out.append((center[0], center[1], radius[0], radius[1]))
        return out

    def __call__(self, img):
        return _aid_apply(img, [self.draw(img.shape[0], img.shape[1])], None)


class HideAndSeek:
    """lib/utils/transforms.py:144-183, same constructor.  `draw(height, width)` consumes numpy's global
    generator as the reference does (gate, randint for the grid -- which never picks the last size, :163 --
    then one rand() per cell, outer loop over the width) and returns (grid_size, [hidden?, ...])."""

    def __init__(self, prob=1.0, prob_hiding_patches=0.5, grid_sizes=(0, 16, 32, 44, 56)):
        self.prob, self.prob_hiding_patches, self.grid_sizes = prob, prob_hiding_patches, grid_sizes

    def draw(self, height, width):
        if not np.random.rand() < self.prob:
            return 0, []
        g = self.grid_sizes[np.random.randint(0, len(self.grid_sizes) - 1)]
        bits = []
        if g != 0:
            for _x in range(0, width, g):
                for _y in range(0, height, g):
                    bits.append(bool(np.random.rand() <= self.prob_hiding_patches))
        return g, bits

    def __call__(self, img):
        return _aid_apply(img, None, [self.draw(img.shape[0], img.shape[1])])


def _aid_pack(cut, hide, n):
    """Drawn parameters of n persons -> the arrays udp_aid_erase takes."""
    ell = ne = gs = mask = None
    min_grid = 0
    if cut is not None:
        P = max(1, max(len(c) for c in cut))
        ell, ne = np.ones((n, P, 4)), np.zeros(n, np.int32)
        for i, c in enumerate(cut):
            ne[i] = len(c)
            if c:
                ell[i, :len(c)] = np.asarray(c, np.float64)
    if hide is not None:
        gs, mask = np.zeros(n, np.int32), np.zeros((n, D.AID_MASK_WORDS), np.uint32)
        for i, (g, bits) in enumerate(hide):
            if g and any(bits):
                if len(bits) > 32 * D.AID_MASK_WORDS:
                    raise ValueError("HideAndSeek: more than %d cells (grid %d)" % (32 * D.AID_MASK_WORDS, g))
                gs[i] = g
                for c, b in enumerate(bits):
                    if b:
                        mask[i, c >> 5] |= np.uint32(1 << (c & 31))
        used = gs[gs > 0]
        if len(used):
            min_grid, mask = int(used.min()), mask.view(np.int32)
        else:
            gs = mask = None      # nobody drew a grid: nothing to hide
    return ell, ne, gs, mask, min_grid


def _aid_apply(img, cut, hide):
    """One uint8 HWC image (numpy) through the device erase; returns a new array (the reference
    writes into its argument and returns it)."""
    t = torch.from_numpy(np.ascontiguousarray(img)[None]).cuda()
    ell, ne, gs, mask, mg = _aid_pack(cut, hide, 1)
    D.aid_erase(t, ell, ne, gs, mask, mg)
    return _np(t)[0]


def _own_aid(obj, cls):
    """The reference's Cutout / HideAndSeek objects (lib/utils/transforms.py:144-223) carry the same
    constructor attributes but no draw(): rebuild ours from them (get_batch on a reference dataset)."""
    if obj is None or obj is False or hasattr(obj, "draw"):
        return obj or None
    if cls is Cutout:
        return Cutout(obj.prob, obj.radius_factor, obj.num_patch)
    return HideAndSeek(obj.prob, obj.prob_hiding_patches, obj.grid_sizes)


def half_body_transform(ds, joints, joints_vis):
    """JointsDataset.half_body_transform (:124-167): host-side box selection from the visible
    upper- or lower-body joints; uses np.random exactly once, like the reference."""
    vis = joints_vis[:, 0] > 0
    upper = [joints[k] for k in range(ds.num_joints) if vis[k] and k in ds.upper_body_ids]
    lower = [joints[k] for k in range(ds.num_joints) if vis[k] and k not in ds.upper_body_ids]
    pick = upper if (np.random.randn() < 0.5 and len(upper) > 2) else (lower if len(lower) > 2 else upper)
    if len(pick) < 2:
        return None, None
    pts = np.array(pick, dtype=np.float32)
    center = pts.mean(axis=0)[:2]
    lo, hi = np.amin(pts, axis=0), np.amax(pts, axis=0)
    bw, bh = hi[0] - lo[0], hi[1] - lo[1]
    if bw > ds.aspect_ratio * bh:
        bh = bw * 1.0 / ds.aspect_ratio
    elif bw < ds.aspect_ratio * bh:
        bw = bh * ds.aspect_ratio
    scale = np.array([bw * 1.0 / ds.pixel_std, bh * 1.0 / ds.pixel_std], dtype=np.float32) * 1.5
    return center, scale


class JointsDataset(torch.utils.data.Dataset):
    """deep_hrnet/lib/dataset/JointsDataset.py:75-385 with the crop, the joint map and the
    target generation on the GPU.  Constructor, attributes, `generate_target` and
    `__getitem__` follow the reference; `get_batch` is the batched form (one launch per
    stage for a whole list of records) that a B200 input pipeline should call instead."""

    def __init__(self, cfg, root, image_set, is_train, transform=None):
        self.num_joints = 0
        self.pixel_std = 200
        self.flip_pairs = []
        self.parent_ids = []
        self.is_train = is_train
        self.root = root
        self.image_set = image_set
        self.output_path = getattr(cfg, "OUTPUT_DIR", "")
        ds = cfg.DATASET
        self.data_format = getattr(ds, "DATA_FORMAT", "jpg")
        self.scale_factor = ds.SCALE_FACTOR
        self.rotation_factor = ds.ROT_FACTOR
        self.flip = ds.FLIP
        self.num_joints_half_body = ds.NUM_JOINTS_HALF_BODY
        self.prob_half_body = ds.PROB_HALF_BODY
        self.color_rgb = ds.COLOR_RGB
        # AID (JointsDataset.py:96-106): three config values = constructor arguments, anything else truthy = defaults
        cc, ch = getattr(ds, "CUTOUT", None), getattr(ds, "HIDE_AND_SEEK", None)
        self.cutout = (Cutout(*cc) if len(cc) == 3 else Cutout()) if cc else None
        self.hide_and_seek = (HideAndSeek(*ch) if len(ch) == 3 else HideAndSeek()) if ch else None
        if self.hide_and_seek:  # fail here, not at a random draw during training: every grid must fit the device mask
            W_, H_ = int(cfg.MODEL.IMAGE_SIZE[0]), int(cfg.MODEL.IMAGE_SIZE[1])
            for g in self.hide_and_seek.grid_sizes[:-1]:  # (the last size is never drawn, transforms.py:163)
                if g and -(-W_ // g) * -(-H_ // g) > 32 * D.AID_MASK_WORDS:
                    raise ValueError("HideAndSeek: grid %d on a %dx%d crop has more than %d cells" % (g, H_, W_, 32 * D.AID_MASK_WORDS))
        self.target_type = cfg.MODEL.TARGET_TYPE
        self.image_size = np.array(cfg.MODEL.IMAGE_SIZE)
        self.heatmap_size = np.array(cfg.MODEL.HEATMAP_SIZE)
        self.sigma = cfg.MODEL.SIGMA
        self.use_different_joints_weight = cfg.LOSS.USE_DIFFERENT_JOINTS_WEIGHT
        self.joints_weight = 1
        self.kpd = cfg.LOSS.KPD
        self.transform = transform
        self.db = []

    def _get_db(self):
        raise NotImplementedError

    def evaluate(self, cfg, preds, output_dir, *args, **kwargs):
        raise NotImplementedError

    def __len__(self):
        return len(self.db)

    def half_body_transform(self, joints, joints_vis):
        return half_body_transform(self, joints, joints_vis)

    # -- a5 / a6 -----------------------------------------------------------------------
    def generate_target_batch(self, joints, joints_vis):
        """joints, joints_vis [N,K,3] -> CUDA tensors (target [N,C,h,w], target_weight [N,K,1])."""
        assert self.target_type in ("gaussian", "offset"), "Only support gaussian map and offset now!"
        jw = None
        if self.use_different_joints_weight:
            jw = np.ascontiguousarray(np.asarray(self.joints_weight, np.float32).reshape(-1))
        img, hm = tuple(int(v) for v in self.image_size), tuple(int(v) for v in self.heatmap_size)
        if self.target_type == "gaussian":
            return D.encode_gaussian(joints, joints_vis, img, hm, float(self.sigma), jw)
        return D.encode_offset(joints, joints_vis, img, hm, float(self.kpd), jw)

    def generate_target(self, joints, joints_vis):
        """JointsDataset.py:291-385 -> (target float32 [K or 3K,h,w], target_weight float32 [K,1])."""
        t, tw = self.generate_target_batch(np.asarray(joints, np.float64)[None], np.asarray(joints_vis, np.float64)[None])
        return _np(t)[0], _np(tw)[0]

    # -- a8 ----------------------------------------------------------------------------
    def _read(self, image_file):
        import cv2
        if self.data_format == "zip":
            raise NotImplementedError("zip data format is outside the codec path")
        img = cv2.imread(image_file, cv2.IMREAD_COLOR | cv2.IMREAD_IGNORE_ORIENTATION)
        if img is None:
            raise ValueError("Fail to read {}".format(image_file))
        if self.color_rgb:
            img = cv2.cvtColor(img, cv2.COLOR_BGR2RGB)
        return img

    def _augment(self, rec, img):
        """The reference's random draws, in the reference's order (JointsDataset.py:206-224)."""
        joints, joints_vis = rec["joints_3d"], rec["joints_3d_vis"]
        c, s, r, flip_w = rec["center"], rec["scale"], 0, 0
        if self.is_train:
            if np.sum(joints_vis[:, 0]) > self.num_joints_half_body and np.random.rand() < self.prob_half_body:
                ch, sh = self.half_body_transform(joints, joints_vis)
                if ch is not None and sh is not None:
                    c, s = ch, sh
            sf, rf = self.scale_factor, self.rotation_factor
            s = s * np.clip(np.random.randn() * sf + 1, 1 - sf, 1 + sf)
            r = np.clip(np.random.randn() * rf, -rf * 2, rf * 2) if random.random() <= 0.6 else 0
            if self.flip and random.random() <= 0.5:
                img = img[:, ::-1, :]
                flip_w = img.shape[1]  # fliplr_joints (:222) runs on the device for the whole batch, see get_batch
                c = c.copy()
                c[0] = img.shape[1] - c[0] - 1
        return img, joints, joints_vis, c, s, r, flip_w

    def get_batch(self, indices, images=None, normalize=None, want_flipped=False):
        """Batched __getitem__: returns (input uint8 CUDA [N,H,W,3], target CUDA, target_weight
        CUDA, metas).  `images` (optional) supplies decoded uint8 arrays instead of files.
        normalize=(mean, std): the reference's `transforms.Compose([ToTensor(), Normalize(mean, std)])`
        (tools/train.py:127-143) is fused into the crop and `input` is float32 CUDA [N,3,H,W]
        (SURVEY 8f-1); with want_flipped it is the pair (input, np.flip(input, 3)) that the flip
        test feeds the model (core/function.py:154-155)."""
        W, H = int(self.image_size[0]), int(self.image_size[1])
        recs, imgs, js, vs, cs, ss, rs, fws = [], [], [], [], [], [], [], []
        cutout, hide = _own_aid(getattr(self, "cutout", None), Cutout), _own_aid(getattr(self, "hide_and_seek", None), HideAndSeek)
        aid = self.is_train and (cutout or hide)
        if aid and normalize is not None:
            raise NotImplementedError("AID erases the uint8 crop (JointsDataset.py:230-235): use the uint8 path, then normalise")
        cuts, hides = [], []
        for n, idx in enumerate(indices):
            rec = copy.deepcopy(self.db[idx])
            img = images[n] if images is not None else self._read(rec["image"])
            img, j, v, c, s, r, fw = self._augment(rec, img)
            if aid:  # drawn per record, after the warp parameters and in the reference's order (:231-235)
                cuts.append(cutout.draw(H, W) if cutout else [])
                hides.append(hide.draw(H, W) if hide else (0, []))
            recs.append(rec); imgs.append(torch.from_numpy(np.ascontiguousarray(img)))
            js.append(j); vs.append(v); cs.append(c); ss.append(s); rs.append(float(r)); fws.append(fw)
        c, s = _csN(np.stack(cs), np.stack(ss))
        rot = np.asarray(rs, np.float64)
        M = D.warp_matrices(c, s, rot, (W, H))
        src = D.SourceImages(imgs, np.arange(len(imgs)))
        if normalize is not None:
            inp = D.crop_normalize_batch(src, M, (W, H), normalize[0], normalize[1], want_flipped=want_flipped)
        else:
            inp = D.crop_batch(src, M, (W, H))
            if aid:
                ell, ne, gs, mask, mg = _aid_pack(cuts if cutout else None, hides if hide else None, len(recs))
                D.aid_erase(inp, ell, ne, gs, mask, mg)
        jd = torch.from_numpy(np.ascontiguousarray(np.stack(js), np.float64)).cuda()
        vd = torch.from_numpy(np.ascontiguousarray(np.stack(vs), np.float64)).cuda()
        fws = np.asarray(fws, np.int32)
        if fws.any():  # the flip branch (:220-224): fliplr_joints of the mirrored records, one launch
            sel = torch.from_numpy(np.nonzero(fws)[0]).cuda()
            jf, vf = D.fliplr_joints_batch(jd[sel], vd[sel], torch.from_numpy(fws[fws != 0]), self.flip_pairs, mul_vis=True)
            jd[sel], vd[sel] = jf, vf
        D.rotate_points_batch(jd, rot, c, s, (W, H), do_clip=False, out=jd)
        target, tw = self.generate_target_batch(jd, vd)
        jh, v = _np(jd), _np(vd)
        metas = []
        for n, rec in enumerate(recs):
            metas.append({"image": rec["image"], "filename": rec.get("filename", ""),
                          "imgnum": rec.get("imgnum", ""), "joints": jh[n], "joints_vis": v[n],
                          "center": cs[n], "scale": ss[n], "rotation": rs[n] if rs[n] else 0,
                          "score": rec.get("score", 1)})
        return inp, target, tw, metas

    def __getitem__(self, idx, image=None):
        inp, target, tw, metas = self.get_batch([idx], None if image is None else [image])
        x = _np(inp)[0]
        if self.transform:
            x = self.transform(x)
        return x, target[0].cpu(), tw[0].cpu(), metas[0]


class RSNJointsEncoder:
    """The codec half of RSN/dataset/JointsDataset.py (:193-220, :276-291): joint map with
    visibility cull, then the five full-map 255*exp() labels."""

    def __init__(self, keypoint_num, input_shape, output_shape, kernels=((15, 15), (11, 11), (9, 9), (7, 7), (5, 5))):
        self.keypoint_num = keypoint_num
        self.input_shape, self.output_shape = tuple(input_shape), tuple(output_shape)  # (H,W), (h,w)
        self.gaussian_kernels = list(kernels)

    def _sizes(self):
        return (self.input_shape[1], self.input_shape[0]), (self.output_shape[1], self.output_shape[0])

    def generate_heatmap(self, joints, valid, kernel=(7, 7)):
        """-> float32 [K,h,w]; valid [K] or [K,1] (drawn iff >= 1)."""
        img, hm = self._sizes()
        v = np.asarray(valid, np.float32).reshape(1, -1)[:, :self.keypoint_num]
        lab = D.encode_rsn(np.asarray(joints, np.float64)[None], v, img, hm, [RSN_KERNEL_SIGMA[kernel[0]]])
        return _np(lab)[0, 0]

    def labels(self, joints, valid):
        """All label scales at once -> float32 [S,K,h,w] (the :214-220 loop)."""
        img, hm = self._sizes()
        v = np.asarray(valid, np.float32).reshape(1, -1)[:, :self.keypoint_num]
        sig = [RSN_KERNEL_SIGMA[k[0]] for k in self.gaussian_kernels]
        return _np(D.encode_rsn(np.asarray(joints, np.float64)[None], v, img, hm, sig))[0]

    def transform_joints(self, joints, joints_vis, rotation, center, scale):
        """get_warpmatrix_inverse + affine_transform + cull (:203-211) -> (joints, joints_vis, trans_inv)."""
        img, _ = self._sizes()
        j, v, Mi = D.rsn_transform_joints_batch(np.asarray(joints, np.float64)[None], np.asarray(joints_vis, np.float64)[None],
                                                np.array([float(rotation)]), np.asarray(center, np.float32).reshape(1, 2),
                                                np.asarray(scale, np.float32).reshape(1, 2), img)
        return _np(j)[0], _np(v)[0], _np(Mi)[0]


# ----------------------------------------------------------------------------- decode side
def _check4(a, what):
    assert isinstance(a, np.ndarray), "%s should be numpy.ndarray" % what
    assert a.ndim == 4, "%s should be 4-ndim" % what


def flip_back(output_flipped, matched_parts):
    """deep_hrnet/lib/utils/transforms.py:15-29 (RSN/lib/utils/transforms.py:59-67).
    Returns a new array; the reference returns a view of its (mutated) argument."""
    assert output_flipped.ndim == 4, "output_flipped should be [batch_size, num_joints, height, width]"
    return _np(D.flip_back(np.ascontiguousarray(output_flipped, np.float32), matched_parts, False))


def flip_back_offset(output_flipped, matched_parts):
    """transforms.py:31-47: mirror, negate the x-offset channels, swap joint triplets."""
    assert output_flipped.ndim == 4, "output_flipped should be [batch_size, num_joints, height, width]"
    return _np(D.flip_back(np.ascontiguousarray(output_flipped, np.float32), matched_parts, True))


def get_max_preds(batch_heatmaps):
    """deep_hrnet/lib/core/inference.py:30-58 -> (preds float32 [B,K,2], maxvals float32 [B,K,1])."""
    _check4(batch_heatmaps, "batch_heatmaps")
    r = D.decode_gaussian(np.ascontiguousarray(batch_heatmaps, np.float32), None, None, post_process=False)
    return _np(r["preds"]), _np(r["maxvals"])


def post(coords, batch_heatmaps):
    """inference.py:60-145 -> float64 [B,K,2].  As in the reference the blurred, renormalised
    maps are written back into `batch_heatmaps`.  `coords` must be get_max_preds(batch_heatmaps)[0]
    -- the only way the reference calls it (:153-155); anything else raises ValueError."""
    _check4(batch_heatmaps, "batch_heatmaps")
    hm = np.ascontiguousarray(batch_heatmaps, np.float32)
    with np.errstate(all="ignore"):
        r = D.decode_gaussian(hm, None, None, post_process=True, want_idx=True, want_blurred=True)
        idx = _np(r["idx"])
        mv = _np(r["maxvals"])[..., 0]
        w = hm.shape[3]
        mine = np.stack([idx % w, idx // w], -1).astype(np.float32) * (mv > 0.0)[..., None]
        if not np.array_equal(mine, np.asarray(coords, np.float32)):
            raise ValueError("post(): coords is not get_max_preds(batch_heatmaps)[0]")
    if batch_heatmaps.dtype == np.float32:
        batch_heatmaps[...] = _np(r["blurred"])
    return _np(r["preds"])


def transform_preds(coords, center, scale, output_size):
    """inference.py:20-27 (= utils/transforms.py:67-74) -> float64 [K,2]."""
    co = np.asarray(coords)
    if co.dtype not in (np.float32, np.float64):
        co = co.astype(np.float64)
    c, s = _cs1(center, scale)
    return _np(D.transform_preds_batch(np.ascontiguousarray(co)[None], c, s, (int(output_size[0]), int(output_size[1]))))[0]


def get_final_preds(config, batch_heatmaps, center, scale, output_flipped=None, flip_pairs=None):
    """inference.py:149-186 -> (preds, maxvals, preds_in_input_space); dtypes as the reference
    (float64 coordinates with TEST.POST_PROCESS in gaussian mode, float32 otherwise).

    Extensions (both optional): `output_flipped` + `flip_pairs` fuse the flip test of
    lib/core/function.py:154-171 (flip_back[_offset] + average) into the same launch, and
    CUDA tensors are accepted for the heat maps, in which case nothing is copied to the host
    except the three small results."""
    if isinstance(batch_heatmaps, np.ndarray):
        assert batch_heatmaps.ndim == 4, "batch_images should be 4-ndim"
    c, s = _csN(center, scale)
    tt = config.MODEL.TARGET_TYPE
    if tt == "offset":
        r = D.decode_offset(batch_heatmaps, c, s, float(config.LOSS.KPD), output_flipped, flip_pairs)
    elif tt == "gaussian":
        r = D.decode_gaussian(batch_heatmaps, c, s, output_flipped, flip_pairs,
                              post_process=bool(config.TEST.POST_PROCESS))
    else:
        raise UnboundLocalError("local variable 'coords' referenced before assignment")  # what the reference does
    return _np(r["preds"]), _np(r["maxvals"]), _np(r["preds_in"])


def rsn_get_results(outputs, center, scale, output_flipped=None, flip_pairs=None):
    """RSN prm decode glue (RSN/exps/RSN18.coco.e1.se.36x8x132000_prm/test.py:183-196):
    outputs/255 -> get_max_preds -> post -> transform_preds; returns (preds f64 [B,K,2], maxvals [B,K,1])."""
    c, s = _csN(center, scale)
    r = D.decode_gaussian(outputs, c, s, output_flipped, flip_pairs, post_process=True, pre_div=255.0)
    return _np(r["preds"]), _np(r["maxvals"])


# --------------------------------------------------------------------------------- install
def accuracy(output, target, hm_type='gaussian', thr=0.5):
    """lib/core/evaluate.py:41-71 -> (acc float64 [K+1], avg_acc, cnt, pred float32 [N,K,2]).
    Accepts numpy arrays (as the reference) or CUDA tensors (skipping the .cpu().numpy() of
    function.py:76-80,187-192); strided [:, ::3] views of the offset mode are gathered on the
    device side.  Only hm_type='gaussian' defines `pred` in the reference (:51-55).  As in the reference,
    `thr` is accepted and ignored: accuracy() never passes it on to dist_acc (:63), whose default 0.5 applies."""
    if hm_type != 'gaussian':
        raise ValueError("accuracy(): the reference defines pred only for hm_type='gaussian' (evaluate.py:51-55)")
    o = output if isinstance(output, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(output, np.float32))
    t = target if isinstance(target, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(target, np.float32))
    # evaluate.py:63 calls dist_acc(dists[idx[i]]) without `thr`: the threshold is always dist_acc's default 0.5
    r = D.accuracy_pck(o.contiguous(), t.contiguous(), 1, 0.5)
    acc = _np(r["acc"])
    return acc, float(_np(r["avg_acc"])[0]) if int(_np(r["cnt"])[0]) else 0, int(_np(r["cnt"])[0]), _np(r["pred"])


def demo_get_final_preds(heatmaps, boxes):
    """tools/infer_utils/decode.py:19-40 `get_final_preds(heatmaps: Tensor, boxes: Tensor)` -- the
    quarter-offset decoder of the demo engines (named demo_* here because lib/core/inference.py owns
    the name get_final_preds).  heatmaps [B,C,H,W] and boxes [B,4] = (cx, cy, sx, sy): torch tensors
    (CPU or CUDA) or numpy arrays; returns int64 numpy [B,C,2] as the reference does."""
    hm = heatmaps if isinstance(heatmaps, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(heatmaps, np.float32))
    bx = boxes if isinstance(boxes, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(boxes))
    out = _np(D.decode_quarter(hm.float(), bx)).astype(np.int64)
    out[out == np.iinfo(np.int32).min] = np.iinfo(np.int64).min
    return out


def rsn_legacy_get_results(outputs, centers, scales, kernel=11, shifts=(0.25,), input_shape=None,
                           outputs_flipped=None, flip_pairs=None):
    """RSN/exps/RSN18.coco/test.py:25-71 `get_results(outputs, centers, scales, kernel, shifts)` (the
    non-prm RSN experiments).  The reference reads cfg.INPUT_SHAPE / OUTPUT_SHAPE / KEYPOINT.NUM from its
    experiment config: here the output shape and joint count come from `outputs`, and input_shape
    defaults to 4x the output shape (RSN/dataset/attribute.py:19-20, 62-63).  Only one shift ratio is
    supported (cfg.TEST.SHIFT_RATIOS = [0.25] in every experiment).  Unlike the reference (:26) the
    caller's `scales` array is not multiplied by 200 in place.  With outputs_flipped (already run through
    the model, NOT yet flipped back) the flip average of compute_on_dataset (:88-96) is fused.
    Returns (preds float64 [N,K,2], maxvals float64 [N,K,1])."""
    if len(shifts) != 1:
        raise NotImplementedError("get_results: one shift ratio (cfg.TEST.SHIFT_RATIOS = [0.25])")
    o = outputs if isinstance(outputs, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(outputs, np.float32))
    f = outputs_flipped
    if f is not None and not isinstance(f, torch.Tensor):
        f = torch.from_numpy(np.ascontiguousarray(f, np.float32))
    h, w = o.shape[2], o.shape[3]
    ish = tuple(input_shape) if input_shape is not None else (4 * h, 4 * w)
    c = np.ascontiguousarray(_np(centers) if isinstance(centers, torch.Tensor) else centers, np.float32)
    s = np.ascontiguousarray(_np(scales) if isinstance(scales, torch.Tensor) else scales, np.float32)
    preds, maxvals = D.rsn_get_results(o, c, s, ish, kernel, shifts[0], f, flip_pairs)
    return _np(preds), _np(maxvals)


def _nms_one(kpts_db, thresh, sigmas, in_vis_thre, soft):
    if len(kpts_db) == 0:
        return []
    preds = np.stack([np.asarray(k['keypoints'], np.float32).reshape(-1, 3) for k in kpts_db])
    area = np.array([k['area'] for k in kpts_db], np.float64)
    score = np.array([k['score'] for k in kpts_db], np.float64)
    n = len(kpts_db)
    r = D.oks_nms_batch(preds, area, score, np.array([0, n], np.int32), n, thresh, soft=soft,
                        sigmas=sigmas if isinstance(sigmas, np.ndarray) else None, in_vis_thre=in_vis_thre)
    cnt = int(_np(r["keep_cnt"])[0])
    if cnt < 0:
        raise D._lib.UdpCodecError("udp_oks_nms: image larger than max_per_image")
    return _np(r["keep"])[:cnt].astype(np.intp)


def oks_nms(kpts_db, thresh, sigmas=None, in_vis_thre=None):
    """lib/nms/nms.py:97-124: kpts_db = list of dicts with 'keypoints' [K,3], 'area', 'score';
    returns the kept indices in keep order ([] for an empty list; an index array otherwise, where the
    reference builds a python list of the same numbers).  As in the reference, a `sigmas` that is not
    an ndarray means COCO's defaults (:76-77)."""
    r = _nms_one(kpts_db, thresh, sigmas, in_vis_thre, False)
    return [int(i) for i in r]


def soft_oks_nms(kpts_db, thresh, sigmas=None, in_vis_thre=None):
    """lib/nms/nms.py:139-180: at most 20 kept indices (np.intp array, as the reference returns)."""
    return _nms_one(kpts_db, thresh, sigmas, in_vis_thre, True)


def coco_rescore_nms(preds, all_boxes, image_ids, in_vis_thre, oks_thre, soft_nms=False):
    """The rescoring + NMS block of COCODataset.evaluate (lib/dataset/coco.py:308-356) for the whole
    result set in two launches.  preds [N,K,3] (validate()'s all_preds), all_boxes [N,6]
    (cx, cy, sx, sy, area, box score), image_ids [N] ints.  Persons are grouped by image in
    first-appearance order, rows in input order (what the reference's defaultdict does, :317-319).
    Returns (score f64 [N], list of kept row lists, one per image in that order)."""
    preds = np.ascontiguousarray(preds, np.float32)
    all_boxes = np.asarray(all_boxes, np.float64)
    ids = np.asarray(image_ids)
    _, first, inv = np.unique(ids, return_index=True, return_inverse=True)
    rank = np.argsort(np.argsort(first))            # image -> position by first appearance
    grp = rank[inv]
    rows = np.argsort(grp, kind="stable")           # grouped row order
    counts = np.bincount(grp, minlength=len(first))
    off = np.zeros(len(first) + 1, np.int32)
    off[1:] = np.cumsum(counts)
    p = D._dev(preds[rows], torch.float32)
    score = D.coco_rescore(p, all_boxes[rows, 5], in_vis_thre)
    r = D.oks_nms_batch(p, all_boxes[rows, 4], score, off, int(counts.max()) if len(counts) else 0, oks_thre, soft=soft_nms)
    keep, cnt, sc = _np(r["keep"]), _np(r["keep_cnt"]), _np(score)
    out_score = np.empty(len(preds), np.float64)
    out_score[rows] = sc
    kept = [[int(rows[j]) for j in keep[off[i]:off[i] + cnt[i]]] for i in range(len(cnt))]
    return out_score, kept


def batched_loader(dataset, batch_size, shuffle=False, drop_last=False, sampler=None, normalize=None,
                   want_flipped=False):
    """The main-process replacement for `DataLoader(dataset, batch_size, num_workers=WORKERS)` of
    tools/train.py:145-158 / tools/test.py:122-128: yields (input, target, target_weight, meta) per batch
    with ONE launch per stage (JointsDataset.get_batch) instead of one __getitem__ per person in worker
    processes.  `dataset` is a reference_api.JointsDataset or an instance of the reference's own
    JointsDataset subclasses (COCODataset, MPIIDataset: same attribute names).  input/target/target_weight
    are CUDA tensors; meta is collated as torch's default_collate would (lists of strings, stacked arrays).

    Why not workers: the per-sample functions need the GPU, the reference starts its workers by fork
    (WORKERS = 4..24) after the parent has initialised CUDA, and CUDA cannot be re-initialised in a forked
    child.  Image decoding (cv2.imread) is the only host work left per record."""
    from torch.utils.data import BatchSampler, RandomSampler, SequentialSampler
    from torch.utils.data._utils.collate import default_collate
    if sampler is None:
        sampler = RandomSampler(dataset) if shuffle else SequentialSampler(dataset)
    for idx in BatchSampler(sampler, batch_size, drop_last):
        inp, target, tw, metas = JointsDataset.get_batch(dataset, idx, normalize=normalize, want_flipped=want_flipped)
        yield inp, target, tw, default_collate(metas)


def install(inference=None, transforms=None, joints_dataset=None, evaluate=None, nms=None, per_sample_encode=False):
    """Substitute the CUDA codec into already-imported reference modules
    (lib.core.inference, lib.utils.transforms, lib.dataset.JointsDataset, lib.core.evaluate, and
    lib.nms.nms and/or lib.dataset.coco for oks_nms / soft_oks_nms -- `nms` may be a module or a list).
    Returns a dict {(module_name, attr): original} so the caller can undo it.

    joints_dataset: by default only ADDS the batched methods (`get_batch`, `generate_target_batch`) to the
    reference's JointsDataset class -- use them from the main process, e.g. through `batched_loader(ds, B)`.
    The per-sample functions the reference calls inside __getitem__ (get_warpmatrix, rotate_points,
    generate_target, and utils.transforms.fliplr_joints) run in DataLoader workers that the reference forks
    after CUDA is up (tools/test.py:122-128, train.py:145-158); a forked child cannot use CUDA, and one
    launch per person would be slower than numpy anyway.  per_sample_encode=True rebinds them all the same
    (debugging, parity checks): the DataLoader must then be built with num_workers=0 or
    multiprocessing_context='spawn'."""
    saved = {}

    def put(mod, name, fn):
        if mod is None:
            return
        saved[(mod.__name__, name)] = getattr(mod, name, None)
        setattr(mod, name, fn)

    for nm in ("get_max_preds", "post", "get_final_preds", "transform_preds"):
        put(inference, nm, globals()[nm])
    for nm in ("flip_back", "flip_back_offset", "transform_preds"):
        put(transforms, nm, globals()[nm])
    if per_sample_encode:
        put(transforms, "fliplr_joints", fliplr_joints)
    put(evaluate, "accuracy", accuracy)
    for mod in (nms if isinstance(nms, (list, tuple)) else [nms]):
        put(mod, "oks_nms", oks_nms)
        put(mod, "soft_oks_nms", soft_oks_nms)
    if joints_dataset is not None:
        cls = joints_dataset.JointsDataset
        for nm in ("get_batch", "generate_target_batch", "_augment", "_read"):
            saved[(joints_dataset.__name__, "JointsDataset." + nm)] = cls.__dict__.get(nm)
            setattr(cls, nm, JointsDataset.__dict__[nm])
        if per_sample_encode:
            put(joints_dataset, "get_warpmatrix", get_warpmatrix)
            put(joints_dataset, "rotate_points", rotate_points)
            put(joints_dataset, "fliplr_joints", fliplr_joints)
            saved[(joints_dataset.__name__, "JointsDataset.generate_target")] = cls.generate_target
            cls.generate_target = JointsDataset.generate_target
    return saved


def uninstall(saved, modules):
    """Undo install(): `modules` is the list of module objects that were passed to it."""
    by_name = {m.__name__: m for m in modules if m is not None}
    for (mod_name, attr), orig in saved.items():
        mod = by_name.get(mod_name)
        if mod is None:
            continue
        target, name = (mod.JointsDataset, attr.split(".", 1)[1]) if attr.startswith("JointsDataset.") else (mod, attr)
        if orig is None:
            if name in getattr(target, "__dict__", {}):
                delattr(target, name)
        else:
            setattr(target, name, orig)
