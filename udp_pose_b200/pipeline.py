"""Host-buffer entry point: the whole codec pass (crop -> target encode -> flip-fused decode)
for a batch of persons whose inputs and outputs live in HOST memory.

This is the call a user of the reference makes when the surrounding program still keeps its
data in numpy (the reference's own calling convention): inputs are copied
host->device, the kernels run through the C ABI, results are copied device->host.  The batch
is cut into chunks that flow through three CUDA streams (upload / compute / download) with
double-buffered device staging, so PCIe transfers in both directions overlap the kernels.

bench.py times this as `e2e`.  Nothing here falls back to the CPU.
"""
import numpy as np
import torch

from . import device_api as D


def pinned(shape, dtype):
    return torch.empty(shape, dtype=dtype, pin_memory=True)


class HostBatch:
    """Pinned host buffers for one batch of `n` persons of a CodecConfig (inputs + outputs)."""

    def __init__(self, cfg, n, n_images, src_hw=(480, 640)):
        K, C = cfg.num_joints, cfg.channels
        (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
        cs = torch.float64 if cfg.cs_dtype == "float64" else torch.float32
        self.cfg, self.n = cfg, n
        # inputs
        self.images = pinned((n_images, src_hw[0], src_hw[1], 3), torch.uint8)
        self.image_index = np.zeros(n, np.int64)
        self.center = pinned((n, 2), cs)
        self.scale = pinned((n, 2), cs)
        self.rot = pinned((n,), torch.float64)
        self.joints = pinned((n, K, 3), torch.float64)      # source-image coordinates
        self.joints_vis = pinned((n, K, 3), torch.float64)
        self.out = pinned((n, C, h, w), torch.float32)      # network output
        self.out_flipped = pinned((n, C, h, w), torch.float32)
        # outputs
        self.crop = pinned((n, H, W, 3), torch.uint8)
        tc = (len(D_RSN) * K) if cfg.target_type == "rsn" else C
        self.target = pinned((n, tc, h, w), torch.float32)
        self.target_weight = pinned((n, K, 1), torch.float32)
        self.keypoints = pinned((n, K, 3), torch.float32)   # x, y, maxval (function.py:208-209)

    def h2d_bytes(self):
        t = [self.images, self.center, self.scale, self.rot, self.joints, self.joints_vis, self.out,
             self.out_flipped]
        return int(sum(x.numel() * x.element_size() for x in t)) + 8 * self.n  # + the image pointer table

    def d2h_bytes(self):
        t = [self.crop, self.target, self.target_weight, self.keypoints]
        return int(sum(x.numel() * x.element_size() for x in t))


from .configs import RSN_SIGMAS as D_RSN  # noqa: E402


class HostPipeline:
    """Chunked, stream-overlapped codec pass over a HostBatch."""

    def __init__(self, cfg, chunk, n_images, src_hw=(480, 640), device="cuda", n_slots=3, split_offsets=None):
        """split_offsets (offset-target configs only; default on): upload only the K classification channels of the
        two network outputs (one strided 2-D copy per tensor and chunk) and let the decoder gather its 7x7 offset
        windows straight from the caller's pinned buffers over the host link (udp_decode_offset_split) -- the
        voting decoder never reads the rest of the offset maps (inference.py:166-167,173-174), so two thirds of
        the upload of lib/core/function.py:154-169 disappear.  Results are bit-identical."""
        self.cfg, self.chunk, self.dev = cfg, chunk, torch.device(device)
        self.split = cfg.target_type == "offset" and (split_offsets is None or bool(split_offsets))
        K, C = cfg.num_joints, cfg.channels
        (W, H), (w, h) = cfg.image_size, cfg.heatmap_size
        cs = torch.float64 if cfg.cs_dtype == "float64" else torch.float32
        dev = self.dev
        self.s_up, self.s_run, self.s_down = (torch.cuda.Stream(dev) for _ in range(3))
        self.images = torch.empty((n_images, src_hw[0], src_hw[1], 3), dtype=torch.uint8, device=dev)
        tc = (len(D_RSN) * K) if cfg.target_type == "rsn" else C

        def slot():
            return dict(
                center=torch.empty((chunk, 2), dtype=cs, device=dev),
                scale=torch.empty((chunk, 2), dtype=cs, device=dev),
                rot=torch.empty((chunk,), dtype=torch.float64, device=dev),
                joints=torch.empty((chunk, K, 3), dtype=torch.float64, device=dev),
                vis=torch.empty((chunk, K, 3), dtype=torch.float64, device=dev),
                jcrop=torch.zeros((chunk, K, 3), dtype=torch.float64, device=dev),
                out=torch.empty((chunk, C, h, w), dtype=torch.float32, device=dev),
                fl=torch.empty((chunk, C, h, w), dtype=torch.float32, device=dev),
                crop=torch.empty((chunk, H, W, 3), dtype=torch.uint8, device=dev),
                ptrs=torch.empty((chunk,), dtype=torch.int64, device=dev),
                target=torch.empty((chunk, tc, h, w), dtype=torch.float32, device=dev),
                up=torch.cuda.Event(), ran=torch.cuda.Event(), down=torch.cuda.Event(),
            )
        self.slots = [slot() for _ in range(max(2, n_slots))]
        # per-person image tables: the pool is dense, so height / width / pitch are constants and
        # only the pointer table travels (pinned, with each chunk's other inputs)
        self.src_h = torch.full((chunk,), src_hw[0], dtype=torch.int32, device=dev)
        self.src_w = torch.full((chunk,), src_hw[1], dtype=torch.int32, device=dev)
        self.src_st = torch.full((chunk,), src_hw[1] * 3, dtype=torch.int32, device=dev)
        self.img_bytes = src_hw[0] * src_hw[1] * 3
        self.ptr_host = [None, None]          # two tables: run k+1 (sync=False) fills one while run k still uploads from the other
        self.ptr_done = [torch.cuda.Event(), torch.cuda.Event()]
        self.runs = 0
        self.chunks_done = 0
        self.launches = 0
        # the fills above (jcrop zeros, the image tables) ran on the caller's current stream; run() reads them from
        # s_run, a non-blocking stream with no dependency on it
        torch.cuda.current_stream(dev).synchronize()

    def _schedule(self, n):
        """(offset, size) of the chunks of one pass.  The first upload has no download to overlap
        and the last download no upload, so the pass ramps up and down through quarter and half
        chunks: the exposed head and tail of the pipeline shrink fourfold."""
        c = self.chunk
        head = [s for s in (c // 4, c // 2) if s >= 16]
        sizes, left = [], n
        if n >= 2 * (sum(head) + c):
            for s in head:
                sizes.append(s)
            left -= 2 * sum(head)
            while left > 0:
                sizes.append(min(c, left))
                left -= sizes[-1]
            sizes += head[::-1]
        else:
            while left > 0:
                sizes.append(min(c, left))
                left -= sizes[-1]
        out, lo = [], 0
        for s in sizes:
            out.append((lo, s))
            lo += s
        return out

    def run(self, hb, sync=True):
        """One pass over the HostBatch `hb`; results land in hb.crop/target/target_weight/keypoints.
        Returns after everything is back on the host; with sync=False it returns once everything is
        enqueued (a following run() then overlaps its uploads with this one's downloads --
        call synchronize() before reading the host buffers)."""
        cfg, n, chunk = self.cfg, hb.n, self.chunk
        par = self.runs & 1
        self.runs += 1
        self.ptr_done[par].synchronize()      # the run before last has finished uploading from this table
        if self.ptr_host[par] is None or self.ptr_host[par].numel() < n:
            self.ptr_host[par] = pinned((n,), torch.int64)
        ptr_host = self.ptr_host[par]
        np.add(self.images.data_ptr(), np.asarray(hb.image_index[:n], np.int64) * self.img_bytes, out=ptr_host.numpy()[:n])
        off_base = (D.host_device_pointer(hb.out), D.host_device_pointer(hb.out_flipped)) if self.split else None
        with torch.cuda.stream(self.s_up):
            self.s_up.wait_stream(self.s_run)  # kernels of an unsynchronised previous run still read the images
            self.images.copy_(hb.images, non_blocking=True)
        for lo, m in self._schedule(n):
            sl = self.slots[self.chunks_done % len(self.slots)]
            self.chunks_done += 1
            sel = slice(lo, lo + m)
            with torch.cuda.stream(self.s_up):
                self.s_up.wait_event(sl["down"])  # slot free again (previous download finished)
                small = (("center", hb.center), ("scale", hb.scale), ("rot", hb.rot), ("joints", hb.joints), ("vis", hb.joints_vis))
                for key, src in small + (() if self.split else (("out", hb.out), ("fl", hb.out_flipped))):
                    sl[key][:m].copy_(src[sel], non_blocking=True)
                if self.split:  # classification channels only: rows of h*w floats, every third map
                    hwb = hb.out.shape[2] * hb.out.shape[3] * 4
                    for key, src in (("out", hb.out), ("fl", hb.out_flipped)):
                        D.copy2d_async(sl[key].data_ptr(), 3 * hwb, src.data_ptr() + lo * cfg.channels * hwb, 3 * hwb, hwb,
                                       m * cfg.num_joints, to_device=True)
                sl["ptrs"][:m].copy_(ptr_host[sel], non_blocking=True)
                sl["up"].record()
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(sl["up"])
                src = D.SourceImages.from_tables(self.images, sl["ptrs"][:m], self.src_h[:m], self.src_w[:m], self.src_st[:m])
                M = D.warp_matrices(sl["center"][:m], sl["scale"][:m], sl["rot"][:m], cfg.image_size)
                D.crop_batch(src, M, cfg.image_size, out=sl["crop"][:m])
                j3 = D.rotate_points_batch(sl["joints"][:m], sl["rot"][:m], sl["center"][:m],
                                           sl["scale"][:m], cfg.image_size, out=sl["jcrop"][:m])
                if cfg.target_type == "rsn":
                    D.encode_batch(cfg, j3, sl["vis"][:m], out=sl["target"][:m].view(
                        m, len(D_RSN), cfg.num_joints, *sl["target"].shape[2:]))
                    tw = sl["vis"][:m, :, 0:1].float()
                else:
                    _, tw = D.encode_batch(cfg, j3, sl["vis"][:m], out=sl["target"][:m])
                kw = {}
                if self.split:  # offset windows: read in place from the caller's pinned buffers
                    pb = lo * cfg.channels * hb.out.shape[2] * hb.out.shape[3] * 4
                    kw["off_ptrs"] = (off_base[0] + pb, off_base[1] + pb)
                r = D.decode_batch(cfg, sl["out"][:m], sl["fl"][:m], sl["center"][:m], sl["scale"][:m],
                                   want_packed=True, **kw)
                self.launches += 4 + D.crop_launches(cfg.image_size)
                sl["ran"].record()
            with torch.cuda.stream(self.s_down):
                self.s_down.wait_event(sl["ran"])
                hb.crop[sel].copy_(sl["crop"][:m], non_blocking=True)
                hb.target[sel].copy_(sl["target"][:m], non_blocking=True)
                hb.target_weight[sel].copy_(tw, non_blocking=True)
                hb.keypoints[sel].copy_(r["packed"], non_blocking=True)
                sl["down"].record()
                # keep the temporaries alive until the download stream is done with them
                tw.record_stream(self.s_down)
                r["packed"].record_stream(self.s_down)
        self.ptr_done[par].record(self.s_up)
        if sync:
            self.synchronize()

    def h2d_bytes(self, hb):
        """Bytes that cross the host link towards the device in one run(hb)."""
        if not self.split:
            return hb.h2d_bytes()
        K, hw4 = self.cfg.num_joints, hb.out.shape[2] * hb.out.shape[3] * 4
        full = 2 * hb.out.numel() * 4
        # classification maps by copy; per joint 2 channels x 2 tensors x 7 window rows, two 32-byte sectors each, read in place
        return hb.h2d_bytes() - full + hb.n * (2 * K * hw4 + K * 2 * 2 * 7 * 64)

    def synchronize(self):
        self.s_down.synchronize()
        self.s_run.synchronize()
