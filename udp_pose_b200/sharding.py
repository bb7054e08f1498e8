"""Person sharding across the GPUs of one box, and the one collective of the path.

Every person instance is independent in every codec stage, so the path shards with no
data-path collective: rank r owns the contiguous block
    [r * ceil(N/R), min(N, (r+1) * ceil(N/R)))
-- the reference's own rule (RSN/lib/utils/dataloader.py:88-93), which keeps the gathered
result in the original order.  Encode needs no exchange at all (targets stay on the GPU that
trains on them).  Decode ends with a single all_gather of the packed keypoints
[ceil(N/R), K, 3] float32 (x, y, maxval: the all_preds layout of lib/core/function.py:208-209);
the last shard is zero-padded for the collective and the padding is sliced off.

One process per GPU (torchrun, env:// rendezvous as RSN/cvpack/.../engine.py:60); backend
"nccl" on GPUs -- NVLink/NVSwitch carries the gather -- and "gloo" in the CPU tests.
"""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_size(n, world_size):
    return -(-n // world_size) if n > 0 else 0


def shard_range(n, rank, world_size):
    """Half-open person range [lo, hi) owned by `rank`."""
    per = shard_size(n, world_size)
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def shard(x, rank=None, world_size=None):
    """The rows of `x` (dim 0 = persons) that this rank owns (a view, no copy)."""
    if rank is None:
        rank, world_size = world()
    lo, hi = shard_range(x.shape[0], rank, world_size)
    return x[lo:hi]


def gather_keypoints(local, n_total, group=None, out=None):
    """all_gather the per-rank packed keypoints [n_local,K,3] -> [n_total,K,3] on every rank.

    `local` holds this rank's shard_range(n_total, ...) rows.  One collective
    (all_gather_into_tensor), no host synchronisation; asynchronous on the current stream
    with NCCL."""
    rank, ws = world()
    if ws == 1:
        return local
    per = shard_size(n_total, ws)
    lo, hi = shard_range(n_total, rank, ws)
    assert local.shape[0] == hi - lo, "local shard has %d rows, expected %d" % (local.shape[0], hi - lo)
    tail = local.shape[1:]
    if local.shape[0] != per:  # last (or empty) shard: pad to the common size
        padded = local.new_zeros((per,) + tuple(tail))
        padded[: local.shape[0]] = local
        local = padded
    if out is None or out.shape[0] != per * ws:
        out = local.new_empty((per * ws,) + tuple(tail))
    dist.all_gather_into_tensor(out, local.contiguous(), group=group)
    return out[:n_total]


def decode_sharded(cfg, out, flipped, center, scale, group=None, decode_fn=None):
    """Decode a replicated batch with each rank doing its own block, then gather.

    out/flipped/center/scale hold all N persons (CUDA tensors or numpy); returns the full
    [N,K,3] float32 keypoints (x, y in source-image pixels, maxval) on every rank.
    decode_fn(cfg, out, flipped, center, scale) -> [n,K,3] replaces the CUDA decoder in the
    CPU (gloo) tests of this host logic; the product path never passes it."""
    rank, ws = world()
    n = out.shape[0]
    lo, hi = shard_range(n, rank, ws)
    fl = None if flipped is None else flipped[lo:hi]
    if decode_fn is None:
        from . import device_api as D
        packed = D.decode_batch(cfg, out[lo:hi], fl, center[lo:hi], scale[lo:hi], want_packed=True)["packed"]
    else:
        packed = decode_fn(cfg, out[lo:hi], fl, center[lo:hi], scale[lo:hi])
    return gather_keypoints(packed, n, group)
