"""Host side of the block-sharded multi-GPU mode (one process per GPU; no reference counterpart,
SURVEY.md 5.8 / 8e).

Partition: a rank owns the blocks whose brick (block coordinates >> shift) hashes to it
(vk_shard_owner in include/vulcan_b200.h) and also stores a one-block halo, so allocation and
integration need no exchange.  The one exchange step of the path is the ray-cast merge: every rank
ray-casts its shard into 64-bit keys (depth bits << 32 | rgb8 << 8 | rank, +inf for "no owned
hit"); a MIN all-reduce over the ranks leaves the nearest hit per pixel.
"""
import numpy as np

MISS_KEY = 0x7F800000 << 32  # +inf depth, no payload


def shard_owner(bx, by, bz, shift, count):
    """numpy restatement of vk::shard_owner (vulcan_b200/csrc/vk_math.cuh) for arrays of block coordinates."""
    m = np.uint64(0xFFFFFFFF)

    def u32(v):
        return (np.asarray(v, dtype=np.int64) >> shift).astype(np.uint64) & m

    h = ((u32(bx) * np.uint64(0x9E3779B1)) & m) ^ ((u32(by) * np.uint64(0x85EBCA77)) & m) ^ \
        ((u32(bz) * np.uint64(0xC2B2AE3D)) & m)
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x7FEB352D)) & m
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x846CA68B)) & m
    h ^= h >> np.uint64(16)
    return (h % np.uint64(count)).astype(np.int64)


def entry_coords(entries):
    """(bx, by, bz) int64 arrays from an (N, 4) int32 hash-entry table (x | y << 16, z | pad << 16, data, next)."""
    e = np.asarray(entries)
    e0 = e[:, 0].astype(np.int64) & 0xFFFFFFFF
    e1 = e[:, 1].astype(np.int64) & 0xFFFF

    def s16(v):
        v = v & 0xFFFF
        return np.where(v >= 0x8000, v - 0x10000, v)

    return s16(e0), s16(e0 >> 16), s16(e1)


def pack_keys(depth, rgb, rank):
    """Merge keys as vk_trace_keys writes them: depth float32 (H, W), rgb float32 (H, W, 3) in [0, 1];
    pixels with depth <= 0 are misses."""
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    bits = depth.view(np.uint32).astype(np.uint64)
    c = np.clip(np.asarray(rgb, dtype=np.float32) * np.float32(255.0) + np.float32(0.5), 0, 255).astype(np.uint64)
    keys = (bits << np.uint64(32)) | (c[..., 0] << np.uint64(24)) | (c[..., 1] << np.uint64(16)) | \
           (c[..., 2] << np.uint64(8)) | np.uint64(rank & 0xFF)
    keys = np.where(depth > 0, keys, np.uint64(MISS_KEY))
    return keys.astype(np.uint64).view(np.int64)  # positive floats: int64 order == depth order


def unpack_keys(keys):
    """Inverse of pack_keys on merged keys: (depth, rgb, winning rank or -1)."""
    k = np.asarray(keys).view(np.uint64)
    hi = (k >> np.uint64(32)).astype(np.uint32)
    hit = hi < np.uint32(0x7F800000)
    depth = np.where(hit, hi.view(np.float32), np.float32(0))
    s = np.float32(1.0 / 255.0)
    rgb = np.stack([((k >> np.uint64(sh)) & np.uint64(0xFF)).astype(np.float32) * s for sh in (24, 16, 8)], axis=-1)
    rgb = np.where(hit[..., None], rgb, np.float32(0))
    rank = np.where(hit, (k & np.uint64(0xFF)).astype(np.int64), -1)
    return depth.astype(np.float32), rgb.astype(np.float32), rank


def merge_keys(keys, group=None):
    """The exchange step: in-place MIN all-reduce of an int64 key tensor over the ranks (NCCL on
    GPUs over NVLink; gloo in the CPU tests)."""
    import torch.distributed as dist
    dist.all_reduce(keys, op=dist.ReduceOp.MIN, group=group)
    return keys
