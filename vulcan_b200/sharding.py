"""Host side of the block-sharded multi-GPU mode (one process per GPU; no reference counterpart,
SURVEY.md 5.8 / 8e).

Partition: a rank owns the blocks whose brick (block coordinates >> shift) hashes to it
(vk_shard_owner in include/vulcan_b200.h) and stores, allocates and integrates exactly those, so
allocation and integration need no exchange and each rank does 1/N of that work.  Ray casting is where
the ranks meet:

  Shard (exact, the mode of bench.py)   the owner's allocation / integration kernels store every hash entry
      and every changed 16-byte piece of block distances into all ranks' copies as they write their own
      (plain stores into peer memory over NVLink / NVSwitch); every rank then ray-casts a band of image
      rows from LOCAL memory only - each ray takes exactly the steps of the unsharded march - and stores
      its band into its peers' images.  Two flag handshakes over peer memory per view, no library
      collective (vk_trace_sharded, include/vulcan_b200.h).
  peer table (exact, no replication)   Tracer.TraceRows with a vkPeerTable: a block outside the local shard
      is read from its owner's tables through peer memory (vk_trace_rows); the long rays' chains of
      dependent look-ups then cross NVLink, which is why the replicated mode replaced it.
  key merge (approximate)  every rank ray-casts its own shard into 64-bit keys (depth bits << 32 |
      rgb8 << 8 | rank, +inf for "no owned hit") and a MIN all-reduce leaves the nearest hit per
      pixel.  No peer mappings needed, but rays step differently where their blocks live on another
      rank (a few percent of the pixels move by more than 0.5 mm), so it does not meet the parity bar.
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


# ---- exact sharded ray casting over peer memory ---------------------------------------------------

def peer_table(rank, entries_ptrs, voxels_ptrs, sdf_ptrs):
    """vkPeerTable from per-rank device pointers (own pointers at index `rank`)."""
    from . import _lib
    t = _lib.vkPeerTable()
    t.count, t.rank = len(entries_ptrs), rank
    for i, (e, v, d) in enumerate(zip(entries_ptrs, voxels_ptrs, sdf_ptrs)):
        t.entries[i] = e
        t.voxels[i] = v
        t.sdf[i] = d
    return t


def row_band(height, rank, count):
    """Rows [begin, end) of the image that `rank` renders: equal bands (the last one takes the rest)."""
    rows = (height + count - 1) // count
    return min(height, rank * rows), min(height, (rank + 1) * rows)


class Shard:
    """One rank of a block-sharded volume with REPLICATED lookup structures (include/vulcan_b200.h,
    vk_shard_*): the rank stores, allocates and integrates only the blocks it owns; its hash entries and
    block distances are copied into every peer as they are written, so every rank ray-casts its band of
    rows from local memory only, bit-identically to an unsharded volume.

        shard = Shard(volume, rank, count, width, height)      # before the volume allocates anything
        shard.connect(group)              # one process per GPU: IPC handles over torch.distributed
        Shard.connect_local(shards)       # or: all ranks in this process (tests / one thread per GPU)
        volume.SetView(frame); integrator.Integrate(frame)      # as usual, on every rank
        shard.Trace(frame)                # collective: fills frame.depth_image / color_image / normal_image
    """

    def __init__(self, volume, rank, count, width, height, shift=3, depth_range=(0.1, 5.0)):
        import ctypes as C
        from . import _lib, api
        self.volume, self.rank, self.count = volume, rank, count
        self.width, self.height = int(width), int(height)
        self.depth_range_ = depth_range
        self._s = _lib.vkShard()
        api.check(_lib.lib.vk_shard_create(C.byref(self._s), C.byref(volume._v), rank, count, shift, self.width,
                                           self.height, api._stream()), "vk_shard_create")
        self._alive = True

    def SetDepthRange(self, dmin, dmax):
        self.depth_range_ = (dmin, dmax)

    def connect(self, group=None):
        """Exchange the CUDA IPC handles of the replicas / output buffers over torch.distributed and map
        the peers' memory (one process per GPU)."""
        import ctypes as C
        import torch.distributed as dist
        from . import _lib, api
        if self.count == 1:
            return
        mine = _lib.vkShardHandles()
        api.check(_lib.lib.vk_shard_export(C.byref(self._s), C.byref(self.volume._v), C.byref(mine)), "vk_shard_export")
        blobs = [None] * self.count
        dist.all_gather_object(blobs, bytes(mine), group=group)
        everyone = (_lib.vkShardHandles * self.count)()
        for r, blob in enumerate(blobs):
            everyone[r] = _lib.vkShardHandles.from_buffer_copy(blob)
        api.check(_lib.lib.vk_shard_connect(C.byref(self._s), everyone), "vk_shard_connect")
        dist.barrier(group=group)

    @staticmethod
    def connect_local(shards):
        import ctypes as C
        from . import _lib, api
        n = len(shards)
        sp = (C.POINTER(_lib.vkShard) * n)(*[C.pointer(s._s) for s in shards])
        vp = (C.POINTER(_lib.vkVolume) * n)(*[C.pointer(s.volume._v) for s in shards])
        api.check(_lib.lib.vk_shard_connect_local(sp, vp, n), "vk_shard_connect_local")

    def TraceStage(self, frame, stage, gather=True):
        """One of the three parts of Trace (vk_trace_sharded_staged) for ranks that share a GPU: every rank
        runs part 1, then every rank part 2, then every rank part 3."""
        import ctypes as C
        from . import _lib, api
        if stage == 1:
            self.volume.SyncDistances()
        f = self._pose_only(frame)
        d, c, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        api.check(_lib.lib.vk_trace_sharded_staged(C.byref(self.volume._v), C.byref(f), self.depth_range_[0],
                                                   self.depth_range_[1], int(bool(gather)), int(stage), C.byref(d),
                                                   C.byref(c), C.byref(n), api._stream()), "vk_trace_sharded_staged")
        h, w = self.height, self.width
        frame.depth_image = api.device_view(d.value, (h, w), "<f4")
        frame.color_image = api.device_view(c.value, (h, w, 3), "<f4")
        frame.normal_image = api.device_view(n.value, (h, w, 3), "<f4")

    def Trace(self, frame, gather=True):
        """Tracer::Trace of the sharded volume (collective).  frame's images become views of this rank's
        output buffer (valid until the second-next Trace); with gather=False only this rank's band of rows
        (self.band()) is rendered and nothing is exchanged but the 38 KB ray-bounds grids."""
        import ctypes as C
        from . import _lib, api
        self.volume.SyncDistances()
        f = self._pose_only(frame)   # the view supplies pose and intrinsics; the images are the shard's
        d, c, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        api.check(_lib.lib.vk_trace_sharded(C.byref(self.volume._v), C.byref(f), self.depth_range_[0], self.depth_range_[1],
                                            int(bool(gather)), C.byref(d), C.byref(c), C.byref(n), api._stream()),
                  "vk_trace_sharded")
        h, w = self.height, self.width
        frame.depth_image = api.device_view(d.value, (h, w), "<f4")
        frame.color_image = api.device_view(c.value, (h, w, 3), "<f4")
        frame.normal_image = api.device_view(n.value, (h, w, 3), "<f4")

    def _pose_only(self, frame):
        from . import _lib
        f = _lib.vkFrame()
        f.depth_w, f.depth_h = self.width, self.height
        f.depth_proj = frame.depth_projection.to_c()
        f.depth_to_world = frame.depth_to_world_transform.to_c()
        return f

    def band(self):
        """Rows [begin, end) this rank rendered in the last Trace (synchronises the stream)."""
        import ctypes as C
        from . import _lib, api
        b0, b1 = C.c_int(0), C.c_int(0)
        api.check(_lib.lib.vk_shard_band(C.byref(self.volume._v), C.byref(b0), C.byref(b1), api._stream()), "vk_shard_band")
        return b0.value, b1.value

    def close(self):
        import ctypes as C
        from . import _lib
        if getattr(self, "_alive", False):
            _lib.lib.vk_shard_destroy(C.byref(self._s), C.byref(self.volume._v))
            self._alive = False
