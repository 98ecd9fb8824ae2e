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


class ShardedPipeline:
    """Host frames in, host image BANDS out, for one rank of a block-sharded volume (the streamed counterpart of
    FramePipeline; the consumer is the host, so no image is exchanged between the GPUs: every rank uploads the depth
    image, ray-casts its band of rows and downloads that band only - the ranks' bands together are one image set).

    Three streams per rank - upload of frame k+1, SetView + Integrate + sharded Trace of frame k, download of the band
    of frame k-1 - over `depth` >= 3 slots.  Which rows a rank renders is decided on the device (the bands are balanced
    from the merged ray bounds), so a frame's band table is fetched asynchronously (vk_shard_band_async) and its band is
    copied one submit later, while the next frame computes; the shard's output images of a view stay valid until the
    second-next Trace, which therefore waits for that copy.

        pipe = ShardedPipeline(shard, integrator, 640, 480)
        k = pipe.submit(pinned_depth, frame)            # every rank, same frames, same order (Trace is collective)
        b0, b1, depth, color, normal = pipe.result(k)   # pinned host images; rows [b0, b1) are this rank's
    """

    class _Slot:
        pass

    def __init__(self, shard, integrator, width, height, depth=3, device=None):
        import torch
        from . import _lib, api
        if depth < 3:
            raise api.VulcanError("a sharded pipeline needs at least 3 slots")
        self.shard_, self.volume_, self.integrator_ = shard, shard.volume, integrator
        self.w_, self.h_ = int(width), int(height)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        f32 = torch.float32
        self.slots_ = []
        for _ in range(depth):
            s = ShardedPipeline._Slot()
            s.d_depth = torch.empty((self.h_, self.w_), dtype=f32, device=dev)
            s.h_depth = torch.zeros((self.h_, self.w_), dtype=f32).pin_memory()
            s.h_color = torch.zeros((self.h_, self.w_, 3), dtype=f32).pin_memory()
            s.h_normal = torch.zeros((self.h_, self.w_, 3), dtype=f32).pin_memory()
            s.h_bands = torch.zeros(17, dtype=torch.int32).pin_memory()
            s.uploaded, s.computed, s.integrated = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
            s.bandinfo, s.downloaded = torch.cuda.Event(), torch.cuda.Event()
            s.frame, s.issued, s.out, s.band = -1, True, None, (0, 0)
            s.cin = _lib.vkFrame()
            s.cin.depth = s.d_depth.data_ptr()
            s.cin.depth_w, s.cin.depth_h = self.w_, self.h_
            self.slots_.append(s)
        self.upload_, self.compute_, self.download_ = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.info_ = torch.cuda.Stream(dev)   # band tables: on the download stream they would queue a band behind the next frame's kernels
        ready = torch.cuda.Event()
        ready.record()
        for st in (self.upload_, self.compute_, self.download_, self.info_):
            st.wait_event(ready)
        self.next_ = 0

    def _slot(self, k):
        return self.slots_[k % len(self.slots_)]

    def _issue_download(self, k):
        """Enqueue the copy of frame k's band (needs its band table on the host: waits for that 68-byte copy)."""
        import torch
        from . import api
        s = self._slot(k)
        if s.frame != k or s.issued:
            return
        s.bandinfo.synchronize()
        bands = s.h_bands.tolist()
        if bands[16] != 0:
            raise api.VulcanError("sharded volume: a peer did not arrive at a handshake (VK_E_TIMEOUT)")
        b0, b1 = bands[self.shard_.rank], bands[self.shard_.rank + 1]
        s.band = (b0, b1)
        with torch.cuda.stream(self.download_):
            # the copy starts when the NEXT frame (if submitted) has finished integrating: allocation and integration
            # store into the peers' replicas, and those stores queue behind download traffic on their way out of the GPU
            if k + 1 < self.next_:
                self.download_.wait_event(self._slot(k + 1).integrated)
            if b1 > b0:
                h, w = self.h_, self.w_
                for host, ptr, shape in ((s.h_depth, s.out[0], (h, w)), (s.h_color, s.out[1], (h, w, 3)),
                                         (s.h_normal, s.out[2], (h, w, 3))):
                    src = api.device_view(ptr, shape, "<f4")
                    host[b0:b1].copy_(src[b0:b1], non_blocking=True)
            s.downloaded.record()
        s.issued = True

    def submit(self, host_depth, frame):
        import ctypes as C
        import torch
        from . import _lib, api
        k = self.next_
        s = self._slot(k)
        if s.frame >= 0:
            self._issue_download(s.frame)
            s.downloaded.synchronize()          # the slot's previous frame has left the device
        # the band of frame k-2 must be on its way before Trace(k) reuses that view's output images
        if k >= 2:
            self._issue_download(k - 2)
        s.frame, s.issued = k, False
        with torch.cuda.stream(self.upload_):
            s.d_depth.copy_(host_depth, non_blocking=True)
            s.uploaded.record()
        cin = s.cin
        cin.depth_proj = frame.depth_projection.to_c()
        cin.depth_to_world = frame.depth_to_world_transform.to_c()
        v = C.byref(self.volume_._v)
        integ, sh = self.integrator_, self.shard_
        d, c, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        with torch.cuda.stream(self.compute_):
            st = C.c_void_p(self.compute_.cuda_stream)
            self.compute_.wait_event(s.uploaded)
            if k >= 2:
                self.compute_.wait_event(self._slot(k - 2).downloaded)
            self.volume_.SyncDistances()
            api.check(_lib.lib.vk_set_view(v, C.byref(cin), st), "SetView")
            api.check(_lib.lib.vk_integrate_depth(v, C.byref(cin), integ.depth_range_[0], integ.depth_range_[1],
                                                  integ.max_distance_weight_, st), "Integrate")
            s.integrated.record()
            api.check(_lib.lib.vk_trace_sharded(v, C.byref(cin), sh.depth_range_[0], sh.depth_range_[1], 0, C.byref(d),
                                                C.byref(c), C.byref(n), st), "vk_trace_sharded")
            s.computed.record()
        s.out = (d.value, c.value, n.value)
        with torch.cuda.stream(self.info_):
            self.info_.wait_event(s.computed)
            api.check(_lib.lib.vk_shard_band_async(v, s.h_bands.data_ptr(), C.c_void_p(self.info_.cuda_stream)),
                      "vk_shard_band_async")
            s.bandinfo.record()
        self.next_ = k + 1
        return k

    def result(self, k):
        """Block until frame k's band is in host memory: (row_begin, row_end, depth, color, normal)."""
        from . import api
        s = self._slot(k)
        if s.frame != k:
            raise api.VulcanError("frame %d is no longer (or not yet) in the pipeline" % k)
        self._issue_download(k)
        s.downloaded.synchronize()
        return s.band[0], s.band[1], s.h_depth, s.h_color, s.h_normal

    def drain(self):
        for j in range(max(0, self.next_ - len(self.slots_)), self.next_):
            self._issue_download(j)
        for st in (self.upload_, self.compute_, self.download_, self.info_):
            st.synchronize()
