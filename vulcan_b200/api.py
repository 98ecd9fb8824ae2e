"""Host-side mirror of the reference's public interface for the TSDF-fusion hot path,
over the C ABI (include/vulcan_b200.h).  Same class and method names, argument meaning
and error behaviour as the reference's C++ classes:

  Volume            include/vulcan/volume.h:16-117
  Integrator etc.   include/vulcan/integrator.h:16-34, depth_/color_/light_integrator.h
  Tracer            include/vulcan/tracer.h:24-76
  Frame             include/vulcan/frame.h:11-32
  DepthTracker      include/vulcan/tracker.h:16-103, depth_tracker.h:9-27
  PyramidTracker    include/vulcan/pyramid_tracker.h
  Transform         include/vulcan/transform.h:9-171
  Projection        include/vulcan/projection.h:11-114

The C++ drop-in headers (include/vulcan/*.h) wrap the same C ABI; this module is what the
parity tests and bench.py use.  torch supplies device memory for images and the stream.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, VulcanError  # noqa: F401


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def SetDefaultMathMode(mode):
    """Math mode of the entry points that take no Volume (Frame.ComputeNormals, ...) and of new volumes."""
    check(lib.vk_set_default_math_mode(int(mode)), "SetDefaultMathMode")


def GetDefaultMathMode():
    return lib.vk_get_default_math_mode()


class _DevView:
    """Zero-copy torch view of library-owned device memory."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2, "strides": None}


def device_view(ptr, shape, typestr):
    if int(np.prod(shape)) == 0:
        dt = {"<i4": torch.int32, "|u1": torch.uint8, "<f4": torch.float32, "<i2": torch.int16}[typestr]
        return torch.empty(tuple(shape), dtype=dt, device="cuda")
    return torch.as_tensor(_DevView(ptr, shape, typestr), device="cuda")


class Transform:
    """Column-major 4x4 matrix plus its cached inverse (include/vulcan/transform.h:166-170)."""

    def __init__(self, matrix=None, inv_matrix=None):
        self.matrix = np.eye(4, dtype=np.float32) if matrix is None else np.asarray(matrix, dtype=np.float32)
        self.inv_matrix = np.eye(4, dtype=np.float32) if inv_matrix is None else np.asarray(inv_matrix, dtype=np.float32)

    @staticmethod
    def Translate(x, y, z):
        m = np.eye(4, dtype=np.float32)
        m[:3, 3] = (x, y, z)
        i = np.eye(4, dtype=np.float32)
        i[:3, 3] = (-np.float32(x), -np.float32(y), -np.float32(z))
        return Transform(m, i)

    @staticmethod
    def Rotate(R):
        m = np.eye(4, dtype=np.float32)
        m[:3, :3] = np.asarray(R, dtype=np.float32)
        return Transform(m, m.T.copy())

    @staticmethod
    def FromRt(R, t):
        """Rigid transform from float64 R, t; matrix and inverse each rounded once to float32."""
        R = np.asarray(R, dtype=np.float64)
        t = np.asarray(t, dtype=np.float64)
        m = np.eye(4)
        m[:3, :3] = R
        m[:3, 3] = t
        i = np.eye(4)
        i[:3, :3] = R.T
        i[:3, 3] = -R.T @ t
        return Transform(m.astype(np.float32), i.astype(np.float32))

    def Inverse(self):
        return Transform(self.inv_matrix, self.matrix)

    def __mul__(self, other):
        # Transform::operator*(Transform) (include/vulcan/transform.h:61-64), float32 accumulate
        return Transform(_matmul4(self.matrix, other.matrix), _matmul4(other.inv_matrix, self.inv_matrix))

    def raw32(self):
        """32 floats: column-major matrix then inverse (byte image of vulcan::Transform)."""
        return np.concatenate([self.matrix.T.reshape(-1), self.inv_matrix.T.reshape(-1)]).astype(np.float32)

    def to_c(self):
        return _lib.vkTransform.from_buffer_copy(self.raw32().tobytes())


def _matmul4(a, b):
    out = np.zeros((4, 4), dtype=np.float32)
    for p in range(4):
        for m in range(4):
            acc = np.float32(0)
            for n in range(4):
                acc = np.float32(acc + np.float32(a[m, n] * b[n, p]))
            out[m, p] = acc
    return out


class Projection:
    """Pinhole intrinsics; defaults as include/vulcan/projection.h:16-20."""

    def __init__(self, fx=500.0, fy=500.0, cx=320.0, cy=240.0):
        self.fx, self.fy, self.cx, self.cy = (float(np.float32(v)) for v in (fx, fy, cx, cy))

    def raw4(self):
        return np.array([self.fx, self.fy, self.cx, self.cy], dtype=np.float32)

    def to_c(self):
        return _lib.vkProjection(self.fx, self.fy, self.cx, self.cy)


class Frame:
    """include/vulcan/frame.h:11-32; images are CUDA float32 tensors (H,W) / (H,W,3)."""

    def __init__(self):
        self.depth_projection = Projection()
        self.color_projection = Projection()
        self.depth_to_world_transform = Transform()
        self.depth_to_color_transform = Transform()
        self.depth_image = None
        self.color_image = None
        self.normal_image = None

    def to_c(self):
        if self.depth_image is None:
            raise VulcanError("missing depth image")  # frame.cpp:10,23
        f = _lib.vkFrame()
        d = self.depth_image
        assert d.is_cuda and d.dtype == torch.float32 and d.is_contiguous()
        f.depth = d.data_ptr()
        f.depth_h, f.depth_w = d.shape[0], d.shape[1]
        if self.color_image is not None:
            c = self.color_image
            assert c.is_cuda and c.dtype == torch.float32 and c.is_contiguous()
            f.color = c.data_ptr()
            f.color_h, f.color_w = c.shape[0], c.shape[1]
        if self.normal_image is not None:
            f.normal = self.normal_image.data_ptr()
        f.depth_proj = self.depth_projection.to_c()
        f.color_proj = self.color_projection.to_c()
        f.depth_to_world = self.depth_to_world_transform.to_c()
        f.depth_to_color = self.depth_to_color_transform.to_c()
        return f

    def ComputeNormals(self):
        """Frame::ComputeNormals (src/frame.cpp:21-36)."""
        if self.depth_image is None:
            raise VulcanError("missing depth image")
        h, w = self.depth_image.shape
        if self.normal_image is None or tuple(self.normal_image.shape) != (h, w, 3):
            self.normal_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        p = self.depth_projection.to_c()
        check(lib.vk_compute_normals(self.depth_image.data_ptr(), C.byref(p), self.normal_image.data_ptr(), w, h,
                                     _stream()), "ComputeNormals")

    def Prepare(self):
        """FilterDepths + ComputeNormals of a sensor frame in one kernel (vk_prepare_frame); the depth
        image is replaced by the filtered one."""
        if self.depth_image is None:
            raise VulcanError("missing depth image")
        h, w = self.depth_image.shape
        out = torch.empty_like(self.depth_image)
        if self.normal_image is None or tuple(self.normal_image.shape) != (h, w, 3):
            self.normal_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        p = self.depth_projection.to_c()
        check(lib.vk_prepare_frame(w, h, self.depth_image.data_ptr(), C.byref(p), out.data_ptr(),
                                   self.normal_image.data_ptr(), _stream()), "Prepare")
        self.depth_image = out

    def Downsample(self, frame):
        """Frame::Downsample (src/frame.cpp:38-57): depth and normals by nearest sample, colour by 2x2
        average, intrinsics halved, transforms copied - one kernel for the three images."""
        if self.depth_image is None or self.color_image is None or self.normal_image is None:
            raise VulcanError("Downsample: depth, color and normal images required")  # frame.cpp:40-42
        h, w = self.depth_image.shape
        ch, cw = self.color_image.shape[:2]
        if (w | h | cw | ch) & 1:
            raise VulcanError("even image dimensions required")  # src/image.cu:185-186
        if frame.depth_image is None or tuple(frame.depth_image.shape) != (h // 2, w // 2):
            frame.depth_image = torch.empty((h // 2, w // 2), dtype=torch.float32, device="cuda")
        if frame.normal_image is None or tuple(frame.normal_image.shape) != (h // 2, w // 2, 3):
            frame.normal_image = torch.empty((h // 2, w // 2, 3), dtype=torch.float32, device="cuda")
        if frame.color_image is None or tuple(frame.color_image.shape) != (ch // 2, cw // 2, 3):
            frame.color_image = torch.empty((ch // 2, cw // 2, 3), dtype=torch.float32, device="cuda")
        src = self.to_c()
        dst = frame.to_c()
        check(lib.vk_downsample_frame(C.byref(src), C.byref(dst), _stream()), "Downsample")
        frame.depth_projection = Projection(dst.depth_proj.fx, dst.depth_proj.fy, dst.depth_proj.cx, dst.depth_proj.cy)
        frame.color_projection = Projection(dst.color_proj.fx, dst.color_proj.fy, dst.color_proj.cx, dst.color_proj.cy)
        frame.depth_to_world_transform = Transform(self.depth_to_world_transform.matrix,
                                                   self.depth_to_world_transform.inv_matrix)
        frame.depth_to_color_transform = Transform(self.depth_to_color_transform.matrix,
                                                   self.depth_to_color_transform.inv_matrix)

    def FilterDepths(self):
        """Frame::FilterDepths (src/frame.cpp:8-19)."""
        if self.depth_image is None:
            raise VulcanError("missing depth image")
        h, w = self.depth_image.shape
        tmp = torch.empty_like(self.depth_image)
        check(lib.vk_filter_depths(w, h, self.depth_image.data_ptr(), tmp.data_ptr(), _stream()), "FilterDepths")
        self.depth_image.copy_(tmp)


class Volume:
    """vulcan::Volume (include/vulcan/volume.h:16-117, src/volume.cu:370-463)."""

    def __init__(self, main_block_count, excess_block_count):
        self._v = _lib.vkVolume()
        check(lib.vk_volume_create(C.byref(self._v), int(main_block_count), int(excess_block_count), _stream()),
              "Volume")
        self._alive = True

    def __del__(self):
        if getattr(self, "_alive", False):
            lib.vk_volume_destroy(C.byref(self._v))
            self._alive = False

    def GetMainBlockCount(self):
        return self._v.main_blocks

    def GetExcessBlockCount(self):
        return self._v.excess_blocks

    def GetDepthRange(self):
        return (self._v.min_depth, self._v.max_depth)

    def SetDepthRange(self, dmin, dmax):
        self._v.min_depth, self._v.max_depth = dmin, dmax

    def GetVoxelLength(self):
        return self._v.voxel_length

    def SetVoxelLength(self, length):
        self._v.voxel_length = length
        # prove the fast division for the voxel and block length now (vk_div_prepare synchronises the
        # stream on the first use of a value), not inside the first frame
        lib.vk_div_prepare(C.c_float(self._v.voxel_length), _stream())
        lib.vk_div_prepare(C.c_float(np.float32(8) * np.float32(self._v.voxel_length)), _stream())

    def GetTruncationLength(self):
        return self._v.trunc_length

    def SetTruncationLength(self, length):
        self._v.trunc_length = length
        lib.vk_div_prepare(C.c_float(self._v.trunc_length), _stream())

    def GetMathMode(self):
        return self._v.math_mode

    def SetMathMode(self, mode):
        """VK_MATH_REFERENCE (0, default): multiply-adds fused exactly where the reference's release
        build fuses them; VK_MATH_STRICT (1): one rounding per operation (-fmad=false build)."""
        if mode not in (_lib.VK_MATH_REFERENCE, _lib.VK_MATH_STRICT):
            raise VulcanError("invalid math mode")
        self._v.math_mode = mode

    def SetSharding(self, rank, count, shift=3, halo=0):
        """Blocks are owned by brick hash (bricks of 2^shift blocks per edge); halo=1 also stores the
        26 neighbours of every owned block (needed only by the key-merge ray cast, TraceKeys)."""
        self._v.shard_rank, self._v.shard_count, self._v.shard_shift, self._v.shard_halo = rank, count, shift, halo

    def Reset(self):
        check(lib.vk_volume_reset(C.byref(self._v), _stream()), "Volume.Reset")

    def Save(self, path):
        """Snapshot of the complete volume state (vk_volume_save); not in the reference."""
        check(lib.vk_volume_save(C.byref(self._v), str(path).encode(), _stream()), "Volume.Save")

    def Load(self, path):
        """Restores a snapshot into this volume (same block counts); it then continues exactly as the
        saved volume would have."""
        check(lib.vk_volume_load(C.byref(self._v), str(path).encode(), _stream()), "Volume.Load")

    def SetView(self, frame):
        """One call reaches the fixed point of repeated reference SetView calls."""
        f = frame.to_c()
        check(lib.vk_set_view(C.byref(self._v), C.byref(f), _stream()), "SetView")

    def SyncCounters(self):
        n = C.c_int(0)
        rc = lib.vk_volume_sync_counters(C.byref(self._v), C.byref(n), _stream())
        if rc != 0 and rc != _lib.VK_E_OVERFLOW:
            check(rc, "SyncCounters")
        return n.value, [self._v.host_counters[i] for i in range(_lib.VK_CNT_COUNT)]

    def GetVisibleCount(self):
        return self.SyncCounters()[0]

    def GetAllocatedBlocks(self):
        raise VulcanError("not implemented")  # Q21, src/volume.cu:444-448

    @property
    def _B(self):
        return self._v.main_blocks + self._v.excess_blocks

    def GetVisibleBlocks(self):
        n = self.GetVisibleCount()
        return device_view(self._v.visible[self._v.visible_slot], (self._B,), "<i4")[:n]

    def GetHashEntries(self):
        """(B, 4) int32 view: [x | y<<16, z | pad<<16, data, next]."""
        return device_view(self._v.entries, (self._B, 4), "<i4")

    def GetVoxels(self):
        """(B*512, 5) int32 view of the 20-byte voxels: distance, r, g, b bits, dw | cw<<16.  Read it; to
        WRITE voxels use GetVoxelsMutable() (the ray marcher samples a mirror of the distances)."""
        return device_view(self._v.voxels, (self._B * 512, 5), "<i4")

    def GetVoxelsMutable(self):
        """The non-const Volume::GetVoxels() (include/vulcan/volume.h:48-50): marks the distance mirror
        stale; the next Trace rebuilds it (vk_volume_sync_sdf)."""
        self._sdf_stale = True
        return self.GetVoxels()

    def SyncDistances(self):
        if getattr(self, "_sdf_stale", False):
            check(lib.vk_volume_sync_sdf(C.byref(self._v), _stream()), "SyncDistances")
            self._sdf_stale = False

    def GetDistances(self):
        """(B*512,) float32 view of vkVolume::sdf."""
        return device_view(self._v.sdf, (self._B * 512,), "<f4")

    def GetBlockVisibility(self):
        return device_view(self._v.visibility, (self._B,), "|u1")


class Integrator:
    """vulcan::Integrator (include/vulcan/integrator.h:16-45, src/integrator.cu:7-56)."""

    def __init__(self, volume):
        self.volume_ = volume
        self.depth_range_ = (0.1, 5.0)
        self.max_distance_weight_ = 16.0
        self.max_color_weight_ = 16.0

    def GetVolume(self):
        return self.volume_

    def GetDepthRange(self):
        return self.depth_range_

    def SetDepthRange(self, dmin, dmax):
        self.depth_range_ = (dmin, dmax)

    def GetMaxDistanceWeight(self):
        return self.max_distance_weight_

    def SetMaxDistanceWeight(self, weight):
        self.max_distance_weight_ = float(weight)

    def GetMaxColorWeight(self):
        return self.max_color_weight_

    def SetMaxColorWeight(self, weight):
        self.max_color_weight_ = float(weight)

    def Integrate(self, frame):
        raise NotImplementedError


class DepthIntegrator(Integrator):
    def Integrate(self, frame):
        f = frame.to_c()
        check(lib.vk_integrate_depth(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0],
                                     self.depth_range_[1], self.max_distance_weight_, _stream()), "Integrate")


class ColorIntegrator(Integrator):
    def Integrate(self, frame):
        if frame.color_image is None:
            raise VulcanError("missing color image")
        f = frame.to_c()
        check(lib.vk_integrate_color(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0],
                                     self.depth_range_[1], self.max_distance_weight_, self.max_color_weight_,
                                     _stream()), "Integrate")


class LightIntegrator(Integrator):
    def __init__(self, volume):
        super().__init__(volume)
        self.light_ = (1.0, (0.0, 0.0, 0.0))  # include/vulcan/light.h:16-20
        self.depth_threshold_ = 0.2           # src/light_integrator.cu:254-258
        self.frame_mask_ = None

    def GetLight(self):
        return self.light_

    def SetLight(self, intensity, position):
        self.light_ = (float(intensity), tuple(float(p) for p in position))

    def Integrate(self, frame):
        if frame.color_image is None or frame.normal_image is None:
            raise VulcanError("missing color or normal image")
        h, w = frame.depth_image.shape
        if self.frame_mask_ is None or tuple(self.frame_mask_.shape) != (h, w):
            self.frame_mask_ = torch.empty((h, w), dtype=torch.float32, device="cuda")
        light = _lib.vkLight()
        light.intensity = self.light_[0]
        for k in range(3):
            light.position[k] = self.light_[1][k]
        f = frame.to_c()
        check(lib.vk_integrate_light(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0],
                                     self.depth_range_[1], self.max_distance_weight_, self.max_color_weight_,
                                     C.byref(light), self.depth_threshold_, self.frame_mask_.data_ptr(), _stream()),
              "Integrate")


class Tracer:
    """vulcan::Tracer (include/vulcan/tracer.h:24-76, src/tracer.cpp:13-147)."""

    def __init__(self, volume):
        self.volume_ = volume
        self.depth_range_ = (0.1, 5.0)
        self.bounds_ = torch.empty((_lib.BOUNDS_H, _lib.BOUNDS_W, 2), dtype=torch.float32, device="cuda")

    def GetVolume(self):
        return self.volume_

    def GetDepthRange(self):
        return self.depth_range_

    def SetDepthRange(self, dmin, dmax):
        self.depth_range_ = (dmin, dmax)

    def Trace(self, frame):
        """Overwrites frame.depth_image (must be pre-sized, src/tracer.cpp:83-84), resizes
        color_image and creates normal_image like the reference (tracer.cpp:88, frame.cpp:25-28)."""
        if frame.depth_image is None:
            raise VulcanError("missing depth image")
        h, w = frame.depth_image.shape
        if frame.color_image is None or tuple(frame.color_image.shape) != (h, w, 3):
            frame.color_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        if frame.normal_image is None or tuple(frame.normal_image.shape) != (h, w, 3):
            frame.normal_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        f = frame.to_c()
        self.volume_.SyncDistances()
        check(lib.vk_trace(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0], self.depth_range_[1],
                           self.bounds_.data_ptr(), frame.depth_image.data_ptr(), frame.color_image.data_ptr(),
                           frame.normal_image.data_ptr(), _stream()), "Trace")

    def _ensure_outputs(self, frame):
        if frame.depth_image is None:
            raise VulcanError("missing depth image")
        h, w = frame.depth_image.shape
        if frame.color_image is None or tuple(frame.color_image.shape) != (h, w, 3):
            frame.color_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        if frame.normal_image is None or tuple(frame.normal_image.shape) != (h, w, 3):
            frame.normal_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")

    def TraceBounds(self, frame):
        """Ray bounds of the local visible blocks into self.bounds_ (sharded mode: combine them
        over the ranks before TraceRows, see vulcan_b200.sharding.ShardedTracer)."""
        f = frame.to_c()
        check(lib.vk_trace_bounds(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0], self.depth_range_[1],
                                  self.bounds_.data_ptr(), _stream()), "TraceBounds")

    def TraceRows(self, frame, peers, row_begin, row_end):
        """Marches rows [row_begin - 2, row_end + 2) with the given bounds and peer tables (or None)
        and computes the normals of rows [row_begin, row_end)."""
        self._ensure_outputs(frame)
        f = frame.to_c()
        check(lib.vk_trace_rows(C.byref(self.volume_._v), C.byref(peers) if peers is not None else None, C.byref(f),
                                self.bounds_.data_ptr(), int(row_begin), int(row_end), frame.depth_image.data_ptr(),
                                frame.color_image.data_ptr(), frame.normal_image.data_ptr(), _stream()), "TraceRows")

    def TraceKeys(self, frame, keys):
        f = frame.to_c()
        check(lib.vk_trace_keys(C.byref(self.volume_._v), C.byref(f), self.depth_range_[0], self.depth_range_[1],
                                self.bounds_.data_ptr(), keys.data_ptr(), _stream()), "TraceKeys")

    def ResolveKeys(self, frame, keys):
        h, w = frame.depth_image.shape
        if frame.color_image is None or tuple(frame.color_image.shape) != (h, w, 3):
            frame.color_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        if frame.normal_image is None or tuple(frame.normal_image.shape) != (h, w, 3):
            frame.normal_image = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
        f = frame.to_c()
        check(lib.vk_resolve_keys(keys.data_ptr(), C.byref(f), frame.depth_image.data_ptr(),
                                  frame.color_image.data_ptr(), frame.normal_image.data_ptr(), _stream()),
              "ResolveKeys")


def _transform_from_c(t):
    m = np.frombuffer(bytes(t), dtype=np.float32)
    return Transform(m[:16].reshape(4, 4).T.copy(), m[16:].reshape(4, 4).T.copy())


class DepthTracker:
    """Tracker + DepthTracker (include/vulcan/tracker.h:16-103, depth_tracker.h:9-27; src/tracker.cpp,
    src/depth_tracker.{cpp,cu}): projective point-to-plane alignment of a frame to the keyframe.
    Every Gauss-Newton iteration is one kernel (system, solve and pose update on the device)."""

    def __init__(self):
        self.translation_enabled_ = True   # src/tracker.cpp:11-12
        self.max_iterations_ = 20
        self.keyframe_ = None
        self.iterations_ = 0
        self.last_update_norm_ = 0.0
        self._workspace = torch.zeros(lib.vk_track_workspace_floats(), dtype=torch.float32, device="cuda")

    def GetKeyframe(self):
        return self.keyframe_

    def SetKeyframe(self, keyframe):
        self.keyframe_ = keyframe

    def GetTranslationEnabled(self):
        return self.translation_enabled_

    def SetTranslationEnabled(self, enabled):
        self.translation_enabled_ = bool(enabled)

    def GetMaxIterations(self):
        return self.max_iterations_

    def SetMaxIterations(self, iterations):
        if iterations <= 0:
            raise VulcanError("iterations > 0")  # src/tracker.cpp:52
        self.max_iterations_ = int(iterations)

    def _validate(self, frame):
        # Tracker::ValidateKeyframe / ValidateFrame (src/tracker.cpp:83-110)
        if self.keyframe_ is None:
            raise VulcanError("keyframe has not been assigned")
        for f, name in ((self.keyframe_, "keyframe"), (frame, "frame")):
            if f.depth_image is None:
                raise VulcanError("%s missing depth image" % name)
            if f.normal_image is None:
                raise VulcanError("%s missing normal image" % name)
            if tuple(f.normal_image.shape[:2]) != tuple(f.depth_image.shape):
                raise VulcanError("%s image size mismatch" % name)

    def Track(self, frame):
        """Tracker::Track (src/tracker.cpp:57-67): frame.depth_to_world_transform is updated in place."""
        self._validate(frame)
        k = self.keyframe_.to_c()
        f = frame.to_c()
        it = C.c_int(0)
        norm = C.c_float(0)
        check(lib.vk_track_depth(C.byref(k), C.byref(f), int(self.translation_enabled_), self.max_iterations_,
                                 self._workspace.data_ptr(), C.byref(it), C.byref(norm), _stream()), "Track")
        frame.depth_to_world_transform = _transform_from_c(f.depth_to_world)
        self.iterations_, self.last_update_norm_ = it.value, norm.value

    def ComputeResiduals(self, frame):
        """DepthTracker::ComputeResiduals (src/depth_tracker.cu:279-306) -> (H, W) tensor."""
        self._validate(frame)
        h, w = frame.depth_image.shape
        out = torch.empty((h, w), dtype=torch.float32, device="cuda")
        k, f = self.keyframe_.to_c(), frame.to_c()
        check(lib.vk_track_residuals(C.byref(k), C.byref(f), out.data_ptr(), _stream()), "ComputeResiduals")
        return out

    def ComputeJacobian(self, frame):
        """DepthTracker::ComputeJacobian (src/depth_tracker.cu:308-344) -> (H, W, 6) tensor."""
        self._validate(frame)
        h, w = frame.depth_image.shape
        out = torch.empty((h, w, 6), dtype=torch.float32, device="cuda")
        k, f = self.keyframe_.to_c(), frame.to_c()
        check(lib.vk_track_jacobian(C.byref(k), C.byref(f), int(self.translation_enabled_), out.data_ptr(), _stream()),
              "ComputeJacobian")
        return out

    def ComputeSystem(self, frame):
        """DepthTracker::ComputeSystem (src/depth_tracker.cu:346-378) -> (packed Hessian, gradient) tensors."""
        self._validate(frame)
        n = 6 if self.translation_enabled_ else 3
        hessian = torch.zeros(n * (n + 1) // 2, dtype=torch.float32, device="cuda")
        gradient = torch.zeros(n, dtype=torch.float32, device="cuda")
        k, f = self.keyframe_.to_c(), frame.to_c()
        check(lib.vk_track_system(C.byref(k), C.byref(f), int(self.translation_enabled_), self._workspace.data_ptr(),
                                  hessian.data_ptr(), gradient.data_ptr(), _stream()), "ComputeSystem")
        return hessian, gradient


class PyramidTracker:
    """PyramidTracker<DepthTracker> (include/vulcan/pyramid_tracker.h, src/pyramid_tracker.cpp:52-90): both
    pyramids are built, the half-size level is tracked with 15 iterations, the full size with 20."""

    def __init__(self):
        self.tracker_ = DepthTracker()
        self.keyframe_ = None
        self.half_keyframe_, self.quarter_keyframe_ = Frame(), Frame()

    def GetTracker(self):
        return self.tracker_

    def GetKeyframe(self):
        return self.keyframe_

    def SetKeyframe(self, keyframe):
        self.keyframe_ = keyframe

    def Track(self, frame):
        if self.keyframe_ is None:
            raise VulcanError("keyframe has not been assigned")
        half_frame, quarter_frame = Frame(), Frame()
        frame.Downsample(half_frame)
        half_frame.Downsample(quarter_frame)
        self.keyframe_.Downsample(self.half_keyframe_)
        self.half_keyframe_.Downsample(self.quarter_keyframe_)
        t = self.tracker_
        t.SetMaxIterations(15)
        t.SetTranslationEnabled(True)
        half_frame.depth_to_world_transform = quarter_frame.depth_to_world_transform
        t.SetKeyframe(self.half_keyframe_)
        t.Track(half_frame)
        t.SetMaxIterations(20)
        frame.depth_to_world_transform = half_frame.depth_to_world_transform
        t.SetKeyframe(self.keyframe_)
        t.Track(frame)


class Mesh:
    """include/vulcan/mesh.h:11-21: points (N,3), faces (F,3); `colors` (N,3) is an addition."""

    def __init__(self):
        self.points = np.zeros((0, 3), dtype=np.float32)
        self.colors = np.zeros((0, 3), dtype=np.float32)
        self.faces = np.zeros((0, 3), dtype=np.int32)


class Extractor:
    """Extractor (include/vulcan/extractor.h, src/extractor.cu:660-740).  The reference's extractor is
    unfinished (edge midpoints of one block; its index and face kernels are empty): this one returns the
    marching-cubes mesh of the whole volume - interpolated vertices with colours (vk_extract_points) and
    triangles over them (vk_extract_mesh)."""

    def __init__(self, volume):
        self.volume_ = volume

    def GetVolume(self):
        return self.volume_

    def Extract(self, mesh=None, capacity=1 << 20, faces=True):
        mesh = mesh if mesh is not None else Mesh()
        fcap = 2 * capacity
        while True:
            pts = torch.empty((capacity, 3), dtype=torch.float32, device="cuda")
            col = torch.empty((capacity, 3), dtype=torch.float32, device="cuda")
            cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
            if faces:
                tri = torch.empty((fcap, 3), dtype=torch.int32, device="cuda")
                check(lib.vk_extract_mesh(C.byref(self.volume_._v), pts.data_ptr(), col.data_ptr(), capacity,
                                          cnt.data_ptr(), tri.data_ptr(), fcap, cnt[1:].data_ptr(), _stream()), "Extract")
            else:
                check(lib.vk_extract_points(C.byref(self.volume_._v), pts.data_ptr(), col.data_ptr(), capacity,
                                            cnt.data_ptr(), _stream()), "Extract")
            n, nf = (int(x) for x in cnt.tolist())
            if n <= capacity and nf <= fcap:
                break
            capacity, fcap = max(capacity, n), max(fcap, nf)
        mesh.points = pts[:n].cpu().numpy()
        mesh.colors = col[:n].cpu().numpy()
        mesh.faces = tri[:nf].cpu().numpy() if faces else np.zeros((0, 3), dtype=np.int32)
        return mesh


class Exporter:
    """Exporter (include/vulcan/exporter.h, src/exporter.cpp:19-71): ASCII PLY with the reference's header
    (x y z, uchar red green blue, face list).  Colours are the fused voxel colours (the reference writes a
    z ramp marked REMOVE)."""

    def __init__(self, file):
        self.file_ = file

    def GetFile(self):
        return self.file_

    def Export(self, mesh):
        rgb = np.clip(np.rint(mesh.colors * 255.0), 0, 255).astype(np.uint8) if len(mesh.colors) == len(mesh.points) \
            else np.full((len(mesh.points), 3), 128, dtype=np.uint8)
        with open(self.file_, "w") as f:
            f.write("ply\nformat ascii 1.0\nelement vertex %d\n" % len(mesh.points))
            f.write("property float x\nproperty float y\nproperty float z\n")
            f.write("property uchar red\nproperty uchar green\nproperty uchar blue\n")
            f.write("element face %d\nproperty list uchar int vertex_indices\nend_header\n" % len(mesh.faces))
            for p, c in zip(mesh.points, rgb):
                f.write("%.9g %.9g %.9g %d %d %d\n" % (p[0], p[1], p[2], c[0], c[1], c[2]))
            for t in mesh.faces:
                f.write("3 %d %d %d\n" % (t[0], t[1], t[2]))   # src/exporter.cpp:63-69
