"""Streaming front end for host frames: SetView -> Integrate -> Trace with the PCIe copies of
neighbouring frames overlapped.

The reference app (apps/vulcan/vulcan.cu:297-325) uploads a depth image, fuses it, traces the
model view and hands the traced images on, all on the legacy stream.  At 640x480 the traced
depth + colour + normal images are 8.6 MB per frame, i.e. their device->host copy takes about
as long as all of the kernels of a frame.  FramePipeline keeps three streams busy instead:

    upload   : pinned host depth k+1  -> device slot (k+1) % depth
    compute  : SetView, Integrate, Trace of frame k into output slot k % depth
    download : traced images of frame k-1 -> pinned host slot (k-1) % depth

Every frame's copies still happen (nothing is skipped or cached); only their latency is hidden
behind the kernels of the neighbouring frames.  `submit()` enqueues a frame and never blocks
on the GPU unless all slots are in flight; `result(k)` blocks until frame k's images are in
host memory and returns them (valid until frame k + depth is submitted).
"""
import ctypes as C

import torch

from . import _lib
from ._lib import lib, check
from .api import DepthIntegrator, ColorIntegrator, VulcanError


class _Slot:
    def __init__(self, h, w, dev, with_color):
        f32 = torch.float32
        self.d_depth = torch.empty((h, w), dtype=f32, device=dev)
        self.d_color = torch.empty((h, w, 3), dtype=f32, device=dev) if with_color else None
        self.o_depth = torch.zeros((h, w), dtype=f32, device=dev)
        self.o_color = torch.zeros((h, w, 3), dtype=f32, device=dev)
        self.o_normal = torch.zeros((h, w, 3), dtype=f32, device=dev)
        self.h_depth = torch.empty((h, w), dtype=f32).pin_memory()
        self.h_color = torch.empty((h, w, 3), dtype=f32).pin_memory()
        self.h_normal = torch.empty((h, w, 3), dtype=f32).pin_memory()
        self.uploaded = torch.cuda.Event()
        self.computed = torch.cuda.Event()
        self.downloaded = torch.cuda.Event()
        self.frame = -1
        # descriptor of the slot's input frame; poses and intrinsics are filled in per submit
        self.cin = _lib.vkFrame()
        self.cin.depth = self.d_depth.data_ptr()
        self.cin.depth_w, self.cin.depth_h = w, h
        if with_color:
            self.cin.color = self.d_color.data_ptr()
            self.cin.color_w, self.cin.color_h = w, h


class FramePipeline:
    def __init__(self, volume, integrator, tracer, width, height, depth=2, device=None):
        if depth < 1:
            raise VulcanError("pipeline depth must be at least 1")
        if not isinstance(integrator, (DepthIntegrator, ColorIntegrator)):
            raise VulcanError("FramePipeline drives a DepthIntegrator or a ColorIntegrator")
        self.volume_, self.integrator_, self.tracer_ = volume, integrator, tracer
        self.w_, self.h_ = int(width), int(height)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.with_color_ = isinstance(integrator, ColorIntegrator)
        self.slots_ = [_Slot(self.h_, self.w_, dev, self.with_color_) for _ in range(depth)]
        self.upload_ = torch.cuda.Stream(dev)
        self.compute_ = torch.cuda.Stream(dev)
        self.download_ = torch.cuda.Stream(dev)
        # everything enqueued so far on the caller's stream (volume creation, settings) comes first
        ready = torch.cuda.Event()
        ready.record()
        for s in (self.upload_, self.compute_, self.download_):
            s.wait_event(ready)
        self.next_ = 0

    def submit(self, host_depth, frame, host_color=None):
        """Enqueue one frame.  `host_depth` ((H,W) float32, pinned) is this frame's sensor image;
        `frame` carries the projections and poses (its image fields are ignored).  Returns the
        frame's sequence number."""
        k = self.next_
        slot = self.slots_[k % len(self.slots_)]
        if slot.frame >= 0:
            slot.downloaded.synchronize()  # the slot's previous frame has left the device
        slot.frame = k
        with torch.cuda.stream(self.upload_):
            slot.d_depth.copy_(host_depth, non_blocking=True)
            if self.with_color_:
                if host_color is None:
                    raise VulcanError("missing color image")
                slot.d_color.copy_(host_color, non_blocking=True)
            slot.uploaded.record()
        cin = slot.cin
        cin.depth_proj = frame.depth_projection.to_c()
        cin.color_proj = frame.color_projection.to_c()
        cin.depth_to_world = frame.depth_to_world_transform.to_c()
        cin.depth_to_color = frame.depth_to_color_transform.to_c()
        v = C.byref(self.volume_._v)
        integ, tr = self.integrator_, self.tracer_
        with torch.cuda.stream(self.compute_):
            s = C.c_void_p(self.compute_.cuda_stream)
            self.compute_.wait_event(slot.uploaded)
            check(lib.vk_set_view(v, C.byref(cin), s), "SetView")
            if self.with_color_:
                check(lib.vk_integrate_color(v, C.byref(cin), integ.depth_range_[0], integ.depth_range_[1],
                                             integ.max_distance_weight_, integ.max_color_weight_, s), "Integrate")
            else:
                check(lib.vk_integrate_depth(v, C.byref(cin), integ.depth_range_[0], integ.depth_range_[1],
                                             integ.max_distance_weight_, s), "Integrate")
            check(lib.vk_trace(v, C.byref(cin), tr.depth_range_[0], tr.depth_range_[1], tr.bounds_.data_ptr(),
                               slot.o_depth.data_ptr(), slot.o_color.data_ptr(), slot.o_normal.data_ptr(), s), "Trace")
            slot.computed.record()
        with torch.cuda.stream(self.download_):
            self.download_.wait_event(slot.computed)
            slot.h_depth.copy_(slot.o_depth, non_blocking=True)
            slot.h_color.copy_(slot.o_color, non_blocking=True)
            slot.h_normal.copy_(slot.o_normal, non_blocking=True)
            slot.downloaded.record()
        self.next_ = k + 1
        return k

    def result(self, k):
        """Block until frame k's traced images are in host memory; returns (depth, color, normal)
        pinned host tensors, valid until frame k + depth is submitted."""
        slot = self.slots_[k % len(self.slots_)]
        if slot.frame != k:
            raise VulcanError("frame %d is no longer (or not yet) in the pipeline" % k)
        slot.downloaded.synchronize()
        return slot.h_depth, slot.h_color, slot.h_normal

    def drain(self):
        """Wait for everything submitted so far; the caller's stream is ordered after it."""
        for s in (self.upload_, self.compute_, self.download_):
            s.synchronize()
