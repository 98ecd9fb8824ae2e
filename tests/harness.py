"""Test infrastructure: ctypes drivers for the CPU oracle (oracle/vulcan_oracle.cpp) and for
the reference's own CUDA build (oracle/_ref, via oracle/ref_shim.cu), plus helpers that
compare volumes by block coordinates (entry indices and voxel-block ids are not pinned by
the reference)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SRC = os.path.join(ROOT, "oracle", "vulcan_oracle.cpp")
ORACLE_LIB = os.path.join(ROOT, "oracle", "_build", "libvulcan_oracle.so")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libvulcan_ref.so")
REF_LIB_NOFMAD = os.path.join(ROOT, "oracle", "_ref", "libvulcan_ref_nofmad.so")

_F = C.c_float
_I = C.c_int
_P = C.c_void_p


def build_oracle(force=False):
    if force or not os.path.exists(ORACLE_LIB) or os.path.getmtime(ORACLE_LIB) < os.path.getmtime(ORACLE_SRC):
        os.makedirs(os.path.dirname(ORACLE_LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O3", "-march=x86-64-v3", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC",
                               "-o", ORACLE_LIB, ORACLE_SRC])
    return ORACLE_LIB


_oracle = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(build_oracle())
        lib.vo_create.restype = _P
        lib.vo_create.argtypes = [_I, _I]
        lib.vo_destroy.argtypes = [_P]
        lib.vo_set_params.argtypes = [_P, _F, _F, _F, _F]
        lib.vo_set_math_mode.argtypes = [_I]
        lib.vo_get_math_mode.restype = _I
        for n in ("vo_voxels", "vo_entries", "vo_visibility", "vo_visible"):
            getattr(lib, n).restype = _P
            getattr(lib, n).argtypes = [_P]
        for n in ("vo_visible_count", "vo_voxel_pointer", "vo_excess_pointer"):
            getattr(lib, n).restype = _I
            getattr(lib, n).argtypes = [_P]
        for n in ("vo_dropped", "vo_voxel_fetches", "vo_entry_fetches"):
            getattr(lib, n).restype = C.c_longlong
            getattr(lib, n).argtypes = [_P]
        lib.vo_set_view.restype = C.c_longlong
        lib.vo_set_view.argtypes = [_P, _P, _I, _I, _P, _P]
        lib.vo_set_view_converged.restype = _I
        lib.vo_set_view_converged.argtypes = [_P, _P, _I, _I, _P, _P, _I]
        lib.vo_integrate_depth.argtypes = [_P, _P, _I, _I, _P, _P, _F, _F, _F]
        lib.vo_integrate_color.argtypes = [_P, _P, _I, _I, _P, _P, _P, _I, _I, _P, _P, _F, _F, _F, _F]
        lib.vo_frame_mask.argtypes = [_I, _I, _P, _P, _F, _P]
        lib.vo_integrate_light.argtypes = [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _F, _F, _F, _F, _F, _P, _F, _P]
        lib.vo_compute_patches.restype = _I
        lib.vo_compute_patches.argtypes = [_P, _P, _I, _P, _P, _F, _F, _F, _I, _I, _I, _I, _P]
        lib.vo_compute_bounds.argtypes = [_P, _I, _P, _I, _I, _I]
        lib.vo_compute_normals.argtypes = [_P, _P, _P, _I, _I]
        lib.vo_compute_points.argtypes = [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I]
        lib.vo_trace.argtypes = [_P, _P, _P, _I, _I, _F, _F, _P, _P, _P, _P]
        lib.vo_extract_points.restype = C.c_longlong
        lib.vo_extract_points.argtypes = [_P, _P, _P, C.c_longlong]
        lib.vo_downsample.argtypes = [_P, _I, _I, _I, _I, _P]
        lib.vo_track_pixels.argtypes = [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _I, _I, _I, _P, _P]
        lib.vo_track_system.argtypes = [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _I, _I, _I, _P, _P]
        lib.vo_track_depth.argtypes = [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _I, _I, _I, _I]
        lib.vo_track_apply_update.argtypes = [_P, _P]
        _oracle = lib
    return _oracle


def _ptr(a):
    return a.ctypes.data_as(_P)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class OracleVolume:
    """CPU oracle with the reference's call structure (one SetView = one reference pass)."""

    def __init__(self, main_blocks, excess_blocks, voxel_length=0.008, trunc_length=0.04, min_depth=0.1,
                 max_depth=5.0):
        self.lib = oracle_lib()
        self.M, self.E, self.B = main_blocks, excess_blocks, main_blocks + excess_blocks
        self.h = self.lib.vo_create(main_blocks, excess_blocks)
        self.lib.vo_set_params(self.h, voxel_length, trunc_length, min_depth, max_depth)
        self.voxel_length, self.trunc_length = voxel_length, trunc_length

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.vo_destroy(self.h)
            self.h = None

    def set_view(self, depth, proj4, T32):
        depth = _f32(depth)
        return self.lib.vo_set_view(self.h, _ptr(depth), depth.shape[1], depth.shape[0], _ptr(_f32(proj4)),
                                    _ptr(_f32(T32)))

    def set_view_converged(self, depth, proj4, T32, max_passes=64):
        depth = _f32(depth)
        return self.lib.vo_set_view_converged(self.h, _ptr(depth), depth.shape[1], depth.shape[0],
                                              _ptr(_f32(proj4)), _ptr(_f32(T32)), max_passes)

    def integrate_depth(self, depth, proj4, T32, dmin=0.1, dmax=5.0, max_w=16.0):
        depth = _f32(depth)
        self.lib.vo_integrate_depth(self.h, _ptr(depth), depth.shape[1], depth.shape[0], _ptr(_f32(proj4)),
                                    _ptr(_f32(T32)), dmin, dmax, max_w)

    def integrate_color(self, depth, dproj4, T32, color, cproj4, Tcd32, dmin=0.1, dmax=5.0, max_w=16.0,
                        max_cw=16.0):
        depth, color = _f32(depth), _f32(color)
        self.lib.vo_integrate_color(self.h, _ptr(depth), depth.shape[1], depth.shape[0], _ptr(_f32(dproj4)),
                                    _ptr(_f32(T32)), _ptr(color), color.shape[1], color.shape[0],
                                    _ptr(_f32(cproj4)), _ptr(_f32(Tcd32)), dmin, dmax, max_w, max_cw)

    def integrate_light(self, depth, dproj4, T32, color, normal, cproj4, Tcd32, intensity, light_pos,
                        dmin=0.1, dmax=5.0, max_w=16.0, max_cw=16.0, depth_threshold=0.2):
        depth, color, normal = _f32(depth), _f32(color), _f32(normal)
        mask = np.zeros(depth.shape, dtype=np.float32)
        self.lib.vo_integrate_light(self.h, _ptr(depth), depth.shape[1], depth.shape[0], _ptr(_f32(dproj4)),
                                    _ptr(_f32(T32)), _ptr(color), _ptr(normal), _ptr(_f32(cproj4)),
                                    _ptr(_f32(Tcd32)), dmin, dmax, max_w, max_cw, intensity,
                                    _ptr(_f32(light_pos)), depth_threshold, _ptr(mask))
        return mask

    def trace(self, w, h, proj4, T32, dmin=0.1, dmax=5.0, normals=True):
        bounds = np.zeros((60, 80, 2), dtype=np.float32)
        depth = np.zeros((h, w), dtype=np.float32)
        color = np.zeros((h, w, 3), dtype=np.float32)
        normal = np.zeros((h, w, 3), dtype=np.float32) if normals else None
        self.lib.vo_trace(self.h, _ptr(_f32(T32)), _ptr(_f32(proj4)), w, h, dmin, dmax, _ptr(bounds), _ptr(depth),
                          _ptr(color), _ptr(normal) if normals else None)
        return depth, color, normal, bounds

    def fetch_counts(self):
        return self.lib.vo_voxel_fetches(self.h), self.lib.vo_entry_fetches(self.h)

    # zero-copy numpy views of the oracle's state
    def entries(self):
        buf = (C.c_int32 * (self.B * 4)).from_address(self.lib.vo_entries(self.h))
        return np.frombuffer(buf, dtype=np.int32).reshape(self.B, 4)

    def voxels(self):
        buf = (C.c_int32 * (self.B * 512 * 5)).from_address(self.lib.vo_voxels(self.h))
        return np.frombuffer(buf, dtype=np.int32).reshape(self.B * 512, 5)

    def visibility(self):
        buf = (C.c_uint8 * self.B).from_address(self.lib.vo_visibility(self.h))
        return np.frombuffer(buf, dtype=np.uint8)

    def visible(self):
        n = self.lib.vo_visible_count(self.h)
        if n == 0:
            return np.zeros((0,), dtype=np.int32)
        buf = (C.c_int32 * n).from_address(self.lib.vo_visible(self.h))
        return np.frombuffer(buf, dtype=np.int32).copy()

    def dropped(self):
        return self.lib.vo_dropped(self.h)


def have_ref(nofmad=False):
    return os.path.exists(REF_LIB_NOFMAD if nofmad else REF_LIB)


_ref_libs = {}


def ref_lib(nofmad=False):
    path = REF_LIB_NOFMAD if nofmad else REF_LIB
    if os.environ.get("VULCAN_REF_VARIANT") == "asshipped" and not nofmad:
        # the reference's default build: no NDEBUG, a device synchronisation after every launch (timed once
        # for BASELINE.md 2.1, never used for parity)
        path = os.path.join(ROOT, "oracle", "_ref", "libvulcan_ref_asshipped.so")
    if path not in _ref_libs:
        lib = C.CDLL(path)
        lib.ref_create.restype = _P
        lib.ref_create.argtypes = [_I, _I, _F, _F, _F, _F]
        lib.ref_destroy.argtypes = [_P]
        lib.ref_set_ranges.argtypes = [_P, _F, _F, _F, _F, _F, _F]
        lib.ref_set_light.argtypes = [_P, _F, _F, _F, _F]
        lib.ref_set_frame.argtypes = [_P, _P, _I, _I, _P, _I, _I, _P, _P, _P, _P, _P]
        lib.ref_set_frame_device.argtypes = [_P, _P, _I, _I, _P, _P]
        lib.ref_compute_input_normals.argtypes = [_P]
        lib.ref_set_view.argtypes = [_P]
        lib.ref_integrate.argtypes = [_P, _I]
        lib.ref_trace.argtypes = [_P, _I, _I, _P, _P]
        lib.ref_device_ptr.restype = _P
        lib.ref_device_ptr.argtypes = [_P, _I]
        lib.ref_visible_count.argtypes = [_P]
        lib.ref_download.argtypes = [_P, _P, C.c_size_t]
        lib.ref_filter_depths.argtypes = [_I, _I, _P, _P]
        lib.ref_compute_normals.argtypes = [_P, _P, _P, _I, _I]
        lib.ref_track_create.restype = _P
        lib.ref_track_destroy.argtypes = [_P]
        lib.ref_track_set_frame.argtypes = [_P, _I, _P, _P, _P, _I, _I, _P, _P]
        lib.ref_track_residuals.argtypes = [_P, _P]
        lib.ref_track_jacobian.argtypes = [_P, _I, _P]
        lib.ref_track_system.argtypes = [_P, _I, _P, _P]
        lib.ref_track.argtypes = [_P, _I, _I, _P]
        lib.ref_track_downsample.argtypes = [_P, _I, _I, _P, _P, _P, _P]
        lib.ref_track_pyramid.argtypes = [_P, _P]
        _ref_libs[path] = lib
    return _ref_libs[path]


def ref_prepare_frame(depth, K, nofmad=False):
    """The reference's input-side pre-pass on the GPU: vulcan::FilterDepths then vulcan::ComputeNormals of the
    filtered depth (src/frame.cu:124-203).  Returns (filtered depth, normals) as numpy arrays."""
    import torch
    lib = ref_lib(nofmad)
    h, w = depth.shape
    src = torch.from_numpy(np.ascontiguousarray(depth, dtype=np.float32)).cuda()
    dst = torch.zeros_like(src)
    nrm = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    assert lib.ref_filter_depths(w, h, src.data_ptr(), dst.data_ptr()) == 0
    k = np.ascontiguousarray(K, dtype=np.float32)
    assert lib.ref_compute_normals(dst.data_ptr(), k.ctypes.data_as(_P), nrm.data_ptr(), w, h) == 0
    return dst.cpu().numpy(), nrm.cpu().numpy()


class RefVolume:
    """The reference's own classes (Volume + integrators + Tracer) on the GPU.  Only one
    instance may be alive per library at a time: the reference keeps its allocator counters
    in translation-unit __device__ globals (Q8)."""

    def __init__(self, main_blocks, excess_blocks, voxel_length=0.008, trunc_length=0.04, min_depth=0.1,
                 max_depth=5.0, nofmad=False):
        self.lib = ref_lib(nofmad)
        self.M, self.E, self.B = main_blocks, excess_blocks, main_blocks + excess_blocks
        self.h = self.lib.ref_create(main_blocks, excess_blocks, voxel_length, trunc_length, min_depth, max_depth)
        if not self.h:
            raise RuntimeError("ref_create failed")
        self.w = self.hgt = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.ref_destroy(self.h)
            self.h = None

    __del__ = close

    def set_ranges(self, imin=0.1, imax=5.0, max_dw=16.0, max_cw=16.0, tmin=0.1, tmax=5.0):
        self.lib.ref_set_ranges(self.h, imin, imax, max_dw, max_cw, tmin, tmax)

    def set_light(self, intensity, pos):
        self.lib.ref_set_light(self.h, intensity, pos[0], pos[1], pos[2])

    def set_frame(self, depth, dproj4, T32, color=None, cproj4=None, Tcd32=None, normal=None):
        depth = _f32(depth)
        self.hgt, self.w = depth.shape
        color = _f32(color) if color is not None else None
        normal = _f32(normal) if normal is not None else None
        rc = self.lib.ref_set_frame(self.h, _ptr(depth), self.w, self.hgt,
                                    _ptr(color) if color is not None else None,
                                    color.shape[1] if color is not None else 0,
                                    color.shape[0] if color is not None else 0,
                                    _ptr(normal) if normal is not None else None, _ptr(_f32(dproj4)),
                                    _ptr(_f32(cproj4)) if cproj4 is not None else None, _ptr(_f32(T32)),
                                    _ptr(_f32(Tcd32)) if Tcd32 is not None else None)
        assert rc == 0, rc

    def set_frame_device(self, depth_ptr, w, h, dproj4, T32):
        self.w, self.hgt = w, h
        rc = self.lib.ref_set_frame_device(self.h, depth_ptr, w, h, _ptr(_f32(dproj4)), _ptr(_f32(T32)))
        assert rc == 0, rc

    def compute_input_normals(self):
        assert self.lib.ref_compute_input_normals(self.h) == 0

    def set_view(self):
        n = self.lib.ref_set_view(self.h)
        assert n >= 0
        return n

    def set_view_converged(self, max_passes=64):
        """SetView until the visible count stops changing (tests/tracer_test.cu:298-303), then
        once more so that the last pass is a pass at the fixed point."""
        prev = -1
        passes = 0
        while passes < max_passes:
            n = self.set_view()
            passes += 1
            if n == prev:
                break
            prev = n
        return passes

    def integrate(self, kind=0):
        assert self.lib.ref_integrate(self.h, kind) == 0

    def trace(self, w, h, dproj4, T32):
        assert self.lib.ref_trace(self.h, w, h, _ptr(_f32(dproj4)), _ptr(_f32(T32))) == 0

    def ptr(self, which):
        return self.lib.ref_device_ptr(self.h, which)

    def visible_count(self):
        return self.lib.ref_visible_count(self.h)

    def tensor(self, which, shape, typestr):
        from vulcan_b200.api import device_view
        return device_view(self.ptr(which), shape, typestr)

    def entries(self):
        return self.tensor(0, (self.B, 4), "<i4")

    def voxels(self):
        return self.tensor(1, (self.B * 512, 5), "<i4")

    def visibility(self):
        return self.tensor(2, (self.B,), "|u1")

    def visible(self):
        return self.tensor(3, (self.B,), "<i4")[: self.visible_count()]

    def bounds(self):
        """The reference Tracer's own 80 x 60 ray-bounds buffer (Tracer::bounds_, include/vulcan/tracer.h:70)."""
        self.lib.ref_tracer_bounds.restype = _P
        self.lib.ref_tracer_bounds.argtypes = [_P]
        from vulcan_b200.api import device_view
        return device_view(self.lib.ref_tracer_bounds(self.h), (60, 80, 2), "<f4")

    def traced(self, w, h):
        return (self.tensor(4, (h, w), "<f4"), self.tensor(5, (h, w, 3), "<f4"), self.tensor(6, (h, w, 3), "<f4"))


# ---- comparisons by block coordinate (numpy or torch tensors alike) ---------------------

def entry_keys(entries):
    """int64 key per entry from the packed coordinates (x | y<<16, z) of an (N,4) int32 table."""
    e0 = entries[:, 0].astype(np.int64) & 0xFFFFFFFF
    e1 = entries[:, 1].astype(np.int64) & 0xFFFF
    return e0 | (e1 << 32)


def block_hash(keys, K):
    """Q1 hash of packed block keys (src/volume.cu:168-180)."""
    k = np.asarray(keys, dtype=np.int64)
    m = np.int64(0xFFFFFFFF)
    bx, by, bz = k & 0xFFFF, (k >> 16) & 0xFFFF, (k >> 32) & 0xFFFF
    sx = np.where(bx >= 0x8000, bx - 0x10000, bx)
    sy = np.where(by >= 0x8000, by - 0x10000, by)
    sz = np.where(bz >= 0x8000, bz - 0x10000, bz)
    h = ((sx * 73856093) & m) ^ ((sy * 19349669) & m) ^ ((sz * 83492791) & m)
    return (h % K).astype(np.int64)


def reachable_mask(entries, M):
    """Per entry: allocated AND found at its own index by a lookup of its coordinates (bucket
    head, then the `next` chain).  An allocated entry that is not reachable sits in a bucket its
    coordinates do not hash to - only the reference's torn request writes (Q24) produce those."""
    entries = np.asarray(entries)
    keys = entry_keys(entries)
    alloc = entries[:, 2] != -1
    idx = np.nonzero(alloc)[0]
    cur = block_hash(keys[idx], M)
    found = np.full(len(idx), -1, dtype=np.int64)
    active = np.ones(len(idx), dtype=bool)
    while active.any():   # chains can be hundreds of entries long in the collision stress tests
        c = cur[active]
        hit = (keys[c] == keys[idx[active]]) & (entries[c, 2] != -1)
        a_idx = np.nonzero(active)[0]
        found[a_idx[hit]] = c[hit]
        nxt = entries[c, 3].astype(np.int64)
        cont = (~hit) & (nxt != -1)
        active[a_idx[~cont]] = False
        cur[a_idx[cont]] = nxt[cont]
    out = np.zeros(len(entries), dtype=bool)
    out[idx[found == idx]] = True
    return out


def allocated_table(entries, M=None):
    """(sorted keys, data ids, entry indices) of the allocated entries (numpy).  With M given,
    entries that a lookup cannot reach are left out (see reachable_mask)."""
    entries = np.asarray(entries)
    mask = entries[:, 2] != -1
    if M is not None:
        mask = reachable_mask(entries, M)
    keys = entry_keys(entries[mask])
    order = np.argsort(keys, kind="stable")
    return keys[order], entries[mask][order, 2], np.nonzero(mask)[0][order]


def visible_keys(entries, visible):
    entries = np.asarray(entries)
    return np.sort(entry_keys(entries[np.asarray(visible)]))


def gather_blocks(voxels, data_ids):
    """(n, 512, 5) int32 voxel words of the given voxel-block ids."""
    return np.asarray(voxels).reshape(-1, 512, 5)[np.asarray(data_ids)]


def split_voxels(words):
    """(..., 5) int32 words -> distance, rgb, distance_weight, color_weight."""
    words = np.ascontiguousarray(words)
    dist = words[..., 0].view(np.float32)
    rgb = words[..., 1:4].view(np.float32)
    w = words[..., 4]
    dw = (w & 0xFFFF).astype(np.int16)
    cw = ((w >> 16) & 0xFFFF).astype(np.int16)
    return dist, rgb, dw, cw


# ---- scenario runners ------------------------------------------------------------------

def _load_synth():
    import importlib.util
    spec = importlib.util.spec_from_file_location("vk_synth", os.path.join(ROOT, "vulcan_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


synth = _load_synth()


def raw32(R, t):
    """Byte image of vulcan::Transform for a rigid pose (float64 R, t rounded once)."""
    m = np.eye(4)
    m[:3, :3] = R
    m[:3, 3] = t
    i = np.eye(4)
    i[:3, :3] = np.asarray(R).T
    i[:3, 3] = -np.asarray(R).T @ np.asarray(t)
    return np.concatenate([m.T.reshape(-1), i.T.reshape(-1)]).astype(np.float32)


def room_pose(k, step_deg=15.0):
    """Camera k of the C4 sweep: in-place rotation about the vertical axis at the room centre
    (24 poses, 15 degree steps)."""
    return synth.rotation_y(step_deg * k), np.zeros(3)


def make_scenario(name="small", frames=(0,), mode="depth", step_deg=1.2, math=0, trace_each=False):
    """Synthetic scenarios.  Sphere-on-plane: frames = orbit indices; room*: frames =
    indices of the in-place rotation (15 degree steps).  math = 0 (reference contraction pattern) or 1
    (strict).  trace_each: ray-cast the model after EVERY frame from that frame's pose (the per-frame loop
    of BASELINE configs 2 and 5) instead of once after the last one; images are stacked (V, H, W[, 3])."""
    cfg = {
        "micro": dict(w=48, h=36, voxel=0.02, M=2048, E=512),
        "tiny": dict(w=64, h=48, voxel=0.016, M=4096, E=1024),
        "small": dict(w=160, h=120, voxel=0.008, M=32768, E=8192),
        "c1": dict(w=640, h=480, voxel=0.004, M=262144, E=65536),
        "c3": dict(w=1280, h=720, voxel=0.004, M=262144, E=65536),
        # 1080p depth, 2 mm voxels, the int-index ceiling of 4 194 303 blocks (42.9 GB pool), sphere scene
        "c4": dict(w=1920, h=1080, voxel=0.002, M=3145728, E=1048575),
        # BASELINE config 4: the 8 x 8 x 3 m room + 3 boxes seen from its centre, depth range widened to
        # 8 m on Volume, Integrator and Tracer so that the room corners (5.7 m) stay valid
        "room": dict(w=1920, h=1080, voxel=0.002, M=3145728, E=1048575, scene="room", depth_range=(0.1, 8.0)),
        # the same geometry at sizes the CPU oracle handles (walls through many block faces, grazing floor rays)
        "room_small": dict(w=240, h=136, voxel=0.016, M=65536, E=16384, scene="room", depth_range=(0.1, 8.0)),
        "room_vga": dict(w=640, h=360, voxel=0.008, M=524288, E=131072, scene="room", depth_range=(0.1, 8.0)),
    }[name]
    cfg.setdefault("scene", "sphere")
    cfg.setdefault("depth_range", (0.1, 5.0))
    w, h = cfg["w"], cfg["h"]
    K = synth.intrinsics(w, h)
    scn = dict(cfg, name=name, mode=mode, math=math, K=np.array(K, dtype=np.float32), trunc=5 * cfg["voxel"],
               frames=[], trace_each=bool(trace_each))
    for k in frames:
        if cfg["scene"] == "room":
            R, t = room_pose(k, step_deg if step_deg != 1.2 else 15.0)
            depth, color = synth.render_room(w, h, K, R, t, max_depth=cfg["depth_range"][1]), None
            if mode != "depth":
                yy, xx = np.mgrid[0:h, 0:w]
                color = np.stack([0.2 + 0.6 * ((xx // 16 + yy // 16) % 2), np.full((h, w), 0.5),
                                  0.2 + 0.6 * (yy / float(h))], axis=-1).astype(np.float32)
        else:
            if k == 0 and step_deg == 1.2:
                R, t = np.eye(3), np.zeros(3)   # C1: identity pose
            else:
                R, t = synth.orbit_pose(k, step_deg=step_deg)
            if mode == "depth":
                depth, color = synth.render_sphere_plane(w, h, K, R, t), None
            else:
                depth, color = synth.render_sphere_plane(w, h, K, R, t, with_color=True)
        scn["frames"].append(dict(depth=depth, color=color, T=raw32(R, t)))
    # depth -> colour camera: 25 mm baseline (App. D, C3)
    scn["Tcd"] = raw32(np.eye(3), np.array([0.025, 0.0, 0.0]))
    scn["light"] = (2.0, (0.025, 0.08, 0.0))
    return scn


def _stack(views):
    """Traced images of one or several views: a single view keeps its (H, W[, 3]) shape."""
    if len(views) == 1:
        return views[0]
    return tuple(np.stack([v[i] for v in views]) for i in range(3))


def _pack_result(entries, voxels, visibility, visible, traced, M):
    keys, data, idx = allocated_table(entries, M)
    return dict(keys=keys, voxels=gather_blocks(voxels, data), vis_flags=np.asarray(visibility)[idx],
                visible_keys=visible_keys(entries, visible), depth=traced[0], color=traced[1], normal=traced[2],
                unreachable=int((np.asarray(entries)[:, 2] != -1).sum() - len(keys)))


def run_oracle(scn):
    dmin, dmax = scn.get("depth_range", (0.1, 5.0))
    ov = OracleVolume(scn["M"], scn["E"], scn["voxel"], scn["trunc"], dmin, dmax)
    ov.lib.vo_set_math_mode(scn.get("math", 0))
    K = scn["K"]
    views = []
    for fr in scn["frames"]:
        passes = ov.set_view_converged(fr["depth"], K, fr["T"], scn.get("max_passes", 4096))
        assert passes < scn.get("max_passes", 4096), "oracle allocation did not converge"
        if scn["mode"] == "depth":
            ov.integrate_depth(fr["depth"], K, fr["T"], dmin, dmax)
        elif scn["mode"] == "color":
            ov.integrate_color(fr["depth"], K, fr["T"], fr["color"], K, scn["Tcd"], dmin, dmax)
        else:
            nrm = np.zeros(fr["depth"].shape + (3,), dtype=np.float32)
            ov.lib.vo_compute_normals(_ptr(_f32(fr["depth"])), _ptr(_f32(K)), _ptr(nrm), scn["w"], scn["h"])
            ov.integrate_light(fr["depth"], K, fr["T"], fr["color"], nrm, K, scn["Tcd"], scn["light"][0],
                               np.array(scn["light"][1], dtype=np.float32), dmin, dmax)
        if scn.get("trace_each"):
            views.append(ov.trace(scn["w"], scn["h"], K, fr["T"], dmin, dmax)[:3])
    if not views:
        last = scn["frames"][-1]
        d, c, n, bounds = ov.trace(scn["w"], scn["h"], K, last["T"], dmin, dmax)
        views.append((d, c, n))
    else:
        bounds = ov.trace(scn["w"], scn["h"], K, scn["frames"][-1]["T"], dmin, dmax)[3]
    res = _pack_result(ov.entries(), ov.voxels(), ov.visibility(), ov.visible(), _stack(views), scn["M"])
    res["bounds"] = bounds
    res["fetches"] = ov.fetch_counts()
    assert ov.dropped() == 0
    return res


def run_ref(scn, nofmad=None):
    """The reference's own CUDA build: release (-fmad=true) for math mode 0, -fmad=false for 1."""
    if nofmad is None:
        nofmad = bool(scn.get("math", 0))
    dmin, dmax = scn.get("depth_range", (0.1, 5.0))
    rv = RefVolume(scn["M"], scn["E"], scn["voxel"], scn["trunc"], dmin, dmax, nofmad=nofmad)
    try:
        K = scn["K"]
        kind = {"depth": 0, "color": 1, "light": 2}[scn["mode"]]
        rv.set_ranges(dmin, dmax, 16.0, 16.0, dmin, dmax)
        rv.set_light(*scn["light"])
        views = []
        for fr in scn["frames"]:
            rv.set_frame(fr["depth"], K, fr["T"], color=fr["color"], cproj4=K, Tcd32=scn["Tcd"])
            if kind == 2:
                rv.compute_input_normals()
            rv.set_view_converged()
            rv.integrate(kind)
            if scn.get("trace_each"):
                rv.trace(scn["w"], scn["h"], K, fr["T"])
                views.append(tuple(x.cpu().numpy() for x in rv.traced(scn["w"], scn["h"])))
        if not views:
            last = scn["frames"][-1]
            rv.trace(scn["w"], scn["h"], K, last["T"])
            views.append(tuple(x.cpu().numpy() for x in rv.traced(scn["w"], scn["h"])))
        d, c, n = _stack(views)
        entries = rv.entries().cpu().numpy()
        keys, data, idx = allocated_table(entries, scn["M"])
        import torch
        vox = rv.voxels().view(-1, 512, 5)[torch.as_tensor(data, device="cuda").long()].cpu().numpy()
        res = dict(keys=keys, voxels=vox, vis_flags=rv.visibility().cpu().numpy()[idx],
                   visible_keys=visible_keys(entries, rv.visible().cpu().numpy()), depth=d, color=c, normal=n,
                   unreachable=int((entries[:, 2] != -1).sum() - len(keys)))
        if hasattr(rv.lib, "ref_tracer_bounds"):
            res["bounds"] = rv.bounds().cpu().numpy()   # the reference Tracer's 80 x 60 grid of the last view
        return res
    finally:
        rv.close()


def run_ours(scn, shard=None):
    import torch
    import vulcan_b200 as vb
    vb.SetDefaultMathMode(scn.get("math", 0))
    dmin, dmax = scn.get("depth_range", (0.1, 5.0))
    vol = vb.Volume(scn["M"], scn["E"])
    vol.SetVoxelLength(scn["voxel"])
    vol.SetTruncationLength(scn["trunc"])
    vol.SetDepthRange(dmin, dmax)
    if shard:
        vol.SetSharding(*shard)
    integ = {"depth": vb.DepthIntegrator, "color": vb.ColorIntegrator, "light": vb.LightIntegrator}[scn["mode"]](vol)
    integ.SetDepthRange(dmin, dmax)
    if scn["mode"] == "light":
        integ.SetLight(*scn["light"])
    tracer = vb.Tracer(vol)
    tracer.SetDepthRange(dmin, dmax)
    K = scn["K"]

    def transform(raw):
        return vb.Transform(raw[:16].reshape(4, 4).T.copy(), raw[16:].reshape(4, 4).T.copy())

    def trace_from(frame):
        out = vb.Frame()
        out.depth_projection = frame.depth_projection
        out.depth_to_world_transform = frame.depth_to_world_transform
        out.depth_image = torch.zeros((scn["h"], scn["w"]), dtype=torch.float32, device="cuda")
        tracer.Trace(out)
        return (out.depth_image.cpu().numpy(), out.color_image.cpu().numpy(), out.normal_image.cpu().numpy())

    frame = vb.Frame()
    views = []
    for fr in scn["frames"]:
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*K)
        frame.color_projection = vb.Projection(*K)
        frame.depth_to_world_transform = transform(fr["T"])
        frame.depth_to_color_transform = transform(scn["Tcd"])
        frame.depth_image = torch.from_numpy(fr["depth"]).cuda()
        if fr["color"] is not None:
            frame.color_image = torch.from_numpy(fr["color"]).cuda()
        if scn["mode"] == "light":
            frame.ComputeNormals()
        vol.SetView(frame)
        integ.Integrate(frame)
        if scn.get("trace_each"):
            views.append(trace_from(frame))
    if not views:
        views.append(trace_from(frame))
    d, c, n = _stack(views)
    n_vis, counters = vol.SyncCounters()
    assert counters[4] == 0, "pool overflow"
    entries = vol.GetHashEntries().cpu().numpy()
    keys, data, idx = allocated_table(entries, scn["M"])
    vox = vol.GetVoxels().view(-1, 512, 5)[torch.as_tensor(data, device="cuda").long()].cpu().numpy()
    res = dict(keys=keys, voxels=vox, vis_flags=vol.GetBlockVisibility().cpu().numpy()[idx],
               visible_keys=visible_keys(entries, vol.GetVisibleBlocks().cpu().numpy()),
               depth=d, color=c, normal=n, counters=counters,
               bounds=tracer.bounds_.cpu().numpy(),
               unreachable=int((entries[:, 2] != -1).sum() - len(keys)))
    del vol
    return res


def _first_index(keys):
    """(unique keys, index of the first occurrence of each, number of duplicate entries)."""
    uniq, first = np.unique(keys, return_index=True)
    return uniq, first, int(len(keys) - len(uniq))


def compare(a, b, trunc):
    """Parity report between two results (dicts of _pack_result).

    Voxel data is compared on the blocks both sides allocated.  The reference can allocate
    blocks no ray asked for: its request kernel stores the 3 x int16 block coordinates of a
    request with three unsynchronised 2-byte stores (src/volume.cu:199,236), so two pixels
    racing on one bucket can leave a torn mix of their coordinates (quirk Q24) - possibly a
    block that hashes to another bucket, i.e. a duplicate of a real block.  Those extras are
    counted in only_b / dup_b and differ from run to run of the reference itself."""
    rep = {}
    ka, ia, dupa = _first_index(a["keys"])
    kb, ib, dupb = _first_index(b["keys"])
    common, ca_, cb_ = np.intersect1d(ka, kb, assume_unique=True, return_indices=True)
    sel_a, sel_b = ia[ca_], ib[cb_]
    rep["alloc_equal"] = bool(len(a["keys"]) == len(b["keys"]) and np.array_equal(a["keys"], b["keys"]))
    rep["alloc_counts"] = (int(len(a["keys"])), int(len(b["keys"])))
    rep["alloc_only_a"] = int(len(ka) - len(common))
    rep["alloc_only_b"] = int(len(kb) - len(common))
    rep["alloc_dup_a"], rep["alloc_dup_b"] = dupa + a.get("unreachable", 0), dupb + b.get("unreachable", 0)
    va, vb_ = np.unique(a["visible_keys"]), np.unique(b["visible_keys"])
    rep["visible_equal"] = bool(len(a["visible_keys"]) == len(b["visible_keys"]) and
                                np.array_equal(a["visible_keys"], b["visible_keys"]))
    rep["visible_counts"] = (int(len(a["visible_keys"])), int(len(b["visible_keys"])))
    rep["visible_only_a"] = int(len(np.setdiff1d(va, vb_, assume_unique=True)))
    rep["visible_only_b"] = int(len(np.setdiff1d(vb_, va, assume_unique=True)))
    # visible entries of b that are not extras of b (blocks a knows) must be visible in a too
    rep["visible_only_b_known"] = int(len(np.intersect1d(np.setdiff1d(vb_, va, assume_unique=True), ka)))
    if len(common):
        xa, xb = a["voxels"][sel_a], b["voxels"][sel_b]
        rep["vis_flags_equal"] = bool(np.array_equal(a["vis_flags"][sel_a], b["vis_flags"][sel_b]))
        da, ca, dwa, cwa = split_voxels(xa)
        db, cb, dwb, cwb = split_voxels(xb)
        rep["voxels"] = int(da.size)
        rep["dw_mismatch"] = int((dwa != dwb).sum())
        rep["cw_mismatch"] = int((cwa != cwb).sum())
        rep["dist_bits_mismatch"] = int((xa[..., 0] != xb[..., 0]).sum())
        same_w = dwa == dwb
        rep["dist_max_abs"] = float(np.abs(da - db)[same_w].max()) if same_w.any() else 0.0
        rep["color_bits_mismatch"] = int((xa[..., 1:4] != xb[..., 1:4]).sum())
        same_c = cwa == cwb
        rep["color_max_abs"] = float(np.abs(ca - cb)[same_c].max()) if same_c.any() else 0.0
    ha, hb = a["depth"] > 0, b["depth"] > 0
    both = ha & hb
    rep["pixels"] = int(ha.size)
    rep["hit_mismatch_frac"] = float((ha != hb).mean())
    rep["depth_bits_mismatch"] = int((a["depth"].view(np.int32) != b["depth"].view(np.int32)).sum())
    rep["depth_max_abs_mm"] = float(np.abs(a["depth"] - b["depth"])[both].max() * 1000) if both.any() else 0.0
    rep["depth_gt_half_mm_frac"] = float((np.abs(a["depth"] - b["depth"])[both] > 0.0005).mean()) if both.any() else 0.0
    rep["color_img_max_abs"] = float(np.abs(a["color"] - b["color"]).max())
    rep["normal_img_max_abs"] = float(np.abs(a["normal"] - b["normal"])[both].max()) if both.any() else 0.0
    rep["normal_bits_mismatch"] = int((a["normal"].view(np.int32) != b["normal"].view(np.int32)).sum())
    if "bounds" in a and "bounds" in b:
        # the 80 x 60 ray-bounds grid of the last traced view (Tracer::bounds_), bit for bit
        rep["bounds_bits_mismatch"] = int((np.ascontiguousarray(a["bounds"]).view(np.int32) !=
                                           np.ascontiguousarray(b["bounds"]).view(np.int32)).sum())
    return rep


# ---- image pyramid + depth tracker -------------------------------------------------

def track_scenario(w=160, h=120, rot_deg=0.8, shift=(0.015, -0.01, 0.02), seed=7, dropouts=0.01):
    """Keyframe and frame of the room scene (walls along all three axes + boxes: every degree of freedom
    of the pose is observable) a small rigid motion apart.  Normals come from the CPU oracle's
    ComputeNormals, so every implementation is fed the same arrays.  Returns a dict of float32 arrays:
    K, key/frame depth, normal, colour, raw32 poses (key, true frame pose)."""
    K = np.array(synth.intrinsics(w, h), dtype=np.float32)
    R0 = synth.rotation_y(25.0)
    t0 = np.array([0.3, 0.2, -0.5])
    a = np.deg2rad(rot_deg)
    Rx = np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    R1 = synth.rotation_y(25.0 + rot_deg) @ Rx
    t1 = t0 + np.asarray(shift)
    rng = np.random.RandomState(seed)
    out = {"K": K, "w": w, "h": h}
    lib = oracle_lib()
    mode = lib.vo_get_math_mode()
    lib.vo_set_math_mode(0)  # the input normals are part of the scenario: one arithmetic mode, always
    for name, (R, t) in (("key", (R0, t0)), ("frame", (R1, t1))):
        d = synth.render_room(w, h, K.astype(np.float64), R, t)
        d = np.ascontiguousarray(d, dtype=np.float32)
        d[rng.rand(h, w) < dropouts] = 0.0
        n = np.zeros((h, w, 3), dtype=np.float32)
        lib.vo_compute_normals(_ptr(d), _ptr(K), _ptr(n), w, h)
        yy, xx = np.mgrid[0:h, 0:w]
        c = np.stack([(xx % 17) / 17.0, (yy % 13) / 13.0, ((xx + yy) % 7) / 7.0], axis=-1).astype(np.float32)
        c = np.ascontiguousarray(c * rng.uniform(0.5, 1.0, size=(h, w, 1)).astype(np.float32))
        out[name + "_depth"], out[name + "_normal"], out[name + "_color"] = d, n, c
        out[name + "_pose"] = raw32(R, t)
    lib.vo_set_math_mode(mode)
    return out


def pose_error(T32, truth32):
    """(rotation error in degrees, translation error in metres) between two raw32 poses."""
    A = np.asarray(T32[:16], dtype=np.float64).reshape(4, 4).T
    B = np.asarray(truth32[:16], dtype=np.float64).reshape(4, 4).T
    dR = A[:3, :3].T @ B[:3, :3]
    # sin(angle) from the antisymmetric part: exact for small angles where arccos(trace) loses half the digits
    ang = np.degrees(np.arcsin(np.clip(np.linalg.norm(dR - dR.T) / (2 * np.sqrt(2)), 0, 1)))
    return float(ang), float(np.linalg.norm(A[:3, 3] - B[:3, 3]))


def oracle_downsample(img, nearest):
    img = np.ascontiguousarray(img, dtype=np.float32)
    h, w = img.shape[:2]
    ch = 1 if img.ndim == 2 else 3
    out = np.zeros((h // 2, w // 2) + (() if ch == 1 else (3,)), dtype=np.float32)
    oracle_lib().vo_downsample(_ptr(img), w, h, ch, int(nearest), _ptr(out))
    return out


def oracle_track_pixels(s, pose32, translation=1):
    w, h = s["w"], s["h"]
    r = np.zeros((h, w), dtype=np.float32)
    J = np.zeros((h, w, 6), dtype=np.float32)
    pose = np.ascontiguousarray(pose32, dtype=np.float32)
    oracle_lib().vo_track_pixels(_ptr(pose), _ptr(s["key_pose"]), _ptr(s["key_depth"]), _ptr(s["key_normal"]),
                                 _ptr(s["K"]), w, h, _ptr(s["frame_depth"]), _ptr(s["frame_normal"]), _ptr(s["K"]),
                                 w, h, int(translation), _ptr(r), _ptr(J))
    return r, J


def oracle_track_system(s, pose32, translation=1):
    w, h = s["w"], s["h"]
    H = np.zeros(21, dtype=np.float64)
    g = np.zeros(6, dtype=np.float64)
    pose = np.ascontiguousarray(pose32, dtype=np.float32)
    oracle_lib().vo_track_system(_ptr(pose), _ptr(s["key_pose"]), _ptr(s["key_depth"]), _ptr(s["key_normal"]),
                                 _ptr(s["K"]), w, h, _ptr(s["frame_depth"]), _ptr(s["frame_normal"]), _ptr(s["K"]),
                                 w, h, int(translation), _ptr(H), _ptr(g))
    n = 6 if translation else 3
    return H[:n * (n + 1) // 2], g[:n]


def oracle_track(s, pose32, translation=1, max_iterations=20):
    w, h = s["w"], s["h"]
    pose = np.array(pose32, dtype=np.float32)
    oracle_lib().vo_track_depth.restype = _I
    it = oracle_lib().vo_track_depth(_ptr(pose), _ptr(s["key_pose"]), _ptr(s["key_depth"]), _ptr(s["key_normal"]),
                                     _ptr(s["K"]), w, h, _ptr(s["frame_depth"]), _ptr(s["frame_normal"]),
                                     _ptr(s["K"]), w, h, int(translation), int(max_iterations))
    return pose, it


class RefTracker:
    """The reference's DepthTracker / Frame::Downsample (oracle/_ref) on a track_scenario."""

    def __init__(self, s, pose32, nofmad=False):
        self.lib = ref_lib(nofmad)
        self.s = s
        self.h = self.lib.ref_track_create()
        assert self.h
        w, h = s["w"], s["h"]
        pose = np.ascontiguousarray(pose32, dtype=np.float32)
        assert self.lib.ref_track_set_frame(self.h, 0, _ptr(s["key_depth"]), _ptr(s["key_normal"]),
                                            _ptr(s["key_color"]), w, h, _ptr(s["K"]), _ptr(s["key_pose"])) == 0
        assert self.lib.ref_track_set_frame(self.h, 1, _ptr(s["frame_depth"]), _ptr(s["frame_normal"]),
                                            _ptr(s["frame_color"]), w, h, _ptr(s["K"]), _ptr(pose)) == 0

    def close(self):
        if self.h:
            self.lib.ref_track_destroy(self.h)
            self.h = None

    def residuals(self):
        r = np.zeros((self.s["h"], self.s["w"]), dtype=np.float32)
        assert self.lib.ref_track_residuals(self.h, _ptr(r)) == 0
        return r

    def jacobian(self, translation=1):
        J = np.zeros((self.s["h"], self.s["w"], 6), dtype=np.float32)
        assert self.lib.ref_track_jacobian(self.h, int(translation), _ptr(J)) == 0
        return J

    def system(self, translation=1):
        H = np.zeros(21, dtype=np.float32)
        g = np.zeros(6, dtype=np.float32)
        assert self.lib.ref_track_system(self.h, int(translation), _ptr(H), _ptr(g)) == 0
        n = 6 if translation else 3
        return H[:n * (n + 1) // 2], g[:n]

    def track(self, translation=1, max_iterations=20):
        pose = np.zeros(32, dtype=np.float32)
        assert self.lib.ref_track(self.h, int(translation), int(max_iterations), _ptr(pose)) == 0
        return pose

    def track_pyramid(self):
        pose = np.zeros(32, dtype=np.float32)
        assert self.lib.ref_track_pyramid(self.h, _ptr(pose)) == 0
        return pose

    def downsample(self, which, level):
        w, h = self.s["w"] >> level, self.s["h"] >> level
        d = np.zeros((h, w), dtype=np.float32)
        c = np.zeros((h, w, 3), dtype=np.float32)
        n = np.zeros((h, w, 3), dtype=np.float32)
        k = np.zeros(4, dtype=np.float32)
        assert self.lib.ref_track_downsample(self.h, which, level, _ptr(d), _ptr(c), _ptr(n), _ptr(k)) == 0
        return d, c, n, k


def ours_track_frames(s, pose32):
    """(keyframe, frame) as vulcan_b200 Frames on the GPU."""
    import torch
    import vulcan_b200 as vb
    frames = []
    for name, pose in (("key", s["key_pose"]), ("frame", pose32)):
        f = vb.Frame()
        f.depth_projection = vb.Projection(*s["K"])
        f.color_projection = vb.Projection(*s["K"])
        f.depth_image = torch.from_numpy(s[name + "_depth"]).cuda()
        f.normal_image = torch.from_numpy(s[name + "_normal"]).cuda()
        f.color_image = torch.from_numpy(s[name + "_color"]).cuda()
        p = np.asarray(pose, dtype=np.float32)
        f.depth_to_world_transform = vb.Transform(p[:16].reshape(4, 4).T.copy(), p[16:].reshape(4, 4).T.copy())
        frames.append(f)
    return frames


# ---- surface points ---------------------------------------------------------------

def oracle_volume_for(scn):
    """The oracle volume after the scenario's frames (run_oracle without the trace)."""
    dmin, dmax = scn.get("depth_range", (0.1, 5.0))
    ov = OracleVolume(scn["M"], scn["E"], scn["voxel"], scn["trunc"], dmin, dmax)
    ov.lib.vo_set_math_mode(scn.get("math", 0))
    for fr in scn["frames"]:
        ov.set_view_converged(fr["depth"], scn["K"], fr["T"])
        if scn["mode"] == "color":
            ov.integrate_color(fr["depth"], scn["K"], fr["T"], fr["color"], scn["K"], scn["Tcd"], dmin, dmax)
        else:
            ov.integrate_depth(fr["depth"], scn["K"], fr["T"], dmin, dmax)
    assert ov.dropped() == 0
    return ov


def oracle_extract_points(ov):
    n = ov.lib.vo_extract_points(ov.h, None, None, 0)
    pts = np.zeros((n, 3), dtype=np.float32)
    col = np.zeros((n, 3), dtype=np.float32)
    assert ov.lib.vo_extract_points(ov.h, _ptr(pts), _ptr(col), n) == n
    return pts, col


def oracle_extract_mesh(ov):
    """(points, colours, faces) of the oracle's marching-cubes mesh (vo_extract_mesh)."""
    lib = ov.lib
    lib.vo_extract_mesh.restype = C.c_longlong
    lib.vo_extract_mesh.argtypes = [_P, _P, _P, C.c_longlong, _P, C.c_longlong, _P]
    nf = C.c_longlong(0)
    n = lib.vo_extract_mesh(ov.h, None, None, 0, None, 0, C.byref(nf))
    pts = np.zeros((n, 3), dtype=np.float32)
    col = np.zeros((n, 3), dtype=np.float32)
    faces = np.zeros((nf.value, 3), dtype=np.int32)
    assert lib.vo_extract_mesh(ov.h, _ptr(pts), _ptr(col), n, _ptr(faces), nf.value, C.byref(nf)) == n
    return pts, col, faces


def canonical_triangles(pts, faces):
    """Triangles as rows of 9 vertex-coordinate bit patterns, each rotated so that its smallest corner comes
    first (winding kept), rows sorted: equal for two meshes iff they have the same oriented triangles."""
    tri = np.ascontiguousarray(pts, dtype=np.float32).view(np.uint32)[faces]          # (F, 3, 3)
    keys = tri[..., 0].astype(np.uint64) * np.uint64(1 << 42) + tri[..., 1].astype(np.uint64) * np.uint64(1 << 21) \
        + tri[..., 2].astype(np.uint64) % np.uint64(1 << 21)
    first = np.argmin(keys, axis=1)
    idx = (first[:, None] + np.arange(3)[None, :]) % 3
    rows = np.take_along_axis(tri, idx[:, :, None], axis=1).reshape(len(faces), 9)
    return rows[np.lexsort(rows.T[::-1])]


def mesh_edge_report(faces):
    """Directed-edge statistics of a triangle mesh: a closed, consistently oriented surface uses every
    directed edge once and always together with its opposite."""
    f = np.asarray(faces, dtype=np.int64)
    a = np.concatenate([f[:, 0], f[:, 1], f[:, 2]])
    b = np.concatenate([f[:, 1], f[:, 2], f[:, 0]])
    m = int(max(a.max(), b.max())) + 1
    fwd = a * m + b
    uniq, counts = np.unique(fwd, return_counts=True)
    rev = np.isin(b * m + a, fwd)
    return dict(edges=int(len(fwd)), repeated_directed=int((counts > 1).sum()), boundary=int((~rev).sum()),
                degenerate=int(((f[:, 0] == f[:, 1]) | (f[:, 1] == f[:, 2]) | (f[:, 0] == f[:, 2])).sum()))


def sorted_rows(pts, col):
    """Rows of [points | colours] in a canonical order (bit patterns), for set comparison."""
    rows = np.concatenate([pts, col], axis=1).astype(np.float32).view(np.uint32)
    order = np.lexsort(rows.T[::-1])
    return rows[order]


def sphere_plane_distance(pts, sphere_c=(0.0, 0.0, 2.0), sphere_r=0.5, plane_y=0.5):
    """Distance of points to the analytic sphere-on-plane surface of synth.render_sphere_plane."""
    p = pts.astype(np.float64)
    ds = np.abs(np.linalg.norm(p - np.asarray(sphere_c), axis=1) - sphere_r)
    dp = np.abs(p[:, 1] - plane_y)
    return np.minimum(ds, dp)
