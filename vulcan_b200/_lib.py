"""ctypes binding of the C ABI declared in include/vulcan_b200.h.

The product path is the CUDA library only: if libvulcan_b200.so is missing this module
raises ImportError (there is no CPU fallback and nothing here touches oracle/).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvulcan_b200.so")

VK_CNT_VISIBLE0 = 0
VK_CNT_VOXEL_PTR = 2
VK_CNT_EXCESS_PTR = 3
VK_CNT_OVERFLOW = 4
VK_CNT_ALLOCATED = 6
VK_CNT_COUNT = 16
VK_E_OVERFLOW = -3
VK_E_TOO_LARGE = -6
VK_MESH_STATIC_MAX_POOL, VK_MESH_STATIC_MAX_BLOCKS = 131072, 2048
VK_MATH_REFERENCE, VK_MATH_STRICT = 0, 1
BOUNDS_W, BOUNDS_H = 80, 60


class vkProjection(C.Structure):
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


class vkTransform(C.Structure):
    _fields_ = [("m", C.c_float * 16), ("inv", C.c_float * 16)]


class vkLight(C.Structure):
    _fields_ = [("intensity", C.c_float), ("position", C.c_float * 3)]


class vkVolume(C.Structure):
    _fields_ = [
        ("voxels", C.c_void_p), ("entries", C.c_void_p), ("free_list", C.c_void_p),
        ("visibility", C.c_void_p), ("visible", C.c_void_p * 2), ("stamps", C.c_void_p),
        ("locks", C.c_void_p), ("lock_tags", C.c_void_p), ("bucket_bits", C.c_void_p), ("space_bits", C.c_void_p), ("sdf", C.c_void_p),
        ("counters", C.c_void_p), ("host_counters", C.POINTER(C.c_int32)),
        ("main_blocks", C.c_int32), ("excess_blocks", C.c_int32),
        ("voxel_length", C.c_float), ("trunc_length", C.c_float),
        ("min_depth", C.c_float), ("max_depth", C.c_float),
        ("shard_rank", C.c_int32), ("shard_count", C.c_int32), ("shard_shift", C.c_int32), ("shard_halo", C.c_int32),
        ("visible_slot", C.c_int32), ("view_id", C.c_int32), ("math_mode", C.c_int32), ("has_color", C.c_int32),
        ("shard", C.c_void_p),
        ("aux_stream", C.c_void_p), ("ev_view", C.c_void_p), ("ev_bounds", C.c_void_p), ("ev_trace", C.c_void_p),
        ("aux_valid", C.c_int32), ("reserved0", C.c_int32),
    ]


MAX_SHARDS = 8


class vkPeerTable(C.Structure):
    _fields_ = [("count", C.c_int32), ("rank", C.c_int32), ("entries", C.c_void_p * MAX_SHARDS),
                ("voxels", C.c_void_p * MAX_SHARDS), ("sdf", C.c_void_p * MAX_SHARDS)]


class vkShard(C.Structure):
    _fields_ = [("count", C.c_int32), ("rank", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("epoch", C.c_int32), ("main_blocks", C.c_int32), ("excess_blocks", C.c_int32), ("connected", C.c_int32),
                ("entries", C.c_void_p * MAX_SHARDS), ("sdf", C.c_void_p * MAX_SHARDS), ("voxels", C.c_void_p * MAX_SHARDS),
                ("push_entries", C.c_void_p * MAX_SHARDS), ("push_sdf", C.c_void_p * MAX_SHARDS),
                ("exchange", C.c_void_p * MAX_SHARDS)]


class vkShardHandles(C.Structure):
    _fields_ = [("voxels", C.c_ubyte * 64), ("exchange", C.c_ubyte * 64),
                ("rep_entries", (C.c_ubyte * 64) * MAX_SHARDS), ("rep_sdf", (C.c_ubyte * 64) * MAX_SHARDS)]


class vkIpcHandles(C.Structure):
    _fields_ = [("entries", C.c_ubyte * 64), ("voxels", C.c_ubyte * 64), ("sdf", C.c_ubyte * 64)]


class vkFrame(C.Structure):
    _fields_ = [
        ("depth", C.c_void_p), ("color", C.c_void_p), ("normal", C.c_void_p),
        ("depth_w", C.c_int32), ("depth_h", C.c_int32), ("color_w", C.c_int32), ("color_h", C.c_int32),
        ("depth_proj", vkProjection), ("color_proj", vkProjection),
        ("depth_to_world", vkTransform), ("depth_to_color", vkTransform),
    ]


# every symbol include/vulcan_b200.h declares: name -> argtypes
_P = C.c_void_p
_F = C.c_float
_I = C.c_int
SYMBOLS = {
    "vk_volume_create": [C.POINTER(vkVolume), _I, _I, _P],
    "vk_volume_reset": [C.POINTER(vkVolume), _P],
    "vk_volume_destroy": [C.POINTER(vkVolume)],
    "vk_set_view": [C.POINTER(vkVolume), C.POINTER(vkFrame), _P],
    "vk_volume_sync_counters": [C.POINTER(vkVolume), C.POINTER(_I), _P],
    "vk_volume_sync_sdf": [C.POINTER(vkVolume), _P],
    "vk_volume_save": [C.POINTER(vkVolume), C.c_char_p, _P],
    "vk_volume_load": [C.POINTER(vkVolume), C.c_char_p, _P],
    "vk_integrate_depth": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _F, _P],
    "vk_integrate_color": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _F, _F, _P],
    "vk_compute_frame_mask": [C.POINTER(vkFrame), _F, _P, _P],
    "vk_integrate_light": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _F, _F, C.POINTER(vkLight), _F, _P, _P],
    "vk_trace": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _P, _P, _P, _P, _P],
    "vk_trace_keys": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _P, _P, _P],
    "vk_resolve_keys": [_P, C.POINTER(vkFrame), _P, _P, _P, _P],
    "vk_compute_patches": [_P, _P, C.POINTER(vkTransform), C.POINTER(vkProjection), _F, _F, _F, _I, _I, _I, _I, _I, _P, _P, _P],
    "vk_compute_bounds": [_P, _P, _I, _I, _P],
    "vk_reset_bounds": [_P, _I, _P],
    "vk_compute_points": [_P, _P, _P, _I, _F, _F, _F, C.POINTER(vkTransform), C.POINTER(vkProjection), _P, _P, _I, _I, _I, _I, _P],
    "vk_compute_normals": [_P, C.POINTER(vkProjection), _P, _I, _I, _P],
    "vk_filter_depths": [_I, _I, _P, _P, _P],
    "vk_prepare_frame": [_I, _I, _P, C.POINTER(vkProjection), _P, _P, _P],
    "vk_shard_owner": [_I, _I, _I, _I, _I],
    "vk_trace_bounds": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _P, _P],
    "vk_trace_rows": [C.POINTER(vkVolume), C.POINTER(vkPeerTable), C.POINTER(vkFrame), _P, _I, _I, _P, _P, _P, _P],
    "vk_ipc_export": [C.POINTER(vkVolume), C.POINTER(vkIpcHandles)],
    "vk_ipc_open": [C.POINTER(vkIpcHandles), C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)],
    "vk_ipc_close": [_P, _P, _P],
    "vk_shard_create": [C.POINTER(vkShard), C.POINTER(vkVolume), _I, _I, _I, _I, _I, _P],
    "vk_shard_export": [C.POINTER(vkShard), C.POINTER(vkVolume), C.POINTER(vkShardHandles)],
    "vk_shard_connect": [C.POINTER(vkShard), C.POINTER(vkShardHandles)],
    "vk_shard_connect_local": [C.POINTER(C.POINTER(vkShard)), C.POINTER(C.POINTER(vkVolume)), _I],
    "vk_shard_destroy": [C.POINTER(vkShard), C.POINTER(vkVolume)],
    "vk_trace_sharded": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _I, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), _P],
    "vk_trace_sharded_staged": [C.POINTER(vkVolume), C.POINTER(vkFrame), _F, _F, _I, _I, C.POINTER(_P), C.POINTER(_P),
                                C.POINTER(_P), _P],
    "vk_shard_band": [C.POINTER(vkVolume), C.POINTER(_I), C.POINTER(_I), _P],
    "vk_shard_band_async": [C.POINTER(vkVolume), _P, _P],
    "vk_downsample_frame": [C.POINTER(vkFrame), C.POINTER(vkFrame), _P],
    "vk_downsample_image": [_P, _I, _I, _I, _I, _P, _P],
    "vk_track_residuals": [C.POINTER(vkFrame), C.POINTER(vkFrame), _P, _P],
    "vk_track_jacobian": [C.POINTER(vkFrame), C.POINTER(vkFrame), _I, _P, _P],
    "vk_track_workspace_floats": [],
    "vk_track_system": [C.POINTER(vkFrame), C.POINTER(vkFrame), _I, _P, _P, _P, _P],
    "vk_track_depth": [C.POINTER(vkFrame), C.POINTER(vkFrame), _I, _I, _P, C.POINTER(_I), C.POINTER(_F), _P],
    "vk_extract_points": [C.POINTER(vkVolume), _P, _P, _I, _P, _P],
    "vk_extract_mesh": [C.POINTER(vkVolume), _P, _P, _I, _P, _P, _I, _P, _P],
    "vk_extract_mesh_static": [C.POINTER(vkVolume), _P, _P, _I, _P, _P, _I, _P, _P],
    "vk_version": [],
    "vk_set_default_math_mode": [_I],
    "vk_get_default_math_mode": [],
    "vk_div_prepare": [_F, _P],
    "vk_launch_count": [],
}


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "vulcan_b200: %s is missing - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.argtypes = argtypes
        fn.restype = (C.c_ulonglong if name == "vk_launch_count" else
                      C.c_int)
    return lib


lib = load()


class VulcanError(RuntimeError):
    """Mirror of vulcan::Exception (include/vulcan/exception.h:33-81) for status codes."""


def check(status, what):
    if status == 0:
        return
    if status > 0:
        raise VulcanError("%s: cuda error %d" % (what, status))
    names = {-1: "invalid argument", -2: "no CUDA device", -3: "voxel/excess pool exhausted",
             -4: "snapshot file missing, truncated or of another version",
             -5: "a peer of the sharded volume did not arrive at a handshake"}
    raise VulcanError("%s: %s" % (what, names.get(status, str(status))))
