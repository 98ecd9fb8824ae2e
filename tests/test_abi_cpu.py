"""CPU-only checks of the drop-in boundary: the C ABI header is plain C, the built library loads
without a GPU and exports every function the header declares (no compute calls here), struct
layouts match the ctypes mirror, and entry points fail loudly - never fall back - without a device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np

import harness as H

HEADER = os.path.join(H.ROOT, "include", "vulcan_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vk_[a-z0-9_]+)\s*\(", src)))


def test_header_is_plain_c(tmp_path):
    c = tmp_path / "abi.c"
    c.write_text('#include "vulcan_b200.h"\n'
                 "int probe(void) { return (int)sizeof(vkVolume) + (int)sizeof(vkFrame); }\n")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only",
                           "-I", os.path.join(H.ROOT, "include"), str(c)])


def test_library_exports_every_declared_symbol():
    import vulcan_b200 as vb
    names = _declared_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(vb.lib, n), "libvulcan_b200.so does not export " + n
        assert n in vb._lib.SYMBOLS, "ctypes mirror misses " + n
    assert vb.lib.vk_version() >= 100


def test_struct_layouts_match_header(tmp_path):
    """sizeof / offsetof of the C structs as the C compiler sees them vs the ctypes mirror."""
    import vulcan_b200 as vb
    c = tmp_path / "layout.c"
    c.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "vulcan_b200.h"\n'
                 "int main(void) { printf(\"%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n\", sizeof(vkVoxel), sizeof(vkHashEntry),"
                 " sizeof(vkPatch), sizeof(vkTransform), sizeof(vkProjection), sizeof(vkLight), sizeof(vkVolume),"
                 " sizeof(vkFrame), offsetof(vkVolume, math_mode), offsetof(vkFrame, depth_to_color),"
                 " sizeof(vkPeerTable), sizeof(vkShard), offsetof(vkShard, push_entries), sizeof(vkIpcHandles));"
                 " return 0; }\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(H.ROOT, "include"), "-o", str(exe), str(c)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    L = vb._lib
    assert got[:6] == [20, 16, 16, 128, 16, 16]   # Appendix B of SURVEY.md: the reference's value types
    assert got[6] == C.sizeof(L.vkVolume) and got[7] == C.sizeof(L.vkFrame)
    assert got[8] == L.vkVolume.math_mode.offset and got[9] == L.vkFrame.depth_to_color.offset
    assert got[10] == C.sizeof(L.vkPeerTable) and got[11] == C.sizeof(L.vkShard)
    assert got[12] == L.vkShard.push_entries.offset and got[13] == C.sizeof(L.vkIpcHandles) == 192


def test_no_device_is_an_error_not_a_fallback():
    """Without a CUDA device the product path reports VK_E_NODEVICE / a CUDA error."""
    import torch
    import vulcan_b200 as vb
    if torch.cuda.is_available():
        return
    v = vb._lib.vkVolume()
    rc = vb.lib.vk_volume_create(C.byref(v), 64, 16, None)
    assert rc != 0
    try:
        vb.api.check(rc, "vk_volume_create")
        raise AssertionError("check() must raise")
    except vb.VulcanError:
        pass
    assert vb.lib.vk_set_view(None, None, None) == -1   # VK_E_INVALID


def test_product_does_not_touch_the_oracle():
    """Nothing under vulcan_b200/ or include/ may reference oracle/ (the oracle is test infrastructure)."""
    bad = []
    for root in ("vulcan_b200", "include"):
        for dp, _, files in os.walk(os.path.join(H.ROOT, root)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"oracle/|vulcan_oracle|libvulcan_ref|import harness", txt):
                        # comments that merely name the oracle are fine; loading it is not
                        if re.search(r"CDLL\([^)]*oracle|dlopen\([^)]*oracle|#include\s+[\"<][^\">]*oracle", txt):
                            bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_transform_mirror():
    """Transform mirror (include/vulcan/transform.h): inverse is a swap, products multiply both."""
    import vulcan_b200 as vb
    a = vb.Transform.Translate(0.3, -1.3, 3.7)
    th = 0.4
    R = np.array([[np.cos(th), 0, np.sin(th)], [0, 1, 0], [-np.sin(th), 0, np.cos(th)]])
    b = vb.Transform.Rotate(R)
    ab = a * b
    assert np.allclose(ab.matrix @ ab.inv_matrix, np.eye(4), atol=1e-6)
    assert np.array_equal(ab.Inverse().matrix, ab.inv_matrix)
    raw = ab.raw32()
    assert raw.shape == (32,) and raw.dtype == np.float32
    assert np.array_equal(raw[:16].reshape(4, 4).T, ab.matrix)
