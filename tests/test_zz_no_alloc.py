"""vk_extract_mesh_static (include/vulcan_b200.h): the mesh extraction without any allocation - its workspace
is a file-scope device array instead of a cudaMalloc per call.  Called through the C-ABI and compared with
vk_extract_mesh on the same volume: the same points (bit for bit, as sorted sets - the order of appended
blocks is arbitrary in both) and the same oriented triangles, on two calls in a row (the array keeps the
previous call's contents), then on a second volume; volumes beyond the stated limits are refused.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _volume(scn):
    import torch
    import vulcan_b200 as vb
    vol = vb.Volume(scn["M"], scn["E"])
    vol.SetVoxelLength(scn["voxel"])
    vol.SetTruncationLength(scn["trunc"])
    integ = (vb.ColorIntegrator if scn["mode"] == "color" else vb.DepthIntegrator)(vol)
    for fr in scn["frames"]:
        f = vb.Frame()
        f.depth_projection = vb.Projection(*scn["K"])
        f.color_projection = vb.Projection(*scn["K"])
        f.depth_to_world_transform = vb.Transform(fr["T"][:16].reshape(4, 4).T.copy(), fr["T"][16:].reshape(4, 4).T.copy())
        f.depth_to_color_transform = vb.Transform(scn["Tcd"][:16].reshape(4, 4).T.copy(),
                                                  scn["Tcd"][16:].reshape(4, 4).T.copy())
        f.depth_image = torch.from_numpy(fr["depth"]).cuda()
        if fr["color"] is not None:
            f.color_image = torch.from_numpy(fr["color"]).cuda()
        vol.SetView(f)
        integ.Integrate(f)
    return vol


def _extract(fn, vol, capacity, poison):
    """One call of an extraction entry point through the C-ABI; returns (rc, points, colors, faces)."""
    import torch
    import vulcan_b200 as vb
    pts = torch.full((capacity, 3), poison, dtype=torch.float32, device="cuda")
    col = torch.full((capacity, 3), poison, dtype=torch.float32, device="cuda")
    tri = torch.full((2 * capacity, 3), -7, dtype=torch.int32, device="cuda")
    cnt = torch.full((2,), -1, dtype=torch.int32, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(C.byref(vol._v), pts.data_ptr(), col.data_ptr(), capacity, cnt.data_ptr(), tri.data_ptr(), 2 * capacity,
            cnt[1:].data_ptr(), stream)
    torch.cuda.synchronize()
    if rc != 0:
        return rc, pts, col, tri, cnt
    n, nf = (int(x) for x in cnt.tolist())
    assert 0 < n <= capacity and 0 < nf <= 2 * capacity
    return rc, pts[:n].cpu().numpy(), col[:n].cpu().numpy(), tri[:nf].cpu().numpy(), cnt


def _same_mesh(a, b):
    _, pa, ca, fa, _ = a
    _, pb, cb, fb, _ = b
    assert len(pa) == len(pb) and len(fa) == len(fb)
    assert np.array_equal(H.sorted_rows(pa, ca), H.sorted_rows(pb, cb))                        # bit for bit
    assert fa.min() >= 0 and fa.max() < len(pa)
    assert np.array_equal(H.canonical_triangles(pa, fa), H.canonical_triangles(pb, fb))


def test_static_mesh_matches_the_allocating_entry_point():
    import torch
    import vulcan_b200 as vb
    from vulcan_b200 import _lib
    lib = vb.lib
    scn = H.make_scenario("tiny", (0, 6), "depth", 5.0)
    vol = _volume(scn)
    allocated = int(vol.SyncCounters()[1][_lib.VK_CNT_ALLOCATED])
    assert 0 < allocated <= _lib.VK_MESH_STATIC_MAX_BLOCKS
    cap = 1 << 17
    want = _extract(lib.vk_extract_mesh, vol, cap, 0.0)
    assert want[0] == 0 and len(want[1]) > 1000 and len(want[3]) > len(want[1])
    # two calls in a row: the second one finds the first one's slots and vertex indices in the array
    got1 = _extract(lib.vk_extract_mesh_static, vol, cap, 1.0)
    got2 = _extract(lib.vk_extract_mesh_static, vol, cap, 2.0)
    assert got1[0] == 0 and got2[0] == 0
    _same_mesh(got1, want)
    _same_mesh(got2, want)
    # a different, smaller model afterwards: nothing of the larger one left in the shared array shows through
    scn_b = H.make_scenario("micro", (0,), "depth", 5.0)
    vol_b = _volume(scn_b)
    want_b = _extract(lib.vk_extract_mesh, vol_b, cap, 0.0)
    got_b = _extract(lib.vk_extract_mesh_static, vol_b, cap, 3.0)
    assert want_b[0] == 0 and got_b[0] == 0 and len(want_b[1]) < len(want[1])
    _same_mesh(got_b, want_b)
    # and the first volume again
    _same_mesh(_extract(lib.vk_extract_mesh_static, vol, cap, 4.0), want)


def test_static_mesh_refuses_what_does_not_fit():
    import torch
    import vulcan_b200 as vb
    from vulcan_b200 import _lib
    lib = vb.lib
    # more allocated blocks than the array has rows for
    scn = H.make_scenario("small", tuple(range(4)), "depth", 45.0)
    vol = _volume(scn)
    assert int(vol.SyncCounters()[1][_lib.VK_CNT_ALLOCATED]) > _lib.VK_MESH_STATIC_MAX_BLOCKS
    rc, pts, col, tri, cnt = _extract(lib.vk_extract_mesh_static, vol, 1 << 12, 5.0)
    assert rc == _lib.VK_E_TOO_LARGE
    assert bool((pts == 5.0).all()) and bool((tri == -7).all()) and cnt.tolist() == [-1, -1]   # nothing written
    del vol
    # a pool larger than the block -> slot table
    big = vb.Volume(_lib.VK_MESH_STATIC_MAX_POOL, 16)
    big.SetVoxelLength(0.008)
    big.SetTruncationLength(0.032)
    rc = _extract(lib.vk_extract_mesh_static, big, 1 << 10, 6.0)[0]
    assert rc == _lib.VK_E_TOO_LARGE
    # bad arguments like the allocating entry point
    assert lib.vk_extract_mesh_static(None, None, None, 0, None, None, 0, None, None) == -1
