"""GPU parity of the STEP-WISE tracer entry points - the free functions of include/vulcan/tracer.cuh:14-32
(ComputePatches, ComputeBounds, ResetBoundsBuffer, ComputePoints), which the reference tests call directly
(tests/tracer_test.cu:22-244,326-337) - against the reference's own functions run on the SAME device
buffers (the layouts are byte images of each other) and against the CPU oracle; plus an explicit check of
the fused 80 x 60 ray-bounds grid against the reference Tracer's own buffer."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu

PATCH = np.dtype([("origin", "<i2", 2), ("size", "<i2", 2), ("bounds", "<f4", 2)])
FMAX = np.finfo(np.float32).max


def _sorted_patches(raw):
    """Patches as rows of 4 int32 words in a canonical order (order is unspecified, Q7)."""
    rows = np.ascontiguousarray(raw).view(np.int32).reshape(-1, 4)
    return rows[np.lexsort(rows.T[::-1])]


def _vk_transform(vb, raw):
    return vb.Transform(raw[:16].reshape(4, 4).T.copy(), raw[16:].reshape(4, 4).T.copy()).to_c()


def _inverse32(raw):
    return np.concatenate([raw[16:], raw[:16]]).astype(np.float32)


def _fused_volume(scn):
    """Our volume after the scenario's frames (kept alive) + the last frame."""
    import torch
    import vulcan_b200 as vb
    vb.SetDefaultMathMode(scn.get("math", 0))
    dmin, dmax = scn["depth_range"]
    vol = vb.Volume(scn["M"], scn["E"])
    vol.SetVoxelLength(scn["voxel"])
    vol.SetTruncationLength(scn["trunc"])
    vol.SetDepthRange(dmin, dmax)
    integ = vb.DepthIntegrator(vol)
    integ.SetDepthRange(dmin, dmax)
    frame = None
    for fr in scn["frames"]:
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*scn["K"])
        T = fr["T"]
        frame.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
        frame.depth_image = torch.from_numpy(fr["depth"]).cuda()
        vol.SetView(frame)
        integ.Integrate(frame)
    return vol, frame


@pytest.mark.parametrize("math", [0, 1])
@pytest.mark.parametrize("name,frames", [("small", (0, 10)), ("room_small", (0, 1))])
def test_stepwise_tracer_matches_reference_and_oracle(name, frames, math):
    import torch
    import vulcan_b200 as vb
    lib = vb.lib
    if not H.have_ref(nofmad=bool(math)):
        pytest.fail("oracle/_ref/libvulcan_ref*.so missing: run oracle/build_ref.sh")
    ref = H.ref_lib(nofmad=bool(math))
    for fn in ("ref_compute_patches", "ref_compute_bounds", "ref_reset_bounds", "ref_compute_points"):
        getattr(ref, fn).restype = C.c_int
    scn = H.make_scenario(name, frames, "depth", math=math)
    w, h = scn["w"], scn["h"]
    dmin, dmax = scn["depth_range"]
    vol, frame = _fused_volume(scn)
    stream = vb.api._stream()
    n_vis = vol.GetVisibleCount()
    assert n_vis > 100
    T = scn["frames"][-1]["T"]
    Tcw = _vk_transform(vb, _inverse32(T))
    Twc = _vk_transform(vb, T)
    proj = vb.Projection(*scn["K"]).to_c()
    L = np.float32(8) * np.float32(scn["voxel"])
    v = vol._v
    visible_ptr = v.visible[v.visible_slot]
    cap = 64 * n_vis

    # ---- ComputePatches (tracer.cuh:14-18) ------------------------------------------------------
    ours_p = torch.zeros((cap, 4), dtype=torch.int32, device="cuda")
    ours_n = torch.zeros(1, dtype=torch.int32, device="cuda")
    vb.api.check(lib.vk_compute_patches(visible_ptr, v.entries, C.byref(Tcw), C.byref(proj), float(L), dmin, dmax, n_vis,
                                        w, h, 80, 60, ours_p.data_ptr(), ours_n.data_ptr(), stream), "vk_compute_patches")
    ref_p = torch.zeros((cap, 4), dtype=torch.int32, device="cuda")
    ref_n = torch.zeros(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    Tcw32 = _inverse32(T)
    K = np.ascontiguousarray(scn["K"], dtype=np.float32)
    assert ref.ref_compute_patches(C.c_void_p(visible_ptr), C.c_void_p(v.entries), Tcw32.ctypes.data_as(C.c_void_p),
                                   K.ctypes.data_as(C.c_void_p), C.c_float(L), C.c_float(dmin), C.c_float(dmax), n_vis,
                                   w, h, 80, 60, C.c_void_p(ref_p.data_ptr()), C.c_void_p(ref_n.data_ptr())) == 0
    n_ours, n_ref = int(ours_n.item()), int(ref_n.item())
    assert n_ours == n_ref and 0 < n_ours <= cap
    a = _sorted_patches(ours_p[:n_ours].cpu().numpy())
    b = _sorted_patches(ref_p[:n_ref].cpu().numpy())
    assert np.array_equal(a, b), "patch multiset differs from the reference's ComputePatches"
    # the CPU oracle on its own volume of the same scenario (entry indices differ, patches must not)
    ov = H.oracle_volume_for(scn)
    ovis = np.ascontiguousarray(ov.visible())
    o_p = np.zeros(cap, dtype=PATCH)
    o_n = ov.lib.vo_compute_patches(ov.h, ovis.ctypes.data_as(C.c_void_p), len(ovis), Tcw32.ctypes.data_as(C.c_void_p),
                                    K.ctypes.data_as(C.c_void_p), float(L), dmin, dmax, w, h, 80, 60,
                                    o_p.ctypes.data_as(C.c_void_p))
    assert o_n == n_ours
    assert np.array_equal(_sorted_patches(o_p[:o_n]), a), "patch multiset differs from the oracle"

    # ---- ResetBoundsBuffer + ComputeBounds (tracer.cuh:20-21,32) -----------------------------------
    ours_b = torch.zeros((60, 80, 2), dtype=torch.float32, device="cuda")
    vb.api.check(lib.vk_reset_bounds(ours_b.data_ptr(), 80 * 60, stream), "vk_reset_bounds")
    assert (ours_b[..., 0] == FMAX).all() and (ours_b[..., 1] == -FMAX).all()
    vb.api.check(lib.vk_compute_bounds(ours_p.data_ptr(), ours_b.data_ptr(), 80, n_ours, stream), "vk_compute_bounds")
    ref_b = torch.zeros((60, 80, 2), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    assert ref.ref_reset_bounds(C.c_void_p(ref_b.data_ptr()), 80 * 60) == 0
    assert ref.ref_compute_bounds(C.c_void_p(ref_p.data_ptr()), C.c_void_p(ref_b.data_ptr()), 80, n_ref) == 0
    ob, rb = ours_b.cpu().numpy(), ref_b.cpu().numpy()
    assert np.array_equal(ob.view(np.int32), rb.view(np.int32)), "bounds grid differs from the reference's ComputeBounds"
    o_b = np.zeros((60, 80, 2), dtype=np.float32)
    ov.lib.vo_compute_bounds(o_p.ctypes.data_as(C.c_void_p), o_n, o_b.ctypes.data_as(C.c_void_p), 80, 60, 1)
    assert np.array_equal(ob.view(np.int32), o_b.view(np.int32)), "bounds grid differs from the oracle"
    assert (ob[..., 0] < ob[..., 1]).sum() > 50   # not vacuous
    # the FUSED kernel of vk_trace (no patch list) leaves the same grid
    tracer = vb.Tracer(vol)
    tracer.SetDepthRange(dmin, dmax)
    out = vb.Frame()
    out.depth_projection, out.depth_to_world_transform = frame.depth_projection, frame.depth_to_world_transform
    out.depth_image = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    tracer.Trace(out)
    fused = tracer.bounds_.cpu().numpy()
    assert np.array_equal(fused.view(np.int32), ob.view(np.int32)), "fused bounds kernel differs from patches + bounds"

    # ---- ComputePoints (tracer.cuh:23-28) on raw pointers -----------------------------------------
    d_ours = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    c_ours = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
    vb.api.check(lib.vk_compute_points(v.entries, v.voxels, ours_b.data_ptr(), v.main_blocks, float(L), scn["voxel"],
                                       scn["trunc"], C.byref(Twc), C.byref(proj), d_ours.data_ptr(), c_ours.data_ptr(),
                                       w, h, 80, 60, stream), "vk_compute_points")
    d_ref = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    c_ref = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    Twc32 = np.ascontiguousarray(T, dtype=np.float32)
    assert ref.ref_compute_points(C.c_void_p(v.entries), C.c_void_p(v.voxels), C.c_void_p(ref_b.data_ptr()), v.main_blocks,
                                  C.c_float(L), C.c_float(scn["voxel"]), C.c_float(scn["trunc"]),
                                  Twc32.ctypes.data_as(C.c_void_p), K.ctypes.data_as(C.c_void_p),
                                  C.c_void_p(d_ref.data_ptr()), C.c_void_p(c_ref.data_ptr()), w, h, 80, 60) == 0
    do, dr = d_ours.cpu().numpy(), d_ref.cpu().numpy()
    assert (do > 0).mean() > 0.2
    assert np.array_equal(do.view(np.int32), dr.view(np.int32)), "ComputePoints depth differs from the reference"
    assert np.array_equal(c_ours.cpu().numpy().view(np.int32), c_ref.cpu().numpy().view(np.int32))
    # and the volume-level trace (filters, caches, free-space runs) equals the plain step-wise call
    assert np.array_equal(out.depth_image.cpu().numpy().view(np.int32), do.view(np.int32))
    assert np.array_equal(out.color_image.cpu().numpy().view(np.int32), c_ours.cpu().numpy().view(np.int32))


def test_patches_reference_scenario():
    """tests/tracer_test.cu:22-166: five blocks placed under given pixels / depths, a rotated and
    translated camera, f = (346.723, 353.914), c = (321.294, 239.052): patch origin / size exact, bounds
    to 1e-4, order free - here bit for bit against the reference's own function and the oracle."""
    import torch
    import vulcan_b200 as vb
    lib = vb.lib
    ref = H.ref_lib()
    ref.ref_compute_patches.restype = C.c_int
    block_length = np.float32(0.008)
    q = np.array([0.7474, 0.3438, -0.3884, 0.4152])
    q /= np.linalg.norm(q)
    qw, qx, qy, qz = q
    R = np.array([[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qz * qw), 2 * (qx * qz + qy * qw)],
                  [2 * (qx * qy + qz * qw), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qx * qw)],
                  [2 * (qx * qz - qy * qw), 2 * (qy * qz + qx * qw), 1 - 2 * (qx * qx + qy * qy)]])
    Tcw = H.raw32(R, np.array([0.3, -1.3, 3.7]))
    K = np.array([346.723, 353.914, 321.294, 239.052], dtype=np.float32)
    pts = [(320.0, 240.0, 2.5), (120.0, 340.0, 1.5), (420.0, 240.0, 0.6), (-20.0, -40.0, 1.0), (720.0, 580.0, 1.0)]
    Minv = Tcw[16:].reshape(4, 4).T.astype(np.float64)
    entries = np.zeros((len(pts), 4), dtype=np.int32)
    for i, (u, v, d) in enumerate(pts):
        Xc = np.array([(u - K[2]) / K[0] * d, (v - K[3]) / K[1] * d, d, 1.0])
        Xw = Minv @ Xc
        b = (Xw[:3] / float(block_length)).astype(np.int64)   # truncation, like the test's short conversion
        packed = np.array([(int(b[0]) & 0xFFFF) | ((int(b[1]) & 0xFFFF) << 16), int(b[2]) & 0xFFFF], dtype=np.uint32)
        entries[i, :2] = packed.view(np.int32)
        entries[i, 2:] = (0, -1)
    d_entries = torch.from_numpy(entries).cuda()
    d_idx = torch.arange(len(pts), dtype=torch.int32, device="cuda")
    cap = 4096
    for math in (0, 1):
        vb.SetDefaultMathMode(math)
        rl = H.ref_lib(nofmad=bool(math))
        rl.ref_compute_patches.restype = C.c_int
        op = torch.zeros((cap, 4), dtype=torch.int32, device="cuda")
        on = torch.zeros(1, dtype=torch.int32, device="cuda")
        t = _vk_transform(vb, Tcw)
        proj = vb.Projection(*K).to_c()
        vb.api.check(lib.vk_compute_patches(d_idx.data_ptr(), d_entries.data_ptr(), C.byref(t), C.byref(proj),
                                            float(block_length), 0.1, 5.0, len(pts), 640, 480, 80, 60, op.data_ptr(),
                                            on.data_ptr(), vb.api._stream()), "vk_compute_patches")
        rp = torch.zeros((cap, 4), dtype=torch.int32, device="cuda")
        rn = torch.zeros(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        assert rl.ref_compute_patches(C.c_void_p(d_idx.data_ptr()), C.c_void_p(d_entries.data_ptr()),
                                      Tcw.ctypes.data_as(C.c_void_p), K.ctypes.data_as(C.c_void_p),
                                      C.c_float(block_length), C.c_float(0.1), C.c_float(5.0), len(pts), 640, 480, 80,
                                      60, C.c_void_p(rp.data_ptr()), C.c_void_p(rn.data_ptr())) == 0
        n = int(on.item())
        assert n == int(rn.item()) and n >= 3   # blocks 4 and 5 project outside the image
        assert np.array_equal(_sorted_patches(op[:n].cpu().numpy()), _sorted_patches(rp[:n].cpu().numpy()))
    vb.SetDefaultMathMode(0)


def test_bounds_known_answer_gpu():
    """tests/tracer_test.cu:168-244: hand-written patches -> per-cell min / max, exact."""
    import torch
    import vulcan_b200 as vb
    patches = np.zeros(6, dtype=PATCH)
    patches[0] = ((2, 3), (4, 2), (1.0, 2.0))
    patches[1] = ((4, 3), (3, 3), (0.5, 1.5))
    patches[2] = ((0, 0), (1, 1), (3.0, 4.0))
    patches[3] = ((64, 44), (16, 16), (0.25, 4.5))
    patches[4] = ((70, 50), (10, 10), (0.2, 0.3))
    patches[5] = ((79, 59), (1, 1), (4.0, 4.75))
    exp = np.empty((60, 80, 2), dtype=np.float32)
    exp[..., 0], exp[..., 1] = FMAX, -FMAX
    for (ox, oy), (sx, sy), (lo, hi) in patches:
        exp[oy:oy + sy, ox:ox + sx, 0] = np.minimum(exp[oy:oy + sy, ox:ox + sx, 0], lo)
        exp[oy:oy + sy, ox:ox + sx, 1] = np.maximum(exp[oy:oy + sy, ox:ox + sx, 1], hi)
    dp = torch.from_numpy(patches.view(np.int32).reshape(-1, 4)).cuda()
    b = torch.zeros((60, 80, 2), dtype=torch.float32, device="cuda")
    s = vb.api._stream()
    vb.api.check(vb.lib.vk_reset_bounds(b.data_ptr(), 80 * 60, s), "vk_reset_bounds")
    vb.api.check(vb.lib.vk_compute_bounds(dp.data_ptr(), b.data_ptr(), 80, len(patches), s), "vk_compute_bounds")
    assert np.array_equal(b.cpu().numpy(), exp)
