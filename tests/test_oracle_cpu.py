"""CPU-only tests (-m "not gpu"): the oracle pinned against golden vectors produced by the
reference's own CUDA build on a B200 (tests/golden/, made by tests/golden/make_golden.py +
pack_golden.py), the reference's known-answer scenarios re-expressed on the oracle
(SURVEY.md 8c), and the harness' own helpers."""
import os
import sys

import numpy as np
import pytest

import harness as H

sys.path.insert(0, os.path.join(H.ROOT, "tests", "golden"))
from pack_golden import block_checksums  # noqa: E402

GOLDEN = os.path.join(H.ROOT, "tests", "golden")
CASES = [
    ("micro", (0,), "depth", 1.2),
    ("micro", (0, 8), "depth", 5.0),
    ("micro", (0, 3), "color", 5.0),
    ("micro", (0, 3), "light", 5.0),
]


@pytest.mark.parametrize("math", [0, 1])
@pytest.mark.parametrize("name,frames,mode,step", CASES)
def test_oracle_matches_reference_golden(name, frames, mode, step, math):
    """Bit-exact: block set, visible set, visibility flags, every voxel word (by checksum, and
    word by word on every 8th block), traced depth / colour / normal images.
    math 0 <-> the reference's release build, math 1 <-> the reference built with -fmad=false."""
    tag = "%s_%s_%d_%g_%s.npz" % (name, mode, len(frames), step, "nofmad" if math else "release")
    g = np.load(os.path.join(GOLDEN, tag))
    r = H.run_oracle(H.make_scenario(name, frames, mode, step, math=math))
    assert np.array_equal(r["keys"], g["keys"])
    assert np.array_equal(r["visible_keys"], g["visible_keys"])
    assert np.array_equal(r["vis_flags"], g["vis_flags"])
    assert np.array_equal(block_checksums(r["voxels"]), g["checksums"])
    assert np.array_equal(r["voxels"][::8], g["sample_voxels"])
    for k in ("depth", "color", "normal"):
        assert np.array_equal(r[k].view(np.int32), g[k].view(np.int32)), k
    assert (r["depth"] > 0).mean() > 0.2   # the fixture is not trivially empty


def test_modes_differ_somewhere():
    """The two arithmetic modes are not the same function (else one golden set would be vacuous)."""
    a = H.run_oracle(H.make_scenario("micro", (0, 8), "depth", 5.0, math=0))
    b = H.run_oracle(H.make_scenario("micro", (0, 8), "depth", 5.0, math=1))
    assert not np.array_equal(a["depth"].view(np.int32), b["depth"].view(np.int32))


# ---- the reference's own known-answer scenarios (tests/*.cpp, *.cu in the reference) ---------

def test_layouts():
    """tests/hash_test.cpp:9-12, tests/block_test.cpp:9-12, tests/voxel_test.cpp:9-19."""
    ov = H.OracleVolume(64, 16)
    assert ov.entries().shape == (80, 4) and ov.entries().itemsize * 4 == 16
    assert ov.voxels().shape == (80 * 512, 5)
    d, rgb, dw, cw = H.split_voxels(ov.voxels())
    assert (d == 1.0).all() and (rgb == 0).all() and (dw == 0).all() and (cw == 0).all()


def test_volume_constructor():
    """tests/volume_test.cpp:35-98: entries invalid, visibility FALSE, free list = identity,
    voxel pointer B-1, excess pointer M."""
    ov = H.OracleVolume(512, 256)
    e = ov.entries()
    assert (e[:, 2] == -1).all() and (e[:, 3] == -1).all() and (e[:, :2] == 0).all()
    assert (ov.visibility() == 1).all()
    assert ov.lib.vo_voxel_pointer(ov.h) == 767 and ov.lib.vo_excess_pointer(ov.h) == 512
    assert len(ov.visible()) == 0


def _wall_scene(w, h, f, depth_value):
    K = np.array([f, f, w / 2.0, h / 2.0], dtype=np.float32)
    depth = np.full((h, w), depth_value, dtype=np.float32)
    T = H.raw32(np.eye(3), np.zeros(3))
    return K, depth, T


@pytest.mark.parametrize("math", [0, 1])
def test_integration_known_answer(math):
    """tests/integrator_test.cu:82-221: 160x120, f=80, constant depth 1.5, tau=0.02, voxel 0.008,
    Volume(512, 256): every visible voxel equals a per-voxel host recomputation (|d| <= 1e-5, weight
    exact); integrating the same frame again doubles the weight."""
    w, h = 160, 120
    K, depth, T = _wall_scene(w, h, 80.0, 1.5)
    ov = H.OracleVolume(512, 256, 0.008, 0.02)
    ov.lib.vo_set_math_mode(math)
    ov.set_view_converged(depth, K, T)
    ov.integrate_depth(depth, K, T)
    ent = ov.entries()
    vis = ov.visible()
    assert len(vis) > 0
    vox = ov.voxels().reshape(-1, 512, 5)
    zz, yy, xx = np.meshgrid(np.arange(8), np.arange(8), np.arange(8), indexing="ij")
    checked = 0
    for e in vis[:64]:
        bx = np.int16(ent[e, 0] & 0xFFFF).astype(np.float64)
        by = np.int16((ent[e, 0] >> 16) & 0xFFFF).astype(np.float64)
        bz = np.int16(ent[e, 1] & 0xFFFF).astype(np.float64)
        X = 0.064 * bx + 0.008 * (xx + 0.5)
        Y = 0.064 * by + 0.008 * (yy + 0.5)
        Z = 0.064 * bz + 0.008 * (zz + 0.5)
        u = 80.0 * X / Z + 80.0
        v = 80.0 * Y / Z + 60.0
        # voxels projecting (almost) exactly on the image border are exempt (:160-168,203)
        inside = (u > 1e-3) & (u < w - 1e-3) & (v > 1e-3) & (v < h - 1e-3)
        border = (np.abs(u) < 1e-3) | (np.abs(u - w) < 1e-3) | (np.abs(v) < 1e-3) | (np.abs(v - h) < 1e-3)
        s = 1.5 - Z
        upd = inside & (s > -0.02 + 1e-6)
        notupd = (~border) & ((~((u >= -1e-3) & (u <= w + 1e-3) & (v >= -1e-3) & (v <= h + 1e-3))) | (s < -0.02 - 1e-6))
        d, _, dw, _ = H.split_voxels(vox[ent[e, 2]])
        d, dw = d.reshape(8, 8, 8), dw.reshape(8, 8, 8)
        exp = np.minimum(1.0, s / 0.02)
        assert np.abs(d[upd] - exp[upd]).max(initial=0) <= 1e-5
        assert (dw[upd] == 1).all() and (dw[notupd] == 0).all() and (d[notupd] == 1.0).all()
        checked += int(upd.sum())
    assert checked > 1000
    before = ov.voxels().copy()
    ov.integrate_depth(depth, K, T)
    _, _, dw0, _ = H.split_voxels(before)
    _, _, dw1, _ = H.split_voxels(ov.voxels())
    assert ((dw1 == 2 * dw0)).all()


@pytest.mark.parametrize("math", [0, 1])
def test_raycast_wall_known_answer(math):
    """tests/tracer_test.cu:246-390: constant-depth wall at 1.5, colour (0.1, 0.2, 0.3), tau=0.04,
    voxel 0.008, Volume(4096, 2048), SetView iterated to the fixed point: traced depth within
    0.01 m and colour within 0.001 away from a 5 px border."""
    w, h = 160, 120
    K, depth, T = _wall_scene(w, h, 80.0, 1.5)
    color = np.empty((h, w, 3), dtype=np.float32)
    color[...] = (0.1, 0.2, 0.3)
    ov = H.OracleVolume(4096, 2048, 0.008, 0.04)
    ov.lib.vo_set_math_mode(math)
    ov.set_view_converged(depth, K, T)
    ov.integrate_color(depth, K, T, color, K, H.raw32(np.eye(3), np.zeros(3)))
    d, c, n, _ = ov.trace(w, h, K, T)
    inner = (slice(5, h - 5), slice(5, w - 5))
    assert (d[inner] > 0).all()
    assert np.abs(d[inner] - 1.5).max() <= 0.01
    assert np.abs(c[inner] - np.array([0.1, 0.2, 0.3], dtype=np.float32)).max() <= 1e-3
    # a fronto-parallel wall: normals point back at the camera
    assert np.abs(n[inner][..., 2]).min() > 0.99


def test_bounds_known_answer():
    """tests/tracer_test.cu:168-244: hand-written patches -> per-cell min/max, exact."""
    import ctypes as C
    lib = H.oracle_lib()
    patches = np.zeros(3, dtype=[("origin", "<i2", 2), ("size", "<i2", 2), ("bounds", "<f4", 2)])
    patches[0] = ((2, 3), (4, 2), (1.0, 2.0))
    patches[1] = ((4, 3), (3, 3), (0.5, 1.5))
    patches[2] = ((0, 0), (1, 1), (3.0, 4.0))
    bounds = np.zeros((60, 80, 2), dtype=np.float32)
    lib.vo_compute_bounds(patches.ctypes.data_as(C.c_void_p), 3, bounds.ctypes.data_as(C.c_void_p), 80, 60, 1)
    fmax = np.finfo(np.float32).max
    exp = np.empty((60, 80, 2), dtype=np.float32)
    exp[..., 0], exp[..., 1] = fmax, -fmax
    for (ox, oy), (sx, sy), (lo, hi) in patches:
        exp[oy:oy + sy, ox:ox + sx, 0] = np.minimum(exp[oy:oy + sy, ox:ox + sx, 0], lo)
        exp[oy:oy + sy, ox:ox + sx, 1] = np.maximum(exp[oy:oy + sy, ox:ox + sx, 1], hi)
    assert np.array_equal(bounds, exp)


def test_allocation_fixed_point_is_ray_walk():
    """tests/volume_test.cpp:250-431 scenario: 64x48, f=32, depth 1+3*((x+y)%100)/99, tau=0.2, voxel
    0.02, camera translated: the converged block set is exactly the set of blocks a float64 ray
    walk touches, up to knife-edge blocks (rays within 1e-4 blocks of a block face)."""
    w, h = 64, 48
    K = np.array([32.0, 32.0, 32.0, 24.0], dtype=np.float32)
    x, y = np.meshgrid(np.arange(w), np.arange(h))
    depth = (1 + 3 * ((x + y) % 100) / 99.0).astype(np.float32)
    t = np.array([-10.73, 2.11, -33.54])
    T = H.raw32(np.eye(3), t)
    ov = H.OracleVolume(4096, 2048, 0.02, 0.2)
    passes = ov.set_view_converged(depth, K, T)
    assert 1 <= passes < 64 and ov.dropped() == 0
    keys, _, _ = H.allocated_table(ov.entries(), 4096)
    assert len(np.unique(keys)) == len(keys)
    # dense float64 sampling of every pixel's truncation band
    L = 0.16
    u = (x + 0.5 - 32.0) / 32.0
    v = (y + 0.5 - 24.0) / 32.0
    dirs = np.stack([u, v, np.ones_like(u)], -1)
    P = t + dirs * depth[..., None].astype(np.float64)
    nd = dirs / np.linalg.norm(dirs, axis=-1, keepdims=True)
    ts = np.linspace(-0.2, 0.2, 801)
    pts = P[..., None, :] + nd[..., None, :] * ts[:, None]
    b = np.floor(pts / L).astype(np.int64).reshape(-1, 3)
    sampled = np.unique((b[:, 0] & 0xFFFF) | ((b[:, 1] & 0xFFFF) << 16) | ((b[:, 2] & 0xFFFF) << 32))
    missing = np.setdiff1d(sampled, keys)
    extra = np.setdiff1d(keys, sampled)
    # the DDA visits every block the segment crosses; sampling can miss corner-grazing blocks only
    assert len(missing) <= 0.01 * len(keys), (len(missing), len(keys))
    assert len(extra) <= 0.12 * len(keys), (len(extra), len(keys))


def test_reachable_mask_flags_wrong_bucket_entries():
    ov = H.OracleVolume(256, 64, 0.02, 0.1)
    K, depth, T = _wall_scene(32, 24, 16.0, 1.0)
    ov.set_view_converged(depth, K, T)
    e = ov.entries().copy()
    assert H.reachable_mask(e, 256).sum() == (e[:, 2] != -1).sum()
    # move an allocated head entry into an empty bucket it does not hash to: unreachable
    src = int(np.nonzero(e[:256, 2] != -1)[0][0])
    dst = int(np.nonzero(e[:256, 2] == -1)[0][0])
    e[dst] = e[src]
    e[dst, 3] = -1
    assert H.reachable_mask(e, 256).sum() == (e[:, 2] != -1).sum() - 1
