"""Image pyramid + depth tracker on the GPU: Frame::Downsample, DepthTracker's per-pixel
residuals / Jacobians, its normal equations, Tracker::Track and the PyramidTracker schedule against the
reference's own sources (oracle/_ref, both builds) and the CPU oracle on the same arrays.

Bars: Downsample bit-exact; residuals / Jacobians bit-exact in the strict mode against the -fmad=false
build and the oracle, within 2e-6 (and <= 0.1 % of pixels gated differently) in the reference-FMA mode;
normal equations within 1e-4 of the largest entry (the reference adds CTA sums in arrival order);
tracked poses within 0.01 degrees / 0.2 mm (the 6x6 solves differ in rounding)."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu

ROT_TOL_DEG, TRANS_TOL_M = 0.01, 2e-4


def _need_ref(math):
    if not H.have_ref(nofmad=bool(math)):
        pytest.fail("oracle/_ref/libvulcan_ref*.so missing: run oracle/build_ref.sh")


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(params=[0, 1], ids=["fma", "strict"])
def math(request):
    import vulcan_b200 as vb
    vb.SetDefaultMathMode(request.param)
    H.oracle_lib().vo_set_math_mode(request.param)
    yield request.param
    vb.SetDefaultMathMode(0)
    H.oracle_lib().vo_set_math_mode(0)


@pytest.mark.parametrize("size", [(160, 120), (640, 480), (1920, 1080), (36, 20)])
def test_downsample_matches_reference(size, math):
    import vulcan_b200 as vb
    _need_ref(math)
    s = H.track_scenario(*size)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    half, quarter = vb.Frame(), vb.Frame()
    frame.Downsample(half)
    half.Downsample(quarter)
    ref = H.RefTracker(s, s["key_pose"], nofmad=bool(math))
    for lvl, f in ((1, half), (2, quarter)):
        d, c, n, k = ref.downsample(1, lvl)
        assert np.array_equal(_bits(f.depth_image.cpu().numpy()), _bits(d)), "depth level %d" % lvl
        assert np.array_equal(_bits(f.color_image.cpu().numpy()), _bits(c)), "colour level %d" % lvl
        assert np.array_equal(_bits(f.normal_image.cpu().numpy()), _bits(n)), "normal level %d" % lvl
        assert np.array_equal(f.depth_projection.raw4(), k)
    ref.close()
    # oracle restatement on the same arrays
    assert np.array_equal(_bits(half.depth_image.cpu().numpy()), _bits(H.oracle_downsample(s["frame_depth"], True)))
    assert np.array_equal(_bits(half.color_image.cpu().numpy()), _bits(H.oracle_downsample(s["frame_color"], False)))
    assert np.array_equal(_bits(half.normal_image.cpu().numpy()), _bits(H.oracle_downsample(s["frame_normal"], True)))


def test_downsample_rejects_odd_sizes():
    import torch
    import vulcan_b200 as vb
    f = vb.Frame()
    f.depth_image = torch.zeros((5, 8), device="cuda")
    f.color_image = torch.zeros((5, 8, 3), device="cuda")
    f.normal_image = torch.zeros((5, 8, 3), device="cuda")
    with pytest.raises(vb.VulcanError):
        f.Downsample(vb.Frame())


@pytest.mark.parametrize("size", [(160, 120), (640, 480), (333, 201), (1920, 1080)])  # ragged: partial tiles
def test_residuals_and_jacobian_match_reference(size, math):
    import vulcan_b200 as vb
    _need_ref(math)
    s = H.track_scenario(*size)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    t = vb.DepthTracker()
    t.SetKeyframe(key)
    ref = H.RefTracker(s, s["key_pose"], nofmad=bool(math))
    o_r, o_J = H.oracle_track_pixels(s, s["key_pose"], 1)
    r = t.ComputeResiduals(frame).cpu().numpy()
    ref_r = ref.residuals()
    assert (r != 0).mean() > 0.5
    for trans in (1, 0):
        t.SetTranslationEnabled(trans)
        J = t.ComputeJacobian(frame).cpu().numpy()
        ref_J = ref.jacobian(trans)
        if trans == 0:
            assert not J[..., 3:].any() and not ref_J[..., 3:].any()
        if math == 1:
            assert np.array_equal(_bits(J), _bits(ref_J)), "Jacobian differs from the -fmad=false reference build"
        else:
            gate = ((J != 0).any(-1) != (ref_J != 0).any(-1)).mean()
            assert gate <= 1e-3
            both = (J != 0).any(-1) & (ref_J != 0).any(-1)
            assert np.abs(J - ref_J)[both].max() <= 2e-6 * max(1.0, np.abs(ref_J).max())
    if math == 1:
        assert np.array_equal(_bits(r), _bits(ref_r)), "residuals differ from the -fmad=false reference build"
    else:
        assert ((r != 0) != (ref_r != 0)).mean() <= 1e-3
        both = (r != 0) & (ref_r != 0)
        assert np.abs(r - ref_r)[both].max() <= 2e-6
    # the oracle is the same function as the CUDA path in both modes
    assert np.array_equal(_bits(r), _bits(o_r))
    t.SetTranslationEnabled(True)
    assert np.array_equal(_bits(t.ComputeJacobian(frame).cpu().numpy()), _bits(o_J))
    ref.close()


@pytest.mark.parametrize("trans", [1, 0])
@pytest.mark.parametrize("size", [(160, 120), (640, 480), (333, 201), (1920, 1080)])
def test_system_matches_reference(size, trans, math):
    import vulcan_b200 as vb
    _need_ref(math)
    s = H.track_scenario(*size)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    t = vb.DepthTracker()
    t.SetKeyframe(key)
    t.SetTranslationEnabled(trans)
    Hs, g = (x.cpu().numpy() for x in t.ComputeSystem(frame))
    H2, g2 = (x.cpu().numpy() for x in t.ComputeSystem(frame))
    assert np.array_equal(_bits(Hs), _bits(H2)) and np.array_equal(_bits(g), _bits(g2)), "sums must be deterministic"
    ref = H.RefTracker(s, s["key_pose"], nofmad=bool(math))
    rH, rg = ref.system(trans)
    ref.close()
    oH, og = H.oracle_track_system(s, s["key_pose"], trans)
    for mine, theirs in ((Hs, rH), (Hs, oH)):
        assert np.abs(mine - theirs).max() <= 1e-4 * np.abs(theirs).max()
    for mine, theirs in ((g, rg), (g, og)):
        assert np.abs(mine - theirs).max() <= 1e-4 * max(np.abs(theirs).max(), 1.0)


@pytest.mark.parametrize("trans", [1, 0])
def test_track_matches_reference_and_oracle(trans, math):
    import vulcan_b200 as vb
    _need_ref(math)
    s = H.track_scenario(160, 120)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    t = vb.DepthTracker()
    t.SetKeyframe(key)
    t.SetTranslationEnabled(trans)
    t.SetMaxIterations(20)
    t.Track(frame)
    mine = frame.depth_to_world_transform.raw32()
    ref = H.RefTracker(s, s["key_pose"], nofmad=bool(math))
    theirs = ref.track(trans, 20)
    ref.close()
    orc, orc_it = H.oracle_track(s, s["key_pose"], trans, 20)
    for other, name in ((theirs, "reference"), (orc, "oracle")):
        rot, tr = H.pose_error(mine, other)
        assert rot <= ROT_TOL_DEG and tr <= TRANS_TOL_M, "%s: %g deg, %g m" % (name, rot, tr)
    assert 1 <= t.iterations_ <= 20
    if trans:
        rot, tr = H.pose_error(mine, s["frame_pose"])
        assert rot < 0.05 and tr < 1e-3, "did not converge to the true pose: %g deg, %g m" % (rot, tr)
    # matrix * inverse == identity
    m = frame.depth_to_world_transform
    assert np.abs(m.matrix.astype(np.float64) @ m.inv_matrix.astype(np.float64) - np.eye(4)).max() < 1e-3


@pytest.mark.parametrize("size", [(160, 120), (640, 480)])
def test_pyramid_tracker(size, math):
    import vulcan_b200 as vb
    _need_ref(math)
    s = H.track_scenario(*size)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    p = vb.PyramidTracker()
    p.SetKeyframe(key)
    p.Track(frame)
    mine = frame.depth_to_world_transform.raw32()
    ref = H.RefTracker(s, s["key_pose"], nofmad=bool(math))
    theirs = ref.track_pyramid()
    ref.close()
    rot, tr = H.pose_error(mine, theirs)
    assert rot <= ROT_TOL_DEG and tr <= TRANS_TOL_M, "reference: %g deg, %g m" % (rot, tr)
    rot, tr = H.pose_error(mine, s["frame_pose"])
    assert rot < 0.05 and tr < 1e-3, "did not converge to the true pose: %g deg, %g m" % (rot, tr)


def test_track_without_valid_pixels_keeps_pose():
    """No overlap (all depths 0): the system is empty, the update 0, the loop stops after one solve."""
    import vulcan_b200 as vb
    s = H.track_scenario(80, 60)
    s["frame_depth"] = np.zeros_like(s["frame_depth"])
    key, frame = H.ours_track_frames(s, s["key_pose"])
    t = vb.DepthTracker()
    t.SetKeyframe(key)
    t.Track(frame)
    assert t.iterations_ == 1
    rot, tr = H.pose_error(frame.depth_to_world_transform.raw32(), s["key_pose"])
    assert rot < 1e-3 and tr < 1e-6


def test_tracker_validation_errors():
    import vulcan_b200 as vb
    t = vb.DepthTracker()
    with pytest.raises(vb.VulcanError):
        t.Track(vb.Frame())          # keyframe has not been assigned
    s = H.track_scenario(80, 60)
    key, frame = H.ours_track_frames(s, s["key_pose"])
    t.SetKeyframe(key)
    frame.normal_image = None
    with pytest.raises(vb.VulcanError):
        t.Track(frame)               # frame missing normal image
    with pytest.raises(vb.VulcanError):
        t.SetMaxIterations(0)


def test_frame_to_model_tracking_closes_the_loop():
    """The reference app's loop with tracking on (apps/vulcan/vulcan.cu:296-325): every pose after the
    first comes from the PyramidTracker against the ray-cast model.  The drift over the trajectory stays
    a small fraction of the motion."""
    import os
    import sys
    sys.path.insert(0, os.path.join(H.ROOT, "tools"))
    import tracked_orbit
    r = tracked_orbit.run(frames=12, w=320, h=240)
    assert r["pool_overflow"] == 0
    assert r["trajectory_deg_m"][0] > 5.0 and r["trajectory_deg_m"][1] > 0.1
    assert r["max_error_deg"] < 0.3 and r["max_error_m"] < 0.01, r
