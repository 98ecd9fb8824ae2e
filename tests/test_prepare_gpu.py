"""Input-side pre-pass: FilterDepths + ComputeNormals.  The separate kernels and the
fused vk_prepare_frame must agree bit for bit with each other and with the reference's own kernels
(both arithmetic modes), on a noisy depth image with invalid pixels."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _noisy_depth(w, h, seed=12345):
    K = np.array(H.synth.intrinsics(w, h), dtype=np.float32)
    R, t = H.synth.orbit_pose(3)
    d = H.synth.render_sphere_plane(w, h, K, R, t).astype(np.float32)
    rng = np.random.RandomState(seed)
    d = d + (d > 0) * rng.normal(0.0, 0.003, size=d.shape).astype(np.float32)
    d[rng.rand(h, w) < 0.02] = 0.0   # dropouts
    return K, np.ascontiguousarray(d, dtype=np.float32)


@pytest.mark.parametrize("math", [0, 1])
@pytest.mark.parametrize("size", [(160, 120), (640, 480), (333, 201)])
def test_prepare_frame_matches_reference(size, math):
    import torch
    import vulcan_b200 as vb
    if not H.have_ref(nofmad=bool(math)):
        pytest.fail("oracle/_ref/libvulcan_ref*.so missing: run oracle/build_ref.sh")
    w, h = size
    K, depth = _noisy_depth(w, h)
    vb.SetDefaultMathMode(math)
    try:
        # separate calls
        f = vb.Frame()
        f.depth_projection = vb.Projection(*K)
        f.depth_image = torch.from_numpy(depth).cuda()
        f.FilterDepths()
        f.ComputeNormals()
        sep_d, sep_n = f.depth_image.cpu().numpy(), f.normal_image.cpu().numpy()
        # fused kernel
        g = vb.Frame()
        g.depth_projection = vb.Projection(*K)
        g.depth_image = torch.from_numpy(depth).cuda()
        g.Prepare()
        fus_d, fus_n = g.depth_image.cpu().numpy(), g.normal_image.cpu().numpy()
    finally:
        vb.SetDefaultMathMode(0)
    ref_d, ref_n = H.ref_prepare_frame(depth, K, nofmad=bool(math))
    assert np.array_equal(sep_d.view(np.uint32), fus_d.view(np.uint32)), "fused depth differs from separate calls"
    assert np.array_equal(sep_n.view(np.uint32), fus_n.view(np.uint32)), "fused normals differ from separate calls"
    bad_d = int((fus_d.view(np.uint32) != ref_d.view(np.uint32)).sum())
    bad_n = int((fus_n.view(np.uint32) != ref_n.view(np.uint32)).any(axis=-1).sum())
    assert bad_d == 0, "filtered depth differs from the reference in %d pixels (max abs %g)" % (
        bad_d, float(np.abs(fus_d - ref_d).max()))
    assert bad_n == 0, "normals differ from the reference in %d pixels" % bad_n
    assert (fus_d > 0).sum() > 0.5 * w * h
