"""Surface points on the GPU: vk_extract_points against the CPU oracle as sorted sets,
bit for bit (positions and colours), points on the analytic surface, PLY export."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _ours_volume(scn):
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


@pytest.mark.parametrize("name,frames,mode,step", [("tiny", (0, 6), "depth", 5.0), ("small", (0, 3), "color", 5.0),
                                                   ("c1", (0,), "depth", 1.2)])
def test_points_match_oracle(name, frames, mode, step, tmp_path):
    import vulcan_b200 as vb
    scn = H.make_scenario(name, frames, mode, step)
    vol = _ours_volume(scn)
    mesh = vb.Extractor(vol).Extract(capacity=1 << 12)   # too small on purpose: Extract grows the buffer
    ov = H.oracle_volume_for(scn)
    pts, col = H.oracle_extract_points(ov)
    assert len(mesh.points) == len(pts) and len(pts) > 1000
    assert np.array_equal(H.sorted_rows(mesh.points, mesh.colors), H.sorted_rows(pts, col))
    dist = H.sphere_plane_distance(mesh.points)
    footprint = 2.5 / float(scn["K"][0])   # pixel footprint at the scene's depth: the accuracy of the fused TSDF
    assert np.median(dist) < 0.25 * footprint and np.quantile(dist, 0.95) < 1.0 * footprint
    assert dist.max() < scn["trunc"]
    # PLY export (src/exporter.cpp:19-71 header)
    path = str(tmp_path / "out.ply")
    vb.Exporter(path).Export(mesh)
    lines = open(path).read().splitlines()
    assert lines[0] == "ply" and lines[2] == "element vertex %d" % len(pts)
    assert lines[9] == "element face 0" and lines[11] == "end_header"
    assert len(lines) == 12 + len(pts)
    x, y, z, r, g, b = lines[12].split()
    assert abs(float(x) - mesh.points[0][0]) < 1e-6 and 0 <= int(r) <= 255


def test_extract_empty_volume():
    import vulcan_b200 as vb
    mesh = vb.Extractor(vb.Volume(4096, 1024)).Extract()
    assert mesh.points.shape == (0, 3) and mesh.faces.shape == (0, 3)
