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
    mesh = vb.Extractor(vol).Extract(capacity=1 << 12)   # too small on purpose: Extract grows the buffers
    ov = H.oracle_volume_for(scn)
    pts, col, faces = H.oracle_extract_mesh(ov)
    assert len(mesh.points) == len(pts) and len(pts) > 1000
    assert np.array_equal(H.sorted_rows(mesh.points, mesh.colors), H.sorted_rows(pts, col))
    # the triangles: the same oriented triangles as the oracle's, compared by the bits of their corner positions
    assert len(mesh.faces) == len(faces) and len(faces) > len(pts)
    assert mesh.faces.min() >= 0 and mesh.faces.max() < len(pts)
    assert np.array_equal(H.canonical_triangles(mesh.points, mesh.faces), H.canonical_triangles(pts, faces))
    rep = H.mesh_edge_report(mesh.faces)
    assert rep["degenerate"] == 0 and rep["repeated_directed"] == 0, rep
    points_only = vb.Extractor(vol).Extract(faces=False)
    assert np.array_equal(H.sorted_rows(points_only.points, points_only.colors), H.sorted_rows(pts, col))
    dist = H.sphere_plane_distance(mesh.points)
    footprint = 2.5 / float(scn["K"][0])   # pixel footprint at the scene's depth: the accuracy of the fused TSDF
    assert np.median(dist) < 0.25 * footprint and np.quantile(dist, 0.95) < 1.0 * footprint
    assert dist.max() < scn["trunc"]
    # PLY export (src/exporter.cpp:19-71 header)
    path = str(tmp_path / "out.ply")
    vb.Exporter(path).Export(mesh)
    lines = open(path).read().splitlines()
    assert lines[0] == "ply" and lines[2] == "element vertex %d" % len(pts)
    assert lines[9] == "element face %d" % len(faces) and lines[11] == "end_header"
    assert len(lines) == 12 + len(pts) + len(faces)
    x, y, z, r, g, b = lines[12].split()
    assert abs(float(x) - mesh.points[0][0]) < 1e-6 and 0 <= int(r) <= 255
    assert lines[12 + len(pts)].split() == ["3"] + [str(i) for i in mesh.faces[0]]


def test_mesh_of_closed_surfaces_is_watertight():
    """Small spheres written into the voxels of a fused volume, centred ON the scene's surface (so that they
    lie inside the allocated band, every voxel around them observed) and placed so that they straddle block
    faces, edges and corners: the extracted mesh is CLOSED - every directed edge exactly once, its opposite
    present - oriented outwards, and encloses the spheres' volume."""
    import torch
    import vulcan_b200 as vb
    from vulcan_b200 import sharding
    scn = H.make_scenario("small", tuple(range(8)), "depth", 45.0)
    vol = _ours_volume(scn)
    entries = vol.GetHashEntries().cpu().numpy()
    alloc = np.nonzero(entries[:, 2] != -1)[0]
    bx, by, bz = (a.astype(np.float64) for a in sharding.entry_coords(entries[alloc]))
    data = entries[alloc, 2].astype(np.int64)
    l, L, tau = scn["voxel"], 8 * scn["voxel"], scn["trunc"]
    t = np.arange(512)
    ox, oy, oz = (t & 7) + 0.5, ((t >> 3) & 7) + 0.5, (t >> 6) + 0.5
    X = bx[:, None] * L + ox[None, :] * l
    Y = by[:, None] * L + oy[None, :] * l
    Z = bz[:, None] * L + oz[None, :] * l
    # centres on the scene sphere (centre (0, 0, 2), radius 0.5), snapped to block corners / faces nearby
    r = 0.022
    centres = []
    for az in range(0, 360, 45):
        for el in (-20.0, 10.0):
            d = np.array([np.sin(np.radians(az)) * np.cos(np.radians(el)), np.sin(np.radians(el)),
                          -np.cos(np.radians(az)) * np.cos(np.radians(el))])
            c = np.array([0.0, 0.0, 2.0]) + 0.5 * d
            snapped = np.round(c / L) * L          # a block corner ...
            if np.linalg.norm(snapped - c) < 0.012:  # ... if one is close enough to stay inside the band
                c = snapped
            centres.append(c)
    sdf = np.full(X.shape, 1.0)
    for c in centres:
        sdf = np.minimum(sdf, (np.sqrt((X - c[0]) ** 2 + (Y - c[1]) ** 2 + (Z - c[2]) ** 2) - r) / tau)
    sdf = np.clip(sdf, -1.0, 1.0).astype(np.float32)
    vox = vol.GetVoxelsMutable().view(-1, 512, 5)
    idx = torch.as_tensor(data, device="cuda")
    words = vox[idx]
    words[:, :, 0] = torch.from_numpy(sdf.view(np.int32)).cuda()
    words[:, :, 4] = 1   # distance_weight 1, color_weight 0
    vox[idx] = words
    vol.SyncDistances()
    mesh = vb.Extractor(vol).Extract()
    assert len(mesh.faces) > 100 * len(centres)
    rep = H.mesh_edge_report(mesh.faces)
    assert rep == dict(edges=3 * len(mesh.faces), repeated_directed=0, boundary=0, degenerate=0), rep
    p = mesh.points[mesh.faces].astype(np.float64)
    volume = np.einsum("ij,ij->i", p[:, 0], np.cross(p[:, 1], p[:, 2])).sum() / 6.0
    want = len(centres) * 4.0 / 3.0 * np.pi * r ** 3
    assert 0.8 * want < volume < 1.05 * want, (volume, want)


def test_extract_empty_volume():
    import vulcan_b200 as vb
    mesh = vb.Extractor(vb.Volume(4096, 1024)).Extract()
    assert mesh.points.shape == (0, 3) and mesh.faces.shape == (0, 3)
