"""Surface points on the CPU oracle: the extracted marching-cubes vertices lie on the
analytic surface the depth maps were rendered from.  (The reference's extractor is unfinished - there are
no reference outputs to pin against; see oracle/vulcan_oracle.cpp: vo_extract_points.)"""
import numpy as np

import harness as H


def test_oracle_points_lie_on_the_scene_surface():
    scn = H.make_scenario("tiny", (0, 6), "depth", 5.0)
    ov = H.oracle_volume_for(scn)
    pts, col = H.oracle_extract_points(ov)
    assert len(pts) > 2000
    dist = H.sphere_plane_distance(pts)
    # zero crossings of a projective TSDF fused from nearest-pixel depth samples: accurate to a fraction of
    # the pixel footprint at the scene's depth (2.5 m / f), far inside the truncation band
    footprint = 2.5 / float(scn["K"][0])
    assert np.median(dist) < 0.25 * footprint
    assert np.quantile(dist, 0.95) < 1.0 * footprint
    assert dist.max() < scn["trunc"]
    # one point per owned edge at most: no duplicates
    assert len(np.unique(H.sorted_rows(pts, col), axis=0)) == len(pts)


def test_oracle_extract_capacity_and_empty_volume():
    ov = H.OracleVolume(2048, 512, 0.02, 0.1)
    assert ov.lib.vo_extract_points(ov.h, None, None, 0) == 0
    scn = H.make_scenario("micro", (0,), "color", 1.2)
    ov = H.oracle_volume_for(scn)
    n = ov.lib.vo_extract_points(ov.h, None, None, 0)
    pts = np.zeros((10, 3), dtype=np.float32)
    assert ov.lib.vo_extract_points(ov.h, H._ptr(pts), None, 10) == n and n > 10   # counts on, stores 10
    full, col = H.oracle_extract_points(ov)
    assert np.array_equal(full[:10], pts)
    assert (col >= 0).all() and (col <= 1).all()


def test_oracle_mesh_is_consistent_and_faces_the_camera_side():
    """vo_extract_mesh: the vertices are those of vo_extract_points in the same order; every face refers to
    three distinct existing vertices; no directed edge is used twice (consistent orientation); open edges
    only along the rim of the observed region; on the sphere the normals point away from its centre
    (towards positive distance)."""
    scn = H.make_scenario("tiny", tuple(range(8)), "depth", 45.0)
    scn["M"], scn["E"] = 32768, 8192   # eight views of the scene
    ov = H.oracle_volume_for(scn)
    pts0, col0 = H.oracle_extract_points(ov)
    pts, col, faces = H.oracle_extract_mesh(ov)
    assert np.array_equal(pts, pts0) and np.array_equal(col, col0)
    assert len(faces) > len(pts) and faces.min() >= 0 and faces.max() < len(pts)
    rep = H.mesh_edge_report(faces)
    assert rep["degenerate"] == 0 and rep["repeated_directed"] == 0, rep
    assert rep["boundary"] < 0.08 * rep["edges"], rep
    tri = pts[faces].astype(np.float64)
    normal = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    centre = tri.mean(axis=1)
    radial = centre - np.array([0.0, 0.0, 2.0])
    on_sphere = (np.abs(np.linalg.norm(radial, axis=1) - 0.5) < 0.02) & (centre[:, 1] < 0.3)
    assert on_sphere.sum() > 500
    assert (np.einsum("ij,ij->i", normal, radial)[on_sphere] > 0).mean() > 0.995   # (a few slivers where fused views disagree)
    # every vertex is used by some face except on the rim (cubes with an unobserved corner are not meshed)
    used = np.zeros(len(pts), dtype=bool)
    used[faces.reshape(-1)] = True
    assert used.mean() > 0.9

