#!/usr/bin/env python
"""Generates the marching-cubes triangle table (vulcan_b200/csrc/vk_mc_tables.h) from first principles
instead of typing 4 096 numbers: for each of the 256 corner-sign cases the iso-surface crosses every cube
face in line segments that depend ONLY on that face's four corner signs (ambiguous faces - two diagonal
inside corners - always cut the inside corners off individually), so neighbouring cubes agree on their
shared face and the mesh is watertight by construction.  The segments of a case chain into closed loops over
the crossed cube edges; each loop is oriented so that its normal points from the inside corners to the
outside ones and is triangulated as a fan.

Conventions (shared by vk_extract.cu and oracle/vulcan_oracle.cpp):
  corner c of a cube      offset (c & 1, (c >> 1) & 1, c >> 2) from the cube's minimum voxel
  case index              bit c set <=> corner c is INSIDE, i.e. !(distance > 0)
  cube edge e = 4 * axis + k   the k-th edge parallel to `axis`, from corner kEdgeCorner[e] to that corner
                               with the axis bit set; it is the lattice edge the voxel at kEdgeCorner[e]
                               owns along `axis` (the vertices of vk_extract_points)
  kTriTable[case]         up to 5 triangles as edge triples, -1 terminated

  python tools/gen_mc_tables.py            # writes the header and runs the self-test
"""
import itertools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CORNER = [np.array([c & 1, (c >> 1) & 1, c >> 2]) for c in range(8)]
EDGE_CORNER = []          # edge id -> (corner a, axis)
for axis in range(3):
    for a in range(8):
        if not (a >> axis) & 1:
            EDGE_CORNER.append((a, axis))
EDGE_OF = {}              # frozenset of two corners -> edge id
for e, (a, axis) in enumerate(EDGE_CORNER):
    EDGE_OF[frozenset((a, a | (1 << axis)))] = e

# the six faces: corners in cyclic order
FACES = []
for axis in range(3):
    u, v = [ax for ax in range(3) if ax != axis]
    for side in (0, 1):
        base = side << axis
        FACES.append([base, base | (1 << u), base | (1 << u) | (1 << v), base | (1 << v)])


def face_segments(inside, face):
    """Segments (pairs of cube edge ids) the surface draws on one face."""
    s = [inside[c] for c in face]
    edges = [EDGE_OF[frozenset((face[i], face[(i + 1) % 4]))] for i in range(4)]   # edge i joins corner i and i+1
    n = sum(s)
    if n == 0 or n == 4:
        return []
    crossed = [i for i in range(4) if s[i] != s[(i + 1) % 4]]
    if len(crossed) == 2:
        return [(edges[crossed[0]], edges[crossed[1]])]
    # ambiguous: two diagonal inside corners, each cut off on its own
    out = []
    for i in range(4):
        if s[i]:
            out.append((edges[(i - 1) % 4], edges[i]))
    return out


def edge_point(e):
    a, axis = EDGE_CORNER[e]
    p = CORNER[a].astype(float)
    p[axis] += 0.5
    return p


def triangles_of_case(case):
    inside = [(case >> c) & 1 for c in range(8)]
    segs = []
    for face in FACES:
        segs += face_segments(inside, face)
    # chain the undirected segments into loops: every crossed edge ends exactly two segments
    ends = {}
    for i, (a, b) in enumerate(segs):
        ends.setdefault(a, []).append(i)
        ends.setdefault(b, []).append(i)
    assert all(len(v) == 2 for v in ends.values()), case
    used = [False] * len(segs)
    tris = []
    for start in range(len(segs)):
        if used[start]:
            continue
        loop = [segs[start][0]]
        cur_edge, cur_seg = segs[start][1], start
        used[start] = True
        while cur_edge != loop[0]:
            loop.append(cur_edge)
            nxt = [i for i in ends[cur_edge] if i != cur_seg][0]
            assert not used[nxt], case
            used[nxt] = True
            a, b = segs[nxt]
            cur_edge, cur_seg = (b if a == cur_edge else a), nxt
        # orientation: the loop's normal (Newell) must point from the inside ends of its edges to the outside ends
        pts = [edge_point(e) for e in loop]
        normal = np.zeros(3)
        for i in range(len(pts)):
            p, q = pts[i], pts[(i + 1) % len(pts)]
            normal += np.cross(p, q)
        outward = np.zeros(3)
        for e in loop:
            a, axis = EDGE_CORNER[e]
            d = np.zeros(3)
            d[axis] = 1.0
            outward += d if inside[a] else -d     # from the inside corner of the edge to its outside corner
        if normal @ outward < 0:
            loop = loop[::-1]
        assert abs(normal @ outward) > 1e-9, case
        for i in range(1, len(loop) - 1):
            tris.append((loop[0], loop[i], loop[i + 1]))
    return tris


def build():
    table = [triangles_of_case(c) for c in range(256)]
    assert max(len(t) for t in table) <= 5
    return table


def extract(dist, table):
    """Reference use of the table on a dense grid (numpy, slow): vertices at edge midpoints by linear
    interpolation; returns (points, faces)."""
    nx, ny, nz = dist.shape
    index, pts, faces = {}, [], []
    for x, y, z in itertools.product(range(nx - 1), range(ny - 1), range(nz - 1)):
        case = 0
        for c in range(8):
            o = CORNER[c]
            if not dist[x + o[0], y + o[1], z + o[2]] > 0:
                case |= 1 << c
        for tri in table[case]:
            ids = []
            for e in tri:
                a, axis = EDGE_CORNER[e]
                v0 = np.array([x, y, z]) + CORNER[a]
                key = (tuple(v0), axis)
                if key not in index:
                    v1 = v0.copy()
                    v1[axis] += 1
                    d0, d1 = dist[tuple(v0)], dist[tuple(v1)]
                    t = d0 / (d0 - d1)
                    index[key] = len(pts)
                    pts.append(v0 + t * (v1 - v0))
                ids.append(index[key])
            faces.append(ids)
    return np.array(pts), np.array(faces)


def self_test(table):
    # a sphere and a two-blob field with ambiguous configurations: closed, consistently oriented, outward
    g = np.mgrid[0:14, 0:14, 0:14].astype(float)
    sphere = np.sqrt(((g - 6.3) ** 2).sum(0)) - 4.2
    rng = np.random.RandomState(5)
    noisy = sphere + 0.9 * rng.randn(*sphere.shape) * (np.abs(sphere) < 2.0)   # exercises all 256 cases
    noisy[[0, -1], :, :] = noisy[:, [0, -1], :] = noisy[:, :, [0, -1]] = 5.0    # closed inside the grid
    for name, field in (("sphere", sphere), ("noisy", noisy)):
        pts, faces = extract(field, table)
        directed = {}
        for f in faces:
            for i in range(3):
                k = (f[i], f[(i + 1) % 3])
                directed[k] = directed.get(k, 0) + 1
        assert all(v == 1 for v in directed.values()), name + ": an edge is used twice in the same direction"
        assert all((b, a) in directed for (a, b) in directed), name + ": open edge - not watertight"
        p = pts[faces]
        vol = np.einsum("ij,ij->i", p[:, 0], np.cross(p[:, 1], p[:, 2])).sum() / 6.0
        assert vol > 0, name + ": normals point inwards"
        if name == "sphere":
            assert abs(vol - 4.0 / 3.0 * np.pi * 4.2 ** 3) / vol < 0.06, vol   # (coarse grid: chords cut the sphere short)
        print("self-test %-6s: %d vertices, %d faces, watertight, volume %.1f" % (name, len(pts), len(faces), vol))
    cases = set()
    for x, y, z in itertools.product(range(13), repeat=3):
        c = 0
        for k in range(8):
            o = CORNER[k]
            if not noisy[x + o[0], y + o[1], z + o[2]] > 0:
                c |= 1 << k
        cases.add(c)
    print("cases exercised by the noisy field: %d of 256" % len(cases))


def write_header(table, path):
    with open(path, "w") as f:
        f.write("// vk_mc_tables.h - GENERATED by tools/gen_mc_tables.py (do not edit): marching-cubes triangle table derived\n"
                "// from the face-segment rule described there (watertight by construction).\n"
                "//   corner c = offset (c & 1, (c >> 1) & 1, c >> 2); case bit c set <=> corner c inside (!(distance > 0));\n"
                "//   cube edge e = 4 * axis + k runs from corner kMcEdgeCorner[e] along `axis` - the lattice edge the voxel at\n"
                "//   that corner owns; kMcTriTable[case] = up to 5 triangles as edge triples, -1 terminated.\n"
                "#pragma once\n\n")
        f.write("static const signed char kMcEdgeCorner[12] = {%s};\n" % ", ".join(str(a) for a, _ in EDGE_CORNER))
        f.write("static const signed char kMcEdgeAxis[12] = {%s};\n\n" % ", ".join(str(ax) for _, ax in EDGE_CORNER))
        f.write("static const signed char kMcTriTable[256][16] = {\n")
        for c, tris in enumerate(table):
            flat = [e for t in tris for e in t]
            flat += [-1] * (16 - len(flat))
            f.write("  {%s},\n" % ", ".join("%2d" % v for v in flat))
        f.write("};\n")


if __name__ == "__main__":
    tbl = build()
    self_test(tbl)
    out = os.path.join(ROOT, "vulcan_b200", "csrc", "vk_mc_tables.h")
    write_header(tbl, out)
    print("wrote", out, "max triangles per case", max(len(t) for t in tbl))
