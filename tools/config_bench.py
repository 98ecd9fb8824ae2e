#!/usr/bin/env python
"""Per-phase times of ours and of the reference's CUDA build on the other BASELINE configurations
(the bench line is config 2; these are the secondary numbers quoted in DESIGN.md):

  c1  640x480 depth        c3  1280x720 depth + colour (ColorIntegrator)       light  640x480 LightIntegrator
  c4  1920x1080 depth, 2 mm voxels, 4 194 303-block volume (orbit scene)
  room  BASELINE config 4's scene: the 8 x 8 x 3 m room, 1920x1080, 2 mm, depth range 0.1-8 m, 4 of the 24 poses

Each phase is timed with CUDA events, L2 flushed (256 MiB memset) before it, on frames 4..7 of an orbit
after 4 warm-up frames (median).  The reference's SetView is repeated until the visible count is stable.

  python tools/config_bench.py [c1 c3 light c4]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
import vulcan_b200 as vb  # noqa: E402

CONFIGS = {"c1": ("c1", "depth"), "c3": ("c3", "color"), "light": ("c1", "light"), "c4": ("c4", "depth"),
           "room": ("room", "depth")}


def timed(fn, flush):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e3


def run(name):
    size, mode = CONFIGS[name]
    frames = tuple(range(8))
    scn = H.make_scenario(size, frames, mode, 15.0) if size == "room" else H.make_scenario(size, frames, mode)
    dmin, dmax = scn.get("depth_range", (0.1, 5.0))
    K = scn["K"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rep = {"config": name, "image": [scn["w"], scn["h"]], "mode": mode}

    # ---- ours
    vol = vb.Volume(scn["M"], scn["E"])
    vol.SetVoxelLength(scn["voxel"])
    vol.SetTruncationLength(scn["trunc"])
    vol.SetDepthRange(dmin, dmax)
    integ = {"depth": vb.DepthIntegrator, "color": vb.ColorIntegrator, "light": vb.LightIntegrator}[mode](vol)
    if mode == "light":
        integ.SetLight(*scn["light"])
    integ.SetDepthRange(dmin, dmax)
    tracer = vb.Tracer(vol)
    tracer.SetDepthRange(dmin, dmax)
    t = {"setview": [], "integrate": [], "trace": []}
    for i, fr in enumerate(scn["frames"]):
        f = vb.Frame()
        f.depth_projection = vb.Projection(*K)
        f.color_projection = vb.Projection(*K)
        T = fr["T"]
        f.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
        Tc = scn["Tcd"]
        f.depth_to_color_transform = vb.Transform(Tc[:16].reshape(4, 4).T.copy(), Tc[16:].reshape(4, 4).T.copy())
        f.depth_image = torch.from_numpy(fr["depth"]).cuda()
        if fr["color"] is not None:
            f.color_image = torch.from_numpy(fr["color"]).cuda()
        if mode == "light":
            f.ComputeNormals()
        out = vb.Frame()
        out.depth_projection, out.depth_to_world_transform = f.depth_projection, f.depth_to_world_transform
        out.depth_image = torch.zeros((scn["h"], scn["w"]), dtype=torch.float32, device="cuda")
        a = timed(lambda: vol.SetView(f), flush)
        b = timed(lambda: integ.Integrate(f), flush)
        c = timed(lambda: tracer.Trace(out), flush)
        if i >= 4:
            t["setview"].append(a); t["integrate"].append(b); t["trace"].append(c)
    rep["ours_us"] = {k: round(float(np.median(v)), 1) for k, v in t.items()}
    rep["visible_blocks"] = int(vol.GetVisibleCount())
    del vol, integ, tracer
    torch.cuda.empty_cache()

    # ---- the reference's own CUDA build
    if H.have_ref():
        rv = H.RefVolume(scn["M"], scn["E"], scn["voxel"], scn["trunc"])
        try:
            kind = {"depth": 0, "color": 1, "light": 2}[mode]
            rv.set_light(*scn["light"])
            rv.set_ranges(dmin, dmax, 16.0, 16.0, dmin, dmax)
            t = {"setview": [], "integrate": [], "trace": []}
            passes = []
            for i, fr in enumerate(scn["frames"]):
                rv.set_frame(fr["depth"], K, fr["T"], color=fr["color"], cproj4=K, Tcd32=scn["Tcd"])
                if kind == 2:
                    rv.compute_input_normals()
                n = []
                a = timed(lambda: n.append(rv.set_view_converged()), flush)
                b = timed(lambda: rv.integrate(kind), flush)
                c = timed(lambda: rv.trace(scn["w"], scn["h"], K, fr["T"]), flush)
                if i >= 4:
                    t["setview"].append(a); t["integrate"].append(b); t["trace"].append(c); passes.append(n[0])
            rep["reference_us"] = {k: round(float(np.median(v)), 1) for k, v in t.items()}
            rep["reference_setview_passes"] = float(np.mean(passes))
        finally:
            rv.close()
    return rep


def main():
    names = sys.argv[1:] or ["c1", "c3", "light", "c4", "room"]
    out = [run(n) for n in names]
    for r in out:
        print(json.dumps(r))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "config_bench.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
