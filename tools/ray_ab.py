#!/usr/bin/env python
"""A/B of tuning-knob sets on one phase of the bench workload (knobs change launch shapes and
scheduling, never results: the traced images are hashed and compared with the first set).

  python tools/ray_ab.py --sets "10=1;10=0;10=1,11=8" [--phase trace|integrate|setview|step] [--scene c2|room]

Each repetition is timed alone with CUDA events on the launching stream, L2 flushed (256 MiB memset)
before it; median and minimum in microseconds."""
import argparse
import ctypes as C
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sets", default="")
    ap.add_argument("--phase", default="trace")
    ap.add_argument("--scene", default="c2")
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--out", default="ray_ab.json")
    args = ap.parse_args()
    lib = vb.lib
    lib.vk_tune.argtypes = [C.c_int, C.c_int]
    dev = torch.device("cuda", 0)
    if args.scene == "room":
        W, Hh, voxel, M, E, drange = 1920, 1080, 0.002, 1048576, 262144, (0.1, 8.0)
    else:
        W, Hh, voxel, M, E, drange = 640, 480, 0.004, 262144, 65536, (0.1, 5.0)
    K = np.array(H.synth.intrinsics(W, Hh), dtype=np.float32)
    vol = vb.Volume(M, E)
    vol.SetVoxelLength(voxel)
    vol.SetTruncationLength(5 * voxel)
    vol.SetDepthRange(*drange)
    integ = vb.DepthIntegrator(vol)
    integ.SetDepthRange(*drange)
    tracer = vb.Tracer(vol)
    tracer.SetDepthRange(*drange)
    frame = None
    for k in range(args.frames):
        if args.scene == "room":
            R, t = H.room_pose(k % 2, 15.0)
            depth = H.synth.render_room(W, Hh, K, R, t, max_depth=8.0)
        else:
            R, t = H.synth.orbit_pose(k)
            depth = H.synth.render_sphere_plane(W, Hh, K, R, t)
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*K)
        frame.depth_to_world_transform = vb.Transform.FromRt(R, t)
        frame.depth_image = torch.from_numpy(depth).to(dev)
        vol.SetView(frame)
        integ.Integrate(frame)
    out = vb.Frame()
    out.depth_projection = frame.depth_projection
    out.depth_to_world_transform = frame.depth_to_world_transform
    out.depth_image = torch.zeros((Hh, W), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def run_phase():
        if args.phase == "trace":
            tracer.Trace(out)
        elif args.phase == "integrate":
            integ.Integrate(frame)
        elif args.phase == "setview":
            vol.SetView(frame)
        else:
            vol.SetView(frame)
            integ.Integrate(frame)
            tracer.Trace(out)

    def timed():
        ts = []
        for _ in range(3):
            run_phase()
        for _ in range(args.reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run_phase()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts = np.array(ts)
        return float(np.median(ts)), float(ts.min())

    def digest():
        tracer.Trace(out)
        torch.cuda.synchronize()
        return (int(out.depth_image.view(torch.int32).long().sum().item()),
                int(out.color_image.view(torch.int32).long().sum().item()),
                int(out.normal_image.view(torch.int32).long().sum().item()))

    defaults = {}
    results, ref = [], None
    for spec in [s for s in args.sets.split(";")] or [""]:
        pairs = [tuple(int(x) for x in kv.split("=")) for kv in spec.split(",") if kv]
        for k, v in pairs:
            old = lib.vk_tune(k, v)
            defaults.setdefault(k, old)
        med, mn = timed()
        d = digest()
        ref = ref or d
        results.append(dict(phase=args.phase, scene=args.scene, knobs=spec, med_us=med, min_us=mn, same_images=((d == ref) if args.phase == "trace" else None),  # other phases change the model between sets
                           
                            visible=vol.GetVisibleCount()))
        print(json.dumps(results[-1]), flush=True)
        for k, v in defaults.items():
            lib.vk_tune(k, v)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", args.out), "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
