#!/usr/bin/env python
"""A few frames of BASELINE config 1 / 2 (640x480, 4 mm, Volume(262144, 65536)) through the C ABI:
SetView + DepthIntegrate + Trace per frame.  The payload for compute-sanitizer and for single-kernel ncu
captures (tools/r02_evidence.sh).  --small uses the 160x120 scenario (racecheck is ~100x slower)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import vulcan_b200 as vb  # noqa: E402
from vulcan_b200 import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--small", action="store_true")
ap.add_argument("--collide", action="store_true", help="tiny hash table: every insert goes through the lock / excess path")
args = ap.parse_args()
w, h, voxel, M, E = (160, 120, 0.008, 32768, 8192) if args.small else (640, 480, 0.004, 262144, 65536)
if args.collide:
    M = 64
K = synth.intrinsics(w, h)
vol = vb.Volume(M, E)
vol.SetVoxelLength(voxel)
vol.SetTruncationLength(5 * voxel)
integ, tracer = vb.DepthIntegrator(vol), vb.Tracer(vol)
out = None
for k in range(args.frames):
    R, t = synth.orbit_pose(k, step_deg=5.0)
    f = vb.Frame()
    f.depth_projection = vb.Projection(*K)
    f.depth_to_world_transform = vb.Transform.FromRt(R, t)
    f.depth_image = torch.from_numpy(synth.render_sphere_plane(w, h, K, R, t)).cuda()
    vol.SetView(f)
    integ.Integrate(f)
    out = vb.Frame()
    out.depth_projection, out.depth_to_world_transform = f.depth_projection, f.depth_to_world_transform
    out.depth_image = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    tracer.Trace(out)
torch.cuda.synchronize()
n, counters = vol.SyncCounters()
print("frames %d visible %d allocated %d overflow %d hit %.3f launches %d" % (
    args.frames, n, counters[6], counters[4], float((out.depth_image > 0).float().mean()), vb.lib.vk_launch_count()))
