#!/usr/bin/env python
"""Diagnostics for the one-GPU emulation of the replicated-shard mode: two ranks on two streams, one
frame; prints per-rank event timelines, the band tables, flags and the handshake time-out counter."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
import vulcan_b200 as vb  # noqa: E402
from vulcan_b200 import sharding as S  # noqa: E402

count = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mode = sys.argv[2] if len(sys.argv) > 2 else "interleaved"
scn = H.make_scenario("small", (0, 10), "depth")
vb.lib.vk_tune.argtypes = [C.c_int, C.c_int]
print("pdl knob was", vb.lib.vk_tune(9, 0))
single = H.run_ours(scn)
streams = [torch.cuda.Stream() for _ in range(count)]
vols, shards, integs = [], [], []
for rank in range(count):
    vol = vb.Volume(scn["M"], scn["E"])
    vol.SetVoxelLength(scn["voxel"])
    vol.SetTruncationLength(scn["trunc"])
    shards.append(S.Shard(vol, rank, count, scn["w"], scn["h"], 2))
    vols.append(vol)
    integs.append(vb.DepthIntegrator(vol))
S.Shard.connect_local(shards)
torch.cuda.synchronize()
K = scn["K"]
for fi, fr in enumerate(scn["frames"]):
    depth = torch.from_numpy(fr["depth"]).cuda()
    T = fr["T"]
    torch.cuda.synchronize()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(count)]
    base = torch.cuda.Event(enable_timing=True)
    base.record()
    outs = []
    frames = []
    for rank in range(count):
        with torch.cuda.stream(streams[rank]):
            streams[rank].wait_event(base)
            frame = vb.Frame()
            frame.depth_projection = vb.Projection(*K)
            frame.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
            frame.depth_image = depth
            evs[rank][0].record()
            vols[rank].SetView(frame)
            integs[rank].Integrate(frame)
            evs[rank][1].record()
            frames.append(frame)
    if mode == "phased":
        torch.cuda.synchronize()
    for rank in range(count):
        with torch.cuda.stream(streams[rank]):
            out = vb.Frame()
            out.depth_projection, out.depth_to_world_transform = frames[rank].depth_projection, frames[rank].depth_to_world_transform
            evs[rank][2].record()
            shards[rank].Trace(out)
            evs[rank][3].record()
            outs.append(out)
    torch.cuda.synchronize()
    for rank in range(count):
        t = [base.elapsed_time(e) for e in evs[rank]]
        hdr = vb.api.device_view(shards[rank]._s.exchange[rank], (80,), "<i4").cpu().numpy()
        print("frame %d rank %d: events ms %s  bounds_flags %s bands_flags %s bands %s ticket %s timeouts %d" % (
            fi, rank, ["%.3f" % x for x in t], hdr[0:16].tolist(), hdr[16:32].tolist(), hdr[32:64].tolist(), hdr[64:68].tolist(), hdr[68]))
        d = outs[rank].depth_image.cpu().numpy()
        print("   depth identical to unsharded:", bool(np.array_equal(d.view(np.int32), single["depth"].view(np.int32))) if fi == len(scn["frames"]) - 1 else "-",
              "hit frac %.3f" % float((d > 0).mean()))
