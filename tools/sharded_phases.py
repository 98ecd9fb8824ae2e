#!/usr/bin/env python
"""Where the time of the sharded ray cast goes (run under torchrun, one rank per GPU):
CUDA-event time of each part of ShardedTracer.Trace on the bench workload, every part timed alone
(device synchronised before it), plus the whole call issued asynchronously and the host time of
issuing it.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29513 \
      tools/sharded_phases.py
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
import vulcan_b200 as vb  # noqa: E402
from vulcan_b200 import sharding as S  # noqa: E402

W, Hh = 640, 480


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl")
    K = np.array(H.synth.intrinsics(W, Hh), dtype=np.float32)
    vol = vb.Volume(262144, 65536)
    vol.SetVoxelLength(0.004)
    vol.SetTruncationLength(0.02)
    vol.SetSharding(rank, world, 3)
    integ = vb.DepthIntegrator(vol)
    frame = None
    for k in range(8):
        R, t = H.synth.orbit_pose(k)
        T = H.raw32(R, t)
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*K)
        frame.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
        frame.depth_image = torch.from_numpy(H.synth.render_sphere_plane(W, Hh, K, R, t)).to(dev)
        vol.SetView(frame)
        integ.Integrate(frame)
    tracer = S.ShardedTracer(vol)
    out = vb.Frame()
    out.depth_projection, out.depth_to_world_transform = frame.depth_projection, frame.depth_to_world_transform
    out.depth_image = torch.zeros((Hh, W), dtype=torch.float32, device=dev)
    t = tracer.tracer
    t._ensure_outputs(out)
    y0, y1 = S.row_band(Hh, rank, world)
    images = (out.depth_image, out.color_image, out.normal_image)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def gather():
        with dist._coalescing_manager(device=dev):
            for img in images:
                dist.all_gather_into_tensor(img, img[y0:y1])

    import ctypes as C
    from vulcan_b200 import _lib, api
    x = tracer._ensure_exchange(Hh, W)

    def bounds_exchange():
        x["epoch"] += 1
        api.check(_lib.lib.vk_exchange_bounds(C.byref(x["desc"]), x["epoch"], t.bounds_.data_ptr(), api._stream()), "xb")

    parts = [("bounds", lambda: t.TraceBounds(out)),
             ("allreduce_38KB_nccl", lambda: dist.all_reduce(t.bounds_, op=dist.ReduceOp.MIN)),
             ("bounds_exchange_peer", bounds_exchange),
             ("rows", lambda: t.TraceRows(out, tracer.peers, y0, y1)),
             ("gather_8.6MB_nccl", gather)]
    rep = {"world": world}
    for name, fn in parts:
        ts = []
        for i in range(13):
            flush.zero_()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            if i >= 3:
                ts.append(e0.elapsed_time(e1) * 1e3)
        v = torch.tensor([float(np.median(ts))], device=dev)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        rep[name + "_us"] = round(float(v.item()), 1)
    ts, hs = [], []
    for i in range(13):
        flush.zero_()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        e0.record()
        tracer.Trace(out)
        e1.record()
        h1 = time.perf_counter()
        e1.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1) * 1e3)
            hs.append((h1 - h0) * 1e6)
    v = torch.tensor([float(np.median(ts)), float(np.median(hs))], device=dev)
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    rep["trace_total_us"], rep["trace_host_issue_us"] = round(float(v[0].item()), 1), round(float(v[1].item()), 1)
    dist.barrier()
    tracer.close()
    if rank == 0:
        print(json.dumps(rep))
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "sharded_phases_%d.json" % world), "w") as f:
            json.dump(rep, f)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
