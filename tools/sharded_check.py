#!/usr/bin/env python
"""Multi-GPU parity check of the block-sharded mode (run under torchrun, one rank per GPU):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/sharded_check.py

Every rank integrates its shard of a small orbit; the sharded ray cast (peer memory over NVLink,
NCCL bounds all-reduce + band all-gather) must be bit-identical to an unsharded volume's ray cast.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
import vulcan_b200 as vb  # noqa: E402
from vulcan_b200 import sharding as S  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist.init_process_group("nccl")
    scn = H.make_scenario("small", (0, 10, 20), "depth")
    K = scn["K"]

    def build(shard):
        vol = vb.Volume(scn["M"], scn["E"])
        vol.SetVoxelLength(scn["voxel"])
        vol.SetTruncationLength(scn["trunc"])
        if shard:
            vol.SetSharding(rank, world, 2)
        integ = vb.DepthIntegrator(vol)
        frame = None
        for fr in scn["frames"]:
            frame = vb.Frame()
            frame.depth_projection = vb.Projection(*K)
            T = fr["T"]
            frame.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
            frame.depth_image = torch.from_numpy(fr["depth"]).cuda()
            vol.SetView(frame)
            integ.Integrate(frame)
        return vol, frame

    def out_frame(frame):
        out = vb.Frame()
        out.depth_projection = frame.depth_projection
        out.depth_to_world_transform = frame.depth_to_world_transform
        out.depth_image = torch.zeros((scn["h"], scn["w"]), dtype=torch.float32, device="cuda")
        return out

    full, frame = build(False)
    want = out_frame(frame)
    vb.Tracer(full).Trace(want)
    shard, _ = build(True)
    tracer = S.ShardedTracer(shard)
    got = out_frame(frame)
    tracer.Trace(got)
    torch.cuda.synchronize()
    n_full = full.SyncCounters()[1][6]
    n_shard = shard.SyncCounters()[1][6]
    ok = all(torch.equal(a.view(torch.int32), b.view(torch.int32)) for a, b in (
        (got.depth_image, want.depth_image), (got.color_image, want.color_image), (got.normal_image, want.normal_image)))
    flags = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    rep = dict(world=world, rank=rank, bit_identical_on_all_ranks=bool(flags.item()), blocks_unsharded=int(n_full),
               blocks_this_shard=int(n_shard), hits=int((want.depth_image > 0).sum().item()))
    dist.barrier()
    tracer.close()
    if rank == 0:
        print(json.dumps(rep))
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "sharded_check_%d.json" % world), "w") as f:
            json.dump(rep, f)
    dist.destroy_process_group()
    sys.exit(0 if flags.item() else 1)


if __name__ == "__main__":
    main()
