#!/usr/bin/env python
"""Multi-GPU parity + per-phase timing of the block-sharded mode on REAL GPUs (one rank per GPU, CUDA IPC
over NVLink), run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/sharded_check.py --config small|c2|c4|c5 [--reps R]

  small  160x120 orbit, 3 frames                    c4  BASELINE config 4: the 8 x 8 x 3 m room, 1920x1080, 2 mm,
  c2     BASELINE config 2: 640x480 orbit, 4 mm          Volume(3145728, 1048575) split / N, 3 of the 24 poses
  c5     BASELINE config 5: 8 sensors on the orbit circle into one sharded volume, a ray cast per sensor

Every rank runs the scenario twice: through an UNSHARDED volume (the single-GPU result) and through its shard
of the sharded volume (vk_shard_*: replicated hash entries + block distances, local ray cast of a row band,
bands stored into the peers).  After every frame the model is ray-cast; every rank's gathered depth / colour /
normal images must be bit-identical to the unsharded ones.  Then the same frames are timed per phase with CUDA
events (L2 flushed before every frame; max over ranks, median over repetitions) for both volumes.
Writes gpurun_out/sharded_check_<N>gpu_<config>.json on rank 0."""
import argparse
import ctypes as C
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


def scenario(config):
    if config == "small":
        return H.make_scenario("small", (0, 10, 20), "depth", trace_each=True), 2
    if config == "c2":
        return H.make_scenario("c1", (0, 1, 2, 3, 4, 5), "depth", trace_each=True), 3
    if config == "c4":
        return H.make_scenario("room", (0, 1, 2), "depth", 15.0, trace_each=True), 3
    if config == "c5":
        return H.make_scenario("c1", tuple(range(8)), "depth", 45.0, trace_each=True), 3
    raise SystemExit("unknown config " + config)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="small")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist.init_process_group("nccl")
    scn, shift = scenario(args.config)
    K = scn["K"]
    w, h = scn["w"], scn["h"]
    dmin, dmax = scn["depth_range"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    lib = vb.lib

    def make_volume(sharded):
        # the sharded volume's pools are the unsharded ones split over the ranks (BASELINE config 4)
        M, E = (scn["M"] // world, scn["E"] // world) if (sharded and args.config == "c4") else (scn["M"], scn["E"])
        vol = vb.Volume(M, E)
        vol.SetVoxelLength(scn["voxel"])
        vol.SetTruncationLength(scn["trunc"])
        vol.SetDepthRange(dmin, dmax)
        shard = None
        if sharded:
            shard = S.Shard(vol, rank, world, w, h, shift, (dmin, dmax))
            shard.connect()
        integ = vb.DepthIntegrator(vol)
        integ.SetDepthRange(dmin, dmax)
        tracer = vb.Tracer(vol)
        tracer.SetDepthRange(dmin, dmax)
        return vol, shard, integ, tracer

    frames = []
    for fr in scn["frames"]:
        f = vb.Frame()
        f.depth_projection = vb.Projection(*K)
        T = fr["T"]
        f.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
        f.depth_image = torch.from_numpy(fr["depth"]).cuda()
        frames.append(f)

    def run(vol, shard, integ, tracer, timed):
        """All frames once: SetView + Integrate + Trace per frame; returns traced images (untimed pass) or
        per-frame (setview, integrate, trace) milliseconds."""
        out_imgs, times = [], []
        for f in frames:
            out = vb.Frame()
            out.depth_projection, out.depth_to_world_transform = f.depth_projection, f.depth_to_world_transform
            if shard is None:
                out.depth_image = torch.zeros((h, w), dtype=torch.float32, device="cuda")
            if timed:
                flush.zero_()
                dist.barrier()
                torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            vol.SetView(f)
            ev[1].record()
            integ.Integrate(f)
            ev[2].record()
            if shard is None:
                tracer.Trace(out)
            else:
                shard.Trace(out)
            ev[3].record()
            torch.cuda.synchronize()
            if timed:
                times.append([ev[i].elapsed_time(ev[i + 1]) for i in range(3)])
            else:
                out_imgs.append((out.depth_image.clone(), out.color_image.clone(), out.normal_image.clone()))
        return times if timed else out_imgs

    def phase_stats(all_times):
        """(reps, frames, 3) ms on this rank -> max over ranks, then median over reps and frames."""
        t = torch.tensor(all_times, dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t = t.cpu().numpy()
        med = np.median(t.reshape(-1, 3), axis=0)
        return dict(setview_ms=float(med[0]), integrate_ms=float(med[1]), trace_ms=float(med[2]), step_ms=float(np.median(t.sum(-1))))

    # ---- unsharded: the single-GPU result, on every rank ---------------------------------------------------
    full = make_volume(False)
    want = run(*full, timed=False)
    n_full = full[0].SyncCounters()[1][6]
    vis_full = full[0].GetVisibleCount()
    # steady state: the same frames again (every block allocated, weights keep growing like in a long sequence)
    t_full = [run(*full, timed=True) for _ in range(args.reps)]
    del full
    torch.cuda.empty_cache()

    # ---- sharded -------------------------------------------------------------------------------------------
    sh = make_volume(True)
    got = run(*sh, timed=False)
    torch.cuda.synchronize()
    n_shard, overflow = sh[0].SyncCounters()[1][6], sh[0].SyncCounters()[1][4]
    ok = overflow == 0
    for (gd, gc, gn), (wd, wc, wn) in zip(got, want):
        ok = ok and torch.equal(gd.view(torch.int32), wd.view(torch.int32)) and \
            torch.equal(gc.view(torch.int32), wc.view(torch.int32)) and torch.equal(gn.view(torch.int32), wn.view(torch.int32))
    band = sh[1].band()
    t_sh = [run(*sh, timed=True) for _ in range(args.reps)]
    flags = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    blocks = torch.tensor([float(n_shard)], dtype=torch.float64, device="cuda")
    gathered = [torch.zeros_like(blocks) for _ in range(world)]
    dist.all_gather(gathered, blocks)
    rep = dict(config=args.config, world=world, bit_identical_on_all_ranks=bool(flags.item()), views=len(want),
               image=[w, h], voxel_length=scn["voxel"], blocks_unsharded=int(n_full), visible_blocks_last_view=int(vis_full),
               blocks_per_shard=[int(g.item()) for g in gathered], last_band_rank0=list(band),
               hits_last_view=int((want[-1][0] > 0).sum().item()),
               unsharded=phase_stats(t_full), sharded=phase_stats(t_sh),
               timing="CUDA events per phase, 256 MiB L2 flush + barrier before every frame, max over ranks, median over "
                      "%d repetitions x %d frames (steady state: the frames are integrated again)" % (args.reps, len(frames)))
    rep["speedup_step"] = rep["unsharded"]["step_ms"] / rep["sharded"]["step_ms"]
    dist.barrier()
    sh[1].close()
    if rank == 0:
        print(json.dumps(rep))
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "sharded_check_%dgpu_%s.json" % (world, args.config)), "w") as f:
            json.dump(rep, f, indent=1)
    dist.destroy_process_group()
    sys.exit(0 if flags.item() else 1)


if __name__ == "__main__":
    main()
