#!/usr/bin/env python
"""Per-phase timing of the hot path on the bench workload under different tuning knobs.

  python tools/tune_bench.py [--frames 8] [--reps 20]

Each phase is timed alone with CUDA events on the launching stream, L2 flushed (256 MiB memset)
before every repetition; knob settings must not change results (the traced images are hashed).
"""
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

W, Hh = 640, 480
KNOB = {"ray_threads": 0, "ray_lanes": 1, "ray_wait": 2, "int_ctas": 3, "alloc_mode": 4, "coop": 5, "free_run": 7, "coop_rounds": 8, "sample_steps": 9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--what", default="ray,int,set")
    ap.add_argument("--ctas", default="0,6,7,8,9,10", help="integrate CTAs/SM values to sweep (0 = occupancy default)")
    args = ap.parse_args()
    lib = vb.lib
    lib.vk_tune.argtypes = [C.c_int, C.c_int]
    dev = torch.device("cuda", 0)
    K = np.array(H.synth.intrinsics(W, Hh), dtype=np.float32)
    vol = vb.Volume(262144, 65536)
    vol.SetVoxelLength(0.004)
    vol.SetTruncationLength(0.02)
    integ = vb.DepthIntegrator(vol)
    tracer = vb.Tracer(vol)
    frame = None
    for k in range(args.frames):
        R, t = H.synth.orbit_pose(k)
        T = H.raw32(R, t)
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*K)
        frame.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
        frame.depth_image = torch.from_numpy(H.synth.render_sphere_plane(W, Hh, K, R, t)).to(dev)
        vol.SetView(frame)
        integ.Integrate(frame)
    out = vb.Frame()
    out.depth_projection = frame.depth_projection
    out.depth_to_world_transform = frame.depth_to_world_transform
    out.depth_image = torch.zeros((Hh, W), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn):
        ts = []
        for _ in range(3):
            fn()
        for _ in range(args.reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts = np.array(ts)
        return float(np.median(ts)), float(ts.min())

    def digest():
        torch.cuda.synchronize()
        return (int(out.depth_image.view(torch.int32).long().sum().item()),
                int(out.color_image.view(torch.int32).long().sum().item()),
                int(out.normal_image.view(torch.int32).long().sum().item()))

    results = []
    what = args.what.split(",")
    if "timeline" in what:
        lib.vk_tune(KNOB["ray_threads"], 32)
        nwarp = (W // 8) * (Hh // 4)
        tl = torch.zeros((nwarp, 8), dtype=torch.int64, device=dev)
        lib.vk_debug_trace_timeline.argtypes = [C.c_void_p]
        flush.zero_()
        lib.vk_debug_trace_timeline(tl.data_ptr())
        tracer.Trace(out)
        torch.cuda.synchronize()
        lib.vk_debug_trace_timeline(None)
        t = tl.cpu().numpy().astype(np.int64)
        alive_sum = t[:, 3] >> 40
        t[:, 3] &= (1 << 40) - 1
        c_la = t[:, 4] >> 40
        t[:, 4] &= (1 << 40) - 1
        single_alu = (t[:, 5] >> 40) << 4     # single-step passes: cycles up to the filter load (max over lanes), and whole pass
        single_all = (t[:, 6] >> 40) << 4
        t[:, 5] &= (1 << 40) - 1
        t[:, 6] &= (1 << 40) - 1
        t0 = t[:, 0].min()
        dur = (t[:, 1] - t[:, 0]) / 1e3
        end = (t[:, 1] - t0) / 1e3
        start = (t[:, 0] - t0) / 1e3
        print("warps %d  kernel span %.1f us  warp duration us: mean %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f" % (
            nwarp, end.max(), dur.mean(), *np.percentile(dur, [50, 90, 99]), dur.max()))
        for q in (10, 20, 30, 40, 60, 80, 100, 120, 140):
            print("  t=%3d us: warps running %5d, not yet started %5d" % (q, int(((start <= q) & (end > q)).sum()), int((start > q).sum())))
        order = np.argsort(-dur)[:12]
        for w_ in order:
            print("  long warp tile (x=%d,y=%d): start %.1f dur %.1f  sm %d trips %d cycles %d (%.0f cyc/trip)" % (
                (w_ % (W // 8)) * 8, (w_ // (W // 8)) * 4, start[w_], dur[w_], t[w_, 2] >> 32, t[w_, 2] & 0xffffffff,
                t[w_, 3], t[w_, 3] / max(1, t[w_, 2] & 0xffffffff)))
            print("      cycles: free-run %d  of which cooperative runs %d  phase A %d  phase B %d  max-lane free steps %d cooperative ray-runs %d  mean live lanes %.1f" % (
                t[w_, 4], c_la[w_], t[w_, 5], t[w_, 6], t[w_, 7] & 0xffff, (t[w_, 7] >> 16) & 0xffff, alive_sum[w_] / max(1, t[w_, 2] & 0xffffffff)))
            print("      paired passes %d  single-step passes %d (cycles: %d up to the filter load, %d in all)" % ((t[w_, 7] >> 48) & 0xffff, (t[w_, 7] >> 32) & 0xffff, single_alu[w_], single_all[w_]))
        trips = (t[:, 2] & 0xffffffff).astype(np.float64)
        print("  all warps: trips total %d, cycles/trip mean %.0f" % (trips.sum(), t[:, 3].sum() / trips.sum()))
        np.save(os.path.join(ROOT, "gpurun_out", "trace_timeline.npy"), t)
    if "ray" in what:
        ref = None
        # (threads per CTA, cooperative-run lane threshold, phase-B batch lanes, batch wait, free-run steps per trip,
        #  cooperative rounds per trip)
        for threads, ahead, lanes, wait, frun in ((64, 0, 12, 3, 8), (64, 2, 12, 3, 8), (64, 4, 12, 3, 8), (64, 8, 12, 3, 8),
                                                  (64, 4, 12, 3, 4), (64, 4, 12, 3, 16), (64, 4, 6, 2, 8), (64, 4, 1, 0, 8),
                                                  (32, 4, 12, 3, 8)):
            rounds = 4
            if True:
                lib.vk_tune(KNOB["free_run"], frun)
                lib.vk_tune(KNOB["coop_rounds"], rounds)
                lib.vk_tune(KNOB["coop"], ahead)
                lib.vk_tune(KNOB["ray_threads"], threads)
                lib.vk_tune(KNOB["ray_lanes"], lanes)
                lib.vk_tune(KNOB["ray_wait"], wait)
                med, mn = timed(lambda: tracer.Trace(out))
                d = digest()
                ref = ref or d
                results.append(dict(phase="trace", threads=threads, free_run=frun, coop=ahead, rounds=rounds, lanes=lanes, wait=wait, med_us=med, min_us=mn,
                                    same=(d == ref)))
                print(json.dumps(results[-1]), flush=True)
    if "int" in what:
        for ctas in [int(x) for x in args.ctas.split(',')]:
            lib.vk_tune(KNOB["int_ctas"], ctas)
            med, mn = timed(lambda: integ.Integrate(frame))
            results.append(dict(phase="integrate", ctas_per_sm=ctas, med_us=med, min_us=mn))
            print(json.dumps(results[-1]), flush=True)
        lib.vk_tune(KNOB["int_ctas"], 0)
    if "set" in what:
        for mode in (256, 128, 64, 1):
            lib.vk_tune(KNOB["alloc_mode"], mode)
            med, mn = timed(lambda: vol.SetView(frame))
            n = vol.GetVisibleCount()
            results.append(dict(phase="setview", alloc_mode=mode, med_us=med, min_us=mn, visible=n))
            print(json.dumps(results[-1]), flush=True)
        lib.vk_tune(KNOB["alloc_mode"], 0)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "tune.json"), "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
