#!/usr/bin/env python
"""Pyramid + depth tracker, ours against the reference build on the same GPU.

    gpurun -- 'python tools/track_bench.py > gpurun_out/track_bench.json'

Per size: Frame::Downsample (two levels), one DepthTracker::Track of 20 iterations started from the
keyframe pose, and the PyramidTracker schedule.  Host wall clock around synchronous calls (both APIs
return the tracked pose to the host), median of `reps` runs after warm-up.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402


def med(fn, reps=15, warm=3):
    import torch
    ts = []
    for i in range(warm + reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        if i >= warm:
            ts.append(time.perf_counter() - t0)
    return float(np.median(ts) * 1e3)


def main():
    import torch
    import vulcan_b200 as vb
    out = {}
    for size in ((640, 480), (1280, 720)):
        s = H.track_scenario(*size)
        res = {}
        key, frame = H.ours_track_frames(s, s["key_pose"])
        start = frame.depth_to_world_transform
        t = vb.DepthTracker()
        t.SetKeyframe(key)
        t.SetMaxIterations(20)

        def ours_track():
            frame.depth_to_world_transform = start
            t.Track(frame)
        res["ours_track20_ms"] = med(ours_track)
        res["ours_track20_iterations"] = t.iterations_
        res["ours_track20_err"] = H.pose_error(frame.depth_to_world_transform.raw32(), s["frame_pose"])
        # the same call at the C ABI with the descriptors built once (what a C++ caller pays; the Python
        # mirror above rebuilds both vkFrame structs on every call)
        import ctypes as C
        kc, fc = key.to_c(), frame.to_c()
        start_c = start.to_c()
        it, norm = C.c_int(0), C.c_float(0)

        def cabi_track():
            fc.depth_to_world = start_c
            vb.lib.vk_track_depth(C.byref(kc), C.byref(fc), 1, 20, t._workspace.data_ptr(), C.byref(it), C.byref(norm), None)
        res["ours_track20_cabi_ms"] = med(cabi_track)
        p = vb.PyramidTracker()
        p.SetKeyframe(key)

        def ours_pyr():
            frame.depth_to_world_transform = start
            p.Track(frame)
        res["ours_pyramid_ms"] = med(ours_pyr)
        res["ours_pyramid_err"] = H.pose_error(frame.depth_to_world_transform.raw32(), s["frame_pose"])
        half, quarter = vb.Frame(), vb.Frame()

        def ours_down():
            frame.Downsample(half)
            half.Downsample(quarter)
        res["ours_downsample2_ms"] = med(ours_down)
        # one system evaluation, device time
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        t.ComputeSystem(frame)
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(50):
            t.ComputeSystem(frame)
        ev[1].record()
        torch.cuda.synchronize()
        res["ours_system_us"] = ev[0].elapsed_time(ev[1]) * 1e3 / 50

        ref = H.RefTracker(s, s["key_pose"])
        pose = np.ascontiguousarray(s["key_pose"], dtype=np.float32)

        def reset():
            w, h = s["w"], s["h"]
            ref.lib.ref_track_set_frame(ref.h, 1, H._ptr(s["frame_depth"]), H._ptr(s["frame_normal"]),
                                        H._ptr(s["frame_color"]), w, h, H._ptr(s["K"]), H._ptr(pose))
        tmp = np.zeros(32, dtype=np.float32)
        ts = []
        for i in range(12):
            reset()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ref.lib.ref_track(ref.h, 1, 20, H._ptr(tmp))
            if i >= 3:
                ts.append(time.perf_counter() - t0)
        res["ref_track20_ms"] = float(np.median(ts) * 1e3)
        res["ref_track20_err"] = H.pose_error(tmp, s["frame_pose"])
        ts = []
        for i in range(12):
            reset()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ref.lib.ref_track_pyramid(ref.h, H._ptr(tmp))
            if i >= 3:
                ts.append(time.perf_counter() - t0)
        res["ref_pyramid_ms"] = float(np.median(ts) * 1e3)
        res["ref_pyramid_err"] = H.pose_error(tmp, s["frame_pose"])
        ref.close()
        out["%dx%d" % size] = res
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
