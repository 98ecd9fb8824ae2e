#!/usr/bin/env python
"""The reference app's frame loop (apps/vulcan/vulcan.cu:296-325) with TRACKING ON, on synthetic frames:

    frame.ComputeNormals(); frame pose := traced frame's pose; tracker.SetKeyframe(traced frame);
    tracker.Track(frame); SetView; Integrate; trace the model from the tracked pose -> next keyframe

A camera moves through the room scene (walls along all three axes: every pose parameter is observable);
the first frame is placed at its true pose, every later pose comes from the PyramidTracker against the
ray-cast model.  Reports the drift against the ground-truth trajectory and per-frame times.

    gpurun -- 'python tools/tracked_orbit.py --frames 60 > gpurun_out/tracked_orbit.json'
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402


def trajectory(n, step_deg=0.6, step_m=0.01):
    """Slow arc through the room: yaw grows by step_deg, the camera slides by step_m per frame."""
    poses = []
    for k in range(n):
        R = H.synth.rotation_y(25.0 + step_deg * k)
        t = np.array([0.3 + step_m * k, 0.2 - 0.2 * step_m * k, -0.5 + 0.5 * step_m * k])
        poses.append((R, t))
    return poses


def run(frames=30, w=640, h=480, voxel=0.008, trunc=0.04, pyramid=True, log=None, oracle_from=None):
    import torch
    import vulcan_b200 as vb
    K = np.array(H.synth.intrinsics(w, h), dtype=np.float32)
    poses = trajectory(frames)
    depths = [np.ascontiguousarray(H.synth.render_room(w, h, K.astype(np.float64), R, t, max_depth=5.0)) for R, t in poses]
    vol = vb.Volume(262144, 65536)
    vol.SetVoxelLength(voxel)
    vol.SetTruncationLength(trunc)
    integ = vb.DepthIntegrator(vol)
    tracer = vb.Tracer(vol)
    tracker = vb.PyramidTracker() if pyramid else vb.DepthTracker()
    trace_frame = vb.Frame()
    trace_frame.depth_projection = vb.Projection(*K)
    trace_frame.color_projection = vb.Projection(*K)
    trace_frame.depth_image = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    errs, times = [], []
    for k in range(frames):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        frame = vb.Frame()
        frame.depth_projection = vb.Projection(*K)
        frame.color_projection = vb.Projection(*K)
        frame.depth_image = torch.from_numpy(depths[k]).cuda()
        frame.color_image = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
        frame.ComputeNormals()
        if k == 0:
            p = H.raw32(*poses[0])
            frame.depth_to_world_transform = vb.Transform(p[:16].reshape(4, 4).T.copy(), p[16:].reshape(4, 4).T.copy())
        else:
            frame.depth_to_world_transform = trace_frame.depth_to_world_transform
            tracker.SetKeyframe(trace_frame)
            if oracle_from is not None and k >= oracle_from:
                # the same Track on the CPU oracle from the same images (diagnosis of the tracker alone)
                sc = {"K": K, "w": w, "h": h, "key_pose": trace_frame.depth_to_world_transform.raw32(),
                      "key_depth": trace_frame.depth_image.cpu().numpy(), "key_normal": trace_frame.normal_image.cpu().numpy(),
                      "frame_depth": depths[k], "frame_normal": frame.normal_image.cpu().numpy()}
                half = dict(sc)
                for name in ("key", "frame"):
                    half[name + "_depth"] = H.oracle_downsample(sc[name + "_depth"], True)
                    half[name + "_normal"] = H.oracle_downsample(sc[name + "_normal"], True)
                half["K"], half["w"], half["h"] = (K / 2).astype(np.float32), w // 2, h // 2
                p1, _ = H.oracle_track(half, sc["key_pose"], 1, 15)
                p2, _ = H.oracle_track(sc, p1, 1, 20)
                valid_key = float((sc["key_depth"] > 0).mean())
                r0, _ = H.oracle_track_pixels(sc, sc["key_pose"], 1)
                if log:
                    e = H.pose_error(p2, H.raw32(*poses[k]))
                    log("  oracle on the same images: %.4f deg %.2f mm from the truth; keyframe valid %.3f, residuals used %.3f"
                        % (e[0], e[1] * 1e3, valid_key, float((r0 != 0).mean())))
            tracker.Track(frame)
        for _ in range(3):   # apps/vulcan/vulcan.cu:316-318
            vol.SetView(frame)
        integ.Integrate(frame)
        trace_frame.depth_to_world_transform = frame.depth_to_world_transform
        tracer.Trace(trace_frame)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
        errs.append(H.pose_error(frame.depth_to_world_transform.raw32(), H.raw32(*poses[k])))
        if log:
            log("frame %d: %.4f deg, %.2f mm" % (k, errs[-1][0], errs[-1][1] * 1e3))
    n_vis, counters = vol.SyncCounters()
    moved = H.pose_error(H.raw32(*poses[0]), H.raw32(*poses[-1]))
    return {"frames": frames, "size": [w, h], "pyramid": bool(pyramid),
            "trajectory_deg_m": moved, "final_error_deg_m": errs[-1],
            "max_error_deg": max(e[0] for e in errs), "max_error_m": max(e[1] for e in errs),
            "ms_per_frame_median": float(np.median(times[1:]) * 1e3) if frames > 1 else None,
            "pool_overflow": int(counters[4]), "visible_blocks_last": int(n_vis)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=30)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--oracle-from", type=int, default=None)
    a = ap.parse_args()
    print(json.dumps(run(a.frames, a.width, a.height, log=lambda m: print(m, file=sys.stderr), oracle_from=a.oracle_from),
                     indent=1))
