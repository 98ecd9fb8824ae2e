#!/usr/bin/env python
"""Generates tests/golden/tracker_*.npz from the REFERENCE's own tracker sources (src/tracker.cpp,
src/depth_tracker.{cpp,cu}, src/frame.cpp, src/image.cu, built unmodified into oracle/_ref by
oracle/build_ref.sh), run on a B200:

    gpurun -- 'python tests/golden/make_golden_tracker.py gpurun_out/golden'   # then copy into tests/golden/

Per build (release = default -fmad=true, nofmad = -fmad=false): per-pixel residuals and Jacobians, the
normal equations (both parameter counts), the pose after Tracker::Track and after the pyramid schedule,
and Frame::Downsample of the frame (half and quarter size) on harness.track_scenario(80, 60).
Note: the 6x6 solve of the reference build is oracle/ref_stub/Eigen/Eigen's LDL^T (Eigen is absent);
everything up to the normal equations is the reference's own arithmetic.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

SIZE = (80, 60)


def main(out):
    os.makedirs(out, exist_ok=True)
    s = H.track_scenario(*SIZE)
    for nofmad in (True, False):
        t = H.RefTracker(s, s["key_pose"], nofmad=nofmad)
        d = {"residuals": t.residuals(), "jacobian6": t.jacobian(1), "jacobian3": t.jacobian(0)}
        d["hessian6"], d["gradient6"] = t.system(1)
        d["hessian3"], d["gradient3"] = t.system(0)
        for lvl in (1, 2):
            dd, cc, nn, kk = t.downsample(1, lvl)
            d["down%d_depth" % lvl], d["down%d_color" % lvl], d["down%d_normal" % lvl], d["down%d_K" % lvl] = dd, cc, nn, kk
        d["pose_track20"] = t.track(1, 20)
        t.close()
        t = H.RefTracker(s, s["key_pose"], nofmad=nofmad)
        d["pose_rot20"] = t.track(0, 20)
        t.close()
        t = H.RefTracker(s, s["key_pose"], nofmad=nofmad)
        d["pose_pyramid"] = t.track_pyramid()
        t.close()
        tag = "tracker_%dx%d_%s" % (SIZE[0], SIZE[1], "nofmad" if nofmad else "release")
        np.savez_compressed(os.path.join(out, tag + ".npz"), **d)
        print(tag, "valid", float((d["residuals"] != 0).mean()), "err track", H.pose_error(d["pose_track20"], s["frame_pose"]),
              "pyramid", H.pose_error(d["pose_pyramid"], s["frame_pose"]))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
