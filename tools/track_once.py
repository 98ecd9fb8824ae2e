#!/usr/bin/env python
"""One PyramidTracker::Track at 640x480 (for `ncu --metrics gpu__time_duration.sum` launch lists)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
import vulcan_b200 as vb  # noqa: E402

s = H.track_scenario(640, 480)
key, frame = H.ours_track_frames(s, s["key_pose"])
p = vb.PyramidTracker()
p.SetKeyframe(key)
for _ in range(2):
    frame.depth_to_world_transform = key.depth_to_world_transform
    p.Track(frame)
t = vb.DepthTracker()
t.SetKeyframe(key)
for _ in range(3):
    t.ComputeSystem(frame)   # the iteration kernel without the solve (launch list: the difference is the solve)
print("pose error (deg, m):", H.pose_error(frame.depth_to_world_transform.raw32(), s["frame_pose"]))
import torch  # noqa: E402
t.SetMaxIterations(1)
for size_frame in (frame,):
    for _ in range(3):
        size_frame.depth_to_world_transform = key.depth_to_world_transform
        t.Track(size_frame)
        st = t._workspace[66:78].cpu().view(torch.int64).numpy()
        print("timeline ns: first CTA entry -> last CTA entry %d, pixels %d, sums+ticket %d, fold %d, solve %d" % (
            st[0] - st[5], st[1] - st[0], st[2] - st[1], st[3] - st[2], st[4] - st[3]))
