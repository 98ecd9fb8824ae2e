#!/usr/bin/env python
"""Distribution of ray-march trip counts on the bench workload (CPU oracle, diagnostics only)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import harness as H
W, Hh = 640, 480
nfr = int(sys.argv[1]) if len(sys.argv) > 1 else 8
K = np.array(H.synth.intrinsics(W, Hh), dtype=np.float32)
ov = H.OracleVolume(262144, 65536, 0.004, 0.02)
for k in range(nfr):
    R, t = H.synth.orbit_pose(k)
    d = H.synth.render_sphere_plane(W, Hh, K, R, t)
    T = H.raw32(R, t)
    ov.set_view_converged(d, K, T)
    ov.integrate_depth(d, K, T)
it = np.zeros((Hh, W), dtype=np.int32)
ov.lib.vo_set_iter_output.argtypes = [C.c_void_p]
ov.lib.vo_set_iter_output(it.ctypes.data_as(C.c_void_p))
depth, color, normal, bounds = ov.trace(W, Hh, K, T)
ov.lib.vo_set_iter_output(None)
print("rays", it.size, "marching", int((it > 0).sum()), "hit", int((depth > 0).sum()))
print("iters: mean %.1f  p50 %d p90 %d p99 %d max %d  sum %d" % (it[it > 0].mean(), *np.percentile(it[it > 0], [50, 90, 99]), it.max(), it.sum()))
hist = np.bincount(np.minimum(it.ravel(), 500) // 10)
for i, c in enumerate(hist):
    if c: print("%4d-%4d %8d" % (i * 10, i * 10 + 9, c))
# per 8x4 warp tile: max vs mean
t = it.reshape(Hh // 4, 4, W // 8, 8).transpose(0, 2, 1, 3).reshape(-1, 32)
print("warp tiles: sum of max %d (x32 = %d lane-iters) vs sum %d -> lane utilisation %.2f" % (t.max(1).sum(), 32 * t.max(1).sum(), t.sum(), t.sum() / (32.0 * t.max(1).sum())))
print("longest warp tiles (max iters):", np.sort(t.max(1))[-10:])
np.save("/tmp/ray_iters.npy", it)
