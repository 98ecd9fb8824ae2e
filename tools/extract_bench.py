#!/usr/bin/env python
"""Surface extraction (vk_extract_points) on BASELINE config 1 after an 8-frame orbit: time per call (CUDA
events, L2 flushed between calls), points found, and the bytes the kernel has to move (16 B per hash entry
scanned + 10 240 B per allocated block + 24 B per point) against the measured HBM peak.

    gpurun -- 'python tools/extract_bench.py > gpurun_out/extract_bench.json'
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402


def main():
    import torch
    import vulcan_b200 as vb
    from vulcan_b200._lib import check
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_extract_gpu as T
    scn = H.make_scenario("c1", tuple(range(0, 80, 10)), "depth", 1.2)
    vol = T._ours_volume(scn)
    cap = 8 << 20
    pts = torch.empty((cap, 3), dtype=torch.float32, device="cuda")
    col = torch.empty((cap, 3), dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    times = []
    for i in range(13):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(vb.lib.vk_extract_points(C.byref(vol._v), pts.data_ptr(), col.data_ptr(), cap, cnt.data_ptr(), None),
              "extract")
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            times.append(e0.elapsed_time(e1))
    n = int(cnt.item())
    # the whole mesh: vertices + marching-cubes faces (wall clock around the call: it sizes its workspace from the
    # allocated-block counter, i.e. it synchronises)
    import time
    tri = torch.empty((2 * cap, 3), dtype=torch.int32, device="cuda")
    cnt2 = torch.zeros(2, dtype=torch.int32, device="cuda")
    mesh_times = []
    for i in range(8):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        check(vb.lib.vk_extract_mesh(C.byref(vol._v), pts.data_ptr(), col.data_ptr(), cap, cnt2.data_ptr(), tri.data_ptr(),
                                     2 * cap, cnt2[1:].data_ptr(), None), "extract mesh")
        torch.cuda.synchronize()
        if i >= 2:
            mesh_times.append((time.perf_counter() - t0) * 1e3)
    n_mesh, n_faces = (int(x) for x in cnt2.tolist())
    entries = vol.GetHashEntries().cpu().numpy()
    blocks = int((entries[:, 2] != -1).sum())
    ms = float(np.median(times))
    bytes_alg = 16 * len(entries) + 10240 * blocks + 24 * n
    peak = 6535.4
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak)
    except Exception:
        pass
    print(json.dumps({"workload": "c1 640x480 4 mm, 8 orbit frames 12 deg apart", "allocated_blocks": blocks,
                      "points": n, "ms": ms, "points_per_s": n / (ms * 1e-3), "algorithmic_bytes": bytes_alg,
                      "achieved_gbs": bytes_alg / (ms * 1e-3) / 1e9, "hbm_peak_gbs": peak,
                      "frac": bytes_alg / (ms * 1e-3) / 1e9 / peak,
                      "mesh": {"vertices": n_mesh, "faces": n_faces, "ms_wall_incl_workspace": float(np.median(mesh_times))}},
                     indent=1))


if __name__ == "__main__":
    main()
