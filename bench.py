#!/usr/bin/env python
"""bench.py — the reference's headline workload on B200.

Workload (BASELINE.json configs[1], "c2_orbit"): 640x480 synthetic sphere-on-plane depth,
4 mm voxels, Volume(262144, 65536), ground-truth poses on a 1.2 deg/frame orbit.  One step =
one frame of the hot path: Volume::SetView (to its fixed point) + DepthIntegrator::Integrate
+ Tracer::Trace.  `value` = voxel updates (visible blocks x 512) per second over the whole
step with inputs resident in HBM; `e2e` = the same with the depth image coming from pinned
host memory and the traced depth/colour/normal images read back, every step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

--impl reference runs the reference's own CUDA sources (oracle/_ref, built unmodified from
/root/reference by oracle/build_ref.sh) through its own classes on the same frames; the
reference has no CPU implementation of this path, so the CPU number reported beside both
arms (`cpu_baseline`) is the oracle port timed on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H = 640, 480
VOXEL = 0.004
TRUNC = 5 * VOXEL
MAIN, EXCESS = 262144, 65536
METRIC = "voxel-updates/s (visible blocks x 512 per frame / SetView+Integrate+Trace time), 640x480, 4 mm"
# one description of the workload for BOTH arms (the driver compares the dicts)
CONFIG = {"workload": "c2_orbit: 640x480 sphere-on-plane depth, 4 mm voxels, Volume(262144,65536), "
                      "1.2 deg/frame orbit, GT poses; step = SetView (to its fixed point)+DepthIntegrate+Trace",
          "l2": "256 MiB device memset between steps, outside the per-step CUDA events"}


def traffic_from_profiles(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the newest committed
    `ncu --set full` summary under profiles/ (rNN_ncu_full_summary.csv); (bytes, file) or (None, None)."""
    import csv
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_summary.csv")), reverse=True):
        try:
            with open(path) as f:
                for row in csv.DictReader(f):
                    if kernel in row.get("Kernel Name", ""):
                        rd = [v for k, v in row.items() if k.startswith("dram__bytes_read.sum")]
                        wr = [v for k, v in row.items() if k.startswith("dram__bytes_write.sum")]
                        unit = [k for k in row if k.startswith("dram__bytes_read.sum")][0]
                        scale = {"[Mbyte]": 1e6, "[Gbyte]": 1e9, "[Kbyte]": 1e3, "[byte]": 1.0}[unit.split()[-1]]
                        return (float(rd[0]) + float(wr[0])) * scale, os.path.relpath(path, ROOT)
        except Exception:
            continue
    return None, None


SHARDING = ("blocks owned by brick hash (8^3-block bricks): a rank stores, allocates and integrates only its own blocks; "
            "hash entries and block distances are replicated into every rank over peer memory (NVLink) as they are "
            "written; every rank ray-casts a band of rows (bands balanced from the merged ray bounds) from local "
            "memory and stores it into its peers' images; two flag handshakes per view, no library collective")


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs.  NVML is
    initialised in the constructor (it takes tens of ms, longer than a short timed region); the
    thread then polls every 2 ms from start() to finish()."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self._nv = self._h = None
        self._names = {}
        try:
            import pynvml as nv
            nv.nvmlInit()
            self._nv, self._h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM)
            self._names = {
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
        except Exception as e:  # NVML missing: report that instead of inventing numbers
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def _sample(self):
        nv, h = self._nv, self._h
        self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        for bit, name in self._names.items():
            if r & bit:
                self.reasons.add(name)

    def run(self):
        if self._nv is None:
            return
        try:
            while not self._stop_evt.is_set():
                self._sample()
                time.sleep(0.002)
        except Exception as e:
            self.reasons.add("nvml_error:%s" % type(e).__name__)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def make_frames(n, start=0):
    import harness as Hh
    frames = []
    K = Hh.synth.intrinsics(W, H)
    for k in range(start, start + n):
        R, t = Hh.synth.orbit_pose(k)
        frames.append((Hh.synth.render_sphere_plane(W, H, K, R, t), Hh.raw32(R, t)))
    return np.array(K, dtype=np.float32), frames


def algorithmic_bytes_integrate(updated_voxels, visible_blocks):
    """40 B per updated voxel + 20 B per visible block + 4*W*H."""
    return 40.0 * updated_voxels + 20.0 * visible_blocks + 4.0 * W * H


def run_ours(args, rank, world):
    import torch
    import torch.distributed as dist
    import vulcan_b200 as vb
    from vulcan_b200 import sharding

    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    K, frames = make_frames(args.warmup + args.steps)
    nfr = len(frames)

    shards = {}

    def new_volume():
        vol = vb.Volume(MAIN, EXCESS)
        vol.SetVoxelLength(VOXEL)
        vol.SetTruncationLength(TRUNC)
        if world > 1:
            # block-sharded volume with replicated lookup structures (include/vulcan_b200.h, vk_shard_*):
            # created before anything is allocated; IPC handles travel over torch.distributed
            sh = sharding.Shard(vol, rank, world, W, H, 3)
            sh.connect()
            shards[id(vol)] = sh
        return vol

    def drop_volume(vol):
        sh = shards.pop(id(vol), None)
        if sh is not None:
            barrier()
            sh.close()
            barrier()

    def transform(raw):
        return vb.Transform(raw[:16].reshape(4, 4).T.copy(), raw[16:].reshape(4, 4).T.copy())

    poses = [transform(T) for _, T in frames]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    import ctypes as C
    lib = vb.lib
    bounds = torch.empty((60, 80, 2), dtype=torch.float32, device=dev)

    def cframe(k, depth_t):
        f = vb.Frame()
        f.depth_projection = vb.Projection(*K)
        f.depth_to_world_transform = poses[k]
        f.depth_image = depth_t
        return f.to_c()

    sharded_out = [C.c_void_p(), C.c_void_p(), C.c_void_p()]

    def step(vol, cin, cout, outs, ev=None, gather=1):
        """One frame through the C ABI (include/vulcan_b200.h) with prebuilt descriptors."""
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        v = C.byref(vol._v)
        if ev: ev[0].record()
        vb.api.check(lib.vk_set_view(v, C.byref(cin), s), "vk_set_view")
        if ev: ev[1].record()
        vb.api.check(lib.vk_integrate_depth(v, C.byref(cin), 0.1, 5.0, 16.0, s), "vk_integrate_depth")
        if ev: ev[2].record()
        if world > 1:
            # every rank ray-casts its band of rows from local memory; two flag handshakes over peer memory
            vb.api.check(lib.vk_trace_sharded(v, C.byref(cout), 0.1, 5.0, gather, C.byref(sharded_out[0]),
                                              C.byref(sharded_out[1]), C.byref(sharded_out[2]), s), "vk_trace_sharded")
        else:
            vb.api.check(lib.vk_trace(v, C.byref(cout), 0.1, 5.0, bounds.data_ptr(), outs[0].data_ptr(),
                                      outs[1].data_ptr(), outs[2].data_ptr(), s), "vk_trace")
        if ev: ev[3].record()

    def frame_for(k, depth_t):
        f = vb.Frame()
        f.depth_projection = vb.Projection(*K)
        f.depth_to_world_transform = poses[k]
        f.depth_image = depth_t
        return f

    # ---------------- device-resident leg -------------------------------------------------
    # Two passes over the same frames, each on a fresh volume.  Pass A is the headline: CUDA events around the
    # STEP only (the three calls back to back, as a caller issues them).  Pass B records events between the phases
    # as well; those events end the programmatic overlap between the kernels of neighbouring phases and cost a
    # drain + timestamp each, so pass B's step is slower - it only supplies the per-phase split and the roofline's
    # kernel time.
    dev_depth = [torch.from_numpy(d).to(dev) for d, _ in frames]
    outs = (torch.zeros((H, W), dtype=torch.float32, device=dev),
            torch.zeros((H, W, 3), dtype=torch.float32, device=dev),
            torch.zeros((H, W, 3), dtype=torch.float32, device=dev))

    def device_pass(phase_events):
        vol = new_volume()
        cin = [cframe(k, dev_depth[k]) for k in range(nfr)]
        cout = [cframe(k, outs[0]) for k in range(nfr)]
        sampler = ClockSampler(dev.index) if not phase_events else None
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(nfr)]
        count_log = torch.zeros(nfr, dtype=torch.int32, device=dev)
        counters = vb.api.device_view(vol._v.counters, (16,), "<i4")
        launches0 = updated0 = 0
        t_wall0 = time.perf_counter()
        barrier()
        for k in range(nfr):
            if k == args.warmup:
                barrier()
                if sampler: sampler.start()
                launches0 = lib.vk_launch_count()
                updated0 = int(counters[8:10].cpu().view(torch.int64).item())
                t_wall0 = time.perf_counter()
            flush.zero_()  # L2 flush between steps, outside the per-step events
            if phase_events:
                step(vol, cin[k], cout[k], outs, evs[k])
            else:
                evs[k][0].record()
                step(vol, cin[k], cout[k], outs, None)
                evs[k][3].record()
            # visible blocks of THIS rank's shard (a rank lists only the blocks it owns)
            count_log[k:k + 1].copy_(counters[vol._v.visible_slot:vol._v.visible_slot + 1])
        barrier()
        res = {"wall": time.perf_counter() - t_wall0, "launches": lib.vk_launch_count() - launches0,
               "clocks": sampler.finish() if sampler else None,
               "counts": count_log[args.warmup:].cpu().tolist(),
               "updated": float(int(counters[8:10].cpu().view(torch.int64).item()) - updated0)}
        assert vol.SyncCounters()[1][4] == 0, "voxel/excess pool overflow"
        timed = range(args.warmup, nfr)
        res["t_step"] = np.array([evs[k][0].elapsed_time(evs[k][3]) for k in timed])
        if phase_events:
            res["t_set"] = np.array([evs[k][0].elapsed_time(evs[k][1]) for k in timed])
            res["t_int"] = np.array([evs[k][1].elapsed_time(evs[k][2]) for k in timed])
            res["t_trc"] = np.array([evs[k][2].elapsed_time(evs[k][3]) for k in timed])
        drop_volume(vol)
        del vol
        torch.cuda.empty_cache()
        return res

    pa = device_pass(False)
    pb = device_pass(True)
    t_wall, launches, clocks, counts, updated_timed = pa["wall"], pa["launches"], pa["clocks"], pa["counts"], pb["updated"]
    tot = torch.tensor([pa["t_step"].sum(), pb["t_int"].sum(), pb["t_trc"].sum(), pb["t_set"].sum(), pb["t_step"].sum()],
                       dtype=torch.float64, device=dev)
    units = torch.tensor([float(sum(counts)) * 512.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(units, op=dist.ReduceOp.SUM)
    tot_ms, int_ms, trc_ms, set_ms, tot_b_ms = (float(x) for x in tot.tolist())
    total_units = float(units.item())

    # roofline of the dominant streaming kernel (integration), live over the timed region:
    # algorithmic bytes from the kernel's own count of voxels written / the
    # CUDA-event time of the Integrate phase summed over the same steps
    roof = None
    if world == 1:
        peaks, which = load_peaks()
        vis_total = float(sum(counts))
        ab = 40.0 * updated_timed + 20.0 * vis_total + 4.0 * W * H * args.steps
        gbs = ab / (int_ms * 1e-3) / 1e9
        traffic, traffic_src = traffic_from_profiles("vk_integrate_kernel<0")
        roof = {"bound": "hbm", "kernel": "vk_integrate_kernel<MODE_DEPTH>", "achieved": gbs,
                "peak": peaks["hbm_gbs"], "peak_source": which, "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                "frac_of_spec_8000": gbs / 8000.0,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel, read from the
                # committed `ncu --set full` capture of this command
                "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": ab / args.steps,
                "updated_voxels_per_launch": updated_timed / args.steps,
                "launch_ms": int_ms / args.steps}

    # ---------------- end-to-end leg (host buffers, copies inside the timed region) ---------
    vol = new_volume()
    pinned_in = [torch.from_numpy(d).pin_memory() for d, _ in frames]
    d_in = torch.empty((H, W), dtype=torch.float32, device=dev)
    cin = [cframe(k, d_in) for k in range(nfr)]
    cout = [cframe(k, outs[0]) for k in range(nfr)]
    h_depth = torch.empty((H, W), dtype=torch.float32).pin_memory()
    h_color = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    h_normal = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    band_rows = []

    def e2e_sync():
        """One frame at a time: upload, SetView+Integrate+Trace, download, host sync.  Sharded: every rank
        uploads the depth image, renders ITS band of rows (no image exchange: the consumer is the host) and
        downloads that band only - the ranks' downloads together are one image set."""
        for k in range(nfr):
            if k == args.warmup:
                barrier()
                t0 = time.perf_counter()
            d_in.copy_(pinned_in[k], non_blocking=True)
            if world == 1:
                step(vol, cin[k], cout[k], outs)
                h_depth.copy_(outs[0], non_blocking=True)
                h_color.copy_(outs[1], non_blocking=True)
                h_normal.copy_(outs[2], non_blocking=True)
            else:
                step(vol, cin[k], cout[k], outs, gather=0)
                b0, b1 = shards[id(vol)].band()   # synchronises: the split is computed on the device
                if k >= args.warmup:
                    band_rows.append(b1 - b0)
                if b1 > b0:
                    for host, ptr, ch in ((h_depth, sharded_out[0], 1), (h_color, sharded_out[1], 3), (h_normal, sharded_out[2], 3)):
                        src = vb.api.device_view(ptr.value, (H, W) if ch == 1 else (H, W, 3), "<f4")
                        host[b0:b1].copy_(src[b0:b1], non_blocking=True)
            torch.cuda.current_stream().synchronize()   # the caller reads the result every step
        barrier()
        return time.perf_counter() - t0

    host_submit = []

    def e2e_streamed():
        """vulcan::FramePipeline (include/vulcan/frame_pipeline.h, C++ host) through ONE C entry point,
        vulcan_pipeline_run: the same copies and kernels for every frame, the copies of neighbouring frames
        overlapped with the kernels; the host takes frame k-lag's images (and reads a pixel of each) after
        submitting frame k; the call returns when every frame's images are in host memory."""
        cxx = C.CDLL(os.path.join(ROOT, "vulcan_b200", "libvulcan.so"))
        cxx.vulcan_pipeline_create.restype = C.c_void_p
        cxx.vulcan_pipeline_create.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int]
        cxx.vulcan_pipeline_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_int,
                                            C.POINTER(C.c_double), C.POINTER(C.c_double)]
        cxx.vulcan_pipeline_destroy.argtypes = [C.c_void_p]
        cxx.vulcan_pipeline_error.restype = C.c_char_p
        cxx.vulcan_pipeline_error.argtypes = [C.c_void_p]
        depth = args.pipeline_depth
        lag = depth - 1
        h = cxx.vulcan_pipeline_create(MAIN, EXCESS, VOXEL, TRUNC, W, H, depth)
        assert h, "vulcan_pipeline_create failed"
        ptrs = (C.c_void_p * nfr)(*[t.data_ptr() for t in pinned_in])   # page-locked host images
        all_poses = np.ascontiguousarray(np.stack([T for _, T in frames]), dtype=np.float32)
        proj = np.ascontiguousarray(K, dtype=np.float32)
        acc = C.c_double(0)
        sub = (C.c_double * nfr)()

        def run(first, count):
            sub_at = C.cast(C.addressof(sub) + 8 * first, C.POINTER(C.c_double))
            rc = cxx.vulcan_pipeline_run(h, count, C.cast(C.addressof(ptrs) + 8 * first, C.POINTER(C.c_void_p)),
                                         proj.ctypes.data, all_poses[first:].ctypes.data, lag, C.byref(acc), sub_at)
            assert rc == 0, cxx.vulcan_pipeline_error(h)

        run(0, args.warmup)
        barrier()
        t0 = time.perf_counter()
        run(args.warmup, args.steps)
        t = time.perf_counter() - t0
        host_submit.clear()
        host_submit.extend([0.0] * args.warmup + [sub[i] * 1e-6 for i in range(args.warmup, nfr)])
        cxx.vulcan_pipeline_destroy(h)
        return t, acc.value

    def e2e_streamed_sharded():
        """vulcan::ShardedPipeline (include/vulcan/sharded_pipeline.h, C++ host) on every rank through ONE C entry
        point, vulcan_sharded_pipeline_run: upload of frame k+1, SetView + Integrate + sharded Trace of frame k and the
        download of this rank's band of frame k-1 overlap; the host takes frame k-2's band (and reads a pixel of each
        image in it) after submitting frame k.  No image is exchanged between the GPUs; torch.distributed only
        carries the CUDA IPC handles once."""
        cxx = C.CDLL(os.path.join(ROOT, "vulcan_b200", "libvulcan.so"))
        cxx.vulcan_sharded_pipeline_create.restype = C.c_void_p
        cxx.vulcan_sharded_pipeline_create.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int,
                                                       C.c_int, C.c_int]
        cxx.vulcan_sharded_pipeline_export.argtypes = [C.c_void_p, C.c_void_p]
        cxx.vulcan_sharded_pipeline_connect.argtypes = [C.c_void_p, C.c_void_p]
        cxx.vulcan_sharded_pipeline_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_int,
                                                    C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
        cxx.vulcan_sharded_pipeline_destroy.argtypes = [C.c_void_p]
        cxx.vulcan_sharded_pipeline_error.restype = C.c_char_p
        cxx.vulcan_sharded_pipeline_error.argtypes = [C.c_void_p]
        depth = max(3, args.pipeline_depth)
        h = cxx.vulcan_sharded_pipeline_create(MAIN, EXCESS, VOXEL, TRUNC, W, H, depth, rank, world)
        assert h, "vulcan_sharded_pipeline_create failed"
        nb = cxx.vulcan_shard_handles_bytes()
        mine = C.create_string_buffer(nb)
        assert cxx.vulcan_sharded_pipeline_export(h, mine) == 0, cxx.vulcan_sharded_pipeline_error(h)
        blobs = [mine.raw]
        if world > 1:
            blobs = [None] * world
            dist.all_gather_object(blobs, mine.raw)
        everyone = C.create_string_buffer(b"".join(blobs), nb * world)
        assert cxx.vulcan_sharded_pipeline_connect(h, everyone) == 0, cxx.vulcan_sharded_pipeline_error(h)
        barrier()
        ptrs = (C.c_void_p * nfr)(*[t.data_ptr() for t in pinned_in])   # page-locked host images
        all_poses = np.ascontiguousarray(np.stack([T for _, T in frames]), dtype=np.float32)
        proj = np.ascontiguousarray(K, dtype=np.float32)
        acc = C.c_double(0)
        sub = (C.c_double * nfr)()
        rows = (C.c_int * nfr)()

        def run(first, count):
            sub_at = C.cast(C.addressof(sub) + 8 * first, C.POINTER(C.c_double))
            rows_at = C.cast(C.addressof(rows) + 4 * first, C.POINTER(C.c_int))
            rc = cxx.vulcan_sharded_pipeline_run(h, count, C.cast(C.addressof(ptrs) + 8 * first, C.POINTER(C.c_void_p)),
                                                 proj.ctypes.data, all_poses[first:].ctypes.data, 2, C.byref(acc), sub_at, rows_at)
            assert rc == 0, cxx.vulcan_sharded_pipeline_error(h)

        run(0, args.warmup)
        barrier()
        t0 = time.perf_counter()
        run(args.warmup, args.steps)
        barrier()
        t = time.perf_counter() - t0
        band_rows.clear()
        band_rows.extend(int(rows[i]) for i in range(args.warmup, nfr))
        host_submit.clear()
        host_submit.extend([0.0] * args.warmup + [sub[i] * 1e-6 for i in range(args.warmup, nfr)])
        barrier()
        cxx.vulcan_sharded_pipeline_destroy(h)
        barrier()
        return t, acc.value

    barrier()
    t_sync = e2e_sync()
    if world == 1 and os.environ.get("BENCH_FORCE_SHARDED_E2E"):   # diagnostics: a single-rank sharded pipeline
        del vol
        torch.cuda.empty_cache()
        t_e2e, _ = e2e_streamed_sharded()
    elif world == 1:
        del vol
        torch.cuda.empty_cache()
        barrier()
        t_e2e, _ = e2e_streamed()   # the C++ pipeline owns its volume
    else:
        drop_volume(vol)
        del vol
        torch.cuda.empty_cache()
        barrier()
        t_e2e, _ = e2e_streamed_sharded()
    e2e_mode = "streamed"
    t_e2e_t = torch.tensor([t_e2e, t_sync], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e_t, op=dist.ReduceOp.MAX)
    t_e2e, t_sync = (float(x) for x in t_e2e_t.tolist())
    h2d = W * H * 4 * world          # every rank uploads the depth image
    d2h = W * H * 4 * 7              # one image set in total: a rank downloads only the rows it rendered
    if world > 1:
        rows = torch.tensor([float(sum(band_rows))], dtype=torch.float64, device=dev)
        dist.all_reduce(rows, op=dist.ReduceOp.SUM)
        d2h = int(round(float(rows.item()) / max(1, args.steps))) * W * 4 * 7

    line = {
        "metric": METRIC, "value": total_units / (tot_ms * 1e-3), "unit": "voxel-updates/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": dict(CONFIG),
        "sharding": "none" if world == 1 else SHARDING,
        "phases": {"setview_ms": set_ms / args.steps, "integrate_ms": int_ms / args.steps,
                   "trace_ms": trc_ms / args.steps,
                   "integrate_voxel_updates_per_s": total_units / (int_ms * 1e-3),
                   "trace_rays_per_s": W * H * args.steps / (trc_ms * 1e-3),
                   "step_ms_with_phase_events": tot_b_ms / args.steps,
                   "from": "a second pass over the same frames on a fresh volume with CUDA events BETWEEN the phases; "
                           "those events end the programmatic overlap of neighbouring kernels, so that pass's step is "
                           "slower than ms_per_step, which is timed with events around the step only"},
        "visible_blocks_per_frame": float(np.mean(counts)),
        "clocks": clocks,
        "e2e": {"value": total_units / t_e2e, "unit": "voxel-updates/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": t_e2e * 1e3 / args.steps,
                "what": ("vulcan::FramePipeline (C++ host, include/vulcan/frame_pipeline.h; %d frames in flight) through the C "
                         "entry point vulcan_pipeline_run: every frame's pinned host depth -> device, "
                         "SetView+Integrate+Trace, traced depth+colour+normal -> pinned host; copies of neighbouring "
                         "frames overlap the kernels; the host takes (and reads a pixel of) frame k-%d's images after "
                         "submitting frame k" % (args.pipeline_depth, args.pipeline_depth - 1))
                if (e2e_mode == "streamed" and world == 1) else
                ("vulcan::ShardedPipeline (C++ host, include/vulcan/sharded_pipeline.h) on every rank through the C entry point "
                 "vulcan_sharded_pipeline_run (%d frames in flight): every frame's pinned host depth -> device on "
                 "every rank, SetView+Integrate+sharded Trace (each rank renders its band of rows, no image exchange), the "
                 "rank's band of depth+colour+normal -> pinned host; copies of neighbouring frames overlap the kernels; the "
                 "host takes (and reads a pixel of) frame k-2's band after submitting frame k" % max(3, args.pipeline_depth))
                if e2e_mode == "streamed" else
                ("pinned host depth -> device on every rank, SetView+Integrate+Trace (each rank renders its band of rows), "
                 "each rank's band of depth+colour+normal -> pinned host, sync every step" if world > 1 else
                 "pinned host depth -> device, SetView+Integrate+Trace, traced depth+colour+normal -> pinned host, sync every step"),
                "host_submit_us_median": (float(np.median(host_submit[args.warmup:]) * 1e6) if host_submit else None),
                "sync_every_step": {"value": total_units / t_sync, "ms_per_step": t_sync * 1e3 / args.steps}},
        "gpu_launches": int(launches),
        "wall_s_timed_region": t_wall,
    }
    if roof:
        line["roofline"] = roof
    return line


def run_reference(args):
    """The reference's own CUDA sources through its own classes (oracle/_ref), same frames."""
    import torch
    import harness as Hh
    if not Hh.have_ref():
        return {"impl": "reference", "unavailable": "oracle/_ref/libvulcan_ref.so not built (needs /root/reference at build time)"}
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    K, frames = make_frames(args.warmup + args.steps)
    nfr = len(frames)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def new_ref():
        rv = Hh.RefVolume(MAIN, EXCESS, VOXEL, TRUNC)
        rv.set_ranges()
        return rv

    # device-resident leg: pass A with events around the step only (the headline, like our arm), pass B with events
    # between the phases for the per-phase split
    dev_depth = [torch.from_numpy(d).to(dev) for d, _ in frames]

    def device_pass(phase_events):
        rv = new_ref()
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(nfr)]
        counts, passes = [], []
        sampler = ClockSampler(0) if not phase_events else None
        torch.cuda.synchronize()
        for k in range(nfr):
            if k == args.warmup:
                torch.cuda.synchronize()
                if sampler: sampler.start()
            flush.zero_()
            rv.set_frame_device(dev_depth[k].data_ptr(), W, H, K, frames[k][1])
            ev = evs[k]
            ev[0].record()
            p = rv.set_view_converged()
            if phase_events: ev[1].record()
            rv.integrate(0)
            if phase_events: ev[2].record()
            rv.trace(W, H, K, frames[k][1])
            ev[3].record()
            if k >= args.warmup:
                counts.append(rv.visible_count())
                passes.append(p)
        torch.cuda.synchronize()
        res = {"clocks": sampler.finish() if sampler else None, "counts": counts, "passes": passes}
        timed = range(args.warmup, nfr)
        res["tot"] = sum(evs[k][0].elapsed_time(evs[k][3]) for k in timed)
        if phase_events:
            res["set"] = sum(evs[k][0].elapsed_time(evs[k][1]) for k in timed)
            res["int"] = sum(evs[k][1].elapsed_time(evs[k][2]) for k in timed)
            res["trc"] = sum(evs[k][2].elapsed_time(evs[k][3]) for k in timed)
        rv.close()
        return res

    pa = device_pass(False)
    pb = device_pass(True)
    clocks, counts, passes = pa["clocks"], pa["counts"], pa["passes"]
    tot_ms, set_ms, int_ms, trc_ms, tot_b_ms = pa["tot"], pb["set"], pb["int"], pb["trc"], pb["tot"]
    units = float(sum(counts)) * 512.0

    # end-to-end leg: Image upload from host, SetView..Trace, download of the traced images
    rv = new_ref()
    pinned_in = [torch.from_numpy(d).pin_memory() for d, _ in frames]
    h_depth = torch.empty((H, W), dtype=torch.float32).pin_memory()
    h_color = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    h_normal = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    torch.cuda.synchronize()
    for k in range(nfr):
        if k == args.warmup:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        rv.set_frame(pinned_in[k].numpy(), K, frames[k][1])
        rv.set_view_converged()
        rv.integrate(0)
        rv.trace(W, H, K, frames[k][1])
        rv.lib.ref_download(h_depth.data_ptr(), rv.ptr(4), W * H * 4)
        rv.lib.ref_download(h_color.data_ptr(), rv.ptr(5), W * H * 12)
        rv.lib.ref_download(h_normal.data_ptr(), rv.ptr(6), W * H * 12)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    rv.close()
    return {
        "impl": "reference", "device": "cuda (the reference is CUDA-only; built unmodified for sm_100a, -O3 -DNDEBUG)",
        "metric": METRIC, "value": units / (tot_ms * 1e-3), "unit": "voxel-updates/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(CONFIG),
        "setview": "Volume::SetView repeated until the visible count is stable (what the reference's callers do)",
        "phases": {"setview_ms": set_ms / args.steps, "integrate_ms": int_ms / args.steps,
                   "trace_ms": trc_ms / args.steps, "setview_passes_mean": float(np.mean(passes)),
                   "integrate_voxel_updates_per_s": units / (int_ms * 1e-3),
                   "trace_rays_per_s": W * H * args.steps / (trc_ms * 1e-3),
                   "step_ms_with_phase_events": tot_b_ms / args.steps,
                   "from": "a second pass over the same frames with CUDA events between the phases"},
        "visible_blocks_per_frame": float(np.mean(counts)),
        "clocks": clocks,
        "e2e": {"value": units / t_e2e, "unit": "voxel-updates/s", "h2d_bytes_per_step": W * H * 4,
                "d2h_bytes_per_step": W * H * 28, "ms_per_step": t_e2e * 1e3 / args.steps},
    }


def cpu_baseline(sample_frames=6):
    """The oracle port on the host cores: a bounded sample (first frames of the same orbit)."""
    import harness as Hh
    K, frames = make_frames(sample_frames)
    threads = os.cpu_count() or 1
    ov = Hh.OracleVolume(MAIN, EXCESS, VOXEL, TRUNC)
    units = 0
    t_int = t_trc = t_set = 0.0
    ray_bytes = []
    for depth, T in frames:
        t0 = time.perf_counter()
        ov.set_view_converged(depth, K, T)
        t1 = time.perf_counter()
        ov.integrate_depth(depth, K, T)
        t2 = time.perf_counter()
        ov.trace(W, H, K, T)
        t3 = time.perf_counter()
        vox, ent = ov.lib.vo_voxel_fetches(ov.h), ov.lib.vo_entry_fetches(ov.h)
        # requested bytes of the reference algorithm for one traced view
        ray_bytes.append(32.0 * W * H + 8.0 * W * H + 20.0 * vox + 16.0 * ent)
        units += len(ov.visible()) * 512
        t_set += t1 - t0
        t_int += t2 - t1
        t_trc += t3 - t2
    tot = t_set + t_int + t_trc
    return {"value": units / tot, "unit": "voxel-updates/s", "cores": threads, "kind": "port",
            "sample": "first %d frames of the same orbit through oracle/vulcan_oracle.cpp (OpenMP, %d threads; "
                      "allocation sequential)" % (sample_frames, threads),
            "ms_per_step": tot * 1e3 / sample_frames,
            "integrate_voxel_updates_per_s": units / t_int, "trace_rays_per_s": W * H * sample_frames / t_trc,
            "trace_algorithmic_bytes_per_view": float(ray_bytes[-1])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-variant", default="release", choices=["release", "asshipped"],
                    help="reference arm: the -O3 -DNDEBUG build (the bar) or the reference's default build without "
                         "NDEBUG (a device synchronisation after every launch), timed once for the record")
    ap.add_argument("--pipeline-depth", type=int, default=3,
                    help="frames in flight in the end-to-end leg (FramePipeline); the host takes frame k-depth+1")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    sys.setswitchinterval(2e-4)  # lets the clock-sampling thread run while the main thread issues launches
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))

    if args.impl == "reference":
        if rank != 0:
            return
        if args.ref_variant == "asshipped":
            os.environ["VULCAN_REF_VARIANT"] = "asshipped"
        line = run_reference(args)
        line["ref_build"] = args.ref_variant
        if "unavailable" not in line and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line))
        return

    import torch
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl")
    line = run_ours(args, rank, world)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
            # ray cast: requested bytes of the reference algorithm (counted by the oracle on frame 5 of the same
            # orbit) over our Trace time - a latency-bound gather kernel, given for completeness
            rb = line["cpu_baseline"]["trace_algorithmic_bytes_per_view"]
            peaks, _ = load_peaks()
            gbs = rb / (line["phases"]["trace_ms"] * 1e-3) / 1e9
            line["phases"]["trace_algorithmic_gbs"] = gbs
            line["phases"]["trace_frac_of_hbm_peak"] = gbs / peaks["hbm_gbs"]
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
