#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json): Mrays/s and ms/frame of the
Office-class scene (34 000 triangles, median-split BVH) at 1920x1080, 1 spp.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU path (OpenMP)

One "step" = one frame = one pass of the hot path over all pixels of the frame.
  value      whole-job Mrays/s, scene/camera/lights resident in HBM: K frames through the frame pipeline
             (my-raytracer_b200/dist.py: one render-kernel launch per rank and frame, finished pixels stored
             straight into rank 0's frame buffer over NVLink, 2 frames in flight), CUDA-event time from the
             first launch to the last frame's completion signal, max over ranks
  e2e        the same pipeline with HOST buffers: per step H2D of the lights from pinned memory, the C-ABI
             call rt_render_device, D2H of the finished RGBA8 frame into pinned memory on rank 0
             (e2e.sync: the synchronous rt_render call, one frame at a time, N = 1)
  roofline   FP32-issue roofline of the render kernel: algorithmic flops of the box /
             triangle / shading tests actually executed (instrumented launch, untimed) over the timed
             kernel duration, against 148 SMs x 128 lanes x 2 x sm_max_mhz
  cpu_baseline  the oracle timed on the host cores on a bounded row sample of the same frame
Rays are counted as the CPU semantics cast them: primary + one shadow ray per (shadowable hit, light)
+ one reflection ray per mirror hit below max_depth (identical counts on both arms, tests check it).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "office_34k_1920x1080_1spp"
NCU_TRAFFIC_BYTES = 8006656 + 541440   # render_kernel<4>, Office 1080p: dram read + write per launch (ncu --set full, profiles/r01_ncu_render_kernel_v8.md)
METRIC = "Mrays/s (Office 1920x1080 1spp; rays = primary + shadow + reflection as the CPU path casts them)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--topology", default="sah", choices=["sah", "reference"],
                    help="tree ordinary rays walk: the library's SAH tree (default) or the reference's median-split topology")
    ap.add_argument("--no-ab", action="store_true", help="skip the A/B frame on the other tree")
    return ap.parse_args()


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def cpu_leg(desc, seconds_target, kind_pref=("reference_fast", "reference", "port")):
    """Times the oracle on a bounded row sample of the frame (all host threads)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc
    kind = next(k for k in kind_pref if orc.available(k))
    o = orc.OracleScene(desc, kind)
    H = int(desc.camera["height"])
    cores = os.cpu_count() or 1
    probe = o.render(want_ids=False, y0=3, ystep=64, nthreads=cores)     # ~1.6 % of the rows
    rows_probe = len(range(3, H, 64))
    per_row = probe["seconds"] / rows_probe
    step = max(1, int(round(H * per_row / max(seconds_target, 1e-3))))
    r = o.render(want_ids=False, y0=0, ystep=step, nthreads=cores)   # torchrun exports OMP_NUM_THREADS=1
    rows = len(range(0, H, step))
    rays = sum(r["rays"].values())
    return {
        "value": rays / r["seconds"] / 1e6, "unit": "Mrays/s", "cores": cores,
        "kind": "reference" if kind.startswith("reference") else "port",
        "sample": f"{rows} of {H} rows (every {step}th row, full width) of the same frame, {r['seconds']:.2f} s, "
                  f"{rays} rays; OpenMP dynamic rows; library {kind}",
        "ms_per_frame_extrapolated": r["seconds"] * 1e3 * H / rows,
        "seconds": r["seconds"], "rays": rays, "rows": rows,
    }


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (oracle/_ref when it was
    compiled in the build container, else the port) on the host cores, same config and metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import rtpkg
    rt = rtpkg.load()
    desc = rt.scenes.office(width=args.width, height=args.height)
    # each step is a bounded row sample sized so that steps+warmup finish within a few minutes
    budget = 150.0 / max(args.steps + args.warmup, 1)
    legs = [cpu_leg(desc, min(budget, args.cpu_seconds)) for _ in range(args.warmup + args.steps)][args.warmup:]
    rays = sum(l["rays"] for l in legs)
    secs = sum(l["seconds"] for l in legs)
    value = rays / secs / 1e6
    frame_ms = sum(l["ms_per_frame_extrapolated"] for l in legs) / len(legs)
    cb = dict(legs[-1]); cb["value"] = value
    for k in ("seconds", "rays", "rows"):
        cb.pop(k, None)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": frame_ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "triangles": desc.num_triangles, "lights": len(desc.lights),
                   "max_depth": desc.max_depth, "note": "ms_per_step = sampled rows extrapolated to the full frame"},
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def l2_read_gbps(dev):
    """Measured L2 read bandwidth: repeated reduction over a 48 MiB buffer that stays L2-resident (126 MB L2).
    MEASURED_PEAKS.json has no L2 figure; the north star's roofline for an L2-resident scene needs one."""
    import torch
    x = torch.ones(48 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    for _ in range(5):
        x.sum()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(3):
        a.record()
        for _ in range(20):
            x.sum()
        b.record(); b.synchronize()
        best = max(best, 20 * x.numel() * 4 / (a.elapsed_time(b) * 1e-3) / 1e9)
    return best


def run_b200(args):
    import numpy as np
    import torch
    import rtpkg
    rt = rtpkg.load()
    from my_raytracer_b200 import capi
    from my_raytracer_b200 import dist as rdist
    lib = rt.load_library()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    capi.check(lib.rt_init_device(local_rank), lib)
    if world > 1:
        import torch.distributed as dist
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)   # the per-frame signal must not queue
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)      # behind the persistent render grids
    desc = rt.scenes.office(width=args.width, height=args.height)
    W, H = args.width, args.height

    scene_flags = capi.RT_SCENE_REFERENCE_TOPOLOGY if args.topology == "reference" else 0
    # ---- scene: rank 0 builds + flattens, one broadcast replicates the blob -----------------------------
    t0 = time.perf_counter()
    host = rt.HostScene(desc, lib)                       # every rank needs camera/lights; cheap (34 k tris)
    host_ms = (time.perf_counter() - t0) * 1e3
    if world > 1:
        blob = rt.DeviceScene(host, scene_flags=scene_flags).export_blob() if rank == 0 else None
        tb = time.perf_counter()
        t_blob = rdist.broadcast_blob(blob, dev)
        torch.cuda.synchronize()
        bcast_ms = (time.perf_counter() - tb) * 1e3
        scene = rt.DeviceScene(lib=lib, blob_device_ptr=t_blob.data_ptr(), blob_bytes=t_blob.numel())
    else:
        scene = rt.DeviceScene(host, scene_flags=scene_flags)
        bcast_ms = 0.0
    info = scene.info()
    cam = host.camera
    chunk = int(os.environ.get("RT_TILE_CHUNK", "0")) or None
    # frames in flight: 4 full-width grids on one GPU.  Split over N GPUs a frame shrinks towards the time of a
    # single heavy tile (8 GPUs: ~0.2 ms of work per GPU), so 8 frames are kept in flight on narrower grids (2 CTAs
    # per SM each at 2-4 GPUs, 1 at 8: frames resident side by side, every warp renders 2-4x more tiles of its
    # frame) - profiles/r01_variants.md
    depth = int(os.environ.get("RT_FRAMES_IN_FLIGHT", "8" if world >= 2 else "4"))
    dyn_fraction = float(os.environ.get("RT_DYN_FRACTION", "0.15"))
    FLUSH = int(os.environ.get("RT_FLUSH_MB", "160")) * 1024 * 1024             # > 126 MB L2 (env: experiments only)
    # 0 = full-width grids (one GPU); half-width at 2-4 GPUs; quarter-width at 8
    ctas = int(os.environ.get("RT_CTAS_PER_SM", "0" if world == 1 else ("2" if world < 8 else "1")))
    pipe = rdist.FramePipeline(lib, scene.handle, host, dev, W, H, depth=depth, ring=depth + 2, tile_chunk=chunk,
                               dyn_fraction=dyn_fraction, flush_bytes=FLUSH, ctas_per_sm=ctas)
    nl = pipe.nl
    stream = torch.cuda.current_stream()

    # ---- algorithmic work of one frame (instrumented launch, untimed; deterministic) -------------------
    # static-only partition here: the per-rank share of the dynamic tail varies from frame to frame, the sum
    # over ranks does not
    d_tmp = torch.zeros(H * W * 4, dtype=torch.uint8, device=dev)
    fo = capi.FrameOut(d_tmp.data_ptr(), None, None, None)
    cparams = host.params(tile=pipe.tile, flags=capi.RT_FLAG_COUNTERS)
    capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                    C.byref(cparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
    cnt = capi.Counters()
    capi.check(lib.rt_get_counters(scene.handle, C.byref(cnt)), lib)
    cnt = cnt.as_dict()
    ct = torch.tensor([cnt[k] for k in sorted(cnt)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(ct)
    tot = dict(zip(sorted(cnt), [int(v) for v in ct.tolist()]))
    rays_frame = tot["rays_primary"] + tot["rays_shadow"] + tot["rays_reflect"]

    def timed(n_frames, host_io):
        """n_frames through the pipeline; CUDA-event time from the first launch to the last frame's completion
        signal, max over ranks."""
        pipe.drain()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        start = torch.cuda.Event(enable_timing=True)
        start.record(pipe.compute[pipe.submitted % pipe.depth])
        for _ in range(n_frames):
            pipe.submit(host_io=host_io)
        last = pipe.last_event
        pipe.drain()
        ms = start.elapsed_time(last)
        return rdist.max_over_ranks(ms, dev) if world > 1 else ms

    # ---- timed region: value (device-resident inputs, frames in flight) --------------------------------------
    for _ in range(max(args.warmup, 3)):
        pipe.submit()
    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms = timed(args.steps, host_io=False)
    clocks = sampler.result()
    ms_per_step = total_ms / args.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: same pipeline with HOST buffers: H2D of the lights, kernel, D2H of the frame on rank 0 --------
    for _ in range(3):
        pipe.submit(host_io=True)
    e2e_total = timed(args.steps, host_io=True)
    e2e_ms = e2e_total / args.steps
    e2e_value = rays_frame / (e2e_ms * 1e-3) / 1e6

    # ---- one frame at a time (latency): serial launches, CUDA events around the kernel ---------------------
    sparams = host.params(tile=pipe.tile, flags=capi.RT_FLAG_NO_CLEAR)
    capi.check(lib.rt_scene_set_launch_width(scene.handle, 0), lib)     # one frame alone gets the whole machine
    kern_ms = []
    for i in range(8):
        pipe.flush.fill_(1)
        capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                        C.byref(sparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
        kern_ms.append(scene.last_kernel_ms())
    kms_local = sum(kern_ms[2:]) / len(kern_ms[2:])
    rank_ms = [kms_local]
    if world > 1:
        g = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(g, torch.tensor([kms_local], dtype=torch.float64, device=dev))
        rank_ms = [float(x.item()) for x in g]
    # the synchronous public call with host buffers (rt_render), one frame at a time, N = 1 only
    sync_e2e = None
    if world == 1:
        h_rgba = torch.zeros(H * W * 4, dtype=torch.uint8).pin_memory()
        hfo = capi.FrameOut(h_rgba.data_ptr(), None, None, None)
        lp = np.ascontiguousarray(host.lights_pos); lc = np.ascontiguousarray(host.lights_color)
        eparams = host.params(flags=capi.RT_FLAG_NO_CLEAR)
        for i in range(3 + 10):
            if i == 3:
                te = time.perf_counter()
            capi.check(lib.rt_render(scene.handle, C.byref(cam), lp.ctypes.data_as(capi.Vec4_p), lc.ctypes.data_as(capi.Vec4_p),
                                     nl, C.byref(eparams), C.byref(hfo)), lib)
        sync_ms = (time.perf_counter() - te) * 1e3 / 10
        sync_e2e = {"value": rays_frame / (sync_ms * 1e-3) / 1e6, "ms_per_step": sync_ms, "api": "rt_render, one frame at a time"}

    # ---- A/B: the same frame with every ray on the reference's median-split topology (N = 1, serial frames) ----
    other_tree = None
    if world == 1 and not args.no_ab:
        other_flags = 0 if scene_flags else capi.RT_SCENE_REFERENCE_TOPOLOGY
        sc2 = rt.DeviceScene(host, scene_flags=other_flags)
        ms2 = []
        for i in range(8):
            pipe.flush.fill_(1)
            capi.check(lib.rt_render_device(sc2.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                            C.byref(sparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
            ms2.append(sc2.last_kernel_ms())
        k2 = sum(ms2[2:]) / len(ms2[2:])
        other_tree = {"traversal_tree": "sah" if scene_flags else "reference", "kernel_ms_single_frame": k2,
                      "mrays": rays_frame / (k2 * 1e-3) / 1e6,
                      "note": "same frame bit for bit (tests/test_gpu_parity.py); compare with latency_ms_single_frame"}
        sc2.close()
    # ---- A/B: ordinary rays on the binary traversal tree instead of its 4-wide twin (N = 1, serial frames) ----
    binary_nodes = None
    if world == 1 and not args.no_ab and info.get("num_nodes_wide", 0) > 0:
        bparams = host.params(tile=pipe.tile, flags=capi.RT_FLAG_NO_CLEAR | capi.RT_FLAG_BINARY_NODES)
        ms3 = []
        for i in range(8):
            pipe.flush.fill_(1)
            capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                            C.byref(bparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
            ms3.append(scene.last_kernel_ms())
        k3 = sum(ms3[2:]) / len(ms3[2:])
        binary_nodes = {"kernel_ms_single_frame": k3, "mrays": rays_frame / (k3 * 1e-3) / 1e6,
                        "note": "RT_FLAG_BINARY_NODES, same frame bit for bit; compare with latency_ms_single_frame"}

    wide = tot.get("nodes4_visited", 0) > 0      # ordinary rays walked the 4-wide tree (render_kernel<K_WIDE>)
    # ---- roofline of the render kernel -------------------------------------------------------------
    peaks = measured_peaks()
    sm_max = float(peaks.get("sm_max_mhz", 1965.0))
    peak_tflops = 148 * 128 * 2 * sm_max * 1e6 / 1e12 * world     # whole-job flops against the N-GPU peak
    L = nl
    flops = 24 * tot["box_tests"] + 48 * tot["tri_filter"] + (60 * L + 20) * tot["shaded_hits"]
    alg_bytes = 32 * tot["box_tests"] + 48 * tot["tri_filter"] + 64 * tot["shaded_hits"] + 16 * tot["texel_fetches"] + 4 * W * H
    achieved = flops / (ms_per_step * 1e-3) / 1e12
    l2_bw = l2_read_gbps(dev) * world
    t_fp32 = flops / (peak_tflops * 1e12)
    t_l2 = alg_bytes / (l2_bw * 1e9)
    roof = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
            # DRAM bytes per launch of this kernel on this workload from the committed `ncu --set full` capture
            # (profiles/r01_ncu_render_kernel_v8.md: dram__bytes_read.sum + dram__bytes_write.sum); only quoted for
            # the configuration that capture was taken on
            "traffic": NCU_TRAFFIC_BYTES if (world == 1 and (W, H) == (1920, 1080) and args.topology == "sah" and wide) else None,
            "traffic_unit": "bytes per launch (ncu, profiles/r01_ncu_render_kernel_v8.md)",
            "kernel": "rtb::render_kernel<4>" if wide else "rtb::render_kernel<0>",
            "timed_over": "flops of one frame / ms_per_step of the timed region (kernels of successive frames overlap, "
                          "so the per-launch duration is not the divisor; flush and launch gaps are inside)",
            "kernel_ms_single_frame": max(rank_ms), "kernel_ms_per_rank": rank_ms,
            "peak_source": "148 SMs x 128 FP32 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json (FP32 is not in the "
                           "driver-measured set; the scene is L2-resident so HBM does not bound it)",
            "flops_per_launch": flops, "algorithmic_bytes_per_launch": alg_bytes,
            "mem_GBps": alg_bytes / (ms_per_step * 1e-3) / 1e9, "hbm_peak_GBps": peaks.get("hbm_gbs"),
            "north_star": {"definition": "slower of FP32 peak for the counted intersection flops and L2 bandwidth for the "
                                         "counted node/triangle bytes (scene resident in L2)",
                           "l2_read_GBps_measured": l2_bw, "l2_how": "torch.sum over an L2-resident 48 MiB buffer, best of 3 x 20",
                           "t_fp32_ms": t_fp32 * 1e3, "t_l2_ms": t_l2 * 1e3,
                           "frac": max(t_fp32, t_l2) * 1e3 / ms_per_step,
                           "note": "node/triangle bytes are served by L1 (96 % hits), so this fraction can exceed 1; "
                                   "roofline.frac (FP32 issue) is the harder, reported bound"}}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 filter + f64 decide/shade", "data": "synthetic",
            "config": {"workload": WORKLOAD, "triangles": info["num_triangles"], "bvh_nodes": info["num_nodes_ref"],
                       "traversal_tree": ("reference median-split topology" if info["topology"] == 1 else
                                          "library-built SAH tree over the same triangles; reference tree kept for d.x == 0 rays"),
                       "traversal_records": info["num_nodes_traversal"],
                       "node_width": 4 if wide else 2, "wide_records": info.get("num_nodes_wide", 0),
                       "lights": nl, "max_depth": desc.max_depth, "spp": 1,
                       "l2": "flushed before every timed frame (160 MiB fill on the frame's stream)",
                       "frames_in_flight": pipe.depth, "ctas_per_sm_per_frame": pipe.ctas_per_sm or "all",
                       "parallelism": f"tiles{world}" if world > 1 else "single",
                       "tile_chunk": pipe.tile[2], "dyn_tiles": pipe.dyn_tiles, "dyn_chunk": pipe.dyn_chunk,
                       "gather": "fused: peer stores into rank 0's frame buffer over NVLink + 4-byte all-reduce signal"
                                 if world > 1 else "none"},
            "mrays_primary": tot["rays_primary"] / (ms_per_step * 1e-3) / 1e6,
            "latency_ms_single_frame": max(rank_ms),
            "other_tree": other_tree,
            "binary_nodes": binary_nodes,
            "rays_per_frame": {k: tot[k] for k in ("rays_primary", "rays_shadow", "rays_reflect")},
            "counters": tot,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(2 * nl * 32) * world, "d2h_bytes_per_step": W * H * 4,
                    "api": "rt_render_device per rank + pinned-host lights in / frame out on rank 0, frames in flight",
                    "sync": sync_e2e},
            "gpu_launches": args.steps * world,
            "roofline": roof,
            "setup_ms": {"host_scene_build": host_ms, "flatten": info["flatten_ms"], "upload": info["upload_ms"],
                         "broadcast": bcast_ms, "blob_bytes": info["blob_bytes"]},
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = {k: v for k, v in cpu_leg(desc, args.cpu_seconds).items() if k not in ("seconds", "rays", "rows")}
        print(json.dumps(line))
    pipe.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _json_only_stdout():
    """Libraries (NCCL's version banner, OpenMP/torchrun notices) write to fd 1; the contract is ONE JSON
    line on stdout.  Point fd 1 at stderr for the whole run and keep the real stdout for the result."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    a = parse()
    _real_stdout = _json_only_stdout()
    import builtins
    _print = builtins.print

    def _result_print(*args, **kw):          # the only print() calls in this file emit the JSON line
        kw.setdefault("file", _real_stdout)
        kw.setdefault("flush", True)
        _print(*args, **kw)
    print = _result_print                     # noqa: A001
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
