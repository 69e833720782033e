#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json): Mrays/s and ms/frame.

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path, Office 1080p 1 spp
    python bench.py --workload bigscene_10m_4k_4spp ...           # BASELINE config 5: 10 M triangles, 3840x2160, 2x2 spp
    python bench.py --impl reference --gpus N --steps K ...       # the reference's own CPU path (OpenMP)

One "step" = one frame = one pass of the hot path over all pixels of the frame, split over the N GPUs by image tiles.
  value      whole-job Mrays/s, scene/camera/lights resident in HBM: frames go through the frame pipeline
             (my-raytracer_b200/dist.py: one render-kernel launch per rank and frame, finished pixels stored straight
             into rank 0's frame buffer over NVLink, several frames in flight).  Timed as a STEADY-STATE window: the
             pipeline is kept fed (untimed leading and trailing frames) and the time is the CUDA-event interval between
             the completion of the frame before the window and the completion of its K-th frame, max over ranks —
             so --steps 20 and --steps 100 agree; the fill and drain of the pipeline are not part of a step.
  e2e        the same metric with HOST buffers inside the timed region: per frame H2D of the lights and the finished
             RGBA8 frame in host memory.  N = 1: the C ABI's pipelined host-buffer call (rt_frame_submit /
             rt_frame_wait, include/rt_b200.h; D2H of frame f overlaps the kernel of frame f + 1).  N > 1:
             rt_render_device per rank with the frame in ONE pinned host block that all ranks' processes map
             (POSIX shm + rt_host_register): every rank's kernel stores its pixels there over its own PCIe link.
             (e2e.sync: the synchronous rt_render call, one frame at a time, N = 1)
  roofline   FP32-issue roofline of the render kernel: algorithmic flops of the box / triangle / shading
             tests actually executed (instrumented launch, untimed) over the timed duration, against
             148 SMs x 128 lanes x 2 x sm_max_mhz
  parity     untimed: the finished GPU frame against the oracle (N = 1: ids, t, 8-bit pixels on the rows the CPU leg
             rendered) and against the 1-GPU frame (every N: bitwise); a mismatch exits non-zero
  cpu_baseline  the oracle timed on the host cores on a bounded row sample of the same frame
Rays are counted as the CPU semantics cast them: primary + one shadow ray per (shadowable hit, light)
+ one reflection ray per mirror hit below max_depth (identical counts on both arms, tests check it).
"""
from __future__ import annotations

import argparse
import ctypes as C
import glob
import hashlib
import json
import os
import re
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# workload name -> (scene function name, kwargs, width, height, spp_grid, metric)
WORKLOADS = {
    "office_34k_1920x1080_1spp": dict(
        scene="office", kwargs={}, width=1920, height=1080, spp_grid=1,
        metric="Mrays/s (Office 1920x1080 1spp; rays = primary + shadow + reflection as the CPU path casts them)"),
    "bigscene_10m_4k_4spp": dict(
        scene="scatter", kwargs={"n_objects": 10000}, width=3840, height=2160, spp_grid=2,
        metric="Mrays/s (10 M-triangle procedural scene 3840x2160, 2x2 sub-pixel grid = 4 spp; rays = primary + shadow "
               "+ reflection as the CPU path casts them)"),
}
ALIASES = {"office": "office_34k_1920x1080_1spp", "bigscene": "bigscene_10m_4k_4spp"}
# committed `ncu --set full` summaries the roofline's DRAM traffic is read from (newest first)
NCU_SUMMARIES = {
    "office_34k_1920x1080_1spp": ["profiles/r02_ncu_render_kernel_office_final.md", "profiles/r02_ncu_render_kernel_office.md", "profiles/r01_ncu_render_kernel_v8.md"],
    "bigscene_10m_4k_4spp": ["profiles/r02_ncu_render_kernel_10M.md", "profiles/r01_ncu_render_kernel_v7_10M.md"],
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="office_34k_1920x1080_1spp")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--objects", type=int, default=0, help="bigscene: number of 1000-triangle objects (default 10000)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--topology", default="sah", choices=["sah", "reference"],
                    help="tree ordinary rays walk: the library's SAH tree (default) or the reference's median-split topology")
    ap.add_argument("--pipeline", default="megakernel", choices=["megakernel", "wavefront"],
                    help="one persistent kernel per frame (default) or TRACE/SHADE kernel pairs per bounce (RT_FLAG_WAVEFRONT)")
    ap.add_argument("--no-ab", action="store_true", help="skip the A/B frames (other tree, binary nodes, flush modes)")
    a = ap.parse_args()
    a.workload = ALIASES.get(a.workload, a.workload)
    if a.workload not in WORKLOADS:
        ap.error("unknown workload %r (have: %s)" % (a.workload, ", ".join(WORKLOADS)))
    wl = WORKLOADS[a.workload]
    a.width = a.width or wl["width"]
    a.height = a.height or wl["height"]
    return a


def make_desc(rt, args, lite=False):
    """lite: a rank that does not assemble the scene only needs its camera, lights and globals (they do not depend
    on the object count)."""
    wl = WORKLOADS[args.workload]
    kw = dict(wl["kwargs"])
    if args.objects and "n_objects" in kw:
        kw["n_objects"] = args.objects
    if lite and "n_objects" in kw:
        kw["n_objects"] = 1
    return getattr(rt.scenes, wl["scene"])(width=args.width, height=args.height, **kw)


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the render kernel, read from the committed
    `ncu --set full` summary of this workload (profiles/), with the file it came from."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for rel in NCU_SUMMARIES.get(workload, []):
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        tot, found = 0.0, 0
        for line in open(path):
            m = re.match(r"\|\s*dram__bytes_(read|write)\.sum\s*\|\s*([0-9.eE+-]+)\s*\|\s*(\w+)\s*\|", line)
            if m and m.group(3) in unit:
                tot += float(m.group(2)) * unit[m.group(3)]
                found += 1
        if found == 2:
            return int(tot), rel
    return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run (one sample is taken
    synchronously at start and at the end, so even a few-millisecond region has samples)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        self.err = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:          # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def sample(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, name in names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception as e:          # noqa: BLE001
            self.err = repr(e)

    def run(self):
        while not self.stop_flag:
            self.sample()
            time.sleep(0.005)

    def result(self):
        self.sample()
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable: %s" % self.err]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


def cpu_leg(desc, seconds_target, spp_grid=1, want_ids=False, kind_pref=("reference_fast", "reference", "port"), oracle=None):
    """Times the oracle on a bounded row sample of the frame (all host threads).  Returns the timing record and, with
    want_ids, the rendered rows (ids / t / rgb) for the parity block."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc
    kind = next(k for k in kind_pref if orc.available(k))
    o = oracle or orc.OracleScene(desc, kind)
    H = int(desc.camera["height"])
    cores = os.cpu_count() or 1
    probe = o.render(want_ids=False, y0=3, ystep=64, spp_grid=spp_grid, nthreads=cores)     # ~1.6 % of the rows
    rows_probe = len(range(3, H, 64))
    per_row = probe["seconds"] / rows_probe
    step = max(1, int(round(H * per_row / max(seconds_target, 1e-3))))
    r = o.render(want_ids=want_ids, y0=0, ystep=step, spp_grid=spp_grid, nthreads=cores)   # torchrun exports OMP_NUM_THREADS=1
    rows = len(range(0, H, step))
    rays = sum(r["rays"].values())
    rec = {
        "value": rays / r["seconds"] / 1e6, "unit": "Mrays/s", "cores": cores,
        "kind": "reference" if kind.startswith("reference") else "port",
        "sample": f"{rows} of {H} rows (every {step}th row, full width) of the same frame, {r['seconds']:.2f} s, "
                  f"{rays} rays; OpenMP dynamic rows; library {kind}",
        "ms_per_frame_extrapolated": r["seconds"] * 1e3 * H / rows,
        "seconds": r["seconds"], "rays": rays, "rows": rows,
    }
    return rec, (r if want_ids else None), step, o


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (oracle/_ref when it was
    compiled in the build container, else the port) on the host cores, same config and metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import rtpkg
    rt = rtpkg.load()
    wl = WORKLOADS[args.workload]
    desc = make_desc(rt, args)
    # each step is a bounded row sample sized so that steps+warmup finish within a few minutes
    budget = 150.0 / max(args.steps + args.warmup, 1)
    legs, oracle = [], None
    for _ in range(args.warmup + args.steps):
        rec, _, _, oracle = cpu_leg(desc, min(budget, args.cpu_seconds), spp_grid=wl["spp_grid"], oracle=oracle)
        legs.append(rec)
    legs = legs[args.warmup:]
    rays = sum(l["rays"] for l in legs)
    secs = sum(l["seconds"] for l in legs)
    value = rays / secs / 1e6
    frame_ms = sum(l["ms_per_frame_extrapolated"] for l in legs) / len(legs)
    cb = dict(legs[-1]); cb["value"] = value
    for k in ("seconds", "rays", "rows"):
        cb.pop(k, None)
    print(json.dumps({
        "impl": "reference", "metric": wl["metric"], "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": frame_ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "triangles": desc.num_triangles, "lights": len(desc.lights),
                   "max_depth": desc.max_depth, "spp": wl["spp_grid"] ** 2,
                   "note": "ms_per_step = sampled rows extrapolated to the full frame"},
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


def sha1(a):
    return hashlib.sha1(memoryview(a).cast("B")).hexdigest()


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
    wl = WORKLOADS[args.workload]
    big = wl["scene"] == "scatter"
    spp_grid = wl["spp_grid"]
    desc = make_desc(rt, args, lite=rank != 0)
    W, H = args.width, args.height

    scene_flags = capi.RT_SCENE_REFERENCE_TOPOLOGY if args.topology == "reference" else 0
    # ---- scene: rank 0 builds + flattens, ONE device-to-device broadcast replicates the resident blob ----------
    t0 = time.perf_counter()
    host = rt.HostScene(desc, lib, gpu_build=big) if rank == 0 else rdist.FrameInputs(desc, lib)
    host_ms = (time.perf_counter() - t0) * 1e3
    bcast_ms = 0.0
    scene0 = rt.DeviceScene(host, scene_flags=scene_flags) if rank == 0 else None
    if world > 1:
        t_src = rdist.wrap_device_ptr(*scene0.blob_device(), device=dev) if rank == 0 else None
        torch.cuda.synchronize()
        tb = time.perf_counter()
        t_blob = rdist.broadcast_blob(t_src, dev)        # NCCL reads rank 0's blob where it lies (no host round trip)
        torch.cuda.synchronize()
        bcast_ms = (time.perf_counter() - tb) * 1e3
        scene = scene0 if rank == 0 else rt.DeviceScene(lib=lib, blob_device_ptr=t_blob.data_ptr(), blob_bytes=t_blob.numel())
        del t_blob
    else:
        scene = scene0
    info = scene.info()
    cam = host.camera
    chunk = int(os.environ.get("RT_TILE_CHUNK", "0")) or None
    # frames in flight: 4 full-width grids on one GPU.  Split over N GPUs a frame shrinks towards the time of a
    # single heavy tile (8 GPUs: ~0.2 ms of work per GPU), so 8 frames are kept in flight on narrower grids (2 CTAs
    # per SM each at 2-4 GPUs, 1 at 8: frames resident side by side, every warp renders 2-4x more tiles of its
    # frame) - profiles/r01_variants.md.  The 10 M-triangle frame is 20 ms of work: 2 frames in flight suffice.
    # (with 4 in flight and 30 % of the tiles in the cross-GPU dynamic queue its heavy tiles balance: 3.08 -> 2.79 ms at 8 GPUs)
    depth = int(os.environ.get("RT_FRAMES_IN_FLIGHT", "4" if big else ("8" if world >= 2 else "4")))
    # the wavefront shards tiles statically only (no cross-GPU dynamic tail)
    dyn_fraction = float(os.environ.get("RT_DYN_FRACTION", ("0.3" if big else "0.15") if args.pipeline == "megakernel" else "0"))
    # tiles per claim from the cross-GPU dynamic queue: a warp renders its claim tile by tile, so the claim is the tail's granularity
    dyn_chunk = int(os.environ.get("RT_DYN_CHUNK", "8"))
    pflags = capi.RT_FLAG_WAVEFRONT if args.pipeline == "wavefront" else capi.RT_FLAG_MEGAKERNEL
    FLUSH = int(os.environ.get("RT_FLUSH_MB", "160")) * 1024 * 1024             # > 126 MB L2 (env: experiments only)
    # 0 = full-width grids (one GPU); half-width at 2-4 GPUs; quarter-width at 8
    ctas = int(os.environ.get("RT_CTAS_PER_SM", "0" if (world == 1 or big) else ("2" if world < 8 else "1")))
    # Cold L2 at the start of every timed frame (timing rule: flush the L2 or use inputs larger than it).  Default
    # "rotate": the frames cycle through identical COPIES of the flattened scene that add up to 1.4x the L2, so a copy is
    # next touched only after the others have streamed through the cache - inputs larger than L2, no fill traffic inside
    # the timed region.  A scene that is itself larger than the L2 needs neither ("big").  The evicting 160 MiB fill of
    # round 1 stays selectable (RT_L2_MODE=inline: on the frame's stream ahead of its kernel; side: once per frame on a
    # stream of its own) and is A/B-timed in `flush_ab` together with no eviction at all ("off").
    l2_bytes = int(torch.cuda.get_device_properties(dev).L2_cache_size)
    blob_bytes = int(info["blob_bytes"])
    l2_mode = os.environ.get("RT_L2_MODE", "big" if blob_bytes > 1.4 * l2_bytes else "rotate")
    ncopies = max(2, min(32, -(-int(1.4 * l2_bytes) // blob_bytes))) if blob_bytes <= 1.4 * l2_bytes else 1
    copies = [scene]
    if ncopies > 1:
        p0, n0 = scene.blob_device()
        copies += [rt.DeviceScene(lib=lib, blob_device_ptr=p0, blob_bytes=n0) for _ in range(ncopies - 1)]
    shared = rdist.SharedHostFrames(lib, W * H * 4 * (depth + 2), rank, world, dev) if world > 1 else None

    def make_pipe(mode):
        rot = mode == "rotate" and ncopies > 1
        return rdist.FramePipeline(lib, [c.handle for c in copies] if rot else scene.handle, host, dev, W, H, depth=depth,
                                   ring=depth + 2, tile_chunk=chunk, dyn_fraction=dyn_fraction, dyn_chunk=dyn_chunk,
                                   flush_bytes=FLUSH if mode in ("inline", "side") else 0, ctas_per_sm=ctas,
                                   flush_mode=mode if mode in ("inline", "side") else "off",
                                   spp_grid=spp_grid, shared_host=shared, flags=pflags)
    flush_mode = l2_mode
    flush_buf = torch.empty(FLUSH, dtype=torch.uint8, device=dev) if FLUSH else None   # serial single-frame timings
    pipe = make_pipe(flush_mode)
    nl = pipe.nl
    stream = torch.cuda.current_stream()

    # ---- algorithmic work of one frame (instrumented launch, untimed; deterministic) -------------------
    # static-only partition here: the per-rank share of the dynamic tail varies from frame to frame, the sum
    # over ranks does not
    d_tmp = torch.zeros(H * W * 4, dtype=torch.uint8, device=dev)
    fo = capi.FrameOut(d_tmp.data_ptr(), None, None, None)
    cparams = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_COUNTERS)
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

    def steady(p, n_frames, host_io):
        """Steady-state time of n_frames: the pipeline is filled with `lead` untimed frames and kept fed by `depth`
        untimed trailing ones; CUDA-event interval between the completion of the last leading frame and of the
        n_frames-th frame after it, max over ranks.  Returns (ms, ring slot of the last timed frame)."""
        p.drain()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        first = len(p.done_events)
        lead = max(2, p.depth)
        slot = None
        for i in range(lead + n_frames + p.depth):
            s = p.submit(host_io=host_io)
            if i == lead + n_frames - 1:
                slot = s
        p.drain()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev = p.done_events
        ms = ev[first + lead - 1].elapsed_time(ev[first + lead - 1 + n_frames])
        del ev[:]                                        # events of finished runs are not needed again
        return (rdist.max_over_ranks(ms, dev) if world > 1 else ms), slot

    # ---- timed region: value (device-resident inputs, frames in flight) --------------------------------------
    # untimed warm-up: at least W frames, and with the rotating scene copies two rounds over all of them (every copy's
    # pages, TLB entries and launch resources touched) plus one pipeline depth - the first frames on a fresh process run
    # up to 10 % slower than the steady state
    n_warm = max(args.warmup, 3, (2 * ncopies + depth) if (flush_mode == "rotate" and ncopies > 1) else (2 * depth + 2))
    for _ in range(n_warm):
        pipe.submit()
    sampler = ClockSampler(local_rank)
    sampler.sample()
    sampler.start()
    total_ms, _ = steady(pipe, args.steps, host_io=False)
    ms_per_step = total_ms / args.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6
    pipeline_ran = scene.last_pipeline()

    # ---- e2e: HOST buffers inside the timed region ---------------------------------------------------------------
    e2e_frame_sha = None
    if world == 1:
        # the C ABI's pipelined host-buffer call: rt_frame_submit(host lights, pinned host frame) / rt_frame_wait
        qopts = capi.FrameQueueOpts(depth, ctas, 0, 0, FLUSH if flush_mode in ("inline", "side") else 0)
        q = C.c_void_p()
        capi.check(lib.rt_frame_queue_create(scene.handle, W, H, C.byref(qopts), C.byref(q)), lib)
        hbuf = C.c_void_p()
        nring = depth + 2
        capi.check(lib.rt_host_alloc(W * H * 4 * nring, C.byref(hbuf)), lib)
        lp = np.ascontiguousarray(host.lights_pos); lc = np.ascontiguousarray(host.lights_color)
        qparams = host.params(spp_grid=spp_grid, flags=capi.RT_FLAG_NO_CLEAR | pflags)
        tick = C.c_longlong()

        def q_submit():
            k = q_submit.n % nring
            if flush_mode == "rotate" and ncopies > 1:       # the next frame renders the next copy of the scene
                capi.check(lib.rt_frame_queue_bind(q, copies[q_submit.n % ncopies].handle), lib)
            capi.check(lib.rt_frame_submit(q, C.byref(cam), lp.ctypes.data_as(capi.Vec4_p), lc.ctypes.data_as(capi.Vec4_p), nl,
                                           C.byref(qparams), C.c_void_p(hbuf.value + k * W * H * 4), C.byref(tick)), lib)
            q_submit.n += 1
            return tick.value, k
        q_submit.n = 0

        def q_done(t):
            ms = C.c_double()
            capi.check(lib.rt_frame_wait(q, t, C.byref(ms)), lib)
            return ms.value
        for _ in range(3):
            q_submit()
        lead = max(2, depth)
        tk = [q_submit() for _ in range(lead + args.steps + depth)]
        t_a = q_done(tk[lead - 1][0])
        t_b = q_done(tk[lead - 1 + args.steps][0])
        q_done(tk[-1][0])
        e2e_ms = (t_b - t_a) / args.steps
        last_slot = tk[lead - 1 + args.steps][1]
        e2e_frame = np.ctypeslib.as_array((C.c_ubyte * (W * H * 4)).from_address(hbuf.value + last_slot * W * H * 4)).copy()
        e2e_frame_sha = sha1(e2e_frame)
        capi.check(lib.rt_frame_queue_destroy(q), lib)
        capi.check(lib.rt_host_free(hbuf), lib)
        e2e_api = ("rt_frame_submit / rt_frame_wait (C ABI): host lights in, pinned host RGBA8 frame out, %d frames in "
                   "flight, D2H on its own stream" % depth)
    else:
        for _ in range(3):
            pipe.submit(host_io=True)
        e2e_total, last_slot = steady(pipe, args.steps, host_io=True)
        e2e_ms = e2e_total / args.steps
        if rank == 0:
            e2e_frame_sha = sha1(np.ascontiguousarray(pipe.host_frame(last_slot)))
        e2e_api = ("rt_render_device per rank, rgba8 = ONE pinned host frame mapped by every rank's process (POSIX shm + "
                   "rt_host_register): each rank's kernel stores its finished pixels over its own PCIe link; pinned-host "
                   "lights in; 4-byte all-reduce completion signal")
    e2e_value = rays_frame / (e2e_ms * 1e-3) / 1e6
    clocks = sampler.result()

    # ---- A/B of the L2 flush modes (N > 1): what the per-frame 160 MiB fill costs -------------------------------
    flush_ab = None
    if not args.no_ab and not big:
        flush_ab = {}
        n_ab = max(10, args.steps // 2)
        for mode in ("rotate", "inline", "side", "off"):
            if mode == flush_mode:
                flush_ab[mode] = ms_per_step
                continue
            pipe.close()
            pipe = make_pipe(mode)
            for _ in range(3):
                pipe.submit()
            flush_ab[mode] = steady(pipe, n_ab, host_io=False)[0] / n_ab
        pipe.close()
        pipe = make_pipe(flush_mode)
        for _ in range(3):
            pipe.submit()
        pipe.drain()

    # ---- parity (untimed): the pipeline's frame against the 1-GPU frame, bitwise; N = 1 also against the oracle ----
    pipe.drain()
    slot = pipe.submit()
    pipe.drain()
    if world > 1:
        dist.barrier()
    parity = None
    full = None
    if rank == 0:
        frame_n = pipe.frame(slot).cpu().numpy().copy()
        want = ("rgba8", "hit_id", "hit_t") if spp_grid == 1 else ("rgba8",)
        full = scene.render(cam, host.lights_pos, host.lights_color, host.params(spp_grid=spp_grid), want=want)
        parity = {"frame_sha1": sha1(np.ascontiguousarray(full["rgba8"])),
                  "frame_equals_1gpu": bool(np.array_equal(frame_n, full["rgba8"])),
                  "e2e_frame_equals_1gpu": e2e_frame_sha == sha1(np.ascontiguousarray(full["rgba8"]))}

    # ---- one frame at a time (latency): serial launches, CUDA events around the kernel ---------------------
    sparams = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_NO_CLEAR | pflags)
    capi.check(lib.rt_scene_set_launch_width(scene.handle, 0), lib)     # one frame alone gets the whole machine
    kern_ms = []
    def evict():
        if flush_buf is not None:
            flush_buf.fill_(1)
    for i in range(5 if big else 8):
        evict()
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
        eparams = host.params(spp_grid=spp_grid, flags=capi.RT_FLAG_NO_CLEAR)
        nrep = 4 if big else 10
        for i in range(3 + nrep):
            if i == 3:
                te = time.perf_counter()
            capi.check(lib.rt_render(scene.handle, C.byref(cam), lp.ctypes.data_as(capi.Vec4_p), lc.ctypes.data_as(capi.Vec4_p),
                                     nl, C.byref(eparams), C.byref(hfo)), lib)
        sync_ms = (time.perf_counter() - te) * 1e3 / nrep
        sync_e2e = {"value": rays_frame / (sync_ms * 1e-3) / 1e6, "ms_per_step": sync_ms, "api": "rt_render, one frame at a time"}

    # ---- A/B: the same frame with every ray on the reference's median-split topology (N = 1, serial frames) ----
    other_tree = None
    if world == 1 and not args.no_ab and not big:
        other_flags = 0 if scene_flags else capi.RT_SCENE_REFERENCE_TOPOLOGY
        sc2 = rt.DeviceScene(host, scene_flags=other_flags)
        ms2 = []
        for i in range(8):
            evict()
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
        bparams = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_NO_CLEAR | capi.RT_FLAG_BINARY_NODES)
        ms3 = []
        for i in range(4 if big else 8):
            evict()
            capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                            C.byref(bparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
            ms3.append(scene.last_kernel_ms())
        k3 = sum(ms3[2:]) / len(ms3[2:])
        binary_nodes = {"kernel_ms_single_frame": k3, "mrays": rays_frame / (k3 * 1e-3) / 1e6,
                        "note": "RT_FLAG_BINARY_NODES, same frame bit for bit; compare with latency_ms_single_frame"}

    # ---- A/B: ordinary rays on the 8-wide twin of the triangle tree (N = 1, serial frames) ------------------------------
    nodes8 = None
    if world == 1 and not args.no_ab and info.get("num_nodes_wide8", 0) > 0:
        p8 = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_NO_CLEAR | capi.RT_FLAG_NODES8 | capi.RT_FLAG_MEGAKERNEL)
        ms8 = []
        for i in range(4 if big else 8):
            evict()
            capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                            C.byref(p8), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
            ms8.append(scene.last_kernel_ms())
        k8 = sum(ms8[2:]) / len(ms8[2:])
        c8p = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_COUNTERS | capi.RT_FLAG_NODES8 | capi.RT_FLAG_MEGAKERNEL)
        capi.check(lib.rt_render_device(scene.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                        C.byref(c8p), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
        c8 = capi.Counters()
        capi.check(lib.rt_get_counters(scene.handle, C.byref(c8)), lib)
        c8 = c8.as_dict()
        f8 = 24 * c8["box_tests"] + 48 * c8["tri_filter"] + (60 * nl + 20) * c8["shaded_hits"]
        nodes8 = {"kernel_ms_single_frame": k8, "mrays": rays_frame / (k8 * 1e-3) / 1e6, "box_tests": c8["box_tests"],
                  "nodes_visited": c8["nodes_visited"], "flops_per_launch": f8,
                  "fp32_frac_single_frame": f8 / (k8 * 1e-3) / 1e12 / (148 * 128 * 2 * float(measured_peaks().get("sm_max_mhz", 1965.0)) * 1e6 / 1e12),
                  "note": "RT_FLAG_NODES8 (256-byte eight-child records), same frame bit for bit; compare with "
                          "latency_ms_single_frame / roofline.frac_single_frame"}

    # ---- A/B: the traversal tree built by the host threads (round 1's builder) instead of the GPU builder --------------
    host_tree = None
    if world == 1 and not args.no_ab and not big and info.get("sah_builder", 0) == 2:
        sc3 = rt.DeviceScene(host, scene_flags=scene_flags | capi.RT_SCENE_HOST_SAH)
        ms4 = []
        for i in range(8):
            evict()
            capi.check(lib.rt_render_device(sc3.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                            C.byref(sparams), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
            ms4.append(sc3.last_kernel_ms())
        k4 = sum(ms4[2:]) / len(ms4[2:])
        hp = host.params(spp_grid=spp_grid, tile=pipe.tile, flags=capi.RT_FLAG_COUNTERS)
        capi.check(lib.rt_render_device(sc3.handle, C.byref(cam), pipe.d_lpos.data_ptr(), pipe.d_lcol.data_ptr(), nl,
                                        C.byref(hp), C.byref(fo), C.c_void_p(stream.cuda_stream)), lib)
        c4 = capi.Counters()
        capi.check(lib.rt_get_counters(sc3.handle, C.byref(c4)), lib)
        c4 = c4.as_dict()
        f4 = 24 * c4["box_tests"] + 48 * c4["tri_filter"] + (60 * nl + 20) * c4["shaded_hits"]
        i3 = sc3.info()
        host_tree = {"kernel_ms_single_frame": k4, "mrays": rays_frame / (k4 * 1e-3) / 1e6, "box_tests": c4["box_tests"],
                     "nodes_visited": c4["nodes_visited"], "flops_per_launch": f4,
                     "fp32_frac_single_frame": f4 / (k4 * 1e-3) / 1e12 / (148 * 128 * 2 * float(measured_peaks().get("sm_max_mhz", 1965.0)) * 1e6 / 1e12),
                     "sah_build_ms": i3["sah_build_ms"], "flatten_ms": i3["flatten_ms"],
                     "note": "RT_SCENE_HOST_SAH: the host builder's tree (32 bins / exact sweep), same frame bit for bit; the GPU "
                             "builder's tree needs fewer box tests per ray, so the same kernel renders the frame faster at a "
                             "LOWER fraction of the FP32 roofline (fewer counted flops); compare with latency_ms_single_frame / "
                             "roofline.frac_single_frame"}
        sc3.close()

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
    traffic, traffic_file = ncu_traffic(args.workload)
    standard = (world == 1 and (W, H) == (wl["width"], wl["height"]) and args.topology == "sah" and wide and not args.objects
                and pipeline_ran == "megakernel")
    roof = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
            # DRAM bytes per launch of this kernel on this workload, read from the committed `ncu --set full` summary
            # (dram__bytes_read.sum + dram__bytes_write.sum); only quoted for the configuration that capture was taken on
            "traffic": traffic if standard else None,
            "traffic_unit": "bytes per launch (ncu --set full, %s)" % traffic_file,
            "kernel": ("rtb::wf_trace_kernel + rtb::wf_shade_kernel" if pipeline_ran == "wavefront" else
                       "rtb::render_kernel<4>" if wide else "rtb::render_kernel<0>"),
            "timed_over": "flops of one frame / ms_per_step of the steady-state window (kernels of successive frames overlap, "
                          "so the per-launch duration is not the divisor; flush and launch gaps are inside)",
            "kernel_ms_single_frame": max(rank_ms), "kernel_ms_per_rank": rank_ms,
            "frac_single_frame": flops / (max(rank_ms) * 1e-3) / 1e12 / peak_tflops,
            "peak_source": "148 SMs x 128 FP32 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json (FP32 is not in the "
                           "driver-measured set; the scene is L2-resident so HBM does not bound it)",
            "flops_per_launch": flops, "algorithmic_bytes_per_launch": alg_bytes,
            "mem_GBps": alg_bytes / (ms_per_step * 1e-3) / 1e9, "hbm_peak_GBps": peaks.get("hbm_gbs"),
            "north_star": {"definition": "slower of FP32 peak for the counted intersection flops and L2 bandwidth for the "
                                         "counted node/triangle bytes (scene resident in L2)",
                           "l2_read_GBps_measured": l2_bw, "l2_how": "torch.sum over an L2-resident 48 MiB buffer, best of 3 x 20",
                           "t_fp32_ms": t_fp32 * 1e3, "t_l2_ms": t_l2 * 1e3,
                           "frac": max(t_fp32, t_l2) * 1e3 / ms_per_step,
                           "note": "node/triangle bytes are served by L1 (91 % hits), so this fraction can exceed 1; "
                                   "roofline.frac (FP32 issue) is the harder, reported bound"}}

    rc = 0
    if rank == 0:
        line = {
            "metric": wl["metric"], "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32 filter + f64 decide/shade", "data": "synthetic",
            "config": {"workload": args.workload, "pipeline": pipeline_ran,
                       "triangles": info["num_triangles"], "bvh_nodes": info["num_nodes_ref"],
                       "traversal_tree": ("reference median-split topology" if info["topology"] == 1 else
                                          "library-built SAH tree over the same triangles; reference tree kept for d.x == 0 rays"),
                       "traversal_records": info["num_nodes_traversal"],
                       "node_width": 4 if wide else 2, "wide_records": info.get("num_nodes_wide", 0),
                       "lights": nl, "max_depth": desc.max_depth, "spp": spp_grid * spp_grid,
                       "l2": {"rotate": "inputs larger than L2: %d identical copies of the flattened scene (%.0f MB, L2 %.0f MB), "
                                        "frame f renders copy f %% %d, so every frame starts with its scene out of cache; no "
                                        "fill inside the timed region (flush_ab: the 160 MiB fill in-line / on a side stream, and "
                                        "no eviction)" % (ncopies, ncopies * blob_bytes / 1e6, l2_bytes / 1e6, ncopies),
                              "big": "inputs larger than L2: the flattened scene is %.0f MB (L2 %.0f MB); no flush" % (blob_bytes / 1e6, l2_bytes / 1e6),
                              "inline": "flushed before every timed frame (160 MiB fill on the frame's stream)",
                              "side": "flushed once per timed frame: 160 MiB fill on a side stream, paced by the frames",
                              "off": "NOT evicted (experiment)"}[flush_mode],
                       "timing": "steady-state window: value is THROUGHPUT with frames_in_flight frames in flight "
                                 "(one frame alone: latency_ms_single_frame)",
                       "warmup_frames_run": n_warm, "frames_in_flight": pipe.depth, "ctas_per_sm_per_frame": pipe.ctas_per_sm or "all",
                       "parallelism": f"tiles{world}" if world > 1 else "single",
                       "tile_chunk": pipe.tile[2], "dyn_tiles": pipe.dyn_tiles, "dyn_chunk": pipe.dyn_chunk,
                       "gather": "fused: peer stores into rank 0's frame buffer over NVLink + 4-byte all-reduce signal"
                                 if world > 1 else "none"},
            "mrays_primary": tot["rays_primary"] / (ms_per_step * 1e-3) / 1e6,
            "latency_ms_single_frame": max(rank_ms),
            "other_tree": other_tree,
            "binary_nodes": binary_nodes,
            "nodes8": nodes8,
            "host_sah_tree": host_tree,
            "flush_ab": flush_ab,
            "rays_per_frame": {k: tot[k] for k in ("rays_primary", "rays_shadow", "rays_reflect")},
            "counters": tot,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(2 * nl * 32) * world, "d2h_bytes_per_step": W * H * 4,
                    "api": e2e_api, "sync": sync_e2e},
            "gpu_launches": args.steps * world * (1 if pipeline_ran == "megakernel" else 2 * (desc.max_depth + 2) * spp_grid * spp_grid),
            "roofline": roof,
            "setup_ms": {"host_scene_build": host_ms, "flatten": info["flatten_ms"], "upload": info["upload_ms"],
                         "flattener": {0: "host threads (rt_flatten.cpp)", 1: "GPU (rt_flatten_gpu.cu)"}.get(info.get("flattener", 0)),
                         "sah_builder": {0: "none", 1: "host threads", 2: "GPU (rt_sah_gpu.cu)"}.get(info.get("sah_builder", 0)),
                         "sah_build": info.get("sah_build_ms"), "sah_build_device_kernels": info.get("sah_device_ms"),
                         "broadcast": bcast_ms, "blob_bytes": info["blob_bytes"],
                         "broadcast_how": "NCCL from rank 0's resident blob (device to device)" if world > 1 else None},
        }
        if not args.no_cpu_baseline and world == 1:
            # the CPU leg renders rows of the SAME frame: they are also the oracle side of the parity block
            rec, rows, step, oracle = cpu_leg(desc, args.cpu_seconds, spp_grid=spp_grid, want_ids=True)
            line["cpu_baseline"] = {k: v for k, v in rec.items() if k not in ("seconds", "rays", "rows")}
            ys = np.arange(0, H, step)
            d = np.abs(rt.quantize_rgb8(rows["rgb"][ys]).astype(int) - full["rgba8"][ys][..., :3].astype(int))
            parity.update(oracle=line["cpu_baseline"]["kind"], checked_rows=int(len(ys)), rows="every %dth row" % step,
                          pix_within_1lsb=float((d <= 1).mean()), pix_max_lsb=int(d.max()))
            if spp_grid == 1:
                parity.update(ids_equal=float((rows["ids"][ys] == full["hit_id"][ys]).mean()),
                              t_equal=float((rows["t"][ys] == full["hit_t"][ys]).mean()))
            else:
                # hit ids / ray parameters are defined for the centre ray: a second, smaller oracle pass at 1 spp
                full1 = scene.render(cam, host.lights_pos, host.lights_color, host.params(), want=("hit_id", "hit_t"))
                r1 = oracle.render(want_ids=True, y0=1, ystep=max(step * 8, 1), spp_grid=1, nthreads=os.cpu_count() or 1)
                y1 = np.arange(1, H, max(step * 8, 1))
                parity.update(ids_equal=float((r1["ids"][y1] == full1["hit_id"][y1]).mean()),
                              t_equal=float((r1["t"][y1] == full1["hit_t"][y1]).mean()), ids_rows=int(len(y1)))
            ok = parity["pix_within_1lsb"] >= 0.999 and parity["ids_equal"] == 1.0
            parity["ok"] = bool(ok and parity["frame_equals_1gpu"] and parity["e2e_frame_equals_1gpu"])
        else:
            parity["ok"] = bool(parity["frame_equals_1gpu"] and parity["e2e_frame_equals_1gpu"])
        line["parity"] = parity
        print(json.dumps(line))
        if not parity["ok"]:
            sys.stderr.write("bench.py: PARITY MISMATCH %s\n" % json.dumps(parity))
            rc = 3
    pipe.close()
    if shared is not None:
        shared.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rc


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
        sys.exit(0)
    sys.exit(run_b200(a) or 0)
