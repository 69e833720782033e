"""BASELINE.json configs 1-4 at their full sizes on one GPU: kernel ms, Mrays/s, the reference CPU path on a row sample
(time + parity of ids / pixels).  Writes gpurun_out/configs.json and prints one JSON line per config."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi
import oracle_py as orc

lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
kind = next(k for k in ("reference_fast", "reference", "port") if orc.available(k))
cores = os.cpu_count() or 1
res = {}
for name in ("spheres", "office", "molecule", "rings", "textured"):
    desc = rt.scenes.REGISTRY[name]()
    W, H = int(desc.camera["width"]), int(desc.camera["height"])
    t = time.time(); host = rt.HostScene(desc, lib, gpu_build=True); build_s = time.time() - t
    dev = rt.DeviceScene(host)
    ms = []
    for i in range(6):
        out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(), want=("rgba8", "hit_id", "hit_t"))
        ms.append(dev.last_kernel_ms())
    c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_COUNTERS), want=("rgba8",))["counters"]
    rays = c["rays_primary"] + c["rays_shadow"] + c["rays_reflect"]
    o = orc.OracleScene(desc, kind)
    step = max(1, H // 40)
    ref = o.render(y0=0, ystep=step, nthreads=cores)
    rows = np.arange(0, H, step)
    d = np.abs(rt.quantize_rgb8(ref["rgb"][rows]).astype(int) - out["rgba8"][rows][..., :3].astype(int))
    ids_equal = float((ref["ids"][rows] == out["hit_id"][rows]).mean())
    cpu_rays = sum(ref["rays"].values())
    r = dict(config=name, width=W, height=H, triangles=host.num_triangles, objects=host.num_objects, lights=len(desc.lights),
             max_depth=desc.max_depth, kernel_ms=min(ms[1:]), rays=rays, mrays_all=rays / (min(ms[1:]) * 1e-3) / 1e6,
             mrays_primary=c["rays_primary"] / (min(ms[1:]) * 1e-3) / 1e6,
             fp32_frac=(24 * c["box_tests"] + 48 * c["tri_filter"] + (60 * len(desc.lights) + 20) * c["shaded_hits"]) / (min(ms[1:]) * 1e-3) / 74.44992e12,
             cpu=dict(kind=kind, cores=cores, rows=len(rows), seconds=ref["seconds"], mrays=cpu_rays / ref["seconds"] / 1e6,
                      ms_per_frame_extrapolated=ref["seconds"] * 1e3 * H / len(rows)),
             parity=dict(ids_equal=ids_equal, t_equal=float((ref["t"][rows] == out["hit_t"][rows]).mean()), pix_max=int(d.max()),
                         within1=float((d <= 1).mean())), host_build_s=build_s, counters=c)
    res[name] = r
    print(json.dumps(r), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w"), indent=1)
