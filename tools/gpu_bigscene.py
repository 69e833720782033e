"""BASELINE config 5 (scaling scene): n_objects x 1000-triangle spheres (10 M triangles at the default) rendered at
3840x2160 with a 2x2 sub-pixel grid.  Prints build / flatten / render timings, the GPU-vs-host BVH build comparison,
counters, and size-independent parity properties (filter == exact-only on a crop, determinism, oracle on sampled rows).

  python tools/gpu_bigscene.py [n_objects=10000] [width=3840] [height=2160] [spp_grid=2] [--oracle-rows K] [--skip-host-build]
"""
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

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n_obj = int(args[0]) if len(args) > 0 else 10000
W = int(args[1]) if len(args) > 1 else 3840
H = int(args[2]) if len(args) > 2 else 2160
spp = int(args[3]) if len(args) > 3 else 2
oracle_rows = int(sys.argv[sys.argv.index("--oracle-rows") + 1]) if "--oracle-rows" in sys.argv else 0
skip_host = "--skip-host-build" in sys.argv
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
out = {"n_objects": n_obj, "width": W, "height": H, "spp_grid": spp}
t = time.time(); desc = rt.scenes.scatter(width=W, height=H, n_objects=n_obj); out["scene_gen_s"] = time.time() - t
t = time.time(); hg = rt.HostScene(desc, lib, gpu_build=True); out["host_scene_gpu_build_s"] = time.time() - t
out["triangles"] = hg.num_triangles; out["bvh_nodes"] = hg.nodes_used; out["bvh_gpu_device_ms"] = hg.bvh_device_ms
print(json.dumps(out), flush=True)
if not skip_host:
    t = time.time(); hc = rt.HostScene(desc, lib); out["host_scene_cpu_build_s"] = time.time() - t
    same = hc.nodes_used == hg.nodes_used and all(np.array_equal(hc.bvh_arrays[k][:hc.nodes_used], hg.bvh_arrays[k][:hc.nodes_used])
                                                   for k in ("left", "first", "count", "bbmin", "bbmax"))
    same = same and np.array_equal(hc.arrays["vertexIdx"], hg.arrays["vertexIdx"])
    out["gpu_build_equals_host_build"] = bool(same)
    del hc
    print(json.dumps({k: out[k] for k in ("host_scene_cpu_build_s", "gpu_build_equals_host_build")}), flush=True)
host = hg
t = time.time(); dev = rt.DeviceScene(host); out["flatten_upload_s"] = time.time() - t
out["info"] = dev.info()
print(json.dumps(out["info"]), flush=True)
p = host.params(spp_grid=spp)
ms = []
for i in range(4):
    r = dev.render(host.camera, host.lights_pos, host.lights_color, p, want=("rgba8",))
    ms.append(dev.last_kernel_ms())
first = r["rgba8"]
if out["info"].get("num_nodes_wide", 0) > 0:   # A/B: the same frame on the binary records of the traversal tree
    msb = []
    for i in range(3):
        rb = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp, flags=capi.RT_FLAG_BINARY_NODES), want=("rgba8",))
        msb.append(dev.last_kernel_ms())
    assert np.array_equal(rb["rgba8"], first)
    out["kernel_ms_binary_nodes"] = min(msb[1:])
c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp, flags=capi.RT_FLAG_COUNTERS), want=("rgba8",))
cnt = c["counters"]
assert np.array_equal(c["rgba8"], first)                               # determinism / instrumented == plain
rays = cnt["rays_primary"] + cnt["rays_shadow"] + cnt["rays_reflect"]
out.update(kernel_ms=min(ms[1:]), kernel_ms_all=ms, counters=cnt, mrays_per_s=rays / (min(ms[1:]) * 1e-3) / 1e6,
           alg_bytes=32 * cnt["box_tests"] + 48 * cnt["tri_filter"] + 64 * cnt["shaded_hits"],
           flops=24 * cnt["box_tests"] + 48 * cnt["tri_filter"] + 140 * cnt["shaded_hits"])
out["alg_GBps"] = out["alg_bytes"] / (out["kernel_ms"] * 1e-3) / 1e9
out["fp32_TFLOPs"] = out["flops"] / (out["kernel_ms"] * 1e-3) / 1e12
print(json.dumps({k: out.get(k) for k in ("kernel_ms", "kernel_ms_binary_nodes", "kernel_ms_all", "mrays_per_s", "alg_GBps", "fp32_TFLOPs", "counters")}), flush=True)
# properties at full scene size on a reduced frame: filter == exact-only, ids/t against the oracle on sampled rows
small = rt.scenes.scatter(width=480, height=270, n_objects=n_obj)
hs = host
cam_small = rt.scene.look_at(small.camera, lib) if hasattr(rt, "scene") else None
from my_raytracer_b200.scene import look_at
cam_small = look_at(small.camera, lib)
a = dev.render(cam_small, host.lights_pos, host.lights_color, host.params())
b = dev.render(cam_small, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_EXACT_ONLY))
out["filter_equals_exact_only"] = bool(all(np.array_equal(a[k], b[k]) for k in ("rgba8", "hit_id", "hit_t", "pixels")))
print(json.dumps({"filter_equals_exact_only": out["filter_equals_exact_only"]}), flush=True)
if oracle_rows:
    import oracle_py as orc
    kind = "reference" if orc.available("reference") else "port"
    t = time.time(); o = orc.OracleScene(small, kind); out["oracle_build_s"] = time.time() - t
    step = max(1, 270 // oracle_rows)
    ref = o.render(y0=1, ystep=step)
    rows = np.arange(1, 270, step)
    d = np.abs(rt.quantize_rgb8(ref["rgb"][rows]).astype(int) - a["rgba8"][rows][..., :3].astype(int))
    out["oracle"] = dict(kind=kind, rows=len(rows), ids_equal=float((ref["ids"][rows] == a["hit_id"][rows]).mean()),
                         t_equal=float((ref["t"][rows] == a["hit_t"][rows]).mean()), pix_max=int(d.max()),
                         within1=float((d <= 1).mean()), seconds=ref["seconds"])
    print(json.dumps(out["oracle"]), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "bigscene_%d.json" % n_obj), "w"))
print("DONE", flush=True)
