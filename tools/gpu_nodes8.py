"""A/B of the traversal-tree widths on one GPU: 4-wide records (default) vs 8-wide (RT_FLAG_NODES8) vs binary, one frame at
a time, CUDA-event kernel time, min / median of N; frames must be bit-identical; counters per ray.

  python tools/gpu_nodes8.py [scene ...] [--big N_OBJECTS]     # default scenes: office rings textured at BASELINE sizes
  RT_B200_LIB=.../variants/pf.so python tools/gpu_nodes8.py --big 10000        # a kernel-variant build on the 10 M scene
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi

SIZES = {"office": dict(width=1920, height=1080), "rings": dict(width=700, height=500), "textured": dict(width=1920, height=1080)}
argv = sys.argv[1:]
big = int(argv[argv.index("--big") + 1]) if "--big" in argv else 0
names = [a for a in argv if a in SIZES] or ([] if big else list(SIZES))
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
jobs = [(n, rt.scenes.REGISTRY[n](**SIZES[n]), 1, False) for n in names]
if big:
    jobs.append(("scatter_%dk_4k_4spp" % big, rt.scenes.scatter(width=3840, height=2160, n_objects=big), 2, True))
out = {"lib": os.environ.get("RT_B200_LIB", "default")}
for name, desc, spp, gpu_build in jobs:
    host = rt.HostScene(desc, lib, gpu_build=gpu_build)
    dev = rt.DeviceScene(host)
    res = {"info": dev.info()}
    reps = 4 if big and gpu_build else 10
    for label, flags in (("wide4", 0), ("wide8", capi.RT_FLAG_NODES8), ("binary", capi.RT_FLAG_BINARY_NODES)):
        ms = []
        for i in range(reps + 2):
            r = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp, flags=flags), want=("rgba8",))
            ms.append(dev.last_kernel_ms())
        c = dev.render(host.camera, host.lights_pos, host.lights_color,
                       host.params(spp_grid=spp, flags=flags | capi.RT_FLAG_COUNTERS), want=("rgba8",))["counters"]
        rays = c["rays_primary"] + c["rays_shadow"] + c["rays_reflect"]
        ms = sorted(ms[2:])
        flops = 24 * c["box_tests"] + 48 * c["tri_filter"] + (60 * len(desc.lights) + 20) * c["shaded_hits"]
        res[label] = {"ms_min": ms[0], "ms_median": ms[len(ms) // 2], "sha1": hashlib.sha1(r["rgba8"].tobytes()).hexdigest(),
                      "visits_per_ray": c["nodes_visited"] / rays, "box_per_ray": c["box_tests"] / rays,
                      "tri_per_ray": c["tri_filter"] / rays, "grays_per_s": rays / ms[len(ms) // 2] / 1e6,
                      "fp32_frac": flops / (ms[len(ms) // 2] * 1e-3) / 74.45e12, "nodes8_visited": c["nodes8_visited"]}
    res["identical"] = len({res[k]["sha1"] for k in ("wide4", "wide8", "binary")}) == 1
    out[name] = res
    print(name, json.dumps(res), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
tag = os.path.basename(os.environ.get("RT_B200_LIB", "default")).replace(".so", "")
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "nodes8_ab_%s%s.json" % (tag, "_big" if big else "")), "w"), indent=1)
if not all(v["identical"] for k, v in out.items() if isinstance(v, dict)):
    raise SystemExit("frames differ between tree widths")
