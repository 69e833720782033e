"""A/B: the persistent kernel against the ray-parallel persistent kernel (RT_FLAG_RAY_PARALLEL) on the BASELINE-size scenes:
kernel time (median of 10) and a bit-for-bit comparison of the frames.  python tools/gpu_rayparallel.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
JOBS = [("spheres", dict(width=500, height=500)), ("molecule", dict(width=500, height=500)), ("rings", dict(width=700, height=500)),
        ("office", dict(width=640, height=360)), ("office", dict(width=1920, height=1080)), ("textured", dict(width=1920, height=1080)),
        ("office", dict(width=480, height=270)), ("office", dict(width=800, height=450)), ("office", dict(width=960, height=540)),
        ("rings", dict(width=350, height=250)), ("rings", dict(width=1400, height=1000)), ("textured", dict(width=640, height=360)),
        ("mixed_small", dict(width=640, height=480))]
if os.environ.get("RT_JOBS"):
    JOBS = [j for j in JOBS if ("%s:%d" % (j[0], j[1]["width"])) in os.environ["RT_JOBS"].split(",")]
for name, kw in JOBS:
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    args = (host.camera, host.lights_pos, host.lights_color)
    res = {}
    for label, fl in (("persistent", capi.RT_FLAG_MEGAKERNEL), ("rayparallel", capi.RT_FLAG_RAY_PARALLEL), ("auto", 0)):
        ms = []
        for i in range(12):
            out = dev.render(*args, host.params(flags=fl), want=("rgba8", "pixels", "hit_id", "hit_t"))
            ms.append(dev.last_kernel_ms())
        res[label] = (sorted(ms[2:])[5], out, dev.last_pipeline())
    a, b = res["persistent"][1], res["rayparallel"][1]
    same = all(np.array_equal(a[k], b[k]) for k in a)
    rays = dev.render(*args, host.params(flags=capi.RT_FLAG_COUNTERS), want=("rgba8",))["counters"]
    nr = rays["rays_primary"] + rays["rays_shadow"] + rays["rays_reflect"]
    print("%-9s %4dx%-4d lights %d  persistent %.4f ms (%.2f Grays/s)  rayparallel %.4f ms (%.2f Grays/s) [%s]  identical %s" % (
        name, kw["width"], kw["height"], host.lights_pos.shape[0], res["persistent"][0], nr / res["persistent"][0] / 1e6,
        res["rayparallel"][0], nr / res["rayparallel"][0] / 1e6, res["rayparallel"][2], same),
        "| default picks", res["auto"][2], "%.4f ms" % res["auto"][0], flush=True)
