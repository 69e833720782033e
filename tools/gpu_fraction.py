"""Kernel time vs fraction of the frame rendered (1 GPU): exposes the latency-bound drain of the last pixels."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
host = rt.HostScene(rt.scenes.office(), lib)
dev = rt.DeviceScene(host)
ntiles = 240 * 270
for world, ranks in ((1, [0]), (64, [0, 13, 40]), (1024, [0, 300, 700]), (ntiles, [100, 20000, 30000, 40000, 50000, 64000])):
    for r in ranks:
        ms = []
        for i in range(6):
            dev.render(host.camera, host.lights_pos, host.lights_color, host.params(tile=(r, world, 1)), want=("rgba8",))
            ms.append(dev.last_kernel_ms())
        c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(tile=(r, world, 1), flags=capi.RT_FLAG_COUNTERS), want=("rgba8",))["counters"]
        rays = c["rays_primary"] + c["rays_shadow"] + c["rays_reflect"]
        print("1/%d rank %d: kernel ms min %.3f | rays %d nodes %d tri %d exact %d shaded %d" % (world, r, min(ms[1:]), rays, c["nodes_visited"], c["tri_filter"], c["tri_exact"], c["shaded_hits"]), flush=True)
