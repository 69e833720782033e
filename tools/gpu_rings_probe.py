"""Where does the rings frame (BASELINE config 3, 700x500) spend its 0.64 ms?  Kernel time by mirror depth, by light count and
by horizontal band of tile rows (a band rendered alone through the tile-share parameters).  python tools/gpu_rings_probe.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
name = os.environ.get("RT_SCENE", "rings")
host = rt.HostScene(rt.scenes.REGISTRY[name](), lib)
dev = rt.DeviceScene(host)
W, H = host.camera.width, host.camera.height


def ms(params, lights=None, n=6):
    lp, lc = (host.lights_pos, host.lights_color) if lights is None else (host.lights_pos[:lights], host.lights_color[:lights])
    t = []
    for _ in range(n):
        dev.render(host.camera, lp, lc, params, want=("rgba8",))
        t.append(dev.last_kernel_ms())
    return min(t[1:])


print("whole frame", ms(host.params()))
for d in range(0, 6):
    print("max_depth", d, ms(host.params(max_depth=d)))
for l in (0, 1, 2):
    print("lights", l, ms(host.params(), lights=l))
tx, ty = (W + 7) // 8, (H + 3) // 4
rows = 5
bands = (ty + rows - 1) // rows
for b in range(bands):
    p = host.params(tile=(b, bands, tx * rows))
    c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(tile=(b, bands, tx * rows), flags=capi.RT_FLAG_COUNTERS), want=("rgba8",))["counters"]
    print("band", b, "pixel rows", b * rows * 4, "ms", round(ms(p), 4), "rays", c["rays_primary"] + c["rays_shadow"] + c["rays_reflect"], "nodes", c["nodes_visited"], "exact", c["tri_exact"], "filter", c["tri_filter"])
