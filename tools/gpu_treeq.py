"""Tree quality of a builder variant (RT_B200_LIB=variants/sah_*.so or the default library): kernel time on the tree it
builds for Office / textured / rings at BASELINE sizes (and --big N_OBJECTS), node visits and triangle tests per ray."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi

lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
argv = sys.argv[1:]
big = int(argv[argv.index("--big") + 1]) if "--big" in argv else 0
flags = capi.RT_SCENE_HOST_SAH if "--host" in argv else 0
jobs = [("office", rt.scenes.office(), 1, False), ("textured", rt.scenes.textured(), 1, False), ("rings", rt.scenes.rings(), 1, False)]
if big:
    jobs = [("scatter_%dk" % big, rt.scenes.scatter(width=3840, height=2160, n_objects=big), 2, True)]
tag = os.path.basename(os.environ.get("RT_B200_LIB", "default")).replace(".so", "") + ("_host" if flags else "")
row = {"lib": tag}
for name, desc, spp, gb in jobs:
    host = rt.HostScene(desc, lib, gpu_build=gb)
    dev = rt.DeviceScene(host, scene_flags=flags)
    info = dev.info()
    ms = []
    for i in range(4 if big else 10):
        dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp), want=("rgba8",))
        ms.append(dev.last_kernel_ms())
    c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp, flags=capi.RT_FLAG_COUNTERS),
                   want=("rgba8",))["counters"]
    rays = c["rays_primary"] + c["rays_shadow"] + c["rays_reflect"]
    row[name] = {"ms": round(min(ms[2:]), 4), "visits": round(c["nodes_visited"] / rays, 2), "box": round(c["box_tests"] / rays, 1),
                 "tri": round(c["tri_filter"] / rays, 2), "sah_ms": round(info["sah_build_ms"], 1),
                 "sah_dev_ms": round(info["sah_device_ms"], 2), "wide": info["num_nodes_wide"], "depth": info["max_depth_bvh"]}
    dev.close()
print(json.dumps(row), flush=True)
