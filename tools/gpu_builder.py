"""Timing of the scene preparation: records and traversal tree on the GPU (default), records on the host threads with the
GPU-built tree (RT_SCENE_HOST_FLATTEN), everything on the host threads (RT_SCENE_HOST_SAH); flatten and upload, the render
time on each blob and a checksum of the frame.  python tools/gpu_builder.py [--big N_OBJECTS]"""
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi

lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
argv = sys.argv[1:]
big = int(argv[argv.index("--big") + 1]) if "--big" in argv else 0
jobs = [("office", rt.scenes.office(), 1, False), ("textured", rt.scenes.textured(), 1, False)]
if big:
    jobs = [("scatter_%dk" % big, rt.scenes.scatter(width=3840, height=2160, n_objects=big), 2, True)]
out = {}
for name, desc, spp, gpu_build in jobs:
    host = rt.HostScene(desc, lib, gpu_build=gpu_build)
    res = {}
    for label, fl in (("device", 0), ("host_flatten", capi.RT_SCENE_HOST_FLATTEN), ("host", capi.RT_SCENE_HOST_SAH), ("device_again", 0)):
        t = time.perf_counter()
        dev = rt.DeviceScene(host, scene_flags=fl)
        wall = (time.perf_counter() - t) * 1e3
        info = dev.info()
        ms = []
        for i in range(5):
            r = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=spp), want=("rgba8",))
            ms.append(dev.last_kernel_ms())
        res[label] = {"create_wall_ms": wall, "flatten_ms": info["flatten_ms"], "upload_ms": info["upload_ms"],
                      "flattener": info["flattener"], "frame_sha1": hashlib.sha1(r["rgba8"].tobytes()).hexdigest()[:12],
                      "sah_builder": info["sah_builder"], "sah_build_ms": info["sah_build_ms"], "sah_device_ms": info["sah_device_ms"],
                      "kernel_ms": min(ms[1:]), "nodes": info["num_nodes_traversal"], "wide": info["num_nodes_wide"],
                      "wide8": info["num_nodes_wide8"], "depth": info["max_depth_bvh"]}
        dev.close()
    out[name] = res
    print(name, json.dumps(res), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "builder%s.json" % ("_big" if big else "")), "w"), indent=1)
