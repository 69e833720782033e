"""Renders RT_SCENE (default office, 1080p) a few times with the default parameters: the target of ncu captures
(no torch import, so the capture starts within seconds).  Usage: ncu ... python tools/gpu_profile_target.py [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
big = int(os.environ.get("RT_BIG", "0"))       # RT_BIG=10000: the 10 M-triangle scene of BASELINE config 5 (3840x2160, 2x2 spp)
spp = 2 if big else 1
desc = rt.scenes.scatter(width=3840, height=2160, n_objects=big) if big else rt.scenes.REGISTRY[os.environ.get("RT_SCENE", "office")]()
host = rt.HostScene(desc, lib, gpu_build=bool(big))
dev = rt.DeviceScene(host)
fl = int(os.environ.get("RT_VARIANT_FLAGS", "0"))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=fl, spp_grid=spp), want=("rgba8",))
    print(i, dev.last_kernel_ms(), flush=True)
