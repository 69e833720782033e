"""Profile target: N frames of one scene through one pipeline (for ncu launch lists / --set full captures).
  RT_PIPE=wavefront|megakernel RT_SCENE=office RT_FRAMES=3 python tools/gpu_profile_wf.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi

lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
name = os.environ.get("RT_SCENE", "office")
host = rt.HostScene(rt.scenes.REGISTRY[name](), lib)
dev = rt.DeviceScene(host)
flags = capi.RT_FLAG_WAVEFRONT if os.environ.get("RT_PIPE", "wavefront") == "wavefront" else capi.RT_FLAG_MEGAKERNEL
for i in range(int(os.environ.get("RT_FRAMES", "3"))):
    dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=flags), want=("rgba8",))
    print(dev.last_pipeline(), dev.last_kernel_ms(), flush=True)
