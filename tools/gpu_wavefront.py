"""A/B of the two pipelines on one GPU: the persistent megakernel vs the wavefront (RT_FLAG_WAVEFRONT), one frame at a
time, CUDA-event time of the whole frame (rt_last_kernel_ms), min / median of N; frames must be bit-identical.

  python tools/gpu_wavefront.py [scene ...]     # default: office spheres molecule rings textured at BASELINE sizes
  RT_B200_LIB=/path/to/variant.so python tools/gpu_wavefront.py office      # a kernel-variant build
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi

SIZES = {"office": dict(width=1920, height=1080), "spheres": dict(width=500, height=500), "molecule": dict(width=500, height=500),
         "rings": dict(width=700, height=500), "textured": dict(width=1920, height=1080)}
names = [a for a in sys.argv[1:] if not a.startswith("-")] or list(SIZES)
reps = 10
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
out = {"lib": os.environ.get("RT_B200_LIB", "default")}
for name in names:
    host = rt.HostScene(rt.scenes.REGISTRY[name](**SIZES[name]), lib)
    dev = rt.DeviceScene(host)
    res = {}
    for label, flags in (("megakernel", capi.RT_FLAG_MEGAKERNEL), ("wavefront", capi.RT_FLAG_WAVEFRONT)):
        ms = []
        for i in range(reps + 2):
            r = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=flags), want=("rgba8",))
            ms.append(dev.last_kernel_ms())
        assert dev.last_pipeline() == label, (dev.last_pipeline(), label)
        ms = sorted(ms[2:])
        res[label] = {"ms_min": ms[0], "ms_median": ms[len(ms) // 2], "sha1": hashlib.sha1(r["rgba8"].tobytes()).hexdigest()}
    res["identical"] = res["megakernel"]["sha1"] == res["wavefront"]["sha1"]
    res["speedup_median"] = res["megakernel"]["ms_median"] / res["wavefront"]["ms_median"]
    out[name] = res
    print(name, json.dumps(res), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
tag = os.path.basename(os.environ.get("RT_B200_LIB", "default")).replace(".so", "")
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "wavefront_ab_%s.json" % tag), "w"), indent=1)
if not all(out[n]["identical"] for n in names):
    raise SystemExit("wavefront frame differs from the megakernel frame")
