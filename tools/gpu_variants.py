"""Times kernel variants (my-raytracer_b200/variants/*.so) on Office 1080p (or RT_SCENE=<name>) and checks each against the
first one bit for bit.  Usage: python tools/gpu_variants.py [names...]; a name may carry render flags,
e.g. wide:8 = variants/wide.so with RT_FLAG_BINARY_NODES; the name "default" is the in-tree library."""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, json, hashlib
import numpy as np
sys.path.insert(0, %r)
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
big = int(os.environ.get("RT_BIG", "0"))     # RT_BIG=10000: the 10 M-triangle scene (3840x2160, 2x2 spp)
desc = rt.scenes.scatter(width=3840, height=2160, n_objects=big) if big else rt.scenes.REGISTRY[os.environ.get("RT_SCENE", "office")]()
host = rt.HostScene(desc, lib, gpu_build=bool(big))
SPP = 2 if big else 1
dev = rt.DeviceScene(host)
ms = []
FL = int(os.environ.get("RT_VARIANT_FLAGS", "0"))
for i in range(12):
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=FL, spp_grid=SPP), want=("rgba8", "hit_id", "hit_t") if SPP == 1 else ("rgba8",))
    ms.append(dev.last_kernel_ms())
h = hashlib.sha1(b"".join(out[k].tobytes() for k in sorted(out))).hexdigest()
c = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_COUNTERS | FL, spp_grid=SPP), want=("rgba8",))["counters"]
print(json.dumps(dict(ms_min=min(ms[2:]), ms_med=sorted(ms[2:])[5], sha=h, nodes=c["nodes_visited"], filt=c["tri_filter"], exact=c["tri_exact"])))
''' % ROOT
names = sys.argv[1:] or sorted(f[:-3] for f in os.listdir(os.path.join(ROOT, "my-raytracer_b200", "variants")) if f.endswith(".so"))
for n in names:
    base, _, fl = n.partition(":")
    env = dict(os.environ, RT_VARIANT_FLAGS=fl or "0")
    if base != "default":
        env["RT_B200_LIB"] = os.path.join(ROOT, "my-raytracer_b200", "variants", base + ".so")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print(n, r.stdout.strip() or r.stderr[-600:], flush=True)
