"""Quick GPU sanity run: parity vs the oracle on small scenes + Office 1080p timing and counters."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
import oracle_py as orc

lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
kind = "reference" if orc.available("reference") else "port"
res = {}
for name, kw in [("mixed_small", {}), ("spheres", dict(width=200, height=200)), ("office", dict(width=240, height=135)),
                 ("rings", dict(width=140, height=100, nu=64, nv=16)), ("molecule", dict(width=100, height=100, n_spheres=60)),
                 ("textured", dict(width=192, height=108, tex_size=128, slices=48, stacks=24))]:
    desc = rt.scenes.REGISTRY[name](**kw)
    host = rt.HostScene(desc, lib)
    dev = rt.DeviceScene(host)
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_COUNTERS))
    ex = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_EXACT_ONLY))
    ref = orc.OracleScene(desc, kind).render()
    d = np.abs(rt.quantize_rgb8(ref["rgb"]).astype(int) - out["rgba8"][..., :3].astype(int))
    r = dict(ids_equal=float((ref["ids"] == out["hit_id"]).mean()), t_equal=float((ref["t"] == out["hit_t"]).mean()),
             pix_max=int(d.max()), within1=float((d <= 1).mean()), rgb_maxdiff=float(np.abs(ref["rgb"] - out["pixels"][..., :3]).max()),
             exact_only_same=bool((ex["rgba8"] == out["rgba8"]).all() and (ex["hit_id"] == out["hit_id"]).all()),
             rays_ref=ref["rays"], counters=out["counters"], kernel_ms=dev.last_kernel_ms())
    res[name] = r
    print(name, json.dumps(r), flush=True)
# Office 1080p timing
desc = rt.scenes.office()
t0 = time.time(); host = rt.HostScene(desc, lib); t1 = time.time()
dev = rt.DeviceScene(host)
print("office host build %.2fs info %s" % (t1 - t0, dev.info()), flush=True)
p = host.params()
for i in range(5):
    out = dev.render(host.camera, host.lights_pos, host.lights_color, p, want=("rgba8",))
    print("frame", i, "kernel_ms", dev.last_kernel_ms(), flush=True)
out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=capi.RT_FLAG_COUNTERS), want=("rgba8", "hit_id"))
print("counters", out["counters"], "kernel_ms(counting)", dev.last_kernel_ms())
# oracle on a row sample of the 1080p frame for parity
o = orc.OracleScene(desc, kind)
ref = o.render(y0=0, ystep=16)
rows = np.arange(0, 1080, 16)
d = np.abs(rt.quantize_rgb8(ref["rgb"][rows]).astype(int) - out["rgba8"][rows][..., :3].astype(int))
print("1080p rows/16: ids_equal", float((ref["ids"][rows] == out["hit_id"][rows]).mean()), "pix max", int(d.max()),
      "oracle secs", ref["seconds"], "rays", ref["rays"])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.save(os.path.join(ROOT, "gpurun_out", "office_1080p.npy"), out["rgba8"])
