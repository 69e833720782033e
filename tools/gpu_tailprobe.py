"""Tail probe (variant built with -DRT_TAILPROBE): per-warp start/end times of the render kernel.
Prints kernel span, mean warp busy time and the idle share (1 - mean busy / span) = what the drain costs."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("RT_B200_LIB", os.path.join(ROOT, "my-raytracer_b200", "variants", "tailprobe.so"))
import torch
import rtpkg
rt = rtpkg.load()
from my_raytracer_b200 import capi
lib = rt.load_library()
capi.check(lib.rt_init_device(0), lib)
name = os.environ.get("RT_SCENE", "office")                    # RT_SCENE=rings: BASELINE config 3 at 700x500
desc = rt.scenes.REGISTRY[name]()
host = rt.HostScene(desc, lib)
dev = rt.DeviceScene(host)
W, H = int(desc.camera["width"]), int(desc.camera["height"])
print("scene", name, W, H, flush=True)
d_rgba = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
d_lpos = torch.tensor(host.lights_pos, dtype=torch.float64, device="cuda")
d_lcol = torch.tensor(host.lights_color, dtype=torch.float64, device="cuda")
fo = capi.FrameOut(d_rgba.data_ptr(), None, None, None)
params = host.params()
st = torch.cuda.current_stream()
for i in range(6):
    capi.check(lib.rt_render_device(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(), d_lpos.shape[0],
                                    C.byref(params), C.byref(fo), C.c_void_p(st.cuda_stream)), lib)
    cnt = capi.Counters(); capi.check(lib.rt_get_counters(dev.handle, C.byref(cnt)), lib)
    raw = [getattr(cnt, f[0]) for f in cnt._fields_]
    sum_busy, t_max, t_min, n, sum_start = raw[:5]
    span = (t_max - t_min) / 1e6
    print("frame %d kernel_ms %.3f | warps %d span %.3f ms mean busy %.3f ms mean start %.3f ms idle share %.1f %%" % (
        i, dev.last_kernel_ms(), n, span, sum_busy / n / 1e6, (sum_start / n - t_min) / 1e6, 100 * (1 - sum_busy / n / 1e6 / span)), flush=True)
