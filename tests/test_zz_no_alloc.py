"""rt_render_device_static: the entry point that allocates nothing (its wavefront path-state scratch is a file-scope
__device__ array of the library, include/rt_b200.h).  No counterpart in the reference.  Checked through the C ABI
against the existing entry point rt_render_device on the same scene: hit ids, ray parameters and 8-bit pixels must be
IDENTICAL (the tolerance of the other GPU tests is 1 LSB on the pixels and exact ids; the two paths share their
arithmetic, so identical is what is asserted), on two calls in a row (the array keeps the first call's contents) and
with a different scene in between.  Collected last (file name) so that a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_render(lib, gpu, fn, dev, host, params, d_lpos, d_lcol, W, H, stream=None, ids_too=True):
    import torch
    rgba = torch.full((H * W * 4,), 7, dtype=torch.uint8, device="cuda")
    ids = torch.full((H * W,), -5, dtype=torch.int32, device="cuda")
    ts = torch.full((H * W,), -7.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    fo = gpu.FrameOut(rgba.data_ptr(), None, ids.data_ptr() if ids_too else None, ts.data_ptr() if ids_too else None)
    gpu.check(fn(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(), d_lpos.shape[0],
                 C.byref(params), C.byref(fo), stream), lib)
    torch.cuda.synchronize()
    return rgba.cpu().numpy().reshape(H, W, 4), ids.cpu().numpy().reshape(H, W), ts.cpu().numpy().reshape(H, W)


def _same(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


def test_static_scratch_entry_point_matches_rt_render_device(rt, lib, gpu):
    import torch
    W, H = 200, 152
    cases = []
    for name, kw in (("mixed_small", dict(width=W, height=H)), ("office", dict(width=W, height=H))):
        host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
        dev = rt.DeviceScene(host)
        d_lpos = torch.tensor(host.lights_pos, dtype=torch.float64, device="cuda")
        d_lcol = torch.tensor(host.lights_color, dtype=torch.float64, device="cuda")
        cases.append((host, dev, d_lpos, d_lcol))
    refs = [_device_render(lib, gpu, lib.rt_render_device, dev, host, host.params(), lp, lc, W, H)
            for host, dev, lp, lc in cases]
    assert (refs[0][1] >= 0).any() and (refs[1][1] >= 0).any()
    # two calls in a row on the same scene, then the other scene, then the first again: the shared array carries the
    # previous call's path state every time
    for k in (0, 0, 1, 1, 0):
        host, dev, lp, lc = cases[k]
        got = _device_render(lib, gpu, lib.rt_render_device_static, dev, host, host.params(), lp, lc, W, H)
        assert _same(got, refs[k]), k
    # 2x2 sub-pixel samples are passes over the same path slots; on a side stream
    host, dev, lp, lc = cases[0]
    p4 = host.params(spp_grid=2)
    ref4 = _device_render(lib, gpu, lib.rt_render_device, dev, host, p4, lp, lc, W, H, ids_too=False)
    st = torch.cuda.Stream()
    got4 = _device_render(lib, gpu, lib.rt_render_device_static, dev, host, p4, lp, lc, W, H, C.c_void_p(st.cuda_stream),
                          ids_too=False)
    assert np.array_equal(got4[0], ref4[0]) and not np.array_equal(ref4[0], refs[0][0])


def test_static_scratch_entry_point_rejects_what_does_not_fit(rt, lib, gpu):
    import torch
    host = rt.HostScene(rt.scenes.mixed_small(width=264, height=256), lib)      # 33 x 64 tiles x 32 = 67584 > 65536 slots
    dev = rt.DeviceScene(host)
    d_lpos = torch.tensor(host.lights_pos, dtype=torch.float64, device="cuda")
    d_lcol = torch.tensor(host.lights_color, dtype=torch.float64, device="cuda")
    out = torch.zeros(264 * 256 * 4, dtype=torch.uint8, device="cuda")
    fo = gpu.FrameOut(out.data_ptr(), None, None, None)
    args = (dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr())
    st = lib.rt_render_device_static(*args, d_lpos.shape[0], C.byref(host.params()), C.byref(fo), None)
    assert st == gpu.RT_ERR_LIMIT and b"65536" in lib.rt_last_error()
    # half of the tiles (this rank's share of a 2-GPU frame) fits
    half = host.params(tile=(0, 2, 1), flags=gpu.RT_FLAG_NO_CLEAR)
    ref = torch.zeros_like(out)
    gpu.check(lib.rt_render_device(*args, d_lpos.shape[0], C.byref(half), C.byref(gpu.FrameOut(ref.data_ptr(), None, None, None)), None), lib)
    gpu.check(lib.rt_render_device_static(*args, d_lpos.shape[0], C.byref(half), C.byref(fo), None), lib)
    torch.cuda.synchronize()
    assert torch.equal(out, ref) and int(ref.count_nonzero()) > 0
    # too many lights
    many = torch.zeros((9, 4), dtype=torch.float64, device="cuda")
    small = rt.HostScene(rt.scenes.mixed_small(width=64, height=48), lib)
    sdev = rt.DeviceScene(small)
    st = lib.rt_render_device_static(sdev.handle, C.byref(small.camera), many.data_ptr(), many.data_ptr(), 9,
                                     C.byref(small.params()), C.byref(fo), None)
    assert st == gpu.RT_ERR_LIMIT and b"lights" in lib.rt_last_error()
    st = lib.rt_render_device_static(sdev.handle, C.byref(small.camera), many.data_ptr(), many.data_ptr(), 2,
                                     C.byref(small.params(flags=gpu.RT_FLAG_COUNTERS)), C.byref(fo), None)
    assert st == gpu.RT_ERR_INVALID_ARG
    assert lib.rt_render_device_static(None, C.byref(small.camera), many.data_ptr(), many.data_ptr(), 2,
                                       C.byref(small.params()), C.byref(fo), None) == gpu.RT_ERR_INVALID_ARG
