"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL over NVLink; gloo in CPU tests).

The path shards by image tiles (rays are independent):
  * the flattened scene blob is produced once on rank 0 and replicated with ONE broadcast;
  * every rank renders the 8x4-pixel tiles it owns (static chunked interleave, rt_render_params
    tile_rank/tile_world/tile_chunk) — no data-path collective during traversal or shading;
  * the framebuffer is gathered to rank 0 (the only real exchange step of the path).
Everything here is device-agnostic (works on CPU tensors with gloo) so that the N > 1 host logic is
covered by world_size-2 CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def broadcast_blob(blob, device, src: int = 0) -> torch.Tensor:
    """Rank `src` passes the flattened scene — host bytes, or a uint8 tensor already on `device` (a view of the
    resident blob, `wrap_device_ptr(*scene.blob_device())`: then the broadcast reads HBM directly, no host
    round trip); every rank returns a uint8 tensor on `device`."""
    rank = dist.get_rank()
    n = torch.zeros(1, dtype=torch.int64, device=device)
    if rank == src:
        n[0] = blob.numel() if isinstance(blob, torch.Tensor) else len(blob)
    dist.broadcast(n, src)
    if rank == src:
        t = blob if isinstance(blob, torch.Tensor) else torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
    else:
        t = torch.empty(int(n.item()), dtype=torch.uint8, device=device)
    dist.broadcast(t, src)
    return t


def default_chunk(world: int) -> int:
    """Tiles per ownership run.  One image row of 8x4 tiles at 1920 px is 240 tiles; runs of 16 tiles
    keep each rank's rays coherent while interleaving finely enough to balance cheap and expensive
    image regions across ranks."""
    return 16


def tile_share(rank: int, world: int, chunk: int | None = None):
    return (rank, world, chunk or default_chunk(world))


def gather_frame(local: torch.Tensor, scratch: torch.Tensor | None = None, dst: int = 0) -> torch.Tensor:
    """Framebuffer gather to rank `dst`.  `local` holds this rank's tiles and zeros elsewhere, so a
    byte-wise SUM over ranks is the union (each pixel is written by exactly one rank).  Returns the
    full frame on `dst` (in `scratch` when given, so `local` keeps its zeros for the next frame)."""
    if dist.get_rank() == dst:
        out = scratch if scratch is not None else torch.empty_like(local)
        out.copy_(local)
        dist.reduce(out, dst, op=dist.ReduceOp.SUM)
        return out
    dist.reduce(local, dst, op=dist.ReduceOp.SUM)
    return local


def max_over_ranks(value: float, device) -> float:
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ------------------------------------------------------------------------------------------------
# Fused multi-GPU frame: peer-mapped frame buffer + cross-GPU dynamic tail queue + frames in flight
# ------------------------------------------------------------------------------------------------

class _DevPtr:
    """__cuda_array_interface__ carrier so that torch can view memory allocated by the C ABI."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def wrap_device_ptr(ptr: int, nbytes: int, device=None) -> torch.Tensor:
    """uint8 torch view (no copy) of `nbytes` of device memory at `ptr` (e.g. an rt_shared_alloc block)."""
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    return torch.as_tensor(_DevPtr(ptr, nbytes), device=dev)


def dyn_split(width: int, height: int, world: int, fraction: float = 0.15, chunk: int = 8):
    """How many of the frame's 8x4 tiles go to the cross-GPU dynamic tail queue (the last ones in tile
    order) and how many tiles one claim takes (a warp renders the tiles of its claim one after the other, so the claim
    is also the granularity of the frame's tail: 8 for cheap tiles, fewer when a tile is expensive).  world == 1: none."""
    ntiles = ((width + 7) // 8) * ((height + 3) // 4)
    if world <= 1:
        return 0, 0
    dyn_tiles = int(ntiles * fraction)
    return dyn_tiles, max(1, int(chunk))


class SharedHostFrames:
    """One block of host memory that every rank's process maps (a file in /dev/shm: POSIX shared memory),
    registered with the CUDA driver in each process (rt_host_register: pinned + device-mapped).  The end-to-end
    multi-GPU frame lands here: every rank's render kernel stores the pixels it finishes straight into the ONE host
    frame over its own PCIe link, so no rank funnels the whole frame through a single host link."""

    def __init__(self, lib, nbytes, rank, world, device, register=True):
        """register=False (CPU tests): the shared mapping only, no CUDA registration."""
        import ctypes as C
        import mmap
        import os
        from . import capi
        self.lib, self.capi, self.nbytes, self.rank = lib, capi, int(nbytes), rank
        name = [None]
        if rank == 0:
            name[0] = "/dev/shm/rt_b200_frames_%d_%08x" % (os.getpid(), int.from_bytes(os.urandom(4), "little"))
            fd = os.open(name[0], os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
            os.ftruncate(fd, self.nbytes)
        if world > 1:
            dist.broadcast_object_list(name, 0, device=device)
        if rank != 0:
            fd = os.open(name[0], os.O_RDWR)
        self.path = name[0]
        self.map = mmap.mmap(fd, self.nbytes)
        os.close(fd)
        self._anchor = (C.c_char * self.nbytes).from_buffer(self.map)
        self.host_ptr = C.addressof(self._anchor)
        self.registered = bool(register)
        self.dev_ptr = None
        if register:
            d = C.c_void_p()
            capi.check(lib.rt_host_register(C.c_void_p(self.host_ptr), self.nbytes, C.byref(d)), lib)
            self.dev_ptr = d.value
        if world > 1:
            dist.barrier()                 # every rank has the file open and registered ...
        if rank == 0:
            os.unlink(self.path)           # ... so the name can go; the memory lives as long as a mapping does

    def numpy(self):
        import numpy as np
        return np.frombuffer(self.map, dtype=np.uint8)

    def close(self):
        if self.map is not None:
            if self.registered:
                self.lib.rt_host_unregister(self.host_ptr)
            self._anchor = None
            try:
                self.map.close()
            except BufferError:
                pass
            self.map = None


class FrameInputs:
    """What a rank that did not build the scene needs to submit frames: camera, lights, render parameters
    (HostScene minus the meshes; used for large scenes that only rank 0 assembles)."""

    def __init__(self, desc, lib):
        import numpy as np
        from . import scene as _scene
        self.desc, self.lib = desc, lib
        self.camera = _scene.look_at(desc.camera, lib)
        self.lights_pos = _scene._vec4_array([l[0] for l in desc.lights]) if desc.lights else np.zeros((0, 4))
        self.lights_color = _scene._vec4_array([l[1] for l in desc.lights]) if desc.lights else np.zeros((0, 4))

    def params(self, *a, **kw):
        from . import scene as _scene
        return _scene.HostScene.params(self, *a, **kw)


class FramePipeline:
    """One frame = one render-kernel launch per rank (north star (c)):

      * rank 0 owns a ring of frame buffers and of dynamic-queue counters (rt_shared_alloc), exported
        through CUDA IPC; the other ranks map them, and their render kernels store finished RGBA8 pixels
        STRAIGHT into rank 0's frame buffer over NVLink — the gather is fused into the kernel epilogue;
      * tiles: a static chunked interleave for most of the frame, the last `dyn_tiles` tiles from a
        cross-GPU ticket counter in rank 0's memory (system-scope atomics, dyn_chunk tiles per claim);
      * per frame ONE tiny all-reduce on a side stream is the completion signal and the flow control:
        frame f may start once the signal of frame f - ring + 1 has completed on this rank, which implies
        that rank 0 consumed that ring slot and zeroed its counter;
      * successive frames alternate between `depth` compute streams, so the drain of frame f overlaps the
        start of frame f + 1 (the persistent grid of f + 1 moves in as the CTAs of f retire).
    With world == 1 the same object runs without IPC and without the collective.
    """

    def __init__(self, lib, scene_handle, host_scene, device, width, height, depth=2, ring=4, tile_chunk=None,
                 dyn_fraction=0.15, flush_bytes=0, ctas_per_sm=0, flush_mode="inline", spp_grid=1, shared_host=None, flags=0,
                 dyn_chunk=8):
        """flush_mode: "inline" = the L2-evicting fill runs on the frame's own stream ahead of its kernel; "side" = one
        fill per frame on a stream of its own, paced by the frames (frame f + depth waits for fill f) but not
        serialised ahead of the kernel; "off".  shared_host: SharedHostFrames for host_io at world > 1 (zero-copy
        stores into ONE host frame); None = rank 0 copies the gathered device frame to pinned memory."""
        import ctypes as C
        from . import capi
        self.C, self.capi, self.lib = C, capi, lib
        # scene_handle may be a LIST of handles of identical copies of the scene: frame f renders copy f % len, so that a
        # copy is next touched len - 1 frames later (with copies adding up to more than the L2, every frame starts
        # with its scene out of cache - "inputs larger than L2" instead of an evicting fill)
        self.scenes = list(scene_handle) if isinstance(scene_handle, (list, tuple)) else [scene_handle]
        self.scene, self.host, self.dev = self.scenes[0], host_scene, device
        self.W, self.H = width, height
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.depth, self.ring = max(1, min(depth, 8)), max(ring, min(depth, 8) + 1)   # rt_b200.h: launch ring of 8
        self.flush_mode = flush_mode if flush_bytes else "off"
        self.spp_grid = spp_grid
        self.shared_host = shared_host
        nb = width * height * 4
        self.frame_bytes = nb
        self._own = []
        fb, ctr = C.c_void_p(), C.c_void_p()
        if self.rank == 0:
            capi.check(lib.rt_shared_alloc(nb * self.ring, C.byref(fb)), lib)
            capi.check(lib.rt_shared_alloc(64 * self.ring, C.byref(ctr)), lib)
            self._own = [fb, ctr]
        self._opened = []
        if self.world > 1:
            h = torch.zeros(128, dtype=torch.uint8, device=device)
            if self.rank == 0:
                hb = (C.c_ubyte * 64)(); hc = (C.c_ubyte * 64)()
                capi.check(lib.rt_ipc_export(fb, hb), lib)
                capi.check(lib.rt_ipc_export(ctr, hc), lib)
                h.copy_(torch.tensor(list(hb) + list(hc), dtype=torch.uint8))
            dist.broadcast(h, 0)
            if self.rank != 0:
                raw = bytes(h.cpu().numpy())
                hb = (C.c_ubyte * 64).from_buffer_copy(raw[:64]); hc = (C.c_ubyte * 64).from_buffer_copy(raw[64:])
                capi.check(lib.rt_ipc_open(hb, C.byref(fb)), lib)
                capi.check(lib.rt_ipc_open(hc, C.byref(ctr)), lib)
                self._opened = [fb, ctr]
        self.fb_ptr, self.ctr_ptr = fb.value, ctr.value
        self.fb_view = wrap_device_ptr(self.fb_ptr, nb * self.ring, device) if self.rank == 0 else None
        self.ctr_view = wrap_device_ptr(self.ctr_ptr, 64 * self.ring, device) if self.rank == 0 else None
        self.dyn_tiles, self.dyn_chunk = dyn_split(width, height, self.world, dyn_fraction, dyn_chunk)
        self.tile = tile_share(self.rank, self.world, tile_chunk) if self.world > 1 else (0, 1, 1)
        self.compute = [torch.cuda.Stream(device=device) for _ in range(self.depth)]
        self.comm = torch.cuda.Stream(device=device)
        self.comm_done = {}
        self.signal = torch.zeros(1, dtype=torch.int32, device=device)
        self.flush = torch.empty(flush_bytes, dtype=torch.uint8, device=device) if flush_bytes else None
        self.d_lpos = torch.tensor(host_scene.lights_pos, dtype=torch.float64, device=device)
        self.d_lcol = torch.tensor(host_scene.lights_color, dtype=torch.float64, device=device)
        self.nl = int(self.d_lpos.shape[0])
        # host-side buffers of the end-to-end mode (submit(host_io=True)): pinned lights in, pinned frames out
        self.h_lpos = torch.tensor(host_scene.lights_pos, dtype=torch.float64).pin_memory()
        self.h_lcol = torch.tensor(host_scene.lights_color, dtype=torch.float64).pin_memory()
        self.h_frames = [torch.zeros(nb, dtype=torch.uint8).pin_memory() for _ in range(self.ring)] if self.rank == 0 else None
        self.params = []
        for s in range(self.ring):
            self.params.append(host_scene.params(spp_grid=spp_grid, tile=self.tile, flags=capi.RT_FLAG_NO_CLEAR | int(flags),
                                                 dyn=(self.dyn_tiles, self.dyn_chunk,
                                                      (self.ctr_ptr + 64 * s) if self.dyn_tiles else None)))
        self.submitted = 0
        self.done_events = []            # completion event of every submitted frame (timing enabled)
        self.flush_stream = torch.cuda.Stream(device=device) if self.flush_mode == "side" else None
        self.flush_done = {}
        self.kernel_done = {}
        # ctas_per_sm > 0: every frame's persistent grid takes only that many CTA slots per SM, so the frames in
        # flight run side by side (rt_scene_set_launch_width)
        self.ctas_per_sm = int(ctas_per_sm)
        for h in self.scenes:
            capi.check(lib.rt_scene_set_launch_width(h, self.ctas_per_sm), lib)

    # -- one frame --------------------------------------------------------------------------------
    def submit(self, host_io=False):
        """Enqueue one frame (asynchronous).  host_io: lights come from pinned host memory (H2D) and rank 0
        copies the finished frame to pinned host memory (D2H) — the end-to-end mode."""
        C, capi, lib = self.C, self.capi, self.lib
        f = self.submitted
        slot = f % self.ring
        cs = self.compute[f % self.depth]
        gate = f - self.ring + 1
        if gate >= 0:
            cs.wait_event(self.comm_done.pop(gate))
        if self.flush_mode == "side":
            # one L2-evicting fill per frame on its own stream, paced by the frames: it starts once the previous
            # frame's kernel is done and frame f + depth waits for it, but it is not serialised ahead of frame f
            prev = self.kernel_done.pop(f - 1, None)
            if prev is not None:
                self.flush_stream.wait_event(prev)
            with torch.cuda.stream(self.flush_stream):
                self.flush.fill_(1)
                fe = torch.cuda.Event()
                fe.record(self.flush_stream)
            self.flush_done[f] = fe
            old = self.flush_done.pop(f - self.depth, None)
            if old is not None:
                cs.wait_event(old)
        zero_copy = host_io and self.shared_host is not None
        with torch.cuda.stream(cs):
            if self.flush_mode == "inline":
                self.flush.fill_(1)                      # evict the scene from L2 between timed frames
            if host_io:
                self.d_lpos.copy_(self.h_lpos, non_blocking=True)   # copyLights (mytracer.cpp:105-118)
                self.d_lcol.copy_(self.h_lcol, non_blocking=True)
            # finished pixels go to rank 0's device frame (peer stores over NVLink) or, end to end at N > 1, straight
            # into the host frame all ranks share (mapped pinned memory, each rank over its own PCIe link)
            base = self.shared_host.dev_ptr if zero_copy else self.fb_ptr
            fo = capi.FrameOut(base + slot * self.frame_bytes, None, None, None)
            capi.check(lib.rt_render_device(self.scenes[f % len(self.scenes)], C.byref(self.host.camera), self.d_lpos.data_ptr(),
                                            self.d_lcol.data_ptr(), self.nl, C.byref(self.params[slot]), C.byref(fo),
                                            C.c_void_p(cs.cuda_stream)), lib)
            done = torch.cuda.Event()
            done.record(cs)
        if self.flush_mode == "side":
            self.kernel_done[f] = done
        self.comm.wait_event(done)
        with torch.cuda.stream(self.comm):
            if self.world > 1:
                dist.all_reduce(self.signal)             # completion signal + flow control (4 bytes)
            if self.rank == 0:
                if host_io and not zero_copy:            # consumer: the finished frame goes to the host
                    self.h_frames[slot].copy_(self.fb_view[slot * self.frame_bytes:(slot + 1) * self.frame_bytes],
                                              non_blocking=True)
                if self.dyn_tiles:
                    self.ctr_view[64 * slot:64 * slot + 8].zero_()   # the slot's ticket counter for frame f + ring
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(self.comm)
        self.comm_done[f] = ev
        self.done_events.append(ev)
        self.last_event = ev
        self.submitted += 1
        return slot

    def frame(self, slot) -> torch.Tensor:
        """Rank 0: uint8 view (H, W, 4) of the frame in ring slot `slot` (valid after its signal completed)."""
        return self.fb_view[slot * self.frame_bytes:(slot + 1) * self.frame_bytes].view(self.H, self.W, 4)

    def host_frame(self, slot):
        """Rank 0: numpy uint8 view (H, W, 4) of the end-to-end frame in ring slot `slot`."""
        if self.shared_host is not None:
            a = self.shared_host.numpy()[slot * self.frame_bytes:(slot + 1) * self.frame_bytes]
        else:
            a = self.h_frames[slot].numpy()
        return a.reshape(self.H, self.W, 4)

    def drain(self):
        for s in self.compute:
            s.synchronize()
        if self.flush_stream is not None:
            self.flush_stream.synchronize()
        self.comm.synchronize()

    def close(self):
        self.drain()
        for h in self.scenes:
            self.lib.rt_scene_set_launch_width(h, 0)
        self.fb_view = self.ctr_view = None
        for p in self._opened:
            self.lib.rt_ipc_close(p)
        if dist.is_initialized() and self.world > 1:
            dist.barrier()                               # every rank unmapped before rank 0 frees
        for p in self._own:
            self.lib.rt_shared_free(p)
        self._opened, self._own = [], []
