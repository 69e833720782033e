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


def broadcast_blob(blob: bytes | None, device, src: int = 0) -> torch.Tensor:
    """Rank `src` passes the exported blob bytes; every rank returns a uint8 tensor on `device`."""
    rank = dist.get_rank()
    n = torch.zeros(1, dtype=torch.int64, device=device)
    if rank == src:
        n[0] = len(blob)
    dist.broadcast(n, src)
    if rank == src:
        t = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
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


def dyn_split(width: int, height: int, world: int, fraction: float = 0.15):
    """How many of the frame's 8x4 tiles go to the cross-GPU dynamic tail queue (the last ones in tile
    order) and how many tiles one claim takes.  world == 1: none."""
    ntiles = ((width + 7) // 8) * ((height + 3) // 4)
    if world <= 1:
        return 0, 0
    dyn_tiles = int(ntiles * fraction)
    return dyn_tiles, 8


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
                 dyn_fraction=0.15, flush_bytes=0, ctas_per_sm=0):
        import ctypes as C
        from . import capi
        self.C, self.capi, self.lib = C, capi, lib
        self.scene, self.host, self.dev = scene_handle, host_scene, device
        self.W, self.H = width, height
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.depth, self.ring = max(1, depth), max(ring, depth + 1)
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
        self.dyn_tiles, self.dyn_chunk = dyn_split(width, height, self.world, dyn_fraction)
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
            self.params.append(host_scene.params(tile=self.tile, flags=capi.RT_FLAG_NO_CLEAR,
                                                 dyn=(self.dyn_tiles, self.dyn_chunk,
                                                      (self.ctr_ptr + 64 * s) if self.dyn_tiles else None)))
        self.submitted = 0
        # ctas_per_sm > 0: every frame's persistent grid takes only that many CTA slots per SM, so the frames in
        # flight run side by side (rt_scene_set_launch_width)
        self.ctas_per_sm = int(ctas_per_sm)
        capi.check(lib.rt_scene_set_launch_width(scene_handle, self.ctas_per_sm), lib)

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
        with torch.cuda.stream(cs):
            if self.flush is not None:
                self.flush.fill_(1)                      # evict the scene from L2 between timed frames
            if host_io:
                self.d_lpos.copy_(self.h_lpos, non_blocking=True)   # copyLights (mytracer.cpp:105-118)
                self.d_lcol.copy_(self.h_lcol, non_blocking=True)
            fo = capi.FrameOut(self.fb_ptr + slot * self.frame_bytes, None, None, None)
            capi.check(lib.rt_render_device(self.scene, C.byref(self.host.camera), self.d_lpos.data_ptr(),
                                            self.d_lcol.data_ptr(), self.nl, C.byref(self.params[slot]), C.byref(fo),
                                            C.c_void_p(cs.cuda_stream)), lib)
            done = torch.cuda.Event()
            done.record(cs)
        self.comm.wait_event(done)
        with torch.cuda.stream(self.comm):
            if self.world > 1:
                dist.all_reduce(self.signal)             # completion signal + flow control (4 bytes)
            if self.rank == 0:
                if host_io:                              # consumer: the finished frame goes to the host
                    self.h_frames[slot].copy_(self.fb_view[slot * self.frame_bytes:(slot + 1) * self.frame_bytes],
                                              non_blocking=True)
                if self.dyn_tiles:
                    self.ctr_view[64 * slot:64 * slot + 8].zero_()   # the slot's ticket counter for frame f + ring
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(self.comm)
        self.comm_done[f] = ev
        self.last_event = ev
        self.submitted += 1
        return slot

    def frame(self, slot) -> torch.Tensor:
        """Rank 0: uint8 view (H, W, 4) of the frame in ring slot `slot` (valid after its signal completed)."""
        return self.fb_view[slot * self.frame_bytes:(slot + 1) * self.frame_bytes].view(self.H, self.W, 4)

    def drain(self):
        for s in self.compute:
            s.synchronize()
        self.comm.synchronize()

    def close(self):
        self.drain()
        self.lib.rt_scene_set_launch_width(self.scene, 0)
        self.fb_view = self.ctr_view = None
        for p in self._opened:
            self.lib.rt_ipc_close(p)
        if dist.is_initialized() and self.world > 1:
            dist.barrier()                               # every rank unmapped before rank 0 frees
        for p in self._own:
            self.lib.rt_shared_free(p)
        self._opened, self._own = [], []
