"""Multi-GPU frame check (launch with torchrun, one rank per GPU): frames rendered through the fused pipeline
(peer stores into rank 0's frame buffer, cross-GPU dynamic tail queue, frames in flight) must equal the frame one
GPU renders alone, bit for bit, for every frame of a burst.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/gpu_multi_check.py [width height frames]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import rtpkg

rt = rtpkg.load()
from my_raytracer_b200 import capi
from my_raytracer_b200 import dist as rdist


def main():
    W = int(sys.argv[1]) if len(sys.argv) > 1 else 1920
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
    frames = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lib = rt.load_library()
    capi.check(lib.rt_init_device(local), lib)
    dist.init_process_group("nccl", device_id=dev, pg_options=dist.ProcessGroupNCCL.Options(is_high_priority_stream=True))
    desc = rt.scenes.office(width=W, height=H)
    # only rank 0 assembles and flattens the scene; the others get the blob device-to-device and submit frames
    # with the camera / lights / parameters alone
    host = rt.HostScene(desc, lib) if rank == 0 else rdist.FrameInputs(desc, lib)
    scene0 = rt.DeviceScene(host) if rank == 0 else None
    t_blob = rdist.broadcast_blob(rdist.wrap_device_ptr(*scene0.blob_device(), device=dev) if rank == 0 else None, dev)
    scene = scene0 if rank == 0 else rt.DeviceScene(lib=lib, blob_device_ptr=t_blob.data_ptr(), blob_bytes=t_blob.numel())
    ok = True
    shared = rdist.SharedHostFrames(lib, W * H * 4 * 5, rank, world, dev)
    for depth, frac, mode, zc in ((2, 0.15, "inline", True), (1, 0.0, "off", False), (3, 1.0, "side", True)):
        pipe = rdist.FramePipeline(lib, scene.handle, host, dev, W, H, depth=depth, ring=depth + 2, dyn_fraction=frac,
                                   flush_bytes=8 << 20, flush_mode=mode, shared_host=shared if zc else None)
        ref = None
        if rank == 0:
            ref = scene.render(host.camera, host.lights_pos, host.lights_color, host.params(), want=("rgba8",))["rgba8"]
        got = []
        for f in range(frames):
            slot = pipe.submit(host_io=(f % 2 == 1))
            if rank == 0:
                pipe.comm_done[f].synchronize()          # this frame's completion signal
                if not (zc and f % 2 == 1):             # (zero copy: the pixels went to the host frame, not to rank 0's HBM)
                    got.append(pipe.frame(slot).cpu().numpy().copy())
                if f % 2 == 1:                          # the end-to-end frame: shared host block (zero copy) or rank 0's D2H
                    got.append(pipe.host_frame(slot).copy())
        pipe.drain()
        if rank == 0:
            bad = [i for i, g in enumerate(got) if not np.array_equal(g, ref)]
            print(f"world {world} depth {depth} dyn_fraction {frac} flush {mode} zero_copy {zc}: {len(got)} frame copies checked, mismatching: {bad}", flush=True)
            ok = ok and not bad
        pipe.close()
    shared.close()
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    if int(flag.item()):
        raise SystemExit("multi-GPU frame differs from the single-GPU frame")
    if rank == 0:
        print("MULTI-GPU CHECK OK", flush=True)


if __name__ == "__main__":
    main()
