"""N > 1 host logic on CPU: world_size 2, gloo.  The scene blob is broadcast from rank 0, every rank
renders its tile share (with the host build of the kernel core standing in for the GPU), the frame is
gathered to rank 0 and must equal the single-rank frame bit for bit."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import rtpkg
    rt = rtpkg.load()
    import hostsim_py as sim
    from my_raytracer_b200 import dist as rdist
    lib = rt.load_library()
    host = rt.HostScene(rt.scenes.mixed_small(width=90, height=50), lib)      # every rank can describe the scene
    blob = sim.flatten(host) if rank == 0 else None                            # ...but only rank 0 flattens it
    # rank 0 hands the broadcast a tensor view of the resident blob (on the GPU: a view of HBM, no host round trip)
    src = torch.frombuffer(bytearray(blob), dtype=torch.uint8) if rank == 0 else None
    t = rdist.broadcast_blob(src, "cpu")
    blob_r = bytes(t.numpy())
    # static chunked interleave + the dynamic tail (host build: pool chunk q goes to rank q % world)
    dyn_tiles, dyn_chunk = rdist.dyn_split(90, 50, world, 0.3)
    assert dyn_tiles > 0 and dyn_chunk > 0 and rdist.dyn_split(90, 50, 1) == (0, 0)
    part = sim.render(host, blob_r, tile=rdist.tile_share(rank, world, 2), dyn=(dyn_tiles, dyn_chunk))
    local = torch.from_numpy(part["rgba8"].copy())
    full = rdist.gather_frame(local, torch.empty_like(local))
    ms = rdist.max_over_ranks(float(rank + 1), "cpu")
    # the end-to-end frame of the multi-GPU path: ONE host block every rank's process maps (POSIX shm); each rank
    # stores the pixels it rendered (on the GPU: its kernel does, through the registered mapping)
    shm = rdist.SharedHostFrames(lib, 90 * 50 * 4, rank, world, "cpu", register=False)
    mine = part["rgba8"][..., 3] == 255
    shm.numpy().reshape(50, 90, 4)[mine] = part["rgba8"][mine]
    dist.barrier()
    shared_frame = shm.numpy().reshape(50, 90, 4).copy()
    dist.barrier()
    shm.close()
    if rank == 0:
        q.put((full.numpy().copy(), len(blob_r), ms, shared_frame))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_frame_equals_single_rank():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    full, nbytes, ms, shared_frame = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import rtpkg
    rt = rtpkg.load()
    import hostsim_py as sim
    host = rt.HostScene(rt.scenes.mixed_small(width=90, height=50), rt.load_library())
    blob = sim.flatten(host)
    ref = sim.render(host, blob)
    assert nbytes == len(blob) and ms == 2.0
    assert np.array_equal(full, ref["rgba8"])
    assert np.array_equal(shared_frame, ref["rgba8"])
