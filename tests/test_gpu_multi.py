"""Multi-GPU frame (needs >= 2 GPUs on the box, skipped otherwise): tools/gpu_multi_check.py under torchrun —
frames rendered through the fused pipeline (peer stores into rank 0's frame buffer over NVLink, cross-GPU dynamic
tail queue, frames in flight) equal the single-GPU frame bit for bit."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fused_multi_gpu_frame_is_bitwise_identical(lib, gpu):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU on this box (covered at N = 2..8 by tools/gpu_multi_check.py under gpurun --gpus N)")
    n = min(n, 4)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "gpu_multi_check.py"), "642", "363", "8"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI-GPU CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
