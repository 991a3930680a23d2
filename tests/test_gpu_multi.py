"""Multi-GPU tests proper (skipped below 2 devices): one process per GPU under torch.distributed.run, NCCL for the
plumbing, the persistent fit exchanging its sums inside the kernel over NVLink.  Each world size runs
``tools/check_multi_gpu.py``, which asserts (a) peer exchange == two-stage form bit for bit and equal weights on all
ranks, (b) rank-count invariance against a one-GPU run of the same job (V0, phi0, psi0, weights, epochs per fit),
(c) sharded VaR quantiles == one-GPU quantiles, (d) sharded device survivors == the unsharded array."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_check(world):
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "check_multi_gpu.py"), "20", "6"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    print(r.stdout[-4000:])
    assert r.returncode == 0, r.stderr[-4000:]
    assert "MULTI_GPU_CHECK PASS" in r.stdout and "FAIL" not in r.stdout, r.stdout[-4000:]
