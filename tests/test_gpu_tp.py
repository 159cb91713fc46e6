"""Multi-GPU bit-exactness of the tensor-parallel forms on REAL NVLink peer memory: spawns
tests/tp_worker.py under torchrun on 2 / 4 / 8 GPUs of this box; skipped when fewer GPUs are visible (the driver's
1-GPU run).  The builder's runs are kept under profiles/ (r2_tp_check_*.log)."""
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
def test_tp_bit_exact_on_nvlink(cuda, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, {torch.cuda.device_count()} visible")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "tp_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    log = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(log):
        open(os.path.join(log, f"tp_check_{world}.log"), "w").write(out.stdout + "\n---- stderr ----\n" + out.stderr[-4000:])
    assert out.returncode == 0 and f"TP CHECK PASS world={world}" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
