"""Multi-GPU parity on hardware: the sharded handle (one process per GPU, NCCL) against the one-rank handle of the same
problem, through the C-ABI.  Needs at least two GPUs on the box (gpurun --gpus 2); skipped otherwise."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    import torch

    return torch.cuda.device_count()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_handle_matches_one_rank(world):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "sharded_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    if out.returncode != 0:  # keep the whole output where a gpurun call brings it back
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "sharded_worker_%d.log" % world), "w") as f:
            f.write(out.stdout + "\n---- stderr ----\n" + out.stderr)
    errs = [ln for ln in out.stderr.splitlines() if "Error" in ln or "assert" in ln.lower() or "aniso" in ln]
    assert out.returncode == 0, (out.stdout[-1500:], errs[-12:])
    assert out.stdout.count("SHARDED_OK") == world, out.stdout[-3000:]
