"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): the edge-sharded Gauss-Newton solve,
with the NCCL all-gather exchange and with the exchange fused into the kernels over NVLink peer memory, is
bit-identical to the single-GPU solve, also when the ranks are skewed by ~300 us against each other in every
iteration.  One process per GPU under torchrun (tests/mgpu_sharded.py)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_solve_bit_identical_to_single_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29600 + world
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tests", "mgpu_sharded.py")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("MGPU_OK") == world, r.stdout[-2000:] + r.stderr[-2000:]
