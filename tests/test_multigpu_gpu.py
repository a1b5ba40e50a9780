"""Multi-rank correctness inside the pytest tier: spawns tests/multigpu_check.py under torchrun on every GPU count
(2, 4, 8) the box offers.  Checks the reference's oracle (all ranks hold byte-identical chains, localTest.sh:70-85) for the
fused peer-memory path and the NCCL baseline, and that both equal the single-GPU chain.  Skipped on a one-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_and_nccl_chains_equal_single_gpu_chain(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, PYTHONPATH=ROOT, BSC_ITERS="5", BSC_PEERS="64")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29700 + world),
                        os.path.join(ROOT, "tests", "multigpu_check.py")], capture_output=True, text=True, env=env, cwd=ROOT,
                       timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "MULTIGPU_CHECK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
