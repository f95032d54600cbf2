"""GPU (>= 2 devices) parity of the tensor-parallel decoder (one process per GPU, NCCL): attn_output / ffn_down
split along K + all-reduce, everything else split by rows, logits all-gathered -- against the same model on one GPU."""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

import pytest

from koboldcpp_b200.lib import lib

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("world", [2])
def test_tensor_parallel_model_matches_single_gpu(world):
    if lib().b200_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port", str(29540 + world),
           str(ROOT / "tests" / "tp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
