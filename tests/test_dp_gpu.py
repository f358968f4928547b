"""Multi-GPU parity of the data-parallel path: runs tools/dp_parity.py under torchrun on 2 GPUs (and on all of them
when the box has more).  Skipped on a single-GPU box; bench.py --gpus N and the committed
profiles/r02_dp_parity_*.txt are the multi-GPU evidence there."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 8])
def test_dp_gradients_match_single_process(world):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    if world > n:
        pytest.skip(f"needs {world} GPUs, this box has {n}")
    env = dict(os.environ, NCCL_DEBUG="WARN")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
                        str(ROOT / "tools" / "dp_parity.py")], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=900)
    print(r.stdout[-4000:])
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-3000:]
    assert "dp_parity result: PASS" in r.stdout
