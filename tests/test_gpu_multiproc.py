"""Real multi-process runs: `torch.distributed.run` with one rank per visible GPU (2 when the box has that many,
4 / 8 likewise), NCCL process group, every rank's results compared with the CPU oracle on the whole job
(tests/mp_worker.py).  Skipped on a one-GPU box — tests/test_gpu_sharded.py covers the same phases with logical
shards there."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [(1, 60000, {}), (0, 30000, {}), (2, 20000, {}), (4, 50000, {}),
         (2, 12000, dict(word_length=10, genome_len=3000))]


def visible_gpus():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ranks_match_oracle(world, tmp_path):
    if visible_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    out = tmp_path / "result.txt"
    port = 29600 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "mp_worker.py"), str(out),
           json.dumps(CASES)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert out.read_text().strip() == "ok", out.read_text()


@pytest.mark.parametrize("world", [2, 3])
def test_processes_on_one_device_use_peer_memory(world, tmp_path):
    """The library's own transport between PROCESSES (shared host segment + CUDA IPC mappings), on whatever single GPU
    the box has."""
    if visible_gpus() < 1:
        pytest.skip("needs a GPU")
    out = tmp_path / "result.txt"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29650 + world), os.path.join(HERE, "mp_worker.py"), str(out),
           json.dumps(CASES[:3])]
    env = dict(os.environ, HUMID_TEST_ONE_DEVICE="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert out.read_text().strip() == "ok", out.read_text()
