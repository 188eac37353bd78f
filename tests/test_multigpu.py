"""GPU, >= 2 devices: same inputs sharded over the GPUs of the box give bit-identical per-image outputs, and the
optional NCCL normaliser all-reduce reproduces the unsharded losses (tests/tools/multigpu_parity.py under torchrun).
Skipped on a one-GPU box; the CPU (gloo, world size 2) version of the host logic is tests/test_host_dist.py."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_sharded_equals_unsharded_on_all_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run: gpurun --gpus 2 -- python -m pytest tests/test_multigpu.py -m gpu)")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "tools", "multigpu_parity.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["multigpu_parity"] == {"local": "ok", "sync": "ok"} and res["world"] == n
