"""Two real peers on two GPUs through bench.py's public path: DHT rendezvous, matchmaking, CUDA-IPC peer memory,
the fused all-reduce + LAMB kernel with cross-GPU flag barriers, bit-identical parameters afterwards."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("compression", ["FLOAT16", "UNIFORM_8BIT", "BLOCKWISE_FP8"])
def test_two_gpu_collaborative_steps(compression):
    port = 29500 + (os.getpid() % 400)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2",
           "--warmup", "2", "--model", "albert-small", "--micro-batch", "8", "--target-batch-size", "64",
           "--seq-len", "128", "--compression", compression, "--skip-e2e"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["params_in_sync"] is True
    assert out["collab"]["solo_rounds"] == 0 and out["collab"]["failed_rounds"] == 0, out["collab"]
    assert out["local_step"] == 4
    assert out["gpu_launches"] > 0
