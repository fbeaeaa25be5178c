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


def test_late_joiner_and_peer_death(tmp_path):
    """Elasticity on real peer memory: GPU 1 joins two steps late (state download from GPU 0 over CUDA IPC), both run
    joint rounds with bit-identical parameters, then GPU 1 vanishes without shutting down and GPU 0 keeps stepping
    (bounded flag waits / matchmaking timeouts, solo rounds) instead of hanging."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    worker = os.path.join(ROOT, "tests", "elastic_worker.py")
    first = subprocess.Popen([sys.executable, worker, "--role", "first", "--dir", str(tmp_path), "--until", "3"],
                             cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    second = subprocess.Popen([sys.executable, worker, "--role", "second", "--dir", str(tmp_path), "--until", "4"],
                              cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        out2, _ = second.communicate(timeout=300)
        out1, _ = first.communicate(timeout=300)
    finally:
        for p in (first, second):
            if p.poll() is None:
                p.kill()
    assert first.returncode == 0, out1[-3000:]
    assert second.returncode == 0, out2[-3000:]
    a = json.load(open(tmp_path / "first.json"))
    b = json.load(open(tmp_path / "second.json"))
    assert b["joined_at"] is not None and b["joined_at"] >= 2, b          # it adopted the collaboration's step
    assert b["stats"].get("state_downloads", 0) >= 1, b["stats"]
    common = sorted(set(a["digests"]) & set(b["digests"]), key=int)
    assert len(common) >= 2, (a["digests"].keys(), b["digests"].keys())
    for s in common:                                                        # bit-identical master parameters
        assert a["digests"][s] == b["digests"][s], (s, a["digests"][s], b["digests"][s])
    assert a["death_step"] is not None and a["local_step"] >= a["death_step"] + 3, a   # kept going after the peer died
    assert a["stats"]["solo_rounds"] >= 1, a["stats"]


def test_nvls_multimem_allreduce_two_gpus():
    """In-switch reduction (multimem.ld_reduce / multimem.st / multimem.red over a multicast object bound to CUDA VMM
    windows): weighted mean on 2 GPUs against a torch reference. Skipped when the device has no multicast support."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from dedloc_b200.parallel import nvls

    if not nvls.multicast_supported(0):
        pytest.skip("no NVLS / multicast support on this device")
    port = 29900 + (os.getpid() % 90)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "test_nvls.py"), "4,20"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "NVLS OK" in res.stdout


@pytest.mark.parametrize("sig", ["SIGKILL", "SIGSTOP"])
def test_peer_dies_inside_the_kernel(tmp_path, sig):
    """A member is killed (or frozen) while it sits INSIDE the all-reduce kernel, before it has sent anything: the
    survivor's bounded waits expire, its launch returns `abort_before_reduce` within the timeout with parameters and
    moments untouched, and it steps on its own gradient instead of hanging."""
    import signal
    import time

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    worker = os.path.join(ROOT, "tests", "abort_worker.py")
    victim = subprocess.Popen([sys.executable, worker, "--role", "victim", "--dir", str(tmp_path)], cwd=ROOT,
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    survivor = subprocess.Popen([sys.executable, worker, "--role", "survivor", "--dir", str(tmp_path)], cwd=ROOT,
                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        t0 = time.time()
        while not (tmp_path / "victim_launching").exists():
            assert victim.poll() is None, victim.stdout.read()[-3000:]
            assert time.time() - t0 < 120
            time.sleep(0.05)
        time.sleep(1.0)                                   # the victim is parked inside its kernel now
        os.kill(victim.pid, getattr(signal, sig))
        out, _ = survivor.communicate(timeout=120)
        assert survivor.returncode == 0, out[-3000:]
        res = json.load(open(tmp_path / "survivor.json"))
        assert res["status"]["name"] == "abort_before_reduce", res
        assert res["seconds"] < 10.0, res                 # timeout_s = 2: bounded, not a hang
        assert res["untouched"] is True, res
        assert res["solo_status"]["code"] == 0 and res["solo_moved"] is True, res
    finally:
        for p in (victim, survivor):
            if p.poll() is None:
                p.kill()
