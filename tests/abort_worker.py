"""Helper process for tests/test_gpu_multi.py::test_peer_dies_inside_the_kernel (not collected by pytest).

Two members on two GPUs run ONE round of the push-based LAMB kernel. The `victim` parks inside its kernel (test hook:
start delay) and is SIGKILLed (or SIGSTOPped) by the test while it is there; the `survivor` must come back from its
launch within the timeout with status `abort_before_reduce`, its parameters untouched, and then finish a solo
round. Handles travel through files: nothing here may depend on a collective that a dead process would hang.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--role", required=True, choices=["survivor", "victim"])
    ap.add_argument("--dir", required=True)
    args = ap.parse_args()
    from dedloc_b200 import lib
    from dedloc_b200.parallel.flat import FlatLayout
    from dedloc_b200.parallel.fused_allreduce import GroupSpec, OptHyper, PeerBuffers, fused_allreduce, fused_allreduce_v2
    from dedloc_b200.utils.compression import CompressionType

    me = 0 if args.role == "survivor" else 1
    torch.cuda.set_device(me)
    layout = FlatLayout([torch.Size([300, 128]), torch.Size([128]), torch.Size([1024, 64])])
    buf = PeerBuffers(layout, me, with_state=True)
    g = torch.Generator().manual_seed(3)
    params = torch.randn(layout.total, generator=g) * 0.05
    buf.p32.copy_(params.cuda())
    buf.p16.copy_(params.cuda())
    buf.acc.copy_((torch.randn(layout.total, generator=g) * 0.01).cuda())
    torch.cuda.synchronize()
    open(os.path.join(args.dir, f"handle{me}.tmp"), "wb").write(buf.handle)
    os.replace(os.path.join(args.dir, f"handle{me}.tmp"), os.path.join(args.dir, f"handle{me}"))
    other = os.path.join(args.dir, f"handle{1 - me}")
    t0 = time.time()
    while not os.path.exists(other):
        time.sleep(0.01)
        assert time.time() - t0 < 60
    handles = [None, None]
    handles[me], handles[1 - me] = buf.handle, open(other, "rb").read()
    bases = [buf.peer_base(h) for h in handles]
    hyper = OptHyper(mode=lib.AR_LAMB, lr=1e-3, bias_correction=1.0, replicate_state=False)
    group = GroupSpec(bases=bases, me=me, weights=[1.0, 1.0], shards=[(0, 1), (0, 1)], slots=[0, 1], tag=4242,
                      timeout_s=2.0 if me == 0 else 60.0)
    if args.role == "victim":
        # park for ~12 s inside the kernel before doing anything (bit 7: exact delay, units of 16 us)
        open(os.path.join(args.dir, "victim_launching"), "w").write(str(os.getpid()))
        fused_allreduce_v2(buf, group, CompressionType.FLOAT16, hyper, debug_flags=((750_000 << 8) | 128) & 0x7FFFFFFF)
        torch.cuda.synchronize()          # never returns: the test kills this process
        return
    while not os.path.exists(os.path.join(args.dir, "victim_launching")):
        time.sleep(0.01)
    time.sleep(0.3)
    t0 = time.time()
    fused_allreduce_v2(buf, group, CompressionType.FLOAT16, hyper)
    torch.cuda.synchronize()
    dt = time.time() - t0
    status = buf.read_status()
    untouched = bool(torch.equal(buf.p32.cpu(), params)) and float(buf.m.abs().max()) == 0.0
    # the survivor steps on its own gradient (what CollaborativeOptimizer._run_solo_round does)
    from dedloc_b200.ops import flat_ops

    flat_ops.encode(CompressionType.NONE, buf.acc, 1.0, buf.send, buf.aux)
    solo = GroupSpec(bases=[buf.symm.ptr], me=0, weights=[1.0], shards=[(0, layout.n_chunks)], slots=[0], tag=7)
    fused_allreduce(buf, solo, CompressionType.NONE, hyper, xgpu_barriers=False)
    torch.cuda.synchronize()
    solo_status = buf.read_status()
    moved = not torch.equal(buf.p32.cpu(), params) and bool(torch.isfinite(buf.p32).all())
    json.dump({"seconds": dt, "status": status, "untouched": untouched, "solo_status": solo_status, "solo_moved": moved},
              open(os.path.join(args.dir, "survivor.json"), "w"))
    print("survivor done", dt, status, flush=True)
    os._exit(0)     # the victim's mapping may be gone: skip teardown


if __name__ == "__main__":
    main()
