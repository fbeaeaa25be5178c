"""Per-stage timeline of the push-based all-reduce kernel (csrc/fused_allreduce_v2.cu) on N GPUs.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/trace_allreduce.py \
        [size_mb] [mode: MEAN|LAMB] [stages or 0] [debug_flags] [sequential|concurrent]

CTA 0 of every role stamps %globaltimer at the start of a stage, after its work, after its system-scope fence and
when the stage's flag goes out. Prints, per role and stage, the offsets from the first stamp in microseconds, and the
device time of the launch -- the evidence for where a round spends its time.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dedloc_b200 import lib as native
from dedloc_b200.parallel.fused_allreduce import GroupSpec, OptHyper, PeerBuffers, fused_allreduce_v2
from dedloc_b200.utils.compression import CompressionType
from scripts.bench_allreduce import albert_shapes, scaled_layout


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    mb = int(sys.argv[1]) if len(sys.argv) > 1 else 68
    mode = {"MEAN": native.AR_MEAN, "LAMB": native.AR_LAMB}[sys.argv[2] if len(sys.argv) > 2 else "MEAN"]
    if len(sys.argv) > 3 and int(sys.argv[3]) > 0:
        os.environ["DEDLOC_AR2_STAGES"] = sys.argv[3]
    debug = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    if len(sys.argv) > 5:
        os.environ["DEDLOC_AR2_SCHEDULE"] = sys.argv[5]
    layout = scaled_layout(albert_shapes(), mb)
    buf = PeerBuffers(layout, local, with_state=True)
    handles = [None] * world
    dist.all_gather_object(handles, buf.handle)
    bases = [buf.peer_base(h) for h in handles]
    buf.acc.copy_(torch.randn(layout.total, device=dev) * 0.01)
    buf.p32.copy_(torch.randn(layout.total, device=dev) * 0.05)
    n_role = 4 * native.AR2_MAX_STAGES * 4
    trace = torch.zeros(n_role + 1024 * 10, dtype=torch.int64, device=dev)
    tag = 7000
    times = []
    for it in range(6):
        tag += 1
        g = GroupSpec(bases=bases, me=rank, weights=[1.0] * world, shards=[(0, 1)] * world, slots=list(range(world)),
                      tag=tag, timeout_s=20.0)
        dist.barrier()
        torch.cuda.synchronize()
        trace.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fused_allreduce_v2(buf, g, CompressionType.FLOAT16, OptHyper(mode=mode, lr=1e-4, replicate_state=False),
                           mean_out32=False, trace=trace, debug_flags=debug)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    st = buf.read_status()
    if rank == 0:
        full = trace.cpu()
        t = full[:n_role].view(4, native.AR2_MAX_STAGES, 4).numpy()
        per_cta = full[n_role:].view(1024, 10).numpy()
        t0 = t[t > 0].min()
        names = ["PUSH  begin/stored/fenced/flag", "REDUCE wait/flag seen/done/last", "UPDATE begin/done/fenced/flag",
                 "APPLY wait/flags seen/done"]
        print(json.dumps({"size_mb": mb, "gpus": world, "mode": sys.argv[2] if len(sys.argv) > 2 else "MEAN",
                          "launch_ms": [round(x, 4) for x in times], "status": st, "debug_flags": debug,
                          "stages_env": os.environ.get("DEDLOC_AR2_STAGES"),
                          "schedule": os.environ.get("DEDLOC_AR2_SCHEDULE", "sequential")}))
        for r in range(4):
            print(names[r])
            for s in range(native.AR2_MAX_STAGES):
                if t[r, s].max() > 0:
                    print(f"  s{s:02d} " + " ".join(f"{(x - t0) / 1e3:8.1f}" if x > 0 else "       -" for x in t[r, s]))
        import numpy as np

        live = per_cta[per_cta[:, 1] > 0]
        if len(live):
            print(f"PUSH per CTA (warp 0), {len(live)} CTAs, us from first stamp: stage  stored min/med/p90/max   fenced min/med/p90/max")
            for s in range(4):
                st, fe = live[:, 1 + s], live[:, 5 + s]
                if st.max() <= 0:
                    continue
                st, fe = (st[st > 0] - t0) / 1e3, (fe[fe > 0] - t0) / 1e3
                q = lambda a: " ".join(f"{x:7.1f}" for x in (a.min(), np.median(a), np.percentile(a, 90), a.max()))
                print(f"  s{s}  {q(st)}   |  {q(fe)}")
            # slowest CTAs of the last traced stage, with their SM
            last = max(s for s in range(4) if live[:, 5 + s].max() > 0)
            order = np.argsort(live[:, 5 + last])
            fmt = lambda i: f"(sm {int(live[i, 0])}: {(live[i, 5 + last] - t0) / 1e3:.1f})"
            print("  fastest:", " ".join(fmt(i) for i in order[:6]), " slowest:", " ".join(fmt(i) for i in order[-6:]))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
