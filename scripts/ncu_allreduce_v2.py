"""One GPU, P emulated members of the push-based all-reduce + LAMB kernel on the ALBERT-large vector: the launch that
`ncu --set full -k regex:ar2_kernel` captures (a multi-rank command must not run under ncu; the emulated launch is the
production device code with every member on one device)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200 import lib
from dedloc_b200.parallel.fused_allreduce import OptHyper, PeerBuffers, fused_allreduce_v2_emulated
from dedloc_b200.utils.compression import CompressionType
from scripts.bench_allreduce import albert_shapes, scaled_layout


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    mode = {"MEAN": lib.AR_MEAN, "LAMB": lib.AR_LAMB}[sys.argv[2] if len(sys.argv) > 2 else "LAMB"]
    layout = scaled_layout(albert_shapes(), 68)
    bufs = [PeerBuffers(layout, 0, with_state=True) for _ in range(P)]
    for r, b in enumerate(bufs):
        g = torch.Generator(device="cuda").manual_seed(r)
        b.acc.copy_(torch.randn(layout.total, device="cuda", generator=g) * 0.01)
        b.p32.copy_(torch.randn(layout.total, device="cuda", generator=torch.Generator(device="cuda").manual_seed(99)) * 0.05)
    hyper = OptHyper(mode=mode, lr=1e-4, bias_correction=1.0, replicate_state=False)
    ev = []
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fused_allreduce_v2_emulated(bufs, [1.0] * P, [(0, 1)] * P, CompressionType.FLOAT16, hyper, tag=500 + i)
        e1.record()
        torch.cuda.synchronize()
        ev.append(round(e0.elapsed_time(e1), 4))
    assert all(b.read_status()["code"] == 0 for b in bufs)
    assert all(torch.equal(b.p32, bufs[0].p32) for b in bufs)
    print({"emulated_members": P, "elements": layout.total, "launch_ms": ev})


if __name__ == "__main__":
    main()
