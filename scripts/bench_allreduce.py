"""BASELINE.json config 5: butterfly all-reduce sweep (1 MB - 1 GB of fp32 gradient) at 2/4/8 GPUs vs NCCL.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_allreduce.py

Ours = the MEAN mode of csrc/fused_allreduce.cu over CUDA-IPC peer memory (reduce-scatter by P2P loads of the
encoded payload + all-gather by P2P stores of the bf16 result), for several wire codecs; baseline =
torch.distributed.all_reduce (NCCL) on a bf16 / fp32 tensor of the same element count. Device-timed, max over ranks.
Roofline per the profiling guide: bytes that must cross NVLink per GPU and direction / 770 GB/s (measured peer copy).
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dedloc_b200.ops import flat_ops
from dedloc_b200.parallel.flat import FlatLayout
from dedloc_b200.parallel.fused_allreduce import GroupSpec, OptHyper, PeerBuffers, fused_allreduce
from dedloc_b200.utils.compression import CompressionType, payload_bytes_per_elem

LINK_GBS = 770.0


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sizes_mb = [1, 4, 16, 64, 68, 256, 1024] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
    codecs = [CompressionType.FLOAT16, CompressionType.NONE, CompressionType.UNIFORM_8BIT, CompressionType.BLOCKWISE_FP8]
    rows = []
    tag = 100
    for mb in sizes_mb:
        n = mb * (1 << 20) // 4                      # elements of the fp32 gradient vector
        n = (n + 2047) // 2048 * 2048
        layout = FlatLayout([torch.Size([n])])
        buf = PeerBuffers(layout, local, with_state=False)
        handles = [None] * world
        dist.all_gather_object(handles, buf.handle)
        bases = [buf.peer_base(h) for h in handles]
        buf.acc.copy_(torch.randn(n, device=dev) * 0.01)
        shards = layout.shard_ranges([1.0] * world)
        x16 = torch.randn(n, device=dev).to(torch.bfloat16)
        x32 = torch.randn(n, device=dev)

        def timed(fn, iters=10, warmup=3):
            for _ in range(warmup):
                fn()
            ts = []
            for _ in range(iters):
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1)], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ts.append(float(t))
            ts.sort()
            return ts[len(ts) // 2]

        res = {"size_mb_fp32": mb, "elements": n, "gpus": world}
        res["nccl_bf16_ms"] = timed(lambda: dist.all_reduce(x16))
        res["nccl_fp32_ms"] = timed(lambda: dist.all_reduce(x32))
        for codec in codecs:
            def ours(encode):
                nonlocal tag
                tag += 1
                if encode:
                    flat_ops.encode(codec, buf.acc, 1.0, buf.send, buf.aux)
                g = GroupSpec(bases=bases, me=rank, weights=[1.0] * world, shards=shards, slots=list(range(world)),
                              tag=tag, timeout_s=20.0)
                fused_allreduce(buf, g, codec, OptHyper(mode=0), mean_out32=False, write_p16=True)

            flat_ops.encode(codec, buf.acc, 1.0, buf.send, buf.aux)
            t_ar = timed(lambda: ours(False))
            t_full = timed(lambda: ours(True))
            pb = payload_bytes_per_elem(codec)
            frac = (world - 1) / world
            roof_ms = frac * n * max(pb, 2) / (LINK_GBS * 1e9) * 1e3    # inbound payload vs outbound bf16, full duplex
            res[codec.name] = {"allreduce_ms": round(t_ar, 4), "encode_plus_allreduce_ms": round(t_full, 4),
                               "roofline_ms": round(roof_ms, 4), "fraction_of_nvlink_roofline": round(roof_ms / t_ar, 3),
                               "vs_nccl_bf16": round(res["nccl_bf16_ms"] / t_ar, 3)}
            assert buf.read_status()["code"] == 0
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
        del buf
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(rows, open(f"gpurun_out/allreduce_sweep_n{world}.json", "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
