"""BASELINE.json config 5: butterfly all-reduce sweep of the ALBERT-large gradient vector (1 MB - 1 GB of fp32
gradient) at 2/4/8 GPUs vs NCCL, in MEAN mode (plain weighted all-reduce) and in LAMB mode (the production call:
all-reduce + trust-ratio update on the owned shard + all-gather of the result).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_allreduce.py \
        [sizes_mb] [kernels: push,pull,nvls]

The vector keeps ALBERT-large's tensor structure (32 tensors; scaled down below 68 MB, repeated above), because LAMB
needs per-tensor norms. Two kernels: `push` = csrc/fused_allreduce_v2.cu (push-based, stage-pipelined, 2 bytes per
element leave the owner), `pull` = csrc/fused_allreduce.cu (P2P loads, barriers between phases; its LAMB mode sends
fp32 master + bf16 mirror = 6 bytes per element). Baseline = torch.distributed.all_reduce (NCCL) on a bf16 / fp32
tensor of the same element count -- which only averages, without the optimizer. Device-timed, max over ranks.

Roofline (B200_PROFILING.md): bytes that must cross NVLink per GPU and direction / link bandwidth. A butterfly
all-reduce sends (P-1)/P of the vector in the reduce-scatter AND in the all-gather, so per direction it is
(P-1)/P * n * (payload bytes + result bytes). Fractions are given against the measured 770 GB/s peer-copy bandwidth
and against the nominal 900 GB/s.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dedloc_b200 import lib as native
from dedloc_b200.ops import flat_ops
from dedloc_b200.parallel.flat import CHUNK, FlatLayout
from dedloc_b200.parallel.fused_allreduce import (GroupSpec, OptHyper, PeerBuffers, fused_allreduce,
                                                  fused_allreduce_v2)
from dedloc_b200.utils.compression import CompressionType, payload_bytes_per_elem

LINK_MEASURED_GBS, LINK_NOMINAL_GBS = 770.0, 900.0


def albert_shapes():
    from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining

    with torch.device("meta"):
        model = AlbertForPreTraining(AlbertConfig.albert_large_v2(), impl="torch")
    return [int(p.numel()) for p in model.parameters()]


def scaled_layout(numels, size_mb):
    """ALBERT-large's tensor list scaled to `size_mb` MB of fp32: shrunk (every tensor >= one chunk) or repeated."""
    full = sum((n + CHUNK - 1) // CHUNK * CHUNK for n in numels) * 4 / 2 ** 20
    if size_mb >= full * 0.99:
        reps = max(1, int(round(size_mb / full)))
        sizes = list(numels) * reps
    else:
        f = size_mb / full
        sizes = [max(CHUNK, int(n * f) // CHUNK * CHUNK) for n in numels]
        # tiny sizes: drop trailing tensors until the vector fits
        while len(sizes) > 1 and sum(sizes) * 4 / 2 ** 20 > size_mb * 1.25:
            sizes.pop()
    return FlatLayout([torch.Size([n]) for n in sizes])


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sizes_mb = [1, 4, 16, 64, 68, 256, 1024] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
    kernels = (sys.argv[2] if len(sys.argv) > 2 else "push,pull").split(",")
    mean_codecs = [CompressionType.FLOAT16, CompressionType.NONE, CompressionType.UNIFORM_8BIT,
                   CompressionType.BLOCKWISE_FP8]
    numels = albert_shapes()
    rows = []
    nvls_server = [None]
    tag = [1000]
    frac = (world - 1) / world

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

    for mb in sizes_mb:
        layout = scaled_layout(numels, mb)
        n = layout.total
        buf = PeerBuffers(layout, local, with_state=True)
        handles = [None] * world
        dist.all_gather_object(handles, buf.handle)
        bases = [buf.peer_base(h) for h in handles]
        buf.acc.copy_(torch.randn(n, device=dev) * 0.01)
        buf.p32.copy_(torch.randn(n, device=dev) * 0.05)
        buf.p16.copy_(buf.p32)
        x16 = torch.randn(n, device=dev).to(torch.bfloat16)
        x32 = torch.randn(n, device=dev)
        owner_mask = [(0, 1)] * world
        contiguous = layout.shard_ranges([1.0] * world)
        res = {"size_mb_fp32": round(n * 4 / 2 ** 20, 1), "elements": n, "tensors": layout.n_tensors, "gpus": world}
        res["nccl_bf16_ms"] = round(timed(lambda: dist.all_reduce(x16)), 4)
        res["nccl_fp32_ms"] = round(timed(lambda: dist.all_reduce(x32)), 4)

        def entry(t_ms, bytes_per_elem_per_dir, extra=None):
            moved = frac * n * bytes_per_elem_per_dir
            out = {"ms": round(t_ms, 4),
                   "of_nvlink_measured_770": round(moved / (LINK_MEASURED_GBS * 1e9) * 1e3 / t_ms, 3),
                   "of_nvlink_nominal_900": round(moved / (LINK_NOMINAL_GBS * 1e9) * 1e3 / t_ms, 3),
                   "busbw_gbs": round(moved / t_ms / 1e6, 1),
                   "vs_nccl_bf16": round(res["nccl_bf16_ms"] / t_ms, 3)}
            if extra:
                out.update(extra)
            return out

        def run(kernel, mode, codec, encode_first):
            tag[0] += 1
            hyper = OptHyper(mode=mode, lr=1e-4, bias_correction=1.0, replicate_state=False)
            if kernel == "push":
                fly = codec in (CompressionType.NONE, CompressionType.FLOAT16, CompressionType.BFLOAT16)
                if not fly and encode_first:
                    flat_ops.encode(codec, buf.acc, 1.0, buf.send, buf.aux)
                g = GroupSpec(bases=bases, me=rank, weights=[1.0] * world, shards=owner_mask,
                              slots=list(range(world)), tag=tag[0], timeout_s=20.0)
                fused_allreduce_v2(buf, g, codec, hyper, src_f32=True, scale=1.0, mean_out32=False, write_p16=True)
            else:
                if encode_first:
                    flat_ops.encode(codec, buf.acc, 1.0, buf.send, buf.aux)
                g = GroupSpec(bases=bases, me=rank, weights=[1.0] * world, shards=contiguous,
                              slots=list(range(world)), tag=tag[0], timeout_s=20.0)
                fused_allreduce(buf, g, codec, hyper, mean_out32=False, write_p16=True)

        for mode_name, mode, codecs in (("MEAN", native.AR_MEAN, mean_codecs),
                                        ("LAMB", native.AR_LAMB, [CompressionType.FLOAT16, CompressionType.UNIFORM_8BIT])):
            for codec in codecs:
                pb = payload_bytes_per_elem(codec)
                flat_ops.encode(codec, buf.acc, 1.0, buf.send, buf.aux)
                key = f"{mode_name}_{codec.name}"
                fly = codec in (CompressionType.NONE, CompressionType.FLOAT16, CompressionType.BFLOAT16)
                # push kernel: the encoder runs inside the kernel for the 16/32-bit codecs, so there is one number;
                # the 8-bit codecs need their two-pass encoder first
                if "push" in kernels:
                    t_push = timed(lambda: run("push", mode, codec, encode_first=not fly))
                    res[key + "_push"] = entry(t_push, pb + 2, {"includes_encode": True})
                    assert buf.read_status()["code"] == 0, buf.read_status()
                if not fly and "push" in kernels:
                    t_push_pre = timed(lambda: run("push", mode, codec, encode_first=False))
                    res[key + "_push"]["ms_payload_pre_encoded"] = round(t_push_pre, 4)
                if "pull" not in kernels:
                    continue
                out_bytes = 2 if mode == native.AR_MEAN else 6       # pull kernel, LAMB: fp32 master + bf16 mirror
                t_pull = timed(lambda: run("pull", mode, codec, encode_first=False))
                t_pull_enc = timed(lambda: run("pull", mode, codec, encode_first=True))
                res[key + "_pull"] = entry(t_pull_enc, pb + out_bytes,
                                           {"includes_encode": True, "ms_payload_pre_encoded": round(t_pull, 4)})
                assert buf.read_status()["code"] == 0, buf.read_status()
        if "nvls" in kernels:
            # in-switch reduction (csrc/nvls_allreduce.cu): fp32 -> fp16 encode + multimem all-reduce, bf16 result
            from dedloc_b200.parallel import nvls

            if nvls.multicast_supported(local):
                if nvls_server[0] is None:
                    nvls_server[0] = nvls.FdServer()
                ar = nvls.NvlsAllreduce(n, local, nvls_server[0])
                ids = [None] * world
                dist.all_gather_object(ids, nvls_server[0].path + "#" + ar.window.key)
                grp = ar.group([s.split("#")[0] for s in ids], rank, salt="|".join(s.split("#")[1] for s in ids))
                t_nvls = timed(lambda: ar.allreduce(buf.acc, 1.0 / world, grp, out16=buf.p16))
                # per GPU and direction: N/P reduced elements in + the other members' pulls out, and the mirror image
                # for the result: about (P+1)/P * n * 2 B; the fraction below uses the peer-to-peer byte count so that
                # the columns stay comparable (it can exceed 1)
                res["MEAN_FLOAT16_nvls"] = entry(t_nvls, 4, {"includes_encode": True})
                assert int(ar.status[0]) == 0
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
        del buf
        torch.cuda.empty_cache()
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(rows, open(f"gpurun_out/allreduce_sweep_n{world}_{'_'.join(kernels)}.json", "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
