"""NVLS (multimem) all-reduce on N GPUs: correctness against a torch reference and device time against NCCL.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/test_nvls.py [sizes_mb]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from dedloc_b200.parallel import nvls

    sizes = [4, 68, 256] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
    out = {"gpus": world, "multicast_supported": nvls.multicast_supported(local), "rows": []}
    if not out["multicast_supported"]:
        if rank == 0:
            print(json.dumps(out))
        return
    server = nvls.FdServer()
    for mb in sizes:
        n = (mb * (1 << 20) // 4 + 2047) // 2048 * 2048
        ar = nvls.NvlsAllreduce(n, local, server)
        socks = [None] * world
        dist.all_gather_object(socks, server.path + "#" + ar.window.key)
        group = ar.group([s.split("#")[0] for s in socks], rank, salt="|".join(s.split("#")[1] for s in socks))
        torch.manual_seed(100 + rank)
        src = torch.randn(n, device=dev) * 0.01
        weights = [(r + 1.0) for r in range(world)]
        wsum = sum(weights)
        out16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)
        out32 = torch.zeros(n, dtype=torch.float32, device=dev)
        ar.allreduce(src, weights[rank] / wsum, group, out16=out16, out32=out32)
        torch.cuda.synchronize()
        ref = (src * (weights[rank] / wsum)).half().float()
        dist.all_reduce(ref)
        err = float((out32 - ref).abs().max())
        scale = float(ref.abs().max())
        ok = int(ar.status[0]) == 0 and err <= 2e-2 * scale and torch.equal(out16.float(), out32)

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

        x16 = torch.randn(n, device=dev).to(torch.bfloat16)
        t_nccl = timed(lambda: dist.all_reduce(x16))
        t_nvls = timed(lambda: ar.allreduce(src, weights[rank] / wsum, group, out16=out16))
        t_nvls_win = timed(lambda: ar.allreduce(src, weights[rank] / wsum, group))
        row = {"size_mb_fp32": mb, "elements": n, "ok": bool(ok), "max_abs_err": err, "ref_max": scale,
               "status": int(ar.status[0]), "nccl_bf16_ms": round(t_nccl, 4),
               "nvls_encode_allreduce_copy_ms": round(t_nvls, 4), "nvls_encode_allreduce_ms": round(t_nvls_win, 4)}
        out["rows"].append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    ok_all = all(r["ok"] for r in out["rows"])
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(out, open(f"gpurun_out/nvls_n{world}.json", "w"), indent=1)
        print("NVLS", "OK" if ok_all else "FAILED")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
