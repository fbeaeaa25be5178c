#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): ALBERT-large collaborative MLM, seq 512, LAMB, target_batch_size 4096.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run)

One rank = one DeDLOC peer on one B200. A "step" is one COLLABORATIVE optimizer step: the peers accumulate
micro-batches until the collaboration has 4096 samples, form a group through the DHT, run the fused
all-reduce + LAMB kernel over NVLink peer memory and advance together. Rank 0 prints ONE JSON line.
`value` = samples/s of the whole job, device-timed (CUDA events, max over ranks) with inputs resident on the
device; `e2e` = the same metric through the public API with the per-micro-batch host->device copy from pinned
memory and the per-step device->host read of the loss inside the timed region (wall clock, max over ranks).
`--impl reference` reports why the unmodified reference cannot run here; `--impl ddp` times the torch
DDP/NCCL + HF-transformers build of the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp"])
    ap.add_argument("--micro-batch", type=int, default=128,
                    help="samples per device per micro-step (128: +4 %% over 64 at N=1 -- larger GEMM M, fewer launches; 256 makes "
                         "rounds of only 2 micro-steps at N=8 and overshoots the target batch)")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--target-batch-size", type=int, default=4096)
    ap.add_argument("--compression", default="FLOAT16")
    ap.add_argument("--model", default="albert-large-v2",
                    choices=["albert-large-v2", "sahajbert", "albert-small", "albert-tiny"],
                    help="sahajbert = ALBERT-large with the 32 000-entry Bengali vocabulary (BASELINE.json config 3)")
    ap.add_argument("--ops", default="native", choices=["native", "torch"], help="linear layers: tcgen05 or torch")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--padded", action="store_true",
                    help="instances of random length in [S/2, S] with padding and a real attention mask (the masked "
                         "attention kernel) instead of full-length sequences")
    ap.add_argument("--ddp-micro-batch", type=int, default=0,
                    help="micro-batch of the --impl ddp arm (0 = the same as ours, halved until it fits in memory)")
    ap.add_argument("--ddp-variant", default="stock", choices=["stock", "fp16hook", "best"],
                    help="--impl ddp: stock DDP + autocast, DDP + fp16_compress_hook, or the best torch-only configuration")
    ap.add_argument("--no-cuda-graph", action="store_true", help="run the micro-step eagerly")
    ap.add_argument("--overlap-wgrad", action="store_true", help="launch weight/bias gradient kernels on a side stream")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def main_reference(args):
    print(json.dumps({
        "impl": "reference",
        "unavailable": "pip install of /root/reference fails offline (no setup.py/pyproject.toml) and its hard "
                       "dependency hivemind==0.9.9.post1 (plus torch_optimizer) is in neither the image nor /opt/wheelhouse",
    }))
    return 0


def committed_baseline(world: int, args):
    """The reference publishes no throughput, so `vs_baseline` is taken against the committed
    same-box torch DDP/NCCL line of the same config (profiles/ddp_n{N}.json, best torch-only variant)."""
    if args.model != "albert-large-v2" or args.target_batch_size != 4096 or args.seq_len != 512:
        return None
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", f"ddp_n{world}.json")
    try:
        rows = [json.loads(l) for l in open(path) if l.strip().startswith("{")]
        row = max(rows, key=lambda r: r["value"])
        return {"value": row["value"], "impl": "ddp", "variant": row.get("variant"), "file": f"profiles/ddp_n{world}.json"}
    except Exception:
        return None


def main_ours(args):
    import torch
    import torch.distributed as dist

    from dedloc_b200 import lib as native
    from dedloc_b200.dht import DHT
    from dedloc_b200.models.albert import AlbertConfig
    from dedloc_b200.trainers.pretrain import PeerConfig, SyntheticMLMBatches, build_albert_peer

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    native.require_native_on_gpu()
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- rendezvous of the DHT (control plane). torch.distributed only carries the bootstrap port + timings.
    if rank == 0:
        dht = DHT(start=True)
        port = [dht.port]
    else:
        port = [None]
    if world > 1:
        dist.broadcast_object_list(port, src=0)
    if rank != 0:
        dht = DHT(start=True, initial_peers=[f"127.0.0.1:{port[0]}"])

    model_cfg = {"albert-large-v2": AlbertConfig.albert_large_v2, "albert-small": AlbertConfig.small,
                 "albert-tiny": AlbertConfig.tiny,
                 "sahajbert": lambda: AlbertConfig.albert_large_v2(vocab_size=32000)}[args.model]()
    seq = min(args.seq_len, model_cfg.max_position_embeddings)
    cfg = PeerConfig(prefix="bench", target_batch_size=args.target_batch_size, micro_batch=args.micro_batch,
                     seq_len=seq, compression=args.compression, target_group_size=max(world, 2),
                     averaging_expiration=5.0, averaging_timeout=60.0, impl=args.ops,
                     kernel_timeout=30.0,     # no faults in this run: a late member is start-up skew, not a dead peer
                     overlap_wgrad=args.overlap_wgrad)
    trainer = build_albert_peer(dht, cfg, model_cfg, device=device, seed=0)
    data = SyntheticMLMBatches(model_cfg.vocab_size, seq, args.micro_batch, seed=1234 + rank, padded=args.padded)
    resident = [data.to_device(b, device) for b in data.batches]   # device-resident copies for the kernel-timed arm
    graphed = False if args.no_cuda_graph else trainer.enable_cuda_graph(resident[0])
    barrier()

    def run_global_step(e2e: bool) -> int:
        """Run micro-steps until this peer takes part in a collaborative step. Returns samples processed."""
        n0 = trainer.samples_processed
        i = 0
        while True:
            if e2e:
                # H2D from pinned memory, every micro-batch (straight into the graph's static inputs when graphed)
                batch = data.next_host() if graphed else data.to_device(data.next_host(), device)
            else:
                batch = resident[i % len(resident)]
            i += 1
            if trainer.micro_step(batch):
                break
        if e2e:
            trainer.pop_mean_loss()                                 # D2H read of the step's result
        return trainer.samples_processed - n0

    def timed_region(steps: int, e2e: bool):
        barrier()
        launches0 = native.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ev0.record()
        samples = 0
        for _ in range(steps):
            samples += run_global_step(e2e)
        ev1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        dev_ms = ev0.elapsed_time(ev1)
        barrier()
        stats = torch.tensor([dev_ms, wall * 1e3, float(samples), float(native.launch_count() - launches0)],
                             dtype=torch.float64, device=device)
        if world > 1:
            mx = stats.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = stats.clone()
            dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        else:
            mx, sm = stats, stats
        return {"dev_ms": float(mx[0]), "wall_ms": float(mx[1]), "samples": float(sm[2]), "launches": int(sm[3])}

    for _ in range(args.warmup):
        run_global_step(False)
    trainer.pop_mean_loss()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev = timed_region(args.steps, e2e=False)
    clocks = sampler.stop() if rank == 0 else None
    e2e = None
    if not args.skip_e2e:
        run_global_step(True)  # warm the pinned-memory path once
        e2e = timed_region(args.steps, e2e=True)

    collab = trainer.collab
    # all peers must hold bit-identical master parameters after the same sequence of collaborative steps
    torch.cuda.synchronize()
    p32 = trainer.opt.buffers.p32
    digest = torch.stack([p32.double().sum(), p32.double().abs().sum(), (p32.double() ** 2).sum()])
    in_sync = True
    if world > 1:
        gathered = [torch.zeros_like(digest) for _ in range(world)]
        dist.all_gather(gathered, digest)
        in_sync = all(bool(torch.equal(g, gathered[0])) for g in gathered)
    if rank == 0:
        value = dev["samples"] / (dev["dev_ms"] / 1e3)
        baseline = committed_baseline(world, args)
        out = {
            "metric": "ALBERT-large collaborative MLM samples/sec", "value": round(value, 2), "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dev["dev_ms"] / args.steps, 3), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": round(value / baseline["value"], 3) if baseline else None,
            "baseline": baseline, "dtype": "bf16", "data": "synthetic", "impl": "ours",
            "config": {"model": args.model, "global_batch": args.target_batch_size, "seq_len": seq,
                       "parallelism": f"dedloc-dp{world}", "optimizer": "LAMB (fused in all-reduce kernel)",
                       "micro_batch": args.micro_batch, "compression": args.compression, "linear_ops": args.ops,
                       "padded_inputs": bool(args.padded),
                       "cuda_graph_micro_step": bool(graphed),
                       "samples_per_step_measured": round(dev["samples"] / args.steps, 1),
                       "l2": "inputs larger than L2 (activations of one micro-batch >> 126 MB)",
                       "step": "one collaborative optimizer step (accumulate to target_batch_size, fused "
                               "all-reduce + LAMB)"},
            "clocks": clocks, "gpu_launches": dev["launches"], "params_in_sync": in_sync,
            "local_step": collab.local_step,
            "collab": {"rounds": collab.stats["rounds"], "solo_rounds": collab.stats["solo_rounds"],
                       "failed_rounds": collab.stats["failed_rounds"],
                       "matchmaking_ms_per_round": round(1e3 * collab.stats["matchmaking_s"] / max(collab.stats["rounds"], 1), 2),
                       "exposed_averaging_device_ms_per_round": round(collab.stats.get("averaging_device_ms", 0.0) / max(collab.stats.get("fused_rounds", 0), 1), 3),
                       "averaging_host_ms_per_round": round(1e3 * collab.stats["averaging_host_s"] / max(collab.stats["rounds"], 1), 2)},
        }
        if e2e is not None:
            out["e2e"] = {"value": round(e2e["samples"] / (e2e["wall_ms"] / 1e3), 2), "unit": "samples/s",
                          "h2d_bytes_per_step": int(data.bytes_per_batch * e2e["samples"] / args.steps / args.micro_batch / world),
                          "d2h_bytes_per_step": 4, "timing": "wall clock incl. H2D of every micro-batch from pinned "
                                                             "memory and D2H of the loss every step; max over ranks"}
        print(json.dumps(out), flush=True)
    collab.shutdown()
    dht.shutdown()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return main_reference(args)
    if args.impl == "ddp":
        from baseline.ddp_baseline import main_ddp

        return main_ddp(args)
    return main_ours(args)


if __name__ == "__main__":
    sys.exit(main())
