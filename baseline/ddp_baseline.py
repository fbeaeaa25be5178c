"""the torch DDP / NCCL (+ cuBLAS, SDPA) build of the SAME config on the same box.

HF `transformers.AlbertForPreTraining` (the model class the reference instantiates, albert/run_trainer.py:64-68),
the restated LAMB of optim/lamb.py (torch_optimizer is not installable offline), bf16 autocast,
DistributedDataParallel with gradient accumulation under no_sync() up to the global batch of 4096, grad-norm
clipping 1.0. None of this repo's kernels or engine is on this path. Same JSON line as bench.py's own arm with
"impl": "ddp".
"""
from __future__ import annotations

import json
import os
import time
from contextlib import nullcontext


def main_ddp(args) -> int:
    import torch
    import torch.distributed as dist
    import transformers
    from torch.nn.parallel import DistributedDataParallel as DDP

    from dedloc_b200.optim import Lamb, get_linear_schedule_with_warmup
    from dedloc_b200.trainers.pretrain import SyntheticMLMBatches, albert_param_groups

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    hf_cfg = transformers.AlbertConfig(
        vocab_size=30000, embedding_size=128, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
        intermediate_size=4096, hidden_act="gelu_new", hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
        max_position_embeddings=512, type_vocab_size=2)
    torch.manual_seed(0)
    model = transformers.AlbertForPreTraining(hf_cfg).to(device).train()
    opt = Lamb(albert_param_groups(model, 0.01), lr=0.00176, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01,
               clamp_value=10000.0, debias=True)
    sched = get_linear_schedule_with_warmup(opt, 5000, 125000)
    ddp = DDP(model, device_ids=[local_rank]) if world > 1 else model
    # the baseline was measured at 64 samples per micro-step (HF's eager ALBERT keeps far more activations than the
    # native path); larger values are honoured when they are asked for explicitly
    seq, mb = args.seq_len, min(args.micro_batch, getattr(args, "ddp_micro_batch", 64) or 64)
    data = SyntheticMLMBatches(hf_cfg.vocab_size, seq, mb, seed=1234 + rank)
    resident = [data.to_device(b, device) for b in data.batches]
    micro_per_step = max(1, args.target_batch_size // (mb * world))

    def step(e2e: bool):
        loss_acc = torch.zeros((), device=device)
        for i in range(micro_per_step):
            batch = data.to_device(data.next_host(), device) if e2e else resident[i % len(resident)]
            sync = (i == micro_per_step - 1) or world == 1
            ctx = nullcontext() if sync else ddp.no_sync()
            with ctx, torch.autocast("cuda", dtype=torch.bfloat16):
                out = ddp(**batch)
            with ctx:
                (out.loss / micro_per_step).backward()
            loss_acc += out.loss.detach()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        sched.step()
        opt.zero_grad(set_to_none=True)
        if e2e:
            return float(loss_acc.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(steps, e2e):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            step(e2e)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        t = torch.tensor([e0.elapsed_time(e1), wall * 1e3], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    for _ in range(args.warmup):
        step(False)
    dev_ms, _ = timed(args.steps, False)
    _, wall_ms = timed(args.steps, True)
    samples = micro_per_step * mb * world * args.steps
    if rank == 0:
        print(json.dumps({
            "metric": "ALBERT-large collaborative MLM samples/sec", "value": round(samples / (dev_ms / 1e3), 2),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dev_ms / args.steps, 3), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ddp",
            "config": {"model": "albert-large-v2 (HF transformers)", "global_batch": micro_per_step * mb * world,
                       "seq_len": seq, "parallelism": f"ddp{world}", "optimizer": "LAMB (python, restated)",
                       "micro_batch": mb},
            "e2e": {"value": round(samples / (wall_ms / 1e3), 2), "unit": "samples/s"},
            "gpu_launches": 0}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0
