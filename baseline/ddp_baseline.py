"""the torch DDP / NCCL (+ cuBLAS, SDPA) build of the SAME config on the same box.

HF `transformers.AlbertForPreTraining` (the model class the reference instantiates, albert/run_trainer.py:64-68),
the restated LAMB of optim/lamb.py (torch_optimizer is not installable offline), gradient accumulation up to the
global batch of 4096, grad-norm clipping 1.0. None of this repo's kernels or engine is on this path. Same JSON line
as bench.py's own arm with "impl": "ddp". Three variants (`--ddp-variant`):

  stock      fp32 parameters + torch.autocast(bf16), DistributedDataParallel with no_sync() accumulation -- what
             classy_vision/generic/distributed_util.py:219-244 builds (B1)
  fp16hook   stock + DDP's fp16_compress_hook (classy_vision/tasks/classification_task.py:876-886) (B2)
  best       the fairest torch can do without custom kernels: bf16 parameters (no autocast casts) with fp32 master
             copies inside LAMB, SDPA attention, the micro-step captured in a CUDA graph when the HF module allows it,
             one NCCL all-reduce of the flat bf16 gradient per optimizer step
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
    variant = getattr(args, "ddp_variant", "stock")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    vocab = 32000 if getattr(args, "model", "") == "sahajbert" else 30000
    hf_cfg = transformers.AlbertConfig(
        vocab_size=vocab, embedding_size=128, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
        intermediate_size=4096, hidden_act="gelu_new", hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
        max_position_embeddings=512, type_vocab_size=2)
    torch.manual_seed(0)
    model = transformers.AlbertForPreTraining(hf_cfg).to(device).train()
    seq = args.seq_len
    notes = []

    best = variant == "best"
    masters = None
    if best:
        torch.backends.cuda.matmul.allow_tf32 = True
        masters = [p.detach().clone().float().requires_grad_(True) for p in model.parameters()]
        for m, (n, _) in zip(masters, model.named_parameters()):
            m._dedloc_name = n
        model = model.to(torch.bfloat16)
        no_decay = ["bias", "LayerNorm.weight", "layer_norm.weight"]
        named = [(n, m) for (n, _), m in zip(model.named_parameters(), masters)]
        groups = [{"params": [m for n, m in named if not any(nd in n for nd in no_decay)], "weight_decay": 0.01},
                  {"params": [m for n, m in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0}]
    else:
        groups = albert_param_groups(model, 0.01)
    opt = Lamb(groups, lr=0.00176, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, clamp_value=10000.0, debias=True)
    sched = get_linear_schedule_with_warmup(opt, 5000, 125000)
    if best or world == 1:
        ddp = model
    else:
        ddp = DDP(model, device_ids=[local_rank])
        if variant == "fp16hook":
            from torch.distributed.algorithms.ddp_comm_hooks import default_hooks

            ddp.register_comm_hook(None, default_hooks.fp16_compress_hook)

    # micro-batch: the same as our arm when it fits, else halve until it does (and say so)
    mb = int(getattr(args, "ddp_micro_batch", 0) or args.micro_batch)

    def fwd_bwd(batch, scale, sync=True):
        if best:
            out = ddp(**batch)
            (out.loss * scale).backward()
        else:
            with (nullcontext() if (sync or ddp is model) else ddp.no_sync()):
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    out = ddp(**batch)
                (out.loss * scale).backward()
        return out.loss.detach()

    while True:
        try:
            data = SyntheticMLMBatches(hf_cfg.vocab_size, seq, mb, seed=1234 + rank)
            resident = [data.to_device(b, device) for b in data.batches]
            fwd_bwd(resident[0], 1.0)
            torch.cuda.synchronize()
            model.zero_grad(set_to_none=True)
            break
        except torch.OutOfMemoryError:
            model.zero_grad(set_to_none=True)
            torch.cuda.empty_cache()
            notes.append(f"micro-batch {mb} ran out of memory")
            mb //= 2
            if mb < 1:
                raise
    micro_per_step = max(1, args.target_batch_size // (mb * world))

    # best: CUDA-graph the micro-step when the HF module captures
    graph, static = None, None
    if best:
        try:
            static = {k: v.clone() for k, v in resident[0].items()}
            side = torch.cuda.Stream(device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    fwd_bwd(static, 1.0 / micro_per_step)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            loss_static = torch.zeros((), device=device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                loss_static.copy_(fwd_bwd(static, 1.0 / micro_per_step))
            graph = g
            notes.append("micro-step captured in a CUDA graph")
        except Exception as e:  # pragma: no cover
            graph = None
            torch.cuda.synchronize()
            notes.append(f"CUDA-graph capture failed ({type(e).__name__}); eager")
        for p in model.parameters():
            if p.grad is not None:
                p.grad.zero_()

    def step(e2e: bool):
        loss_acc = torch.zeros((), device=device)
        for i in range(micro_per_step):
            batch = data.next_host() if e2e else resident[i % len(resident)]
            if graph is not None:
                for k, v in batch.items():
                    static[k].copy_(v, non_blocking=True)
                graph.replay()
                loss_acc += loss_static
            else:
                if e2e:
                    batch = data.to_device(batch, device)
                loss_acc += fwd_bwd(batch, 1.0 / micro_per_step, sync=(i == micro_per_step - 1))
        if best:
            grads = [p.grad for p in model.parameters()]
            if world > 1:
                flat = torch.cat([g.reshape(-1) for g in grads])
                dist.all_reduce(flat)
                flat.div_(world)
                off = 0
                for g in grads:
                    g.copy_(flat[off:off + g.numel()].view_as(g))
                    off += g.numel()
            for m, g in zip(masters, grads):
                m.grad = g.float()
            torch.nn.utils.clip_grad_norm_(masters, 1.0)
            opt.step()
            with torch.no_grad():
                for p, m in zip(model.parameters(), masters):
                    p.copy_(m)
                    p.grad.zero_()
        else:
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            opt.zero_grad(set_to_none=True)
        sched.step()
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
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ddp", "variant": variant,
            "config": {"model": f"albert-large-v2 (HF transformers {transformers.__version__}), vocab {vocab}",
                       "global_batch": micro_per_step * mb * world,
                       "seq_len": seq, "parallelism": f"ddp{world}", "optimizer": "LAMB (python, restated)",
                       "micro_batch": mb, "attn_implementation": getattr(hf_cfg, "_attn_implementation", None),
                       "notes": notes},
            "e2e": {"value": round(samples / (wall_ms / 1e3), 2), "unit": "samples/s"},
            "gpu_launches": 0}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0
