"""BASELINE.json config 4: SwAV ResNet-50 collaborative training with LARS, synthetic 2x224 + 6x96 px crops, one peer
per GPU. Device-timed images/s over `--steps` training steps (max over ranks), incl. the collaborative optimizer (fused
LARC + SGD-momentum in the all-reduce kernel across the peers; a solo round at one peer) and the Sinkhorn kernel.

    python scripts/bench_swav.py --batch 64 --steps 10                                   # one peer
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_swav.py --gpus N
    ... --impl ddp      the torch-only arm (baseline/ddp_swav_baseline.py: torchvision trunk, PyTorch head / Sinkhorn /
                        loss / LARC, DistributedDataParallel over NCCL) -- BASELINE.md B4

torch.distributed is used by OUR arm only to hand rank 0's DHT address to the other ranks and to take the max of the
timings; the peers find each other, form groups and average through the DHT / matchmaking / fused kernel as in
production. One collaborative step per `--rounds_every` local steps (target batch = gpus x batch x rounds_every).
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def run_ours(args) -> dict:
    import torch.distributed as dist

    from dedloc_b200.dht import DHT
    from dedloc_b200.trainers.swav.config import compose
    from dedloc_b200.trainers.swav.data import synthetic_multicrop_batch
    from dedloc_b200.trainers.swav.train import SelfSupervisionTrainer, standard_train_step

    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    on_gpu = args.device != "cpu"
    device = torch.device("cuda", local) if on_gpu else torch.device("cpu")
    if on_gpu:
        torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl" if on_gpu else "gloo", **({"device_id": device} if on_gpu else {}))
        holder = [None]
        if rank == 0:
            dht = DHT(start=True)
            holder[0] = f"127.0.0.1:{dht.port}"
        dist.broadcast_object_list(holder, src=0)
        if rank != 0:
            dht = DHT(start=True, initial_peers=[holder[0]])
    else:
        dht = DHT(start=True)
    tmp = tempfile.mkdtemp()
    target = args.target_batch_size or world * args.batch * args.rounds_every
    cfg = compose(["config=pretrain/swav/swav_1node_resnet_submit", "config.DATA.TRAIN.DATA_SOURCES=[synthetic]",
                   f"config.DATA.TRAIN.BATCHSIZE_PER_REPLICA={args.batch}", "config.DATA.NUM_DATALOADER_WORKERS=0",
                   "config.HOOKS.LOG_GPU_STATS=false", "config.HOOKS.MEMORY_SUMMARY.PRINT_MEMORY_SUMMARY=false",
                   "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false", f"config.CHECKPOINT.DIR={tmp}",
                   "+config.OPTIMIZER.lr=0.6", f"+config.OPTIMIZER.batch_size_for_tracking={args.batch}",
                   f"+config.OPTIMIZER.target_batch_size={target}", "+config.OPTIMIZER.exp_prefix=benchswav"])
    trainer = SelfSupervisionTrainer(cfg, device=device, dht=dht, dataloader=[],
                                     collab_kwargs=dict(min_refresh_period=0.05, default_refresh_period=0.2,
                                                        max_refresh_period=1.0, target_group_size=max(world, 2)))
    task = trainer.task
    for h in trainer.hooks:
        h.on_start(task)
    batches = [synthetic_multicrop_batch(args.batch, task.device, size_crops=(args.big, args.small)) for _ in range(2)]
    label = torch.arange(args.batch)
    import time

    def step(i):
        standard_train_step(task, {"input": batches[i % 2], "label": label}, trainer.hooks)

    for i in range(args.warmup):
        step(i)
    if world > 1:
        dist.barrier()
    collab = task.optimizer.optimizer
    steps_before = collab.local_step
    if on_gpu:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    if on_gpu:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
    else:
        ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if world > 1:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    out = {"metric": "SwAV RN50 collaborative images/sec", "impl": "ours",
           "value": round(world * args.batch / ms * 1e3, 1), "unit": "images/s", "n_gpus": world,
           "ms_per_step": round(ms, 2), "batch_per_replica": args.batch, "target_batch_size": target,
           "crops": f"2x{args.big} + 6x{args.small}", "optimizer": type(collab.opt).__name__,
           "global_steps_in_window": collab.local_step - steps_before, "loss": float(task.last_loss),
           "collab": {k: v for k, v in collab.stats.items() if isinstance(v, (int, float))}, "data": "synthetic"}
    if world > 1:
        dist.barrier()
    collab.shutdown()
    dht.shutdown()
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "ddp"], default="ours")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--target_batch_size", type=int, default=0, help="0 = gpus x batch x rounds_every")
    ap.add_argument("--rounds_every", type=int, default=4, help="local steps per collaborative step (ours)")
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--big", type=int, default=224)
    ap.add_argument("--small", type=int, default=96)
    ap.add_argument("--sync_bn", action="store_true", help="ddp arm: SyncBatchNorm (the recipe converts BN when it can)")
    args = ap.parse_args()
    if args.impl == "ddp":
        from baseline.ddp_swav_baseline import run

        out = run(args)
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
