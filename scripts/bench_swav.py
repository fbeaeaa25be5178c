"""BASELINE.json config 4: SwAV ResNet-50 collaborative training with LARS, synthetic 2x224 + 6x96 px crops, one peer
per GPU (this script: one peer). Device-timed images/s over `--steps` training steps incl. the collaborative
optimizer (fused LARC + SGD-momentum in the all-reduce kernel, solo round at one peer) and the Sinkhorn kernel.

    python scripts/bench_swav.py --batch 64 --steps 10
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--target_batch_size", type=int, default=256)
    args = ap.parse_args()
    from dedloc_b200.dht import DHT
    from dedloc_b200.trainers.swav.config import compose
    from dedloc_b200.trainers.swav.data import synthetic_multicrop_batch
    from dedloc_b200.trainers.swav.train import SelfSupervisionTrainer, standard_train_step

    dht = DHT(start=True)
    tmp = tempfile.mkdtemp()
    cfg = compose(["config=pretrain/swav/swav_1node_resnet_submit", "config.DATA.TRAIN.DATA_SOURCES=[synthetic]",
                   f"config.DATA.TRAIN.BATCHSIZE_PER_REPLICA={args.batch}", "config.DATA.NUM_DATALOADER_WORKERS=0",
                   "config.HOOKS.LOG_GPU_STATS=false", "config.HOOKS.MEMORY_SUMMARY.PRINT_MEMORY_SUMMARY=false",
                   "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false", f"config.CHECKPOINT.DIR={tmp}",
                   "+config.OPTIMIZER.lr=0.6", f"+config.OPTIMIZER.batch_size_for_tracking={args.batch}",
                   f"+config.OPTIMIZER.target_batch_size={args.target_batch_size}", "+config.OPTIMIZER.exp_prefix=benchswav"])
    trainer = SelfSupervisionTrainer(cfg, dht=dht, dataloader=[], collab_kwargs=dict(min_refresh_period=0.05))
    task = trainer.task
    for h in trainer.hooks:
        h.on_start(task)
    batches = [synthetic_multicrop_batch(args.batch, task.device) for _ in range(2)]
    label = torch.arange(args.batch)

    def step(i):
        standard_train_step(task, {"input": batches[i % 2], "label": label}, trainer.hooks)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    print(json.dumps({"metric": "SwAV RN50 collaborative images/sec (1 peer)", "value": round(args.batch / ms * 1e3, 1),
                      "unit": "images/s", "ms_per_step": round(ms, 2), "batch_per_replica": args.batch,
                      "crops": "2x224 + 6x96", "optimizer": type(task.optimizer.optimizer.opt).__name__,
                      "global_steps": task.optimizer.optimizer.local_step, "loss": float(task.last_loss),
                      "data": "synthetic"}))
    task.optimizer.optimizer.shutdown()
    dht.shutdown()


if __name__ == "__main__":
    main()
