"""Sinkhorn kernel vs the PyTorch restatement; fused LARS path through SGDCollaborative on the GPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,K,n_out,dtype", [(200, 3000, 200, torch.float32), (4040, 3000, 200, torch.bfloat16),
                                             (64, 257, 33, torch.float32), (300, 4096, 300, torch.bfloat16)])
def test_sinkhorn_kernel_matches_reference(native, n, K, n_out, dtype):
    from dedloc_b200.ops.sinkhorn import sinkhorn, sinkhorn_reference

    g = torch.Generator().manual_seed(0)
    emb = torch.nn.functional.normalize(torch.randn(n, 128, generator=g), dim=1)
    proto = torch.nn.functional.normalize(torch.randn(K, 128, generator=g), dim=1)
    scores = (emb @ proto.t()).to(dtype).cuda()
    for eps in (0.03, 0.05):
        got = sinkhorn(scores, eps, 3, n_out)
        ref = sinkhorn_reference(scores.float().cpu(), eps, 3, n_out)
        assert got.shape == (n_out, K)
        assert torch.allclose(got.sum(dim=1).cpu(), torch.ones(n_out), atol=1e-4)
        assert torch.allclose(got.cpu(), ref, rtol=2e-3, atol=1e-7), (got.cpu() - ref).abs().max()
    # two calls in a row give identical results (scratch is re-initialised by every call)
    again = sinkhorn(scores, 0.05, 3, n_out)
    assert torch.allclose(again, got, rtol=1e-4, atol=1e-9)   # fp32 atomics: order-dependent rounding only


def test_swav_gpu_step_with_fused_lars(native, tmp_path):
    from dedloc_b200.dht import DHT
    from dedloc_b200.trainers.swav.config import compose
    from dedloc_b200.trainers.swav.data import synthetic_multicrop_batch
    from dedloc_b200.trainers.swav.train import SelfSupervisionTrainer, standard_train_step

    dht = DHT(start=True)
    cfg = compose(["config=pretrain/swav/swav_1node_resnet_submit", "config.DATA.TRAIN.DATA_SOURCES=[synthetic]",
                   "config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=8", "config.DATA.NUM_DATALOADER_WORKERS=0",
                   "config.HOOKS.LOG_GPU_STATS=false", "config.HOOKS.MEMORY_SUMMARY.PRINT_MEMORY_SUMMARY=false",
                   "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false", f"config.CHECKPOINT.DIR={tmp_path}",
                   "config.LOSS.swav_loss.queue.start_iter=1", "config.LOSS.swav_loss.queue.queue_length=64",
                   "+config.OPTIMIZER.lr=0.6", "+config.OPTIMIZER.batch_size_for_tracking=8",
                   "+config.OPTIMIZER.target_batch_size=16", "+config.OPTIMIZER.exp_prefix=gpuswav"])
    trainer = SelfSupervisionTrainer(cfg, dht=dht, dataloader=[], collab_kwargs=dict(min_refresh_period=0.05))
    task = trainer.task
    from dedloc_b200.optim.fused import FusedLARS

    assert isinstance(task.optimizer.optimizer.opt, FusedLARS) and task.prototypes_normalized_by_kernel
    for h in trainer.hooks:
        h.on_start(task)
    losses = []
    for it in range(4):
        sample = {"input": synthetic_multicrop_batch(8, task.device), "label": torch.arange(8)}
        standard_train_step(task, sample, trainer.hooks)
        losses.append(float(task.last_loss))
    assert all(l == l for l in losses)
    assert task.optimizer.optimizer.local_step == 2            # 16 samples per collaborative step, 8 per iteration
    w = task.model.prototype_weights[0]
    # frozen for the first 5005 iterations: prototypes did not move (they were only normalised at start)
    torch.cuda.synchronize()
    assert torch.allclose(w.float().norm(dim=1), torch.ones(w.shape[0], device=w.device), atol=1e-3)
    task.optimizer.optimizer.shutdown()
    dht.shutdown()
