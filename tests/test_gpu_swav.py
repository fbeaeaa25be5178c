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


def test_l2_normalize128_matches_torch(native):
    import torch.nn.functional as F

    from dedloc_b200.ops.swav_ops import l2_normalize128

    torch.manual_seed(0)
    x = (torch.randn(512, 128, device="cuda") * 3).to(torch.bfloat16).requires_grad_(True)
    y = l2_normalize128(x)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    ref = F.normalize(xr, dim=1, p=2)
    ref.backward(dy.float())
    assert torch.allclose(y.float(), ref, atol=1e-2, rtol=1e-2)
    assert (x.grad.float() - xr.grad).abs().max() <= 2e-2 * xr.grad.abs().max()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_swav_cross_entropy_matches_the_loss_restatement(native, dtype):
    """csrc/swav_ops.cu against the PyTorch restatement of vissl/losses/swav_loss.py:275-302 (fp32)."""
    import torch.nn.functional as F

    from dedloc_b200.ops.swav_ops import swav_cross_entropy

    torch.manual_seed(1)
    num_crops, bs, K, T = 8, 64, 3000, 0.1
    scores = (torch.randn(num_crops * bs, K, device="cuda") * 0.3).to(dtype).requires_grad_(True)
    assign = [0, 1]
    q = torch.softmax(torch.randn(len(assign), bs, K, device="cuda") * 2, dim=-1)
    loss = swav_cross_entropy(scores, q, assign, num_crops, T)
    loss.backward()
    sr = scores.detach().float().requires_grad_(True)
    logp = F.log_softmax(sr / T, dim=1).view(num_crops, bs, K)
    total = 0
    for i, c in enumerate(assign):
        others = [p for p in range(num_crops) if p != c]
        total = total - torch.einsum("bk,pbk->", q[i], logp[others]) / (bs * len(others))
    ref = total / len(assign)
    ref.backward()
    assert abs(float(loss) - float(ref)) <= 2e-3 * abs(float(ref)), (float(loss), float(ref))
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-4
    assert (scores.grad.float() - sr.grad).abs().max() <= tol * sr.grad.abs().max()


def test_head_on_own_kernels_matches_torch_path(native):
    """SwAVPrototypesHead._forward_native (tcgen05 GEMMs + fused L2-normalise) against the module's PyTorch path in
    fp32: embeddings, prototype scores and every parameter gradient (accumulated in main_grad)."""
    from dedloc_b200.models.swav import SwAVPrototypesHead

    torch.manual_seed(2)
    head = SwAVPrototypesHead((2048, 2048, 128), use_bn=True, num_clusters=(3000,)).cuda().train()
    ref = SwAVPrototypesHead((2048, 2048, 128), use_bn=True, num_clusters=(3000,)).cuda().train()
    ref.load_state_dict(head.state_dict())
    for p in head.parameters():
        p.main_grad = torch.zeros_like(p, dtype=torch.float32)
    feats = torch.randn(512, 2048, device="cuda")
    f1 = feats.clone().requires_grad_(True)
    f2 = feats.clone().requires_grad_(True)
    assert head._native_ok(f1)
    emb, scores = head(f1)
    emb_r, scores_r = ref(f2)
    assert torch.allclose(emb.float(), emb_r, atol=2e-2, rtol=5e-2)
    assert torch.allclose(scores.float(), scores_r, atol=3e-2, rtol=5e-2)
    g = torch.randn_like(scores_r) * 0.1
    scores.backward(g.to(scores.dtype))
    scores_r.backward(g)
    assert (f1.grad - f2.grad).abs().max() <= 8e-2 * f2.grad.abs().max()
    for (n, p), (_, pr) in zip(head.named_parameters(), ref.named_parameters()):
        got = p.main_grad if p.grad is None else p.main_grad + p.grad.float()
        # bf16 activations in front of a BatchNorm with a large 1/std: single elements of dW1 move by up to 15 % of
        # the maximum even in a pure-PyTorch bf16 emulation of this path, so the check is on the whole tensor
        if float(pr.grad.abs().max()) < 1e-5:
            # the bias in front of the BatchNorm has a zero gradient (the batch mean is subtracted): only rounding noise
            assert float(got.abs().max()) <= 5e-2, (n, float(got.abs().max()))
            continue
        rel = float((got - pr.grad).norm() / pr.grad.norm().clamp_min(1e-12))
        assert rel <= 6e-2, (n, rel)
        assert (got - pr.grad).abs().max() <= 0.3 * pr.grad.abs().max() + 1e-6, n
