"""SwAV on CPU: Sinkhorn restatement, loss with queue keyed on the collaboration step, hydra-style config,
sgd_collaborative with two peers, VISSL-shaped checkpoints, hooks (prototype normalisation, freeze)."""
import os
import threading

import pytest
import torch

from dedloc_b200.models.swav import SwAVLoss, SwAVModel, normalize_prototypes
from dedloc_b200.ops.sinkhorn import sinkhorn_reference
from dedloc_b200.trainers.swav.config import compose


def literal_reference_sinkhorn(scores, eps, iters, bs):
    """Line-by-line arithmetic of distributed_sinkhornknopp (world_size 1) for cross-checking the restatement."""
    a = scores / eps
    a = a - a.max()
    Q = torch.exp(a).t()
    Q /= Q.sum()
    K, n = Q.shape
    r, c = torch.ones(K) / K, torch.ones(n) / n
    for _ in range(iters):
        u = Q.sum(dim=1)
        Q *= (r / u).unsqueeze(1)
        Q *= (c / Q.sum(dim=0)).unsqueeze(0)
    return (Q / Q.sum(dim=0, keepdim=True)).t()[:bs]


def test_sinkhorn_restatement_matches_literal():
    torch.manual_seed(0)
    s = torch.randn(50, 30) * 0.3
    got = sinkhorn_reference(s, 0.05, 3, 20)
    ref = literal_reference_sinkhorn(s.clone(), 0.05, 3, 20)
    assert torch.allclose(got, ref, rtol=1e-5, atol=1e-8)
    assert torch.allclose(got.sum(dim=1), torch.ones(20), atol=1e-5)


def test_loss_queue_switches_on_collaboration_step():
    torch.manual_seed(0)
    B, crops, K, D = 8, 4, 16, 8
    loss = SwAVLoss(num_crops=crops, num_prototypes=[K], embedding_dim=D, queue_length=32, queue_start_iter=5,
                    epsilon=0.05)
    emb = torch.nn.functional.normalize(torch.randn(B * crops, D), dim=1)
    proto = torch.nn.Linear(D, K, bias=False)
    scores = proto(emb)
    l0 = loss([emb, scores], None, training_iterations=0)
    assert not loss.use_queue and torch.isfinite(l0)
    assert loss.local_emb_queue.abs().sum() == 0
    l1 = loss([emb, scores], None, training_iterations=5)        # the COLLABORATION reached start_iter
    assert loss.use_queue and torch.isfinite(l1)
    assert loss.local_emb_queue.abs().sum() > 0
    head = torch.nn.Module()
    head.nmb_heads = 1
    head.prototypes = lambda i: proto
    loss.compute_queue_scores(head)
    assert loss.local_queue0.abs().sum() > 0
    l1.backward()
    assert proto.weight.grad is not None


def test_config_compose_and_overrides():
    cfg = compose(["config=pretrain/swav/swav_1node_resnet_submit", "hydra.verbose=true",
                   "config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=64", "+config.OPTIMIZER.batch_size_for_tracking=64",
                   "+config.OPTIMIZER.lr=2.4", "+config.OPTIMIZER.warmup_start_lr=0.3",
                   "+config.OPTIMIZER.warmup_epochs=500", "+config.OPTIMIZER.max_epochs=5000",
                   "+config.OPTIMIZER.eta_min=0.0048", "+config.OPTIMIZER.target_group_size=4",
                   "config.OPTIMIZER.dht_listen_on_port=1124", "config.LOSS.swav_loss.queue.start_iter=98000",
                   '+config.OPTIMIZER.dht_initial_peers=["10.0.0.1:1337"]'])
    assert cfg.OPTIMIZER.name == "sgd_collaborative" and cfg.OPTIMIZER.lr == 2.4
    assert cfg.OPTIMIZER.larc_config.trust_coefficient == 0.001 and cfg.OPTIMIZER.larc_config.clip is False
    assert cfg.LOSS.swav_loss.num_prototypes == [3000] and cfg.LOSS.swav_loss.embedding_dim == 128
    assert cfg.LOSS.swav_loss.num_crops == 8 and cfg.LOSS.swav_loss.queue.local_queue_length == 3840
    assert cfg.LOSS.swav_loss.queue.start_iter == 98000 and cfg.LOSS.swav_loss.epsilon == 0.03
    assert cfg.MODEL.TEMP_FROZEN_PARAMS_ITER_MAP == [["module.heads.0.prototypes0.weight", 5005]]
    assert cfg.OPTIMIZER.dht_initial_peers == ["10.0.0.1:1337"]


def _tiny_overrides(tmp, prefix, peers, port_note=""):
    return ["config=pretrain/swav/swav_1node_resnet_submit", "config.MACHINE.DEVICE=cpu",
            "config.MODEL.TRUNK.TRUNK_PARAMS.RESNETS.DEPTH=18",
            'config.MODEL.HEAD.PARAMS=[["swav_head", {"dims": [512, 64, 16], "use_bn": true, "num_clusters": [32]}]]',
            'config.MODEL.TEMP_FROZEN_PARAMS_ITER_MAP=[["module.heads.0.prototypes0.weight", 2]]',
            "config.DATA.TRAIN.DATA_SOURCES=[synthetic]", "config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=4",
            "config.DATA.NUM_DATALOADER_WORKERS=0", "config.DATA.TRAIN.DATA_LIMIT=64",
            'config.DATA.TRAIN.TRANSFORMS=[{"name": "ImgPilToMultiCrop", "total_num_crops": 4, "size_crops": [64, 32], '
            '"num_crops": [2, 2], "crop_scales": [[0.14, 1], [0.05, 0.14]]}, {"name": "RandomHorizontalFlip", "p": 0.5}, '
            '{"name": "ImgPilColorDistortion", "strength": 1.0}, {"name": "ImgPilGaussianBlur", "p": 0.5, '
            '"radius_min": 0.1, "radius_max": 2.0}, {"name": "ToTensor"}, {"name": "Normalize", "mean": [0.485, 0.456, '
            '0.406], "std": [0.229, 0.224, 0.225]}]',
            "config.LOSS.swav_loss.queue.queue_length=16", "config.LOSS.swav_loss.queue.start_iter=2",
            "config.HOOKS.LOG_GPU_STATS=false", "config.HOOKS.MEMORY_SUMMARY.PRINT_MEMORY_SUMMARY=false",
            "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false", "config.HOOKS.PERF_STATS.ROLLING_BTIME_FREQ=2",
            f"config.CHECKPOINT.DIR={tmp}", "config.CHECKPOINT.CHECKPOINT_ITER_FREQUENCY=3", "config.OPTIMIZER.num_epochs=1",
            "+config.OPTIMIZER.lr=0.3", "+config.OPTIMIZER.warmup_epochs=2", "+config.OPTIMIZER.max_epochs=20",
            "+config.OPTIMIZER.batch_size_for_tracking=4", f"+config.OPTIMIZER.exp_prefix={prefix}",
            "+config.OPTIMIZER.target_group_size=2", "+config.OPTIMIZER.target_batch_size=16",
            "+config.OPTIMIZER.compression=NONE", f"+config.OPTIMIZER.dht_initial_peers={peers}"]


def test_two_collaborative_swav_peers(tmp_path):
    from dedloc_b200.dht import DHT
    from dedloc_b200.trainers.swav import run_initial_dht_node
    from dedloc_b200.trainers.swav.train import train_main

    boot = DHT(start=True)
    peers = f'["127.0.0.1:{boot.port}"]'
    fast = dict(min_refresh_period=0.05, default_refresh_period=0.1, max_refresh_period=0.5)
    results, errors = {}, []

    from dedloc_b200.trainers.swav.train import SelfSupervisionTrainer

    # peers are separate processes in production (same SEED_VALUE -> same init); in-process peers are constructed
    # one after the other because the global torch RNG is shared between threads
    trainers = [SelfSupervisionTrainer(compose(_tiny_overrides(str(tmp_path / f"p{i}"), "swavtest", peers)),
                                       collab_kwargs=fast, max_iterations=40, max_global_steps=3) for i in range(2)]

    def run(i):
        try:
            results[i] = trainers[i].train()
        except Exception:  # pragma: no cover
            import traceback

            errors.append(traceback.format_exc())

    threads = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    [t.start() for t in threads]
    [t.join(timeout=300) for t in threads]
    assert not errors, errors[0]
    t0, t1 = results[0], results[1]
    c0, c1 = t0.optimizer.optimizer, t1.optimizer.optimizer
    assert c0.local_step == 3 and c1.local_step == 3
    assert c0.stats["solo_rounds"] == 0 and c1.stats["solo_rounds"] == 0
    for (n, a), (_, b) in zip(t0.model.named_parameters(), t1.model.named_parameters()):
        assert torch.allclose(a, b, atol=1e-6), n
    w = t0.model.prototype_weights[0]
    assert torch.allclose(w.norm(dim=1), torch.ones(w.shape[0]), atol=1e-5)      # NormalizePrototypesHook
    assert t0.loss.use_queue                                                       # queue keyed on the global step
    # VISSL-shaped checkpoints; the optimizer entry is None for sgd_collaborative
    ck = torch.load(tmp_path / "p0" / "checkpoint.torch", weights_only=False)
    assert set(ck) >= {"phase_idx", "iteration", "loss", "classy_state_dict"}
    assert ck["classy_state_dict"]["optimizer"] is None
    assert "local_emb_queue" in ck["classy_state_dict"]["loss"]
    assert (tmp_path / "p0" / "model_iteration3.torch").exists() and (tmp_path / "p0" / "metrics.json").exists()
    assert t0.perf_stats.host["forward"].n == t0.iteration and "optimizer_step" in t0.perf_stats.as_dict()
    for t in (t0, t1):
        t.optimizer.optimizer.shutdown()
        t.optimizer.dht.shutdown()
    boot.shutdown()
    run_initial_dht_node.main(["--listen_on", "127.0.0.1:*", "--refresh_period", "0.01", "--max_iterations", "2"])


def test_model_complexity_and_profiler_hooks():
    """HOOKS.MODEL_COMPLEXITY (#params / #flops / #activations on a dummy forward,
    vissl/hooks/state_update_hooks.py:16-63) and the one-forward ProfilerHook of ClassyVision."""
    import types

    from dedloc_b200.models.swav import SwAVModel
    from dedloc_b200.trainers.swav.hooks import (ProfilerHook, SSLModelComplexityHook, default_hook_generator,
                                                  model_complexity)

    model = SwAVModel(depth=18, head_dims=(512, 64, 16), num_clusters=(32,))
    c = model_complexity(model, (3, 64, 64))
    assert c["params"] == sum(p.numel() for p in model.parameters())
    head_macs = 512 * 64 + 64 * 16 + 16 * 32
    # ResNet-18 at 224 px is 1.814 GMACs; convolutions scale with the pixel count
    trunk = c["macs_per_sample"] - head_macs
    assert abs(trunk / (1.814e9 * (64 / 224) ** 2) - 1) < 0.03, trunk
    assert c["flops_per_sample"] == 2 * c["macs_per_sample"] and c["activations_per_sample"] > 0
    assert model.training, "the hook restores the training mode"

    cfg = compose(["config=pretrain/swav/swav_1node_resnet_submit", "config.HOOKS.MODEL_COMPLEXITY.COMPUTE_COMPLEXITY=true",
                   "config.HOOKS.MODEL_COMPLEXITY.INPUT_SHAPE=[3, 64, 64]", "config.HOOKS.PROFILER.PROFILE_FORWARD=true",
                   "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false"])
    kinds = [type(h) for h in default_hook_generator(cfg)]
    assert SSLModelComplexityHook in kinds and ProfilerHook in kinds
    task = types.SimpleNamespace(model=model, device=torch.device("cpu"))
    for h in default_hook_generator(cfg):
        if isinstance(h, (SSLModelComplexityHook, ProfilerHook)):
            h.on_start(task)
    assert task.model_complexity["params"] == c["params"] and "conv" in task.profile_table.lower()
    off = [type(h) for h in default_hook_generator(compose(["config=pretrain/swav/swav_1node_resnet_submit",
                                                           "config.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD=false"]))]
    assert SSLModelComplexityHook not in off and ProfilerHook not in off


def test_stateful_peer_sampler_and_epoch_hook():
    """Per-peer shuffle (no two peers walk the dataset in the same order), new order every phase, mid-epoch resume."""
    import types

    from dedloc_b200.trainers.swav.data import StatefulPeerSampler, build_dataloader, peer_shuffle_seed
    from dedloc_b200.trainers.swav.hooks import SetDataSamplerEpochHook

    s = StatefulPeerSampler(20, batch_size=4, seed=7)
    first = list(s)
    assert sorted(first) == list(range(20)) and list(s) == first           # a function of (seed, epoch)
    s.set_epoch(1)
    assert list(s) != first
    s.set_epoch(0)
    s.set_start_iter(2)
    assert list(s) == first[8:] and list(s) == first, "only the first epoch after a resume is partial"
    assert list(StatefulPeerSampler(20, 4, seed=8)) != first

    base = ["config=pretrain/swav/swav_1node_resnet_submit", "config.MACHINE.DEVICE=cpu",
            "config.DATA.TRAIN.DATA_SOURCES=[synthetic]", "config.DATA.TRAIN.DATA_LIMIT=32",
            "config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=4", "config.DATA.NUM_DATALOADER_WORKERS=0"]
    a, b = compose(base + ["config.DATA.TRAIN.PEER_SHUFFLE_SALT=peer-a"]), compose(base + ["config.DATA.TRAIN.PEER_SHUFFLE_SALT=peer-b"])
    assert peer_shuffle_seed(a) == peer_shuffle_seed(a) != peer_shuffle_seed(b)
    assert peer_shuffle_seed(compose(base)) != peer_shuffle_seed(compose(base)), "no salt: unique per call / process"
    loader = build_dataloader(a, torch.device("cpu"))
    order0 = list(loader.sampler)
    task = types.SimpleNamespace(dataloader=loader, phase_idx=3)
    SetDataSamplerEpochHook().on_phase_start(task)
    assert loader.sampler.epoch == 3 and list(loader.sampler) != order0
