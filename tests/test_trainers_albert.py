"""ALBERT entry points end to end on CPU: coordinator + trainer + auxiliary peer through their CLIs
(flag names of /root/reference/albert/arguments.py), metrics on the DHT, HF-compatible checkpoints."""
import json
import threading
import time

import pytest
import torch

from dedloc_b200.dht import DHT
from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining
from dedloc_b200.trainers.albert import metrics_utils, run_aux, run_first_peer, run_trainer
from dedloc_b200.trainers.albert.arguments import AlbertTrainingArguments, CollaborationArguments, DatasetArguments
from dedloc_b200.utils.argparse_dataclass import DataclassArgumentParser


def test_reference_flag_defaults():
    p = DataclassArgumentParser((AlbertTrainingArguments, DatasetArguments, CollaborationArguments))
    t, d, c = p.parse_args_into_dataclasses(["--experiment_prefix", "e"])
    assert (c.target_batch_size, c.averaging_expiration, c.averaging_timeout, c.compression) == (4096, 5.0, 30.0, "FLOAT16")
    assert (c.metadata_expiration, c.statistics_expiration, c.target_group_size, c.bandwidth) == (30, 600, 256, 100.0)
    assert (t.per_device_train_batch_size, t.gradient_accumulation_steps, t.seq_length) == (4, 2, 512)
    assert (t.learning_rate, t.warmup_steps, t.adam_epsilon, t.weight_decay, t.clamp_value) == (0.00176, 5000, 1e-6, 0.01, 10000.0)
    assert (t.save_steps, t.save_total_limit, t.max_grad_norm, t.output_dir) == (500, 2, 1.0, "outputs")


def test_coordinator_trainer_aux_roundtrip(tmp_path):
    cfg = AlbertConfig(vocab_size=128, embedding_size=16, hidden_size=32, num_hidden_layers=2, num_attention_heads=2,
                       intermediate_size=64, max_position_embeddings=32)
    cfg_dir = tmp_path / "cfg"
    cfg.save_pretrained(str(cfg_dir))
    repo = tmp_path / "repo"
    repo.mkdir()
    boot = DHT(start=True)   # stands in for "any live peer can be an initial peer"
    peers = f"127.0.0.1:{boot.port}"
    common = ["--experiment_prefix", "t1", "--initial_peers", peers, "--target_batch_size", "16",
              "--averaging_expiration", "1.0", "--averaging_timeout", "10", "--min_refresh_period", "0.05",
              "--default_refresh_period", "0.1", "--max_refresh_period", "0.5", "--compression", "NONE"]
    coord_args = ["--experiment_prefix", "t1", "--initial_peers", peers, "--address", "127.0.0.1",
                  "--dht_listen_on", "127.0.0.1:*", "--refresh_period", "0.3", "--save_checkpoint_step_interval", "1",
                  "--model_config_path", str(cfg_dir), "--repo_path", str(repo), "--max_iterations", "40",
                  "--target_batch_size", "16", "--compression", "NONE", "--averaging_timeout", "10",
                  "--min_refresh_period", "0.05", "--default_refresh_period", "0.1", "--max_refresh_period", "0.5"]
    train_args = common + ["--config_path", str(cfg_dir), "--synthetic_data", "--per_device_train_batch_size", "2",
                           "--gradient_accumulation_steps", "2", "--seq_length", "32", "--output_dir",
                           str(tmp_path / "out"), "--save_steps", "2", "--max_local_steps", "14", "--device", "cpu",
                           "--dht_listen_on", "127.0.0.1:*", "--listen_on", "127.0.0.1:*", "--warmup_steps", "2",
                           "--statistics_expiration", "60"]
    aux_args = common + ["--config_path", str(cfg_dir), "--output_dir", str(tmp_path / "aux_out"), "--max_local_steps", "1",
                         "--device", "cpu", "--dht_listen_on", "127.0.0.1:*", "--listen_on", "127.0.0.1:*"]
    errors = []

    def guard(fn, a):
        try:
            fn(a)
        except Exception as e:  # pragma: no cover
            errors.append(repr(e))
            raise

    tc = threading.Thread(target=guard, args=(run_first_peer.main, coord_args), daemon=True)
    ta = threading.Thread(target=guard, args=(run_aux.main, aux_args), daemon=True)
    tc.start()
    ta.start()
    run_trainer.main(train_args)
    # metrics are signed records under {prefix}_metrics
    reader = DHT(start=True, initial_peers=[peers], record_validators=metrics_utils.make_validators("t1")[0])
    got = reader.get("t1_metrics", latest=True)
    assert got is not None and len(got.value) >= 1
    m = metrics_utils.LocalMetrics.parse_obj(next(iter(got.value.values())).value)
    assert m.step >= 1 and m.mini_steps >= 1
    tc.join(timeout=60)
    assert not errors, errors
    # trainer checkpoints: HF layout + optimizer.pt with the collaborative optimizer's state
    ckpts = sorted((tmp_path / "out").glob("checkpoint-*"))
    assert ckpts, "trainer wrote no checkpoint"
    assert len(ckpts) <= 2
    restored = AlbertForPreTraining.from_pretrained(str(ckpts[-1]))
    assert restored.config.hidden_size == 32
    opt_state = torch.load(ckpts[-1] / "optimizer.pt", weights_only=False)
    assert "local_step" in opt_state["state"]
    # coordinator pulled the state from the trainer and wrote save_pretrained + optimizer_state.pt
    assert (repo / "config.json").exists() and (repo / "optimizer_state.pt").exists()
    coord_model = AlbertForPreTraining.from_pretrained(str(repo))
    assert json.loads((repo / "config.json").read_text())["hidden_size"] == 32
    reader.shutdown()
    boot.shutdown()
