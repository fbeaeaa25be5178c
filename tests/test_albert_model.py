"""ALBERT re-implementation vs Hugging Face's AlbertForPreTraining on CPU fp32 (SURVEY 7.4 oracle) and the
HF checkpoint round trip."""
import pytest
import torch

from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining

transformers = pytest.importorskip("transformers")


def _batch(cfg, B=3, S=32, seed=0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(5, cfg.vocab_size, (B, S), generator=g)
    labels = torch.full((B, S), -100)
    mask = torch.rand(B, S, generator=g) < 0.15
    mask[0, 1] = True
    labels[mask] = ids[mask]
    ids = ids.clone()
    ids[mask] = 4
    att = torch.ones(B, S, dtype=torch.long)
    att[1, S - 5:] = 0
    tt = torch.zeros(B, S, dtype=torch.long)
    tt[:, S // 2:] = 1
    sop = torch.randint(0, 2, (B,), generator=g)
    return dict(input_ids=ids, attention_mask=att, token_type_ids=tt, labels=labels, sentence_order_label=sop)


def _hf_model(cfg):
    hf_cfg = transformers.AlbertConfig(**{k: v for k, v in cfg.to_dict().items() if k not in ("architectures", "model_type")})
    hf_cfg._attn_implementation = "eager"
    return transformers.AlbertForPreTraining(hf_cfg).eval()


def test_matches_hf_forward_and_backward():
    cfg = AlbertConfig.tiny(classifier_dropout_prob=0.0)
    torch.manual_seed(0)
    mine = AlbertForPreTraining(cfg).eval()
    hf = _hf_model(cfg)
    missing, unexpected = hf.load_state_dict(mine.hf_state_dict(), strict=False)
    assert not unexpected and all("position_ids" in m for m in missing), (missing, unexpected)
    batch = _batch(cfg)
    out = mine(**batch, return_logits=True)
    ref = hf(**batch)
    assert torch.allclose(out.prediction_logits, ref.prediction_logits, atol=2e-5, rtol=1e-4)
    assert torch.allclose(out.sop_logits, ref.sop_logits, atol=2e-5, rtol=1e-4)
    assert torch.allclose(out.loss, ref.loss, atol=1e-5, rtol=1e-5)
    out.loss.backward()
    ref.loss.backward()
    hf_grads = dict(hf.named_parameters())
    for name, p in mine.named_parameters():
        if name.startswith("predictions.decoder."):
            continue
        g_ref = hf_grads[name].grad
        assert torch.allclose(p.grad, g_ref, atol=2e-5, rtol=1e-3), name


def test_param_count_and_names_of_albert_large():
    m = AlbertForPreTraining(AlbertConfig.albert_large_v2())
    assert m.num_parameters() == 17_847_474
    assert len(list(m.parameters())) == 32
    assert "albert.encoder.albert_layer_groups.0.albert_layers.0.attention.query.weight" in m.state_dict()


def test_hf_checkpoint_round_trip(tmp_path):
    cfg = AlbertConfig.tiny()
    mine = AlbertForPreTraining(cfg)
    mine.save_pretrained(str(tmp_path / "ckpt"))
    hf = transformers.AlbertForPreTraining.from_pretrained(str(tmp_path / "ckpt"))
    for k, v in mine.hf_state_dict().items():
        assert torch.equal(hf.state_dict()[k], v), k
    hf.save_pretrained(str(tmp_path / "hf"))
    back = AlbertForPreTraining.from_pretrained(str(tmp_path / "hf"))
    for (k, a), (_, b) in zip(mine.state_dict().items(), back.state_dict().items()):
        assert torch.equal(a, b), k


def test_resize_token_embeddings_keeps_tie():
    m = AlbertForPreTraining(AlbertConfig.tiny())
    m.resize_token_embeddings(1100)
    assert m.predictions.decoder.weight is m.albert.embeddings.word_embeddings.weight
    assert m.predictions.bias.shape[0] == 1100
    out = m(**_batch(m.config))
    assert torch.isfinite(out.loss)
