"""csrc/embedding.cu against plain PyTorch fp32 references of the same ops (ALBERT embedding front end + LayerNorm(128))."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _with_main_grad(*params):
    for p in params:
        p.main_grad = torch.zeros_like(p, dtype=torch.float32)


def test_embed_layer_norm_matches_fp32_reference(native):
    from dedloc_b200.models.albert import AlbertConfig, AlbertEmbeddings
    from dedloc_b200.ops import embedding as emb

    torch.manual_seed(0)
    cfg = AlbertConfig(vocab_size=1000, embedding_size=128, max_position_embeddings=64, type_vocab_size=2)
    mod = AlbertEmbeddings(cfg).cuda().to(torch.bfloat16)
    with torch.no_grad():
        mod.LayerNorm.weight.copy_(torch.rand(128) + 0.5)
        mod.LayerNorm.bias.copy_(torch.randn(128) * 0.1)
    params = [mod.word_embeddings.weight, mod.position_embeddings.weight, mod.token_type_embeddings.weight,
              mod.LayerNorm.weight, mod.LayerNorm.bias]
    _with_main_grad(*params)
    B, S = 6, 48
    ids = torch.randint(1, 1000, (B, S), device="cuda")
    ids[:, :5] = 4                                  # repeated ids: many rows add into the same gradient row
    ids[0, 7] = cfg.pad_token_id                    # the padding row receives no gradient
    tts = torch.randint(0, 2, (B, S), device="cuda")
    assert emb.embed_supported(mod, ids)
    y = emb.embed_layer_norm(mod, ids, tts)
    dy = torch.randn_like(y)
    y.backward(dy)

    # fp32 reference on the bf16-rounded tables (sum rounded to bf16 after each add, as torch does for bf16 tensors)
    w, p, t, g, b = [q.detach().float().requires_grad_(True) for q in params]
    x = (w[ids] + t[tts]).bfloat16().float()
    x = (x + p[:S][None]).bfloat16().float()
    # straight-through for the roundings: gradients flow as if the sums were exact
    xs = w[ids] + t[tts] + p[:S][None]
    x = xs + (x - xs).detach()
    ref = F.layer_norm(x, (128,), g, b, cfg.layer_norm_eps)
    ref.backward(dy.float())
    assert torch.allclose(y.float(), ref, atol=2e-2, rtol=2e-2)
    for got, want, name in zip([q.main_grad for q in params], [w.grad, p.grad, t.grad, g.grad, b.grad],
                               ["word", "pos", "type", "gamma", "beta"]):
        want = want.clone()
        if name == "word":
            want[cfg.pad_token_id] = 0
        scale = want.abs().max().clamp_min(1e-6)
        assert (got - want).abs().max() <= 2e-3 * scale, (name, float((got - want).abs().max()), float(scale))


def test_layer_norm128_matches_fp32_reference(native):
    from dedloc_b200.ops import embedding as emb

    torch.manual_seed(1)
    ln = torch.nn.LayerNorm(128, eps=1e-12).cuda().to(torch.bfloat16)
    with torch.no_grad():
        ln.weight.copy_(torch.rand(128) + 0.5)
        ln.bias.copy_(torch.randn(128) * 0.1)
    _with_main_grad(ln.weight, ln.bias)
    x = (torch.randn(777, 128, device="cuda") * 2 + 0.3).to(torch.bfloat16).requires_grad_(True)
    assert emb.ln128_supported(x, ln)
    y = emb.layer_norm128(x, ln)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    g, b = ln.weight.detach().float().requires_grad_(True), ln.bias.detach().float().requires_grad_(True)
    ref = F.layer_norm(xr, (128,), g, b, 1e-12)
    ref.backward(dy.float())
    assert torch.allclose(y.float(), ref, atol=2e-2, rtol=2e-2)
    assert (x.grad.float() - xr.grad).abs().max() <= 2e-2 * xr.grad.abs().max()
    assert (ln.weight.main_grad - g.grad).abs().max() <= 2e-3 * g.grad.abs().max()
    assert (ln.bias.main_grad - b.grad).abs().max() <= 2e-3 * b.grad.abs().max()
