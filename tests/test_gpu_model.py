"""Row-wise kernels, fused FFN autograd node and the whole ALBERT model on the native path vs PyTorch fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).cuda()


def _close(a, b, what, rtol=2e-2, atol=2e-2):
    a, b = a.float(), b.float()
    scale = b.abs().max().item() + 1e-6
    assert torch.allclose(a, b, rtol=rtol, atol=atol * scale), f"{what}: max err {(a - b).abs().max().item():.4g} vs scale {scale:.4g}"


@pytest.mark.parametrize("H", [256, 1024, 2048])
def test_layernorm_fwd_bwd(native, H):
    from dedloc_b200.ops.layernorm import layer_norm

    M = 1000
    ln = torch.nn.LayerNorm(H, eps=1e-12).cuda().to(torch.bfloat16)
    with torch.no_grad():
        ln.weight.copy_(_rand((H,), 1, 0.2) + 1.0)
        ln.bias.copy_(_rand((H,), 2, 0.2))
    ln.weight.main_grad = torch.zeros(H, device="cuda")
    ln.bias.main_grad = torch.zeros(H, device="cuda")
    x = _rand((4, M // 4, H), 3, 2.0).requires_grad_(True)
    dy = _rand((4, M // 4, H), 4, 1.0)
    for rep in range(2):  # second pass checks that parameter gradients ACCUMULATE
        x.grad = None
        y = layer_norm(x, ln, "native")
        y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    wr, br = ln.weight.detach().float().requires_grad_(True), ln.bias.detach().float().requires_grad_(True)
    yr = torch.nn.functional.layer_norm(xr, (H,), wr, br, 1e-12)
    yr.backward(dy.float())
    _close(y, yr, "ln fwd")
    _close(x.grad, xr.grad, "ln dx")
    _close(ln.weight.main_grad, 2 * wr.grad, "ln dgamma", rtol=1e-2, atol=1e-2)
    _close(ln.bias.main_grad, 2 * br.grad, "ln dbeta", rtol=1e-2, atol=1e-2)


def test_colsum_accumulate(native):
    from dedloc_b200.ops.layernorm import colsum_accumulate

    dy = _rand((3000, 1032), 5)
    out = torch.ones(1032, device="cuda")
    colsum_accumulate(dy, out)
    _close(out, 1.0 + dy.float().sum(0), "colsum", rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("side_stream", [False, True])
def test_fused_ffn_autograd(native, side_stream):
    from dedloc_b200.ops import gemm as gemm_ops
    from dedloc_b200.ops.elementwise import gelu_new
    from dedloc_b200.ops.gemm import ffn

    class L:
        pass

    M, K, I = 512, 256, 1024
    l1, l2 = L(), L()
    l1.weight, l1.bias = _rand((I, K), 1, 0.05), _rand((I,), 2, 0.1)
    l2.weight, l2.bias = _rand((K, I), 3, 0.05), _rand((K,), 4, 0.1)
    for l in (l1, l2):
        l.weight.main_grad = torch.zeros(l.weight.shape, device="cuda")
        l.bias.main_grad = torch.zeros(l.bias.shape, device="cuda")
    x = _rand((2, M // 2, K), 5, 0.5).requires_grad_(True)
    dy = _rand((2, M // 2, K), 6, 0.1)
    gemm_ops.overlap_wgrad(side_stream)      # weight / bias gradients on the side stream, joined before they are read
    try:
        y = ffn(x, l1, l2, residual=x)
        y.backward(dy)
    finally:
        gemm_ops.overlap_wgrad(False)
        gemm_ops.join_wgrad_stream()
    xr = x.detach().float().requires_grad_(True)
    ps = [t.float().requires_grad_(True) for t in (l1.weight, l1.bias, l2.weight, l2.bias)]
    yr = gelu_new(xr @ ps[0].t() + ps[1]) @ ps[2].t() + ps[3] + xr
    yr.backward(dy.float())
    _close(y, yr, "ffn fwd")
    _close(x.grad, xr.grad, "ffn dx", 3e-2, 3e-2)
    for got, ref, name in zip((l1.weight.main_grad, l1.bias.main_grad, l2.weight.main_grad, l2.bias.main_grad),
                              (p.grad for p in ps), ("dW1", "db1", "dW2", "db2")):
        _close(got, ref, name, 3e-2, 3e-2)


def test_albert_native_matches_torch_path(native):
    """Same bf16 weights, same batch: the native path (tcgen05 linears, fused FFN, own LayerNorm) must reproduce
    the loss and the parameter gradients of the plain PyTorch path within bf16 tolerance."""
    from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining
    from dedloc_b200.optim.fused import FusedLamb
    from dedloc_b200.trainers.pretrain import SyntheticMLMBatches, albert_param_groups

    cfg = AlbertConfig(vocab_size=2048, embedding_size=128, hidden_size=256, num_hidden_layers=3,
                       num_attention_heads=4, intermediate_size=1024, max_position_embeddings=128,
                       classifier_dropout_prob=0.0)
    torch.manual_seed(0)
    ref = AlbertForPreTraining(cfg, impl="torch").cuda().to(torch.bfloat16).train()
    nat = AlbertForPreTraining(cfg, impl="native").cuda().to(torch.bfloat16).train()
    nat.load_state_dict(ref.state_dict())
    opt = FusedLamb(albert_param_groups(nat, 0.01), lr=1e-3, debias=True)
    opt.materialize()
    data = SyntheticMLMBatches(cfg.vocab_size, 128, 8, pin=False, pool=1)
    batch = data.to_device(data.next_host(), "cuda")
    out_r = ref(**batch)
    out_r.loss.backward()
    out_n = nat(**batch)
    out_n.loss.backward()
    nat.flush_fused_grads()
    assert abs(out_n.loss.item() - out_r.loss.item()) < 0.05, (out_n.loss.item(), out_r.loss.item())
    ref_grads = dict(ref.named_parameters())
    worst = 0.0
    for name, p in nat.named_parameters():
        if name.endswith("key.bias"):
            continue  # softmax is invariant to the key bias: its exact gradient is 0, both paths only hold noise
        total = p.grad.float() + p.main_grad
        g_ref = ref_grads[name].grad.float()
        denom = g_ref.norm().item() + 1e-6
        rel = (total - g_ref).norm().item() / denom
        worst = max(worst, rel)
        assert rel < 0.08, f"{name}: relative gradient error {rel:.3f}"
    # one fused LAMB step keeps the bf16 mirror in sync with the fp32 master copy
    opt.step()
    torch.cuda.synchronize()
    for p, v32 in zip(opt.flat_params, opt.layout.views(opt.buffers.p32)):
        assert torch.equal(p.data, v32.to(torch.bfloat16))
