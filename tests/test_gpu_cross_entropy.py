"""Fused bf16 cross-entropy (csrc/cross_entropy.cu) against the plain PyTorch fp32 computation."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,V", [(64, 30000), (37, 1000), (5, 32008)])
def test_cross_entropy_matches_torch(rows, V):
    from dedloc_b200.ops.cross_entropy import cross_entropy_rows, supported

    torch.manual_seed(0)
    logits = (torch.randn(rows, V, device="cuda") * 3).to(torch.bfloat16).requires_grad_(True)
    labels = torch.randint(0, V, (rows,), device="cuda")
    weight = torch.rand(rows, device="cuda")
    weight[::5] = 0.0
    assert supported(logits)
    loss = cross_entropy_rows(logits, labels)
    (loss * weight).sum().backward()
    ref_logits = logits.detach().float().requires_grad_(True)
    ref = F.cross_entropy(ref_logits, labels, reduction="none")
    (ref * weight).sum().backward()
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-4)
    # the gradient is rounded to bf16 (what the decoder GEMMs consume)
    torch.testing.assert_close(logits.grad.float(), ref_logits.grad, rtol=1e-2, atol=1e-5)
    assert logits.grad.dtype == torch.bfloat16
    assert torch.all(logits.grad[::5] == 0)


def test_cross_entropy_extreme_rows():
    from dedloc_b200.ops.cross_entropy import cross_entropy_rows

    V = 1024
    logits = torch.full((3, V), -1e4, device="cuda", dtype=torch.bfloat16)
    logits[0, 7] = 50.0          # confident and right
    logits[1, 9] = 50.0          # confident and wrong
    logits[2] = 0.0              # uniform
    labels = torch.tensor([7, 3, 0], device="cuda")
    loss = cross_entropy_rows(logits, labels)
    ref = F.cross_entropy(logits.float(), labels, reduction="none")
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-3)
    assert torch.isfinite(loss).all()
