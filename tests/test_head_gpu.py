"""Fused classifier head (csrc/head_ops.cu) vs Linear + log_softmax + nll_loss in float64, and the whole
tensor-core model trained through forward_with_loss vs through the plain torch head."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("B", [128, 8, 1])
def test_head_forward_backward_matches_float64(B):
    from torch_distlearn_b200.ops import head_nll
    g = torch.Generator().manual_seed(B)
    feat = torch.randn(B, 2, 2, 512, generator=g).to(DEV).requires_grad_(True)
    lin = torch.nn.Linear(2048, 10).to(DEV)
    y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    loss, logp = head_nll(feat, lin.weight, lin.bias, y)
    (loss * 1.5).backward()
    f64 = feat.detach().double().requires_grad_(True)
    w64, b64 = lin.weight.detach().double().requires_grad_(True), lin.bias.detach().double().requires_grad_(True)
    out64 = F.log_softmax(F.linear(f64.permute(0, 3, 1, 2).reshape(B, 2048), w64, b64), dim=1)   # (c, h, w) feature order
    l64 = F.nll_loss(out64, y)
    (l64 * 1.5).backward()

    def close(a, b, tol=2e-5):
        err = float((a.detach().double() - b.detach()).abs().max() / (b.detach().abs().max() + 1e-30))
        assert err < tol, err

    close(logp, out64)
    close(loss, l64)
    close(feat.grad, f64.grad)
    close(lin.weight.grad, w64.grad)
    close(lin.bias.grad, b64.grad)
    # second backward: accumulates into the existing .grad inside the kernel
    w1, b1 = lin.weight.grad.clone(), lin.bias.grad.clone()
    loss2, _ = head_nll(feat.detach().requires_grad_(True), lin.weight, lin.bias, y)
    (loss2 * 1.5).backward()
    torch.testing.assert_close(lin.weight.grad, 2 * w1, rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(lin.bias.grad, 2 * b1, rtol=1e-5, atol=1e-7)


def test_model_forward_with_loss_equals_plain_head():
    from torch_distlearn_b200.models import cifar_convnet_tc
    g = torch.Generator().manual_seed(3)
    x = torch.randn(32, 3, 32, 32, generator=g).to(DEV)
    y = torch.randint(0, 10, (32,), generator=g).to(DEV)
    a, b = cifar_convnet_tc(0).to(DEV), cifar_convnet_tc(0).to(DEV)
    out_a, loss_a = a.forward_with_loss(x, y)
    loss_a.backward()
    out_b = b(x)
    loss_b = F.nll_loss(out_b, y)
    loss_b.backward()
    torch.testing.assert_close(out_a, out_b, rtol=1e-4, atol=1e-5)
    assert abs(float(loss_a) - float(loss_b)) < 1e-5
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        torch.testing.assert_close(p.grad, q.grad, rtol=2e-3, atol=1e-5 * float(q.grad.abs().max()) + 1e-7, msg=lambda s, n=n: f"{n}: {s}")
