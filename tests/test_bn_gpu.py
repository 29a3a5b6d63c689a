"""Fused BatchNorm(train) + ReLU + 2x2 max-pool (csrc/bn_ops.cu) vs the same three torch ops in float64."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("B,H,C", [(8, 32, 64), (16, 16, 128), (32, 8, 256), (128, 4, 512), (4, 4, 12)])
def test_fused_bn_relu_pool_matches_float64(B, H, C):
    from torch_distlearn_b200.ops import FusedBNReLUPool
    g = torch.Generator().manual_seed(B + C)
    y = (torch.randn(B, H, H, C, generator=g) * 2.0 + 0.5).to(DEV)
    dp = torch.randn(B, H // 2, H // 2, C, generator=g).to(DEV)
    mod = FusedBNReLUPool(C, eps=1e-3).to(DEV)
    with torch.no_grad():
        mod.weight.copy_(torch.rand(C, generator=g) + 0.5)
        mod.bias.copy_(torch.randn(C, generator=g) * 0.3)
    ref = torch.nn.BatchNorm2d(C, eps=1e-3).to(DEV).double()
    ref.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in mod.state_dict().items()})
    y1 = y.clone().requires_grad_(True)
    p = mod(y1)
    p.backward(dp)
    y64 = y.double().requires_grad_(True)
    p64 = F.max_pool2d(F.relu(ref(y64.permute(0, 3, 1, 2))), 2, 2).permute(0, 2, 3, 1)
    p64.backward(dp.double())

    def close(a, b, tol):
        err = float((a.double() - b).abs().max() / (b.abs().max() + 1e-30))
        assert err < tol, err

    close(p, p64, 2e-6)
    close(y1.grad, y64.grad, 5e-5)
    close(mod.weight.grad, ref.weight.grad, 5e-5)
    close(mod.bias.grad, ref.bias.grad, 5e-5)
    # a second backward accumulates into the existing .grad inside the kernel: exactly twice the first
    g1, b1 = mod.weight.grad.clone(), mod.bias.grad.clone()
    rm, rv = mod.running_mean.clone(), mod.running_var.clone()
    mod(y.clone().requires_grad_(True)).backward(dp)
    torch.testing.assert_close(mod.weight.grad, 2 * g1, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(mod.bias.grad, 2 * b1, rtol=1e-5, atol=1e-5)
    mod.running_mean.copy_(rm); mod.running_var.copy_(rv)
    close(mod.running_mean, ref.running_mean, 1e-5)
    close(mod.running_var, ref.running_var, 1e-5)
    # eval mode uses the running statistics through the plain torch path
    mod.eval(); ref.eval()
    close(mod(y), F.max_pool2d(F.relu(ref(y.double().permute(0, 3, 1, 2))), 2, 2).permute(0, 2, 3, 1), 1e-5)


@pytest.mark.parametrize("B,H,C", [(128, 16, 128), (16, 32, 64), (128, 4, 512)])
def test_conv_bias_gradient_comes_out_of_the_bn_backward(B, H, C):
    """dl_bn_relu_pool_bwd_bias: the column sums of dy (= bias gradient of the convolution in front) from the same pass."""
    from torch_distlearn_b200.ops import FusedBNReLUPool
    from torch_distlearn_b200.ops.bn import _BIAS_DONE
    g = torch.Generator().manual_seed(C)
    y = torch.randn(B, H, H, C, generator=g).to(DEV)
    dp = torch.randn(B, H // 2, H // 2, C, generator=g).to(DEV)
    conv_bias = torch.nn.Parameter(torch.zeros(C, device=DEV))
    conv_bias.grad = torch.full((C,), 0.25, device=DEV)          # the kernel adds into the existing gradient
    mod = FusedBNReLUPool(C).to(DEV)
    mod.attach_conv_bias(conv_bias)
    for p in mod.parameters():
        p.grad = torch.zeros_like(p)                              # direct accumulation needs allocated gradients
    y1 = y.clone().requires_grad_(True)
    mod(y1).backward(dp)
    torch.cuda.synchronize()
    assert id(conv_bias) in _BIAS_DONE
    _BIAS_DONE.discard(id(conv_bias))
    want = y1.grad.double().sum((0, 1, 2)) + 0.25
    # the true value is ~0 (a bias in front of a BatchNorm has no gradient): compare on the scale of the summands
    scale = float(y1.grad.abs().sum((0, 1, 2)).max())
    assert float((conv_bias.grad.double() - want).abs().max()) < 1e-5 * scale + 1e-6
