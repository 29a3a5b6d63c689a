"""tcgen05 / TMEM / TMA GEMM (csrc/tc_gemm.cu) vs a float64 reference and vs torch's strict-fp32 matmul."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _mats(M, N, K, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(M, K, generator=g).to(DEV)
    b = torch.randn(N, K, generator=g).to(DEV)
    return a, b


def _rel_err(got, want):
    return float((got.double() - want).abs().max() / want.abs().max())


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (128, 64, 128), (256, 128, 1600), (384, 256, 96), (2048, 512, 6400),
                                   (200, 72, 75 + 1), (128, 10, 2048)])
def test_3xtf32_matches_fp64_like_fp32_does(M, N, K):
    from torch_distlearn_b200.ops import tc_gemm
    if K % 4:
        K += 4 - K % 4
    a, b = _mats(M, N, K)
    want = a.double() @ b.double().T
    got = tc_gemm(a, b)
    torch.cuda.synchronize()
    torch.backends.cuda.matmul.allow_tf32 = False
    fp32 = a @ b.T
    e_tc, e_fp32 = _rel_err(got, want), _rel_err(fp32, want)
    # 3xTF32 must sit in fp32 territory (plain TF32 would be ~1e-3)
    assert e_tc < 2e-5 and e_tc < 8 * e_fp32 + 2e-6, (e_tc, e_fp32)


def test_bias_splitk_accumulate_and_strided_views():
    from torch_distlearn_b200.ops import tc_gemm
    M, N, K = 256, 128, 4096
    a, b = _mats(M, N, K, 1)
    bias = torch.randn(N, device=DEV)
    want = a.double() @ b.double().T + bias.double()
    for split in (1, 4, 7):
        got = tc_gemm(a, b, bias=bias, split_k=split)
        torch.cuda.synchronize()
        assert _rel_err(got, want) < 2e-5, split
    base = torch.randn(M, N, device=DEV)
    out = base.clone()
    tc_gemm(a, b, out=out, split_k=3, accumulate=True)
    assert _rel_err(out, want - bias.double() + base.double()) < 2e-5
    # operands and output as views with a row pitch larger than the row
    big_a = torch.randn(M, K + 64, device=DEV)
    big_c = torch.zeros(M, N + 32, device=DEV)
    av, cv = big_a[:, :K], big_c[:, :N]
    tc_gemm(av, b, out=cv)
    torch.cuda.synchronize()
    assert _rel_err(cv, av.double() @ b.double().T) < 2e-5
    assert float(big_c[:, N:].abs().max()) == 0.0


def test_plain_tf32_mode_is_coarser():
    from torch_distlearn_b200.ops import tc_gemm
    a, b = _mats(256, 128, 2048, 2)
    want = a.double() @ b.double().T
    e1 = _rel_err(tc_gemm(a, b, terms=1), want)
    e3 = _rel_err(tc_gemm(a, b, terms=3), want)
    assert 1e-5 < e1 < 5e-3 and e3 < e1 / 20, (e1, e3)


@pytest.mark.parametrize("a_mn,b_mn", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 72, 1024), (64, 1600, 4096), (512, 76, 2048)])
def test_mn_major_operands(a_mn, b_mn, M, N, K):
    """Operands stored transposed ([K, M] / [K, N]) go through MN-major UMMA descriptors, no transpose copy."""
    from torch_distlearn_b200.ops import tc_gemm
    a, b = _mats(M, N, K, 3)
    want = a.double() @ b.double().T
    a_in = a.T.contiguous() if a_mn else a
    b_in = b.T.contiguous() if b_mn else b
    got = tc_gemm(a_in, b_in, a_mn=a_mn, b_mn=b_mn, split_k=3 if K >= 2048 else 1)
    torch.cuda.synchronize()
    assert _rel_err(got, want) < 2e-5


def _conv_case(B, cin, cout, H, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cin, H, H, generator=g).to(DEV)
    conv = torch.nn.Conv2d(cin, cout, 5, 1, 2).to(DEV)
    dy = torch.randn(B, cout, H, H, generator=g).to(DEV)
    return x, conv, dy


@pytest.mark.parametrize("implicit", [False, True])
@pytest.mark.parametrize("B,cin,cout,H", [(4, 3, 64, 32), (8, 64, 128, 16), (6, 128, 256, 8), (4, 256, 512, 4), (128, 64, 128, 16), (1, 32, 32, 16), (2, 32, 96, 8)])
def test_conv5x5_forward_backward_vs_float64_conv2d(B, cin, cout, H, implicit):
    """Explicit (im2col + GEMM + col2im) and implicit (4-D TMA halo boxes) paths against float64 conv2d."""
    from torch_distlearn_b200.ops import TCConv5x5
    from torch_distlearn_b200.ops.conv import _implicit_ok
    if implicit and not _implicit_ok(B, H, H, cin, cout):
        pytest.skip("geometry handled by the explicit path")
    x, conv, dy = _conv_case(B, cin, cout, H)
    tc = TCConv5x5(cin, cout, implicit=implicit).to(DEV)
    tc.load_conv2d(conv)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous().requires_grad_(True)
    y = tc(x_nhwc)
    y.backward(dy.permute(0, 2, 3, 1).contiguous())
    # float64 oracle
    x64 = x.double().requires_grad_(True)
    conv64 = torch.nn.Conv2d(cin, cout, 5, 1, 2).to(DEV).double()
    conv64.load_state_dict({k: v.double() for k, v in conv.state_dict().items()})
    y64 = conv64(x64)
    y64.backward(dy.double())
    # strict fp32 cuDNN for scale
    torch.backends.cudnn.allow_tf32 = False
    y32 = conv(x)

    def rel(a, b):
        return float((a.double() - b).abs().max() / b.abs().max())

    e_y, e_y32 = rel(y.permute(0, 3, 1, 2), y64), rel(y32, y64)
    assert e_y < 1e-5 and e_y < 10 * e_y32 + 1e-6, (e_y, e_y32)
    assert rel(x_nhwc.grad.permute(0, 3, 1, 2), x64.grad) < 1e-5
    assert rel(tc.weight.grad.view(cout, 5, 5, cin).permute(0, 3, 1, 2), conv64.weight.grad) < 1e-5
    assert rel(tc.bias.grad, conv64.bias.grad) < 1e-5
    # second backward: .grad exists now, so the backward kernels add into it themselves (ops.bn.grad_sink) —
    # the result must be what autograd's own accumulation would give: twice the first gradient
    w1, b1 = tc.weight.grad.clone(), tc.bias.grad.clone()
    tc(x_nhwc.detach().requires_grad_(True)).backward(dy.permute(0, 2, 3, 1).contiguous())
    torch.testing.assert_close(tc.weight.grad, 2 * w1, rtol=1e-5, atol=1e-6 * float(w1.abs().max()))
    torch.testing.assert_close(tc.bias.grad, 2 * b1, rtol=1e-5, atol=1e-6 * float(b1.abs().max()) + 1e-6)


def test_tc_model_matches_reference_model():
    """CifarConvNetTC == CifarConvNet: logits and parameter gradients are as close to a float64 run of the
    reference network as the strict-fp32 cuDNN network itself is."""
    import copy
    import torch.nn.functional as F
    from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = cifar_convnet(0).to(DEV)
    tc = cifar_convnet_tc(0).to(DEV)
    ref64 = copy.deepcopy(ref).double()
    g = torch.Generator().manual_seed(1)
    x = torch.randn(32, 3, 32, 32, generator=g).to(DEV)
    y = torch.randint(0, 10, (32,), generator=g).to(DEV)
    o_ref, o_tc, o_64 = ref(x), tc(x), ref64(x.double())
    F.nll_loss(o_ref, y).backward(); F.nll_loss(o_tc, y).backward(); F.nll_loss(o_64, y).backward()

    def err(a, b):
        return float((a.detach().double() - b.detach()).abs().max() / b.detach().abs().max())

    e_tc, e_ref = err(o_tc, o_64), err(o_ref, o_64)
    assert e_tc < 10 * e_ref + 1e-5, f"logits: tensor-core model {e_tc:.2e} vs cuDNN fp32 {e_ref:.2e} from float64"
    for i, (c64, cr, ct) in enumerate(zip(ref64.convs, ref.convs, tc.convs)):
        want = c64.weight.grad.permute(0, 2, 3, 1).reshape(ct.weight.shape)
        e_tc, e_ref = err(ct.weight.grad, want), err(cr.weight.grad, c64.weight.grad)
        assert e_tc < 10 * e_ref + 1e-5, f"conv{i + 1} weight grad: {e_tc:.2e} vs cuDNN {e_ref:.2e}"
        assert err(ct.bias.grad, c64.bias.grad) < 10 * err(cr.bias.grad, c64.bias.grad) + 1e-4
    assert err(tc.linear.weight.grad, ref64.linear.weight.grad) < 10 * err(ref.linear.weight.grad, ref64.linear.weight.grad) + 1e-5
