"""Stream-K form of the tensor-core GEMM (csrc/tc_gemm_streamk.inl) against the default one-tile-per-CTA kernel and
float64.  The stream-K kernel is the default (validated on a B200 in round 2); these tests switch between the two
forms explicitly with dl_tc_gemm_streamk().
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture()
def streamk():
    from torch_distlearn_b200 import _lib
    lib = _lib.cuda()

    class Switch:
        def __call__(self, on):
            lib.dl_tc_gemm_streamk(int(on))
    before = lib.dl_tc_gemm_streamk(0)
    yield Switch()
    lib.dl_tc_gemm_streamk(before)


def _rel_err(got, want):
    return float((got.double() - want).abs().max() / want.abs().max())


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 128, 1600), (32768, 128, 1600), (8192, 256, 3200), (2048, 512, 6400),
                                   (200, 72, 76), (128, 10, 2048), (131072, 64, 96), (4096, 4096, 512)])
@pytest.mark.parametrize("terms", [3, 1])
def test_streamk_gemm_matches_default_kernel(streamk, M, N, K, terms):
    from torch_distlearn_b200.ops import tc_gemm
    g = torch.Generator().manual_seed(M + N + K)
    a, b = torch.randn(M, K, generator=g).to(DEV), torch.randn(N, K, generator=g).to(DEV)
    bias = torch.randn(N, generator=g).to(DEV)
    streamk(0)
    want = tc_gemm(a, b, bias=bias, terms=terms)
    streamk(1)
    got = tc_gemm(a, b, bias=bias, terms=terms)
    torch.cuda.synchronize()
    tol = 2e-5 if terms == 3 else 3e-3
    assert _rel_err(got, want.double()) < tol
    if terms == 3:
        assert _rel_err(got, a.double() @ b.double().T + bias.double()) < 2e-5
    # accumulate: C += A B^T, nothing zeroed
    base = torch.randn(M, N, generator=g).to(DEV)
    out = base.clone()
    tc_gemm(a, b, out=out, accumulate=True, terms=terms)
    torch.cuda.synchronize()
    assert _rel_err(out, want.double() - bias.double() + base.double()) < tol


def test_streamk_operand_layouts_and_strided_output(streamk):
    from torch_distlearn_b200.ops import tc_gemm
    g = torch.Generator().manual_seed(7)
    M, N, K = 640, 200, 4096
    a_t, b_t = torch.randn(K, M, generator=g).to(DEV), torch.randn(K, N, generator=g).to(DEV)      # MN-major operands
    want = a_t.double().T @ b_t.double()
    streamk(1)
    got = tc_gemm(a_t, b_t, a_mn=True, b_mn=True)
    torch.cuda.synchronize()
    assert _rel_err(got, want) < 2e-5
    big = torch.zeros(M, N + 56, device=DEV)
    view = big[:, :N]
    tc_gemm(a_t, b_t, a_mn=True, b_mn=True, out=view)
    torch.cuda.synchronize()
    assert _rel_err(view, want) < 2e-5 and float(big[:, N:].abs().max()) == 0.0


@pytest.mark.parametrize("B,H,cin,cout", [(128, 16, 64, 128), (128, 8, 128, 256), (128, 4, 256, 512), (16, 16, 64, 128)])
def test_streamk_convolution_forward_backward(streamk, B, H, cin, cout):
    from torch_distlearn_b200.ops.conv import conv5x5_nhwc
    g = torch.Generator().manual_seed(B + H + cin)
    x = torch.randn(B, H, H, cin, generator=g).to(DEV)
    w = (torch.randn(cout, 25 * cin, generator=g) * 0.05).to(DEV)
    bias = torch.randn(cout, generator=g).to(DEV)
    dy = torch.randn(B, H, H, cout, generator=g).to(DEV)
    outs = []
    for on in (0, 1):
        streamk(on)
        xi, wi, bi = x.clone().requires_grad_(), w.clone().requires_grad_(), bias.clone().requires_grad_()
        y = conv5x5_nhwc(xi, wi, bi)
        y.backward(dy)
        torch.cuda.synchronize()
        outs.append((y.detach(), xi.grad, wi.grad, bi.grad))
    for got, want in zip(outs[1], outs[0]):
        assert _rel_err(got, want.double()) < 2e-5


def test_streamk_whole_model_step(streamk):
    """The flagship step with every GEMM on the stream-K kernel, inside a CUDA graph, against the default kernels."""
    import torch_distlearn_b200 as dl
    from torch_distlearn_b200.models import cifar_convnet_tc
    from torch_distlearn_b200.parallel.trainer import TrainStep
    outs = []
    for on in (0, 1):
        streamk(on)
        comm = dl.P2PCommunicator(rank=0, world=1, device=torch.device(DEV), heap_mb=96)
        model = cifar_convnet_tc().to(DEV)
        flat = dl.FlatParams(comm, model)
        sgd = dl.AllReduceSGD(comm, flat)
        step = TrainStep(model, flat, sgd, (128, 3, 32, 32), lr=0.01, use_graph=True, warmup=2)
        g = torch.Generator().manual_seed(3)
        losses, first = [], None
        for _ in range(4):
            step.step(torch.randn(128, 3, 32, 32, generator=g), torch.randint(0, 10, (128,), generator=g))
            losses.append(step.loss_value())
            if first is None:
                first = flat.params.tensor.clone()
        outs.append((flat.params.tensor.clone(), losses, first))
    # one step apart the two kernels differ by summation order only; after four steps at lr 0.01 on random labels that
    # seed has grown to ~1e-4 (measured 1.4e-4 on a B200; two runs of the SAME kernel drift apart at the same rate
    # because the split-K partial tiles meet in red.add order)
    torch.testing.assert_close(outs[1][2], outs[0][2], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(outs[1][0], outs[0][0], rtol=1e-2, atol=5e-4)
    torch.testing.assert_close(torch.tensor(outs[1][1]), torch.tensor(outs[0][1]), rtol=1e-3, atol=1e-3)

