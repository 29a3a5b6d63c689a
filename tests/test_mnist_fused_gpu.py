"""The one-launch MNIST step (csrc/mnist_step.cu, an 8-CTA cluster with distributed shared memory) against the plain
PyTorch network in float64: loss, log-probabilities, every gradient, and the fused SGD update."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True, params=[2, 1], ids=["form2", "form1"])
def kernel_form(request):
    """Every test runs on both forms of the kernel (csrc/mnist_step.cu: 416 threads, the default; and the first one)."""
    from torch_distlearn_b200 import _lib
    before = _lib.cuda().dl_mnist_form(request.param)
    yield request.param
    _lib.cuda().dl_mnist_form(before)


def _reference(model, x, y):
    ref = copy.deepcopy(model).double().cpu()
    out = ref(x.double().cpu())
    loss = F.nll_loss(out, y.cpu())
    loss.backward()
    return ref, out.detach(), float(loss)


@pytest.mark.parametrize("B", [1, 3, 8, 32, 200])
def test_gradients_match_float64(B):
    from torch_distlearn_b200.models import mnist_convnet
    from torch_distlearn_b200.ops.mnist import PARAM_NAMES, mnist_step, model_tensors
    g = torch.Generator().manual_seed(B)
    model = mnist_convnet().to(DEV)
    x = torch.randn(B, 1024, generator=g).to(DEV)
    y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    for p in model.parameters():
        p.grad = torch.full_like(p, 0.5)            # the kernel ADDS into .grad
    params, grads = model_tensors(model)
    loss = torch.zeros(1, device=DEV)
    logp = torch.zeros(B, 10, device=DEV)
    before = [p.clone() for p in params]
    mnist_step(x, y, params, grads, loss, logp)
    torch.cuda.synchronize()
    ref, out, ref_loss = _reference(model, x, y)
    assert abs(float(loss) - ref_loss) < 1e-5 * max(1.0, abs(ref_loss))
    torch.testing.assert_close(logp.cpu().double(), out, rtol=1e-5, atol=1e-5)
    named = dict(ref.named_parameters())
    for n, gr, p, old in zip(PARAM_NAMES, grads, params, before):
        assert torch.equal(p, old), "parameters must not move without apply_sgd"
        want = named[n].grad
        got = gr.cpu().double() - 0.5
        scale = float(want.abs().max()) + 1e-12
        assert float((got - want).abs().max()) < 2e-5 * scale + 1e-6, (n, float((got - want).abs().max()), scale)


@pytest.mark.parametrize("B", [1, 8])
def test_fused_update_and_determinism(B):
    from torch_distlearn_b200.models import mnist_convnet
    from torch_distlearn_b200.ops.mnist import PARAM_NAMES, mnist_step, model_tensors
    g = torch.Generator().manual_seed(7)
    x = torch.randn(B, 1024, generator=g).to(DEV)
    y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    runs = []
    for _ in range(2):
        model = mnist_convnet().to(DEV)
        for p in model.parameters():
            p.grad = torch.zeros_like(p)
        params, grads = model_tensors(model)
        loss = torch.zeros(1, device=DEV)
        ref, _, _ = _reference(model, x, y)
        for _ in range(1):
            mnist_step(x, y, params, grads, loss, None, lr=0.1, apply_sgd=True)
        torch.cuda.synchronize()
        named = dict(ref.named_parameters())
        for n, p, gr in zip(PARAM_NAMES, params, grads):
            want = named[n].detach() - 0.1 * named[n].grad
            torch.testing.assert_close(p.cpu().double(), want, rtol=1e-5, atol=1e-6)
            assert float(gr.abs().max()) == 0.0, "the fused update leaves .grad alone"
        runs.append([p.clone() for p in params])
    for a, b in zip(*runs):
        assert torch.equal(a, b), "two runs must give the same bits"


def test_training_with_the_fused_step_learns():
    from torch_distlearn_b200.models import mnist_convnet
    from torch_distlearn_b200.ops.mnist import mnist_step, model_tensors
    g = torch.Generator().manual_seed(3)
    model = mnist_convnet().to(DEV)
    for p in model.parameters():
        p.grad = torch.zeros_like(p)
    params, grads = model_tensors(model)
    x = torch.randn(4, 1024, generator=g).to(DEV)
    y = torch.randint(0, 10, (4,), generator=g).to(DEV)
    loss = torch.zeros(1, device=DEV)
    first = None
    for _ in range(60):
        mnist_step(x, y, params, grads, loss, None, lr=0.05, apply_sgd=True)
        first = float(loss) if first is None else first
    assert float(loss) < 0.5 * first
