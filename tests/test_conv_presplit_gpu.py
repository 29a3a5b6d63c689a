"""Weights handed to the tensor-core convolution already split into TF32 hi / lo (ops/conv.py: split_weights,
csrc/tc_gemm.cu: dl_tc_conv5x5_ws) against the same convolution with the split done inside the kernel, and against
float64: same products in the same order, so the two agree to the last bits that stream-K's red.add order leaves open."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture
def presplit_state():
    from torch_distlearn_b200.ops import conv
    before = conv.presplit()
    yield conv
    conv.presplit(before)


def _run(conv, on, x, w, b, dy):
    conv.presplit(on)
    xi = x.clone().requires_grad_(True)
    wi, bi = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = conv.conv5x5_nhwc(xi, wi, bi)
    y.backward(dy)
    torch.cuda.synchronize()
    return y.detach(), xi.grad, wi.grad, bi.grad


@pytest.mark.parametrize("B,H,cin,cout", [(32, 16, 64, 128), (32, 8, 128, 256), (16, 4, 256, 512), (8, 16, 64, 64)])
def test_presplit_matches_in_kernel_split_and_float64(presplit_state, B, H, cin, cout):
    conv = presplit_state
    g = torch.Generator(device=DEV).manual_seed(cin + cout)
    x = torch.randn(B, H, H, cin, device=DEV, generator=g)
    w = torch.randn(cout, 25 * cin, device=DEV, generator=g) * 0.05
    b = torch.randn(cout, device=DEV, generator=g)
    dy = torch.randn(B, H, H, cout, device=DEV, generator=g)
    on = _run(conv, True, x, w, b, dy)
    off = _run(conv, False, x, w, b, dy)
    for a, c, name in zip(on, off, ("y", "dx", "dw", "db")):
        scale = float(c.abs().max())
        assert float((a - c).abs().max()) <= 2e-6 * scale, (name, float((a - c).abs().max()), scale)
    # float64 reference (NCHW, weight [Cout, Cin, 5, 5] from the (kh, kw, cin) column order)
    xd = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    wd = w.double().view(cout, 5, 5, cin).permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yd = F.conv2d(xd, wd, b.double(), padding=2)
    yd.backward(dy.double().permute(0, 3, 1, 2))
    y_ref = yd.detach().permute(0, 2, 3, 1)
    dx_ref = xd.grad.permute(0, 2, 3, 1)
    assert float((on[0].double() - y_ref).abs().max()) <= 3e-6 * float(y_ref.abs().max())
    assert float((on[1].double() - dx_ref).abs().max()) <= 3e-6 * float(dx_ref.abs().max())


def test_split_halves_add_up_and_hi_is_tf32():
    from torch_distlearn_b200.ops.conv import split_weights
    w = torch.randn(128, 1600, device=DEV)
    e = split_weights(w)
    torch.cuda.synchronize()
    hi, lo = e.buf[0], e.buf[1]
    assert torch.equal(hi + lo, w), "lo is the exact remainder"
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0, "hi has the 13 low mantissa bits clear"


def test_buffer_taken_by_a_second_forward_is_rebuilt_for_the_first_backward(presplit_state):
    conv = presplit_state
    conv.presplit(True)
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(8, 8, 8, 128, device=DEV, generator=g, requires_grad=True)
    w = torch.nn.Parameter(torch.randn(128, 3200, device=DEV, generator=g) * 0.05)
    y1 = conv.conv5x5_nhwc(x, w, None)
    with torch.no_grad():
        w_old = w.detach().clone()
    y2 = conv.conv5x5_nhwc(x.detach(), w, None)          # same weight tensor: reuses (and re-makes) the split buffer
    dy = torch.randn_like(y1)
    y1.backward(dy)
    torch.cuda.synchronize()
    conv.presplit(False)
    x2 = x.detach().clone().requires_grad_(True)
    conv.conv5x5_nhwc(x2, w_old.requires_grad_(True), None).backward(dy)
    torch.cuda.synchronize()
    assert float((x.grad - x2.grad).abs().max()) <= 2e-6 * float(x2.grad.abs().max())
    assert y2.shape == y1.shape
