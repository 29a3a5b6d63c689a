"""Step time of the CIFAR convnet: reference modules (cuDNN strict fp32 / TF32) vs tensor-core modules."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc

dev = "cuda"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
x = torch.randn(B, 3, 32, 32, device=dev); y = torch.randint(0, 10, (B,), device=dev)
torch.backends.cudnn.benchmark = True


def bench(model, iters=20, graph=True):
    def body():
        for p in model.parameters():
            p.grad = None
        F.nll_loss(model(x), y).backward()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            body()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    run = body
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        run = g.replay
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for name, mk, tf32 in [("cuDNN fp32", lambda: cifar_convnet(0), False), ("cuDNN tf32", lambda: cifar_convnet(0), True),
                       ("tcgen05 3xTF32, torch BN/ReLU/pool", lambda: cifar_convnet_tc(0, 3, fused_tail=False), False),
                       ("tcgen05 3xTF32 + fused BN/ReLU/pool", lambda: cifar_convnet_tc(0, 3), False),
                       ("tcgen05 TF32 + fused BN/ReLU/pool", lambda: cifar_convnet_tc(0, 1), False)]:
    torch.backends.cudnn.allow_tf32 = tf32
    m = mk().to(dev)
    print(f"{name}: fwd+bwd {bench(m):.3f} ms/step at batch {B}", flush=True)
# per-op breakdown of the tensor-core model (eager, profiler-free: CUDA events around pieces)
from torch_distlearn_b200.ops import im2col5x5, tc_gemm, col2im5x5
for (cin, cout, H) in [(3, 64, 32), (64, 128, 16), (128, 256, 8), (256, 512, 4)]:
    xh = torch.randn(B, H, H, cin, device=dev); M = B * H * H
    ld = (25 * cin + 3) // 4 * 4
    w = torch.randn(cout, ld, device=dev); dy = torch.randn(M, cout, device=dev)
    col = im2col5x5(xh)

    def t(fn, n=10):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3
    from torch_distlearn_b200.ops.conv import _splits
    bn = 64 if cout <= 64 else 128
    t_i2c = t(lambda: im2col5x5(xh, out=col))
    t_fwd = t(lambda: tc_gemm(col, w, split_k=_splits((M + 127) // 128, (cout + bn - 1) // bn, ld)))
    bn2 = 64 if ld <= 64 else 128
    t_wg = t(lambda: tc_gemm(dy, col, a_mn=True, b_mn=True, split_k=_splits((cout + 127) // 128, (ld + bn2 - 1) // bn2, M)))
    dcol = torch.empty(M, ld, device=dev)
    t_dg = t(lambda: tc_gemm(dy, w, b_mn=True, out=dcol))
    t_c2i = t(lambda: col2im5x5(dcol, (B, H, H, cin)))
    print(f"conv {cin}->{cout} @{H}x{H}: im2col {t_i2c:.0f}us  fwd {t_fwd:.0f}us  wgrad {t_wg:.0f}us  dgrad {t_dg:.0f}us  col2im {t_c2i:.0f}us", flush=True)

# implicit-GEMM pieces (no im2col / col2im)
from torch_distlearn_b200.ops.conv import _tc_conv, _implicit_ok, TAPS
for (cin, cout, H) in [(64, 128, 16), (128, 256, 8), (256, 512, 4)]:
    if not _implicit_ok(B, H, H, cin, cout):
        continue
    xh = torch.randn(B, H, H, cin, device=dev); M = B * H * H
    w = torch.randn(cout, 25 * cin, device=dev); dy = torch.randn(B, H, H, cout, device=dev)
    y = torch.empty(M, cout, device=dev); dw = torch.empty(cout, 25 * cin, device=dev); dx = torch.empty(B, H, H, cin, device=dev)
    bn = 64 if cout <= 64 else 128
    bnc = 64 if cin <= 64 else 128
    t_f = t(lambda: _tc_conv(1, xh, w, y, None, B, H, H, cin, cout, _splits((M + 127) // 128, (cout + bn - 1) // bn, TAPS * cin), 3))
    t_w = t(lambda: _tc_conv(3, xh, dy, dw, None, B, H, H, cin, cout, _splits((cout + 127) // 128, (25 * cin + 127) // 128, M), 3))
    t_d = t(lambda: _tc_conv(2, dy, w, dx, None, B, H, H, cin, cout, _splits((M + 127) // 128, (cin + bnc - 1) // bnc, TAPS * cout), 3))
    fl = 2.0 * M * cout * 25 * cin
    print(f"implicit conv {cin}->{cout} @{H}x{H}: fwd {t_f:.0f}us ({fl / t_f / 1e6:.0f}TF)  wgrad {t_w:.0f}us ({fl / t_w / 1e6:.0f}TF)  dgrad {t_d:.0f}us ({fl / t_d / 1e6:.0f}TF)", flush=True)
