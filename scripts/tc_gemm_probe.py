"""First-light probe for the tcgen05 GEMM: tiny problem, prints max error; run under `timeout`."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch_distlearn_b200.ops import tc_gemm
torch.manual_seed(0)
for (M, N, K) in [(128, 64, 32), (128, 128, 32), (128, 128, 128), (256, 256, 1600)]:
    a = torch.randn(M, K, device="cuda"); b = torch.randn(N, K, device="cuda")
    for terms in (1, 3):
        c = tc_gemm(a, b, terms=terms)
        torch.cuda.synchronize()
        want = a.double() @ b.double().T
        err = float((c.double() - want).abs().max() / want.abs().max())
        print(f"M={M} N={N} K={K} terms={terms}: rel err {err:.3e}", flush=True)
        if err > 1e-2:
            d = (c.double() - want).abs()
            bad = (d > 1e-2 * want.abs().max()).nonzero()
            print("  bad count", len(bad), "first", bad[:5].tolist(), "rows bad", sorted(set(bad[:, 0].tolist()))[:16], "cols bad", sorted(set(bad[:, 1].tolist()))[:16])
            print("  c[0,:8]", c[0, :8].tolist(), "want", want[0, :8].tolist())


def bench(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


print("shape: ours 3xTF32 / ours TF32 / torch fp32 / torch tf32  (us, TFLOP/s of 2MNK)")
for (M, N, K) in [(32768, 128, 1600), (8192, 256, 3200), (2048, 512, 6400), (131072, 64, 96), (512, 6400, 2048), (8192, 8192, 8192)]:
    a = torch.randn(M, K, device="cuda"); b = torch.randn(N, K, device="cuda"); c = torch.empty(M, N, device="cuda")
    fl = 2.0 * M * N * K
    t3 = bench(lambda: tc_gemm(a, b, out=c, terms=3))
    want = a[:256].double() @ b.double().T
    err3 = float((c[:256].double() - want).abs().max() / want.abs().max())
    t1 = bench(lambda: tc_gemm(a, b, out=c, terms=1))
    torch.backends.cuda.matmul.allow_tf32 = False
    tf = bench(lambda: torch.matmul(a, b.T, out=c))
    errf = float((c[:256].double() - want).abs().max() / want.abs().max())
    torch.backends.cuda.matmul.allow_tf32 = True
    tt = bench(lambda: torch.matmul(a, b.T, out=c))
    torch.backends.cuda.matmul.allow_tf32 = False
    print(f"{M}x{N}x{K}: {t3:.0f}us {fl / t3 / 1e6:.0f}TF (err {err3:.1e}) / {t1:.0f}us {fl / t1 / 1e6:.0f}TF / "
          f"{tf:.0f}us {fl / tf / 1e6:.0f}TF (err {errf:.1e}) / {tt:.0f}us {fl / tt / 1e6:.0f}TF", flush=True)
