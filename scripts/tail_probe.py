#!/usr/bin/env python
"""Warm, in-graph time of every kernel group of the flagship step at its real shapes (CIFAR convnet, batch 128).

`profiles/step_launch_list.md` is an ncu list: cold caches, serialised launches — good for shares, wrong for absolutes
(the step runs in 1.13 ms as a graph but the list adds up to 1.32 ms).  Here every group is captured into a CUDA graph
of `--reps` back-to-back calls and replayed: launch gaps and cache state are the ones of the real step.

    python scripts/tail_probe.py [--batch 128] [--reps 50] [--out gpurun_out/tail_probe.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch


def timed_graph(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    torch.cuda.current_stream().wait_stream(side)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps       # microseconds per call


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from torch_distlearn_b200 import _lib
    from torch_distlearn_b200.ops import bn as bnops
    from torch_distlearn_b200.ops import conv as convops
    from torch_distlearn_b200.ops.head import head_nll
    lib = _lib.cuda()
    B = args.batch
    st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    rows = []

    def add(name, us, nbytes=0):
        rows.append({"op": name, "us": round(us, 2), "gbs": round(nbytes / us / 1e3, 1) if nbytes else None})
        print(f"{name:44s} {us:8.2f} us" + (f"   {nbytes / us / 1e3:7.0f} GB/s" if nbytes else ""), flush=True)

    chans = (3, 64, 128, 256, 512)
    for li in range(4):
        cin, cout, H = chans[li], chans[li + 1], 32 >> li
        M = B * H * H
        x = torch.randn(B, H, H, cin, device=dev)
        w = torch.randn(cout, 25 * cin, device=dev) * 0.05
        bias = torch.randn(cout, device=dev)
        y = torch.randn(B, H, H, cout, device=dev)
        dy = torch.randn(B, H, H, cout, device=dev)
        dp = torch.randn(B, H // 2, H // 2, cout, device=dev)
        p = torch.empty(B, H // 2, H // 2, cout, device=dev)
        gamma, beta = torch.ones(cout, device=dev), torch.zeros(cout, device=dev)
        mean, invstd = torch.empty(cout, device=dev), torch.empty(cout, device=dev)
        dgamma, dbeta = torch.zeros(cout, device=dev), torch.zeros(cout, device=dev)
        work = torch.zeros(int(lib.dl_bn_work_bytes(cout)) // 8 + 1, dtype=torch.float64, device=dev)
        act_bytes = M * cout * 4
        tag = f"L{li + 1} [{B}x{H}x{H}x{cout}]"
        # convolution: forward / dgrad / wgrad through the autograd function's own launch code
        xi = x.clone().requires_grad_(li > 0)
        wi, bi = torch.nn.Parameter(w.clone()), torch.nn.Parameter(bias.clone())
        wi.grad, bi.grad = torch.zeros_like(wi), torch.zeros_like(bi)

        def conv_fwd():
            return convops.conv5x5_nhwc(xi, wi, bi)

        with torch.no_grad():
            add(f"{tag} conv forward", timed_graph(conv_fwd, args.reps))

        def conv_fwd_bwd():
            out = convops.conv5x5_nhwc(xi, wi, bi)
            out.backward(dy)

        add(f"{tag} conv forward+backward", timed_graph(conv_fwd_bwd, args.reps))

        if li > 0:      # the three implicit GEMMs one by one, fp32-accurate (3xTF32) and plain TF32
            TAPS = 25
            yo = torch.empty(M, cout, device=dev)
            dxo = torch.empty(B, H, H, cin, device=dev)
            dwo = torch.zeros(cout, TAPS * cin, device=dev)
            dyc = dy.contiguous()
            flop = 2.0 * M * cout * TAPS * cin
            for terms in (3, 1):
                sp_f = convops._splits((M + 127) // 128, (cout + 127) // 128, TAPS * cin)
                sp_d = convops._splits((M + 127) // 128, (cin + (63 if cin <= 64 else 127)) // (64 if cin <= 64 else 128), TAPS * cout)
                sp_w = convops._splits((cout + 127) // 128, (TAPS * cin + 127) // 128, M)
                for nm, fn in (("forward", lambda: convops._tc_conv(1, x, w, yo, bias, B, H, H, cin, cout, sp_f, terms)),
                               ("dgrad", lambda: convops._tc_conv(2, dyc, w, dxo, None, B, H, H, cin, cout, sp_d, terms)),
                               ("wgrad", lambda: convops._tc_conv(3, x, dyc, dwo, None, B, H, H, cin, cout, sp_w, terms, accumulate=True))):
                    us = timed_graph(fn, args.reps)
                    rows.append({"op": f"{tag} gemm {nm} terms={terms}", "us": round(us, 2), "tflops": round(flop / us / 1e6, 1)})
                    print(f"{tag + ' gemm ' + nm + f' terms={terms}':44s} {us:8.2f} us   {flop / us / 1e6:7.1f} TFLOP/s (algorithmic)", flush=True)

        def bn_fwd():
            _lib.check(lib.dl_bn_relu_pool_fwd(y.data_ptr(), p.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(),
                                               invstd.data_ptr(), None, None, B, H, H, cout, 1e-3, 0.1, work.data_ptr(), st()))

        add(f"{tag} bn+relu+pool forward", timed_graph(bn_fwd, args.reps), int(act_bytes * 2.25))

        def bn_bwd():
            _lib.check(lib.dl_bn_relu_pool_bwd(y.data_ptr(), dp.data_ptr(), dy.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                               mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), B, H, H, cout, 1,
                                               work.data_ptr(), st()))

        add(f"{tag} bn+relu+pool backward", timed_graph(bn_bwd, args.reps), int(act_bytes * 3.5))

        def cs():
            _lib.check(lib.dl_colsum(dy.data_ptr(), dbeta.data_ptr(), M, cout, 1, work.data_ptr(), st()))

        add(f"{tag} colsum (conv bias gradient)", timed_graph(cs, args.reps), act_bytes)
    # head
    feat = torch.randn(B, 2, 2, 512, device=dev, requires_grad=True)
    hw = torch.nn.Parameter(torch.randn(10, 2048, device=dev) * 0.02)
    hb = torch.nn.Parameter(torch.zeros(10, device=dev))
    hw.grad, hb.grad = torch.zeros_like(hw), torch.zeros_like(hb)
    tgt = torch.randint(0, 10, (B,), device=dev)
    with torch.no_grad():
        add("head forward (linear+logsoftmax+nll)", timed_graph(lambda: head_nll(feat, hw, hb, tgt), args.reps))

    def head_fb():
        loss, _ = head_nll(feat, hw, hb, tgt)
        loss.backward()

    add("head forward+backward", timed_graph(head_fb, args.reps))
    xin = torch.randn(B, 3, 32, 32, device=dev)
    add("input NCHW->NHWC permute", timed_graph(lambda: xin.permute(0, 2, 3, 1).contiguous(), args.reps), B * 3 * 32 * 32 * 8)
    n = 4328992
    pflat, gflat = torch.randn(n, device=dev), torch.randn(n, device=dev)
    add("local sgd + zero grads (world 1)", timed_graph(
        lambda: _lib.check(lib.dl_local_sgd(pflat.data_ptr(), gflat.data_ptr(), None, n, 0.1, 0.0, 0.0, 0, 1.0, 1, st())), args.reps),
        n * 4 * 4)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump({"batch": B, "reps": args.reps, "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
