#!/usr/bin/env python
"""MNIST convnet with AllReduceEA (elastic averaging) — counterpart of examples/mnist-ea.lua.

Every node runs plain local SGD; every ``tau`` steps ``averageParameters`` pulls the node towards the
shared center and the center towards the nodes (tau=10, alpha=0.2 as shipped, examples/mnist-ea.lua:18).
Here that call is one fused kernel (csrc/fused_ea.cu).  Epoch end: ``synchronizeCenter`` (:121).
"""
from common import batches_this_epoch, forward_backward, setup, Throughput, MnistKernelStep

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import mnist_convnet
from torch_distlearn_b200.ops import local_sgd
from torch_distlearn_b200.utils import ConfusionMatrix, mnist_dataset
from torch_distlearn_b200.utils.config import base_parser
from torch_distlearn_b200.utils.data import MNIST_CLASSES


def main(argv=None):
    p = base_parser("MNIST AllReduceEA (examples/mnist-ea.lua)", batch_size=1, learning_rate=0.01)
    p.add_argument("--tau", "--communicationTime", dest="tau", type=int, default=10)
    p.add_argument("--alpha", type=float, default=0.2)
    p.add_argument("--eager", action="store_true", help="stock torch kernels for forward/backward instead of the one-launch step")
    opt = p.parse_args(argv)
    comm, log = setup(opt)
    dev = comm.device
    ds = mnist_dataset(opt.nodeIndex, opt.numNodes, opt.trainSamples, seed=1)
    if opt.pythonLoader:
        get_batch, num_batches = ds.sampledBatcher(opt.batchSize, "permutation", device=dev)
    else:       # C++ worker threads fill pinned batches ahead of the step (utils/native_loader.py)
        get_batch, num_batches = ds.nativeBatcher(opt.batchSize, "permutation", threads=opt.loaderThreads).sampledBatcher(dev)
    model = mnist_convnet(seed=opt.seed + opt.nodeIndex).to(dev)   # nodes start apart; the initial sync aligns them
    flat = dl.FlatParams(comm, model, with_momentum=opt.momentum != 0)
    ea = dl.AllReduceEA(comm, opt.tau, opt.alpha, flat=flat)
    if opt.resume:
        dl.load_checkpoint(opt.resume, flat, center=ea.center)
    ea.synchronizeParameters()                                     # examples/mnist-ea.lua:63
    cm = ConfusionMatrix(MNIST_CLASSES, comm=comm)
    n_batches = batches_this_epoch(num_batches(), opt)
    use_kernel = dev.type == "cuda"
    kstep = None
    if MnistKernelStep.usable(model, dev, opt.batchSize) and not opt.eager:
        kstep = MnistKernelStep(model, opt.batchSize)              # the whole local step in one launch (csrc/mnist_step.cu)
    for epoch in range(1, opt.numEpochs + 1):
        cm.zero()
        thr, loss_sum = Throughput(), 0.0
        for _ in range(n_batches):
            batch = get_batch()
            done = False
            if kstep is not None and batch.input.shape[0] == opt.batchSize:
                plain = opt.momentum == 0 and opt.weightDecay == 0
                out, loss, done = kstep(batch, lr=opt.learningRate if plain else None)
            else:
                out, loss = forward_backward(model, batch)
            if done:                                               # gradient and update happened in the same kernel
                pass
            elif use_kernel:                                       # local SGD (examples/mnist-ea.lua:103-107)
                local_sgd(flat.params.tensor, flat.grads.tensor, flat.momentum.tensor if flat.momentum else None,
                          lr=opt.learningRate, momentum=opt.momentum, weight_decay=opt.weightDecay, zero_grads=True)
            else:
                flat.params.tensor.add_(flat.grads.tensor, alpha=-opt.learningRate)
                flat.zero_grad()
            ea.averageParameters()                                 # :110 — every tau-th call talks to the peers
            cm.batchAdd(out, batch.target)
            loss_sum += float(loss)
            thr.add(batch.batchSize)
        ea.synchronizeCenter()                                     # :121
        cm.allReduce()
        comm.synchronize()
        dist2 = float(((flat.params.tensor - ea.center) ** 2).sum())
        log.print(f"epoch {epoch}: loss {loss_sum / max(n_batches, 1):.4f}  accuracy {100 * cm.totalValid:.2f}%  "
                  f"|x-center|^2 {dist2:.3e}  {thr.rate():.0f} samples/s/node")
        log.metric(epoch=epoch, loss=loss_sum / max(n_batches, 1), accuracy=cm.totalValid, dist2=dist2)
        if opt.save and opt.nodeIndex == 1:
            import os
            os.makedirs(opt.save, exist_ok=True)
            dl.save_checkpoint(os.path.join(opt.save, "mnist_ea.dlb200"), flat, center=ea.center, algo="AllReduceEA",
                               tau=opt.tau, alpha=opt.alpha, step=ea.step, world=comm.world, extra={"epoch": epoch})
    final = float(cm.totalValid)
    shutdown(comm)
    return final


if __name__ == "__main__":
    main()
