#!/usr/bin/env python
"""MNIST convnet with AllReduceSGD — the counterpart of examples/mnist.lua / examples/mnist.sh.

    ./examples/mnist.sh                                   # 4 CPU processes, like the reference
    python -m torch_distlearn_b200.launch -n 4 examples/mnist.py
    torchrun --nproc-per-node 8 examples/mnist.py --cuda  # one process per B200

Per step (examples/mnist.lua:97-116): gradients of one batch, ``sumAndNormalizeGradients``, SGD.
With ``--fused`` (default) those two become ONE launch of the fused sm_100a kernel
(``AllReduceSGD.step``); ``--no-fused`` keeps the reference's call sequence.  Every ``--reportEvery``
iterations the confusion matrix is summed over the nodes (:119-125), at the end of every epoch the
parameters are re-synchronised (:129), which also absorbs uneven partition sizes.
"""
from common import batches_this_epoch, forward_backward, min_over_nodes, setup, Throughput, MnistKernelStep

import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import mnist_convnet
from torch_distlearn_b200.utils import ConfusionMatrix, mnist_dataset
from torch_distlearn_b200.utils.config import base_parser
from torch_distlearn_b200.utils.data import MNIST_CLASSES


def main(argv=None):
    p = base_parser("MNIST AllReduceSGD (examples/mnist.lua)", batch_size=1, learning_rate=0.01)
    p.add_argument("--reportEvery", type=int, default=1000)
    p.add_argument("--eager", action="store_true", help="torch forward/backward instead of the one-launch kernel (GPU)")
    opt = p.parse_args(argv)
    comm, log = setup(opt)
    dev = comm.device
    # partition the data over the nodes (examples/mnist.lua:26-29)
    ds = mnist_dataset(opt.nodeIndex, opt.numNodes, opt.trainSamples, seed=1)
    if opt.pythonLoader:
        get_batch, num_batches = ds.sampledBatcher(opt.batchSize, "permutation", device=dev)
    else:       # C++ worker threads fill pinned batches ahead of the step (utils/native_loader.py)
        get_batch, num_batches = ds.nativeBatcher(opt.batchSize, "permutation", threads=opt.loaderThreads).sampledBatcher(dev)
    model = mnist_convnet(seed=opt.seed).to(dev)
    flat = dl.FlatParams(comm, model, with_momentum=opt.momentum != 0)
    sgd = dl.AllReduceSGD(comm, flat)
    start_epoch = 1
    if opt.resume:
        hdr = dl.load_checkpoint(opt.resume, flat, momentum=flat.momentum)
        start_epoch = int(hdr["extra"].get("epoch", 0)) + 1
    sgd.synchronizeParameters()                                    # examples/mnist.lua:72
    cm = ConfusionMatrix(MNIST_CLASSES, comm=comm)
    n_batches = batches_this_epoch(num_batches(), opt)
    common_batches = min_over_nodes(comm, n_batches)               # periodic reports need every node present
    kstep = None
    if MnistKernelStep.usable(model, dev, opt.batchSize) and not opt.eager:
        kstep = MnistKernelStep(model, opt.batchSize)              # forward + backward in one launch (csrc/mnist_step.cu);
                                                                   # it ADDS into .grad like autograd, the all-reduce follows
    for epoch in range(start_epoch, opt.numEpochs + 1):
        cm.zero()
        thr, loss_sum = Throughput(), 0.0
        for i in range(1, n_batches + 1):
            batch = get_batch()
            if kstep is not None and batch.input.shape[0] == opt.batchSize:
                out, loss, _ = kstep(batch)
            else:
                out, loss = forward_backward(model, batch)
            if opt.fused:
                sgd.step(lr=opt.learningRate, momentum=opt.momentum, weight_decay=opt.weightDecay, variant=opt.variant)
            else:
                sgd.sumAndNormalizeGradients(flat)                 # examples/mnist.lua:109
                with torch.no_grad():                              # :112-116
                    flat.params.tensor.add_(flat.grads.tensor, alpha=-opt.learningRate)
                flat.zero_grad()
            cm.batchAdd(out, batch.target)
            loss_sum += float(loss)
            thr.add(batch.batchSize)
            if i % opt.reportEvery == 0 and i <= common_batches:
                cm.allReduce()                                     # examples/mnist.lua:122
                comm.synchronize()
                log.print(f"epoch {epoch} it {i}: accuracy {100 * cm.totalValid:.2f}% (all nodes)")
                cm.zero()
        winner = sgd.synchronizeParameters()                       # examples/mnist.lua:129
        cm.allReduce()
        comm.synchronize()
        log.print(f"epoch {epoch}: loss {loss_sum / max(n_batches, 1):.4f}  accuracy {100 * cm.totalValid:.2f}%  "
                  f"{thr.rate():.0f} samples/s/node  (params from node {winner + 1})")
        log.metric(epoch=epoch, loss=loss_sum / max(n_batches, 1), accuracy=cm.totalValid, samples_per_s=thr.rate() * comm.world)
        if opt.save and opt.nodeIndex == 1:
            import os
            os.makedirs(opt.save, exist_ok=True)
            dl.save_checkpoint(os.path.join(opt.save, "mnist.dlb200"), flat, momentum=flat.momentum, algo="AllReduceSGD",
                               world=comm.world, extra={"epoch": epoch})
    final = float(cm.totalValid)
    shutdown(comm)
    return final


if __name__ == "__main__":
    main()
