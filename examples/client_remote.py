#!/usr/bin/env python
"""CIFAR-10 elastic averaging across hosts — counterpart of examples/client_remote.lua.

The reference script builds an ``ipc.Tree`` by hand from ``--host/--port/--base/--clientIP``
(:31-41) so that nodes on different machines — and of different kinds: client_remote.sh pairs one
CPU node with one GPU node — can train together, then runs the AllReduceEA flow (local SGD,
``averageParameters`` every tau steps, ``synchronizeCenter`` per epoch; its calls are spelled against
AsyncEA by mistake).  Here ``--backend dist`` rendezvouses on ``--host:--port`` through
torch.distributed and runs the same algorithm classes over library collectives; inside one NVSwitch
domain use ``--backend p2p`` (the default of the other examples) for the fused kernels.
"""
import math

from common import batches_this_epoch, evaluate, forward_backward, setup, Throughput

import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import cifar_convnet
from torch_distlearn_b200.utils import ConfusionMatrix, Data
from torch_distlearn_b200.utils.config import base_parser


def main(argv=None):
    p = base_parser("CIFAR-10 AllReduceEA across hosts (examples/client_remote.lua)", batch_size=32, learning_rate=0.1)
    p.add_argument("--tau", "--communicationTime", dest="tau", type=int, default=10)
    p.add_argument("--alpha", type=float, default=0.2)
    p.add_argument("--base", type=int, default=2, help="accepted for parity: there is no k-ary tree any more")
    p.add_argument("--clientIP", type=str, default="", help="accepted for parity")
    p.set_defaults(backend="dist")
    opt = p.parse_args(argv)
    comm, log = setup(opt)
    dev = comm.device
    batch = math.ceil(opt.batchSize / opt.numNodes)
    data = Data(opt.nodeIndex, opt.numNodes, batch, device=dev, train_samples=opt.trainSamples, test_samples=opt.testSamples,
                native=not opt.pythonLoader, threads=opt.loaderThreads)
    model = cifar_convnet(seed=opt.seed).to(dev)
    flat = dl.FlatParams(comm, model)
    ea = dl.AllReduceEA(comm, opt.tau, opt.alpha, flat=flat)
    ea.synchronizeParameters()                                      # examples/client_remote.lua:158
    cm = ConfusionMatrix(data.classes, comm=comm)
    n_train = batches_this_epoch(data.numTrainingBatches(), opt)
    for epoch in range(1, opt.numEpochs + 1):
        cm.zero()
        thr = Throughput()
        for i in range(1, n_train + 1):
            b = data.getTrainingBatch()
            out, loss = forward_backward(model, b)
            with torch.no_grad():                                   # :204-208
                flat.params.tensor.add_(flat.grads.tensor, alpha=-opt.learningRate)
            flat.zero_grad()
            ea.averageParameters()                                  # :210
            cm.batchAdd(out, b.target)
            thr.add(b.batchSize)
        cm.allReduce()                                              # :223
        ea.synchronizeCenter()                                      # :228
        comm.synchronize()
        log.print(f"epoch {epoch} train: accuracy {100 * cm.totalValid:.2f}%  {thr.rate() * comm.world:.0f} samples/s (all nodes)")
        cm.zero()
        test_loss = evaluate(model, data.getTestBatch, data.numTestBatches(), cm)
        cm.allReduce()                                              # :254
        comm.synchronize()
        log.print(f"epoch {epoch} test : loss {test_loss:.4f}  accuracy {100 * cm.totalValid:.2f}%")
    final = float(cm.totalValid)
    shutdown(comm)
    return final


if __name__ == "__main__":
    main()
