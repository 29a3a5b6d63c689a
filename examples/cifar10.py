#!/usr/bin/env python
"""CIFAR-10 convnet with AllReduceSGD — counterpart of examples/cifar10.lua (and its verbatim copy
cifar10tests.lua) / examples/cifar10.sh / examples/cifar10-cuda.sh.

The global ``--batchSize`` (default 32) is divided over the nodes (examples/cifar10.lua:36); training
batches come from a label-uniform sampler (:53-54); each epoch ends with ``synchronizeParameters`` and
a pass over the test partition whose confusion matrix is summed over the nodes (:203-236).

On GPUs the default path is the flagship one: forward + backward + the fused all-reduce/SGD kernel
captured in ONE CUDA graph per step (``--graph``), optionally launched per gradient bucket under
backward (``--bucketMB``).  ``--no-fused`` runs the reference's call sequence instead.
"""
import math

from common import batches_this_epoch, evaluate, forward_backward, setup, Throughput

import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc
from torch_distlearn_b200.parallel.trainer import TrainStep
from torch_distlearn_b200.utils import ConfusionMatrix, Data
from torch_distlearn_b200.utils.config import base_parser


def main(argv=None):
    p = base_parser("CIFAR-10 AllReduceSGD (examples/cifar10.lua)", batch_size=32, learning_rate=0.1)
    p.add_argument("--graph", action="store_true", help="capture the whole step in a CUDA graph (needs equal step counts)")
    p.add_argument("--bucketMB", type=float, default=0.0, help="overlap per-bucket fused launches with backward")
    p.add_argument("--flatOrder", choices=["module", "execution"], default="module",
                   help="layout of the flat parameter vector ('execution': buckets are complete layer by layer)")
    p.add_argument("--conv", choices=["auto", "tc", "cudnn"], default="auto",
                   help="tc = convolutions as implicit GEMMs on the tensor cores (3xTF32, fp32-level accuracy; CUDA only), "
                        "cudnn = stock torch modules; auto = tc with --cuda")
    opt = p.parse_args(argv)
    comm, log = setup(opt)
    dev = comm.device
    batch = math.ceil(opt.batchSize / opt.numNodes)                 # examples/cifar10.lua:36
    data = Data(opt.nodeIndex, opt.numNodes, batch, device=dev, train_samples=opt.trainSamples, test_samples=opt.testSamples,
                native=not opt.pythonLoader, threads=opt.loaderThreads)
    use_tc = dev.type == "cuda" and opt.conv in ("auto", "tc")
    model = (cifar_convnet_tc(seed=opt.seed) if use_tc else cifar_convnet(seed=opt.seed)).to(dev)
    flat = dl.FlatParams(comm, model, with_momentum=opt.momentum != 0, order=opt.flatOrder)
    sgd = dl.AllReduceSGD(comm, flat)
    if opt.resume:
        dl.load_checkpoint(opt.resume, flat, momentum=flat.momentum)
    sgd.synchronizeParameters()                                     # :139
    cm = ConfusionMatrix(data.classes, comm=comm)
    n_train = batches_this_epoch(data.numTrainingBatches(), opt)
    n_test = data.numTestBatches()
    step = None
    if opt.fused and (opt.graph or opt.bucketMB > 0) and dev.type == "cuda":
        step = TrainStep(model, flat, sgd, (batch, 3, 32, 32), lr=opt.learningRate, momentum=opt.momentum,
                         weight_decay=opt.weightDecay, use_graph=opt.graph, variant=opt.variant, bucket_mb=opt.bucketMB)
    for epoch in range(1, opt.numEpochs + 1):
        cm.zero()
        thr, loss_sum = Throughput(), 0.0
        for i in range(1, n_train + 1):
            b = data.getTrainingBatch()
            if step is not None and b.batchSize == batch:
                step.step(b.input, b.target)
                out, loss = step.logits, step.loss
            else:
                out, loss = forward_backward(model, b)              # :181
                if opt.fused:
                    sgd.step(lr=opt.learningRate, momentum=opt.momentum, weight_decay=opt.weightDecay, variant=opt.variant)
                else:
                    sgd.sumAndNormalizeGradients(flat)              # :184
                    with torch.no_grad():                           # :187-191
                        flat.params.tensor.add_(flat.grads.tensor, alpha=-opt.learningRate)
                    flat.zero_grad()
            cm.batchAdd(out, b.target)                              # :194-196
            if i % 20 == 0 or i == n_train:
                loss_sum += float(loss)
            thr.add(b.batchSize)
            log.progress(i, n_train)
        # The drain first, the metrics after (the reference has them the other way round, :203 then :208): a node whose
        # partition had fewer batches must be able to serve its peers' remaining steps from synchronizeParameters; were
        # it parked in the confusion-matrix collective instead, nobody could make progress until the timeout.
        winner = sgd.synchronizeParameters()                        # :208
        cm.allReduce()                                              # :203
        comm.synchronize()
        log.print(f"epoch {epoch} train: accuracy {100 * cm.totalValid:.2f}%  {thr.rate() * comm.world:.0f} samples/s (all nodes)"
                  f"  params from node {winner + 1}")
        train_acc = cm.totalValid
        cm.zero()
        test_loss = evaluate(model, data.getTestBatch, n_test, cm)  # :213-231
        cm.allReduce()                                              # :234
        comm.synchronize()
        log.print(f"epoch {epoch} test : loss {test_loss:.4f}  accuracy {100 * cm.totalValid:.2f}%")
        log.metric(epoch=epoch, train_accuracy=train_acc, test_accuracy=cm.totalValid, test_loss=test_loss,
                   samples_per_s=thr.rate() * comm.world)
        if opt.save and opt.nodeIndex == 1:
            import os
            os.makedirs(opt.save, exist_ok=True)
            dl.save_checkpoint(os.path.join(opt.save, "cifar10.dlb200"), flat, momentum=flat.momentum, algo="AllReduceSGD",
                               world=comm.world, extra={"epoch": epoch})
    final = float(cm.totalValid)
    shutdown(comm)
    return final


if __name__ == "__main__":
    main()
