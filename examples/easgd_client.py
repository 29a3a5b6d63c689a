#!/usr/bin/env python
"""Asynchronous EASGD — a worker (counterpart of examples/EASGD_client.lua).

Loop order is the reference's (examples/EASGD_client.lua:106-117): gradient at x, then ``syncClient``
(every ``--communicationTime`` steps: ONE kernel that loads the center from the server GPU, applies
d = alpha (x - c), x -= d locally and adds d into the center with NVLink atomics), then the SGD step
with the gradient computed before the sync.

On a GPU the worker runs the flagship path (``AsyncTrainStep``): the network with its convolutions on the tensor cores,
the whole step as a CUDA-graph replay, and on sync steps the exchange kernel on a side stream UNDER forward/backward —
the elastic difference depends on x and the center only, so it can be formed while the gradient is being computed and
the update ``(x - d) - lr g`` applied afterwards: the reference's arithmetic and order of effects, with the exchange
hidden.  ``--eager`` keeps the plain loop (stock cuDNN model, one launch after another); CPU nodes always use it.
"""
from common import batches_this_epoch, easgd_topology, forward_backward, setup, Throughput

import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import cifar_convnet, cifar_convnet_tc
from torch_distlearn_b200.ops import local_sgd
from torch_distlearn_b200.utils import Data
from torch_distlearn_b200.utils.config import easgd_parser


def main(argv=None):
    parser = easgd_parser("EASGD client (examples/EASGD_client.lua)", "client")
    parser.add_argument("--eager", action="store_true", help="plain loop with the stock cuDNN model instead of AsyncTrainStep")
    opt = parser.parse_args(argv)
    world, server_rank, _ = easgd_topology(opt)
    comm, log = setup(opt, rank=opt.nodeIndex, world=world)          # client i is rank i (the server is rank 0)
    dev = comm.device
    data = Data(opt.nodeIndex, opt.numNodes, opt.batchSize, device=dev, train_samples=opt.trainSamples,
                test_samples=opt.testSamples,
                native=not opt.pythonLoader, threads=opt.loaderThreads)                        # partitioned over the clients (:75)
    flagship = dev.type == "cuda" and not opt.eager
    model = (cifar_convnet_tc(seed=opt.seed) if flagship else cifar_convnet(seed=opt.seed)).to(dev)
    flat = dl.FlatParams(comm, model)
    algo = dl.AsyncEA(comm, "client", opt.communicationTime, opt.alpha, server_rank=server_rank,
                      num_clients=opt.numNodes, exclusive=opt.exclusive, verbose=opt.verbose)
    algo.initClient(flat, timeout=600.0)                             # :95 — start from the server's center
    n_train = batches_this_epoch(data.numTrainingBatches(), opt)
    step = None
    if flagship:
        from torch_distlearn_b200.parallel.trainer import AsyncTrainStep
        step = AsyncTrainStep(model, flat, algo, (opt.batchSize, 3, 32, 32), lr=opt.learningRate)
    for epoch in range(1, opt.numEpochs + 1):
        thr, loss_sum, synced = Throughput(), 0.0, 0
        for i in range(1, n_train + 1):
            b = data.getTrainingBatch()
            if step is not None and b.batchSize == opt.batchSize:
                loss = step.step(b.input, b.target)                  # :106-117 as one graph replay (sync steps: two streams)
                synced += algo.step % algo.tau == 0
                thr.add(b.batchSize)
                if i % 20 == 0 or i == n_train:
                    loss_sum = step.loss_value()
                continue
            _, loss = forward_backward(model, b)                     # :106
            synced += bool(algo.syncClient(flat))                    # :109
            if dev.type == "cuda":                                   # :113-117
                local_sgd(flat.params.tensor, flat.grads.tensor, None, lr=opt.learningRate, zero_grads=True)
            else:
                with torch.no_grad():
                    flat.params.tensor.add_(flat.grads.tensor, alpha=-opt.learningRate)
                flat.zero_grad()
            thr.add(b.batchSize)
            if i % 20 == 0 or i == n_train:
                loss_sum = float(loss)
        comm.synchronize()
        print(f"[client {opt.nodeIndex}] epoch {epoch}: loss {loss_sum:.4f}  {synced} syncs  {thr.rate():.0f} samples/s", flush=True)
    algo.finish()
    comm.barrier()
    shutdown(comm)


if __name__ == "__main__":
    main()
