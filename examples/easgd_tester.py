#!/usr/bin/env python
"""Asynchronous EASGD — the evaluator (counterpart of examples/EASGD_tester.lua).

Waits for the server's test request, pulls the center over NVLink into its own model, computes the
error over the whole (unpartitioned) training and test sets (examples/EASGD_tester.lua:109-159), logs
a row to ``<save>/ErrorRate.log`` (:47,:161) and acknowledges (:169).  Unlike the reference, client
syncs are never stalled while the evaluation runs.
"""
import os

from common import easgd_topology, evaluate, setup

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import cifar_convnet
from torch_distlearn_b200.utils import ConfusionMatrix, Data, Logger
from torch_distlearn_b200.utils.config import easgd_parser


def main(argv=None):
    opt = easgd_parser("EASGD tester (examples/EASGD_tester.lua)", "tester").parse_args(argv)
    opt.tester = True
    world, server_rank, tester_rank = easgd_topology(opt)
    comm, log = setup(opt, rank=tester_rank, world=world, loud=True)
    dev = comm.device
    data = Data(1, 1, opt.batchSize, device=dev, train_samples=opt.trainSamples, test_samples=opt.testSamples,
                native=not opt.pythonLoader, threads=opt.loaderThreads)   # full data (:75)
    model = cifar_convnet(seed=opt.seed).to(dev)
    flat = dl.FlatParams(comm, model)          # same allocation sequence on every rank: the heap is symmetric
    algo = dl.AsyncEA(comm, "tester", opt.communicationTime, opt.alpha, server_rank=server_rank,
                      num_clients=opt.numNodes, verbose=opt.verbose)
    algo.initTester(flat)                                            # :92
    save_dir = opt.save or os.path.join("Results", "easgd")
    logger = Logger(os.path.join(save_dir, "ErrorRate.log"))         # :36-47
    cm = ConfusionMatrix(data.classes, device=dev)
    tests = 0
    while algo.startTest(flat, timeout=None):                        # :104
        tests += 1
        cm.zero()
        train_loss = evaluate(model, data.getTrainingBatch, data.numTrainingBatches(), cm, batch_stats=True)
        train_err = 1.0 - cm.totalValid
        cm.zero()
        test_loss = evaluate(model, data.getTestBatch, data.numTestBatches(), cm, batch_stats=True)
        test_err = 1.0 - cm.totalValid
        logger.add({"Training Error": train_err, "Test Error": test_err, "Training Loss": train_loss, "Test Loss": test_loss})
        log.print(f"[tester] test {tests}: train error {100 * train_err:.2f}%  test error {100 * test_err:.2f}%")
        algo.finishTest()                                            # :169
    comm.barrier()
    shutdown(comm)


if __name__ == "__main__":
    main()
