#!/usr/bin/env python
"""Asynchronous EASGD — the center server (counterpart of examples/EASGD_server.lua).

The reference server owns the center variable in host memory, opens 1 + N + 1 sockets
(examples/EASGD_server.lua:67-77) and serves one client at a time.  Here the center is a buffer in
this rank's symmetric heap (GPU memory with ``--cuda``): clients pull it and push their elastic
differences with P2P loads and atomics, so the server does no per-sync work at all.  It keeps the
reference's role: publish the initial center (:108), watch syncs arrive (:118-122), trigger the
tester every ``--testTime`` syncs (:124-126) — and, new here, stop cleanly when all clients are done.
"""
from common import easgd_topology, setup

import torch_distlearn_b200 as dl
from torch_distlearn_b200.launch import shutdown
from torch_distlearn_b200.models import cifar_convnet
from torch_distlearn_b200.utils.config import easgd_parser


def main(argv=None):
    opt = easgd_parser("EASGD center server (examples/EASGD_server.lua)", "server").parse_args(argv)
    world, server_rank, tester_rank = easgd_topology(opt)
    comm, log = setup(opt, rank=server_rank, world=world, loud=True)
    model = cifar_convnet(seed=opt.seed).to(comm.device)
    flat = dl.FlatParams(comm, model)          # same allocation sequence on every rank: the heap is symmetric
    algo = dl.AsyncEA(comm, "server", opt.communicationTime, opt.alpha, server_rank=server_rank,
                      num_clients=opt.numNodes, exclusive=opt.exclusive, verbose=opt.verbose)
    if opt.resume:
        dl.load_checkpoint(opt.resume, flat)
    algo.initServer(flat)                                           # examples/EASGD_server.lua:108
    log.print(f"[server] center published ({flat.numel} parameters), {opt.numNodes} clients"
              + (", tester attached" if tester_rank is not None else ""))
    next_test = opt.testTime
    while algo.syncServer(flat, timeout=None):                      # :118-122
        syncs = algo.syncs
        if tester_rank is not None and syncs >= next_test:          # :124-126
            log.print(f"[server] {syncs} syncs: asking the tester to evaluate the center")
            algo.testNet(block=True)
            next_test = syncs + opt.testTime
    if tester_rank is not None:
        algo.testNet(block=True)                                    # final evaluation of the converged center
    algo.stop()
    log.print(f"[server] all clients finished after {algo.syncs} syncs")
    if opt.save:
        import os
        os.makedirs(opt.save, exist_ok=True)
        dl.save_checkpoint(os.path.join(opt.save, "easgd_center.dlb200"), flat, center=flat.params.tensor, algo="AsyncEA",
                           tau=opt.communicationTime, alpha=opt.alpha, steps=algo.syncs, world=world)
    comm.barrier()
    shutdown(comm)


if __name__ == "__main__":
    main()
