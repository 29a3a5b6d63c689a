"""Command-line flags mirroring the reference's ``lapp`` blocks (examples/cifar10.lua:1-10,
examples/EASGD_server.lua:1-23, examples/EASGD_client.lua:1-17, examples/EASGD_tester.lua:1-22),
plus torchrun-style environment variables (RANK / WORLD_SIZE / LOCAL_RANK) and the knobs that are
new here (--alpha, --momentum, --variant, --backend)."""
from __future__ import annotations

import argparse
import os


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def base_parser(description: str, batch_size: int = 32, learning_rate: float = 0.1) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--nodeIndex", type=int, default=_env_int("RANK", 0) + 1, help="node index (1-based)")
    p.add_argument("--numNodes", type=int, default=_env_int("WORLD_SIZE", 1), help="num nodes spawned in parallel")
    p.add_argument("--batchSize", type=int, default=batch_size, help="batch size (global for cifar10: split over nodes)")
    p.add_argument("--learningRate", type=float, default=learning_rate, help="learning rate")
    p.add_argument("--cuda", action="store_true", help="use cuda")
    p.add_argument("--gpu", type=int, default=_env_int("LOCAL_RANK", 0) + 1, help="which gpu to use (1-based, only with --cuda)")
    p.add_argument("--host", type=str, default="127.0.0.1", help="rendezvous host")
    p.add_argument("--port", type=int, default=8080, help="rendezvous port")
    p.add_argument("--numEpochs", type=int, default=2, help="epochs to run (the reference hard-codes 100)")
    p.add_argument("--epochSize", type=int, default=0, help="cap on batches per epoch (0 = whole partition)")
    p.add_argument("--trainSamples", type=int, default=4096, help="size of the synthetic training set")
    p.add_argument("--testSamples", type=int, default=1024, help="size of the synthetic test set")
    p.add_argument("--momentum", type=float, default=0.0)
    p.add_argument("--weightDecay", type=float, default=0.0)
    p.add_argument("--variant", choices=["auto", "ll", "oneshot", "twoshot", "nvls", "hybrid", "light"], default="auto",
                   help="collective algorithm (auto = measured variant table)")
    p.add_argument("--backend", choices=["auto", "p2p", "host", "dist"], default="auto",
                   help="p2p = CUDA symmetric heap over NVLink, host = CPU shared memory, "
                        "dist = torch.distributed collectives (several hosts, mixed CPU/GPU nodes)")
    p.add_argument("--heapMB", type=float, default=256.0, help="symmetric heap size per node")
    p.add_argument("--fused", action=argparse.BooleanOptionalAction, default=True,
                   help="use the fused all-reduce+SGD kernel (off = reference-style sumAndNormalizeGradients + local update)")
    p.add_argument("--save", type=str, default="", help="checkpoint / results directory")
    p.add_argument("--resume", type=str, default="", help="checkpoint file to resume from")
    p.add_argument("--metrics", type=str, default="", help="JSONL metrics file (node 1)")
    p.add_argument("--verbose", action="store_true", help="protocol tracing")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--pythonLoader", action="store_true",
                   help="assemble batches in Python instead of the native loader threads (csrc/loader.cpp)")
    p.add_argument("--loaderThreads", type=int, default=2, help="worker threads of the native batch loader")
    return p


def easgd_parser(description: str, role: str) -> argparse.ArgumentParser:
    p = base_parser(description, batch_size=128, learning_rate=0.01)
    p.add_argument("--communicationTime", type=int, default=10, help="tau: local steps between syncs with the center")
    p.add_argument("--alpha", type=float, default=0.2, help="elastic moving rate")
    p.add_argument("--testTime", type=int, default=20, help="server: trigger the tester every this many syncs")
    p.add_argument("--exclusive", action="store_true", help="serialise client syncs (the reference's server mutex)")
    p.add_argument("--tester", action="store_true", help="the job has a tester rank (the last one)")
    p.add_argument("--server", action="store_true", help="(parity flag) this process is the server")
    p.add_argument("--visualize", "--visualise", dest="visualize", action="store_true", help="accepted; plotting is not built")
    p.set_defaults(role=role)
    return p
