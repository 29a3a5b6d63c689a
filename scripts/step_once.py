"""A few eager training steps of the flagship model (tensor-core convs, fused BN tail, fused all-reduce+SGD) for
`ncu --metrics gpu__time_duration.sum`: the launch list of one step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch_distlearn_b200 as dl
from torch_distlearn_b200.models import cifar_convnet_tc
from torch_distlearn_b200.parallel.trainer import TrainStep

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
comm = dl.P2PCommunicator(rank=0, world=1, device=dev, heap_mb=128)
model = cifar_convnet_tc().to(dev)
flat = dl.FlatParams(comm, model)
sgd = dl.AllReduceSGD(comm, flat)
step = TrainStep(model, flat, sgd, (128, 3, 32, 32), lr=0.01, use_graph=False, warmup=1)
x = torch.randn(128, 3, 32, 32, device=dev); y = torch.randint(0, 10, (128,), device=dev)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    step.step(x, y)
    torch.cuda.synchronize()
    print("STEP", i, step.loss_value(), "kernels/step", step.kernels_per_step, flush=True)
