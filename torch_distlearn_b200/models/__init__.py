"""Model families of the reference's examples: the MNIST convnet (examples/mnist.lua:53-81) and
the CIFAR-10 convnet (examples/Model.lua:20-66)."""
from .cifar import CifarConvNet, cifar_convnet
from .mnist import MnistConvNet, mnist_convnet

__all__ = ["CifarConvNet", "cifar_convnet", "MnistConvNet", "mnist_convnet"]
