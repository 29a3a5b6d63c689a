"""Model families of the reference's examples: the MNIST convnet (examples/mnist.lua:53-81) and
the CIFAR-10 convnet (examples/Model.lua:20-66)."""
from .cifar import CifarConvNet, CifarConvNetTC, cifar_convnet, cifar_convnet_tc
from .mnist import MnistConvNet, mnist_convnet

__all__ = ["CifarConvNet", "CifarConvNetTC", "cifar_convnet", "cifar_convnet_tc", "MnistConvNet", "mnist_convnet"]
