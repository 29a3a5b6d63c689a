"""Auxiliary subsystems: checkpoint format, synthetic partitioned data, confusion matrix, loggers,
CLI flags, measurement helpers, coloured protocol tracing."""
from .checkpoint import load_checkpoint, read_checkpoint, save_checkpoint
from .color_print import print_client, print_server
from .data import Batch, Data, SyntheticDataset, cifar_dataset, mnist_dataset
from .logging import Logger, RankLogger
from .metrics import ConfusionMatrix
from .native_loader import NativeBatcher, read_cifar10_bin
