"""B200-native data-parallel CNN trainer.

Same capabilities as dat-rohit/distributed-neural-network (the CIFAR-10 CNN, the
three training entry points, timing logs, failure simulation), re-designed for
Blackwell: one process per GPU, hand-written sm_100a kernels for the whole
training step, and a gradient exchange that runs over NVLink peer memory fused
into the weight-gradient/optimizer kernel.

    from distributed_neural_network_b200 import Network, Trainer, TrainConfig
"""
from .models.network import Network  # noqa: F401
from .ops import flat, functional, native  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):  # lazy: the trainer pulls in torch.distributed
    if name in ("Trainer", "TrainConfig"):
        from . import trainer as _t
        return getattr(_t, name)
    raise AttributeError(name)
