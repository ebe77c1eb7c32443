"""b200net: a B200-native (sm_100a) implementation of the ResNet-50 data-parallel hot path of
VaibhavKamat/Performance-Comparison-of-TensorFlow-PyTorch-and-their-Distributed-Counterparts.

    import b200net
    model = b200net.models.resnet50(pretrained=False)     # instead of torchvision.models.resnet50
    model.fc = torch.nn.Sequential(...)                   # the reference's head, stock modules are fine
    optimizer = b200net.optim.Adam(model.fc.parameters(), lr=0.003)
    model = b200net.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)

The rest of the reference's train / eval / inference loops runs unchanged.  All arithmetic on the path is done by
hand-written CUDA kernels in libb200net.so (C-ABI: include/b200net.h); importing this package without the built
library raises.
"""
from . import _C  # noqa: F401  (fails loudly when libb200net.so is missing)
from . import ops, functional, nn, models, optim, utils, parallel  # noqa: F401
from .functional import set_precision, get_precision, set_cuda_graphs  # noqa: F401

__version__ = "0.1.0"
