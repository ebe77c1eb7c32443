"""b200net: a B200-native (sm_100a) implementation of the data-parallel hot paths of
VaibhavKamat/Performance-Comparison-of-TensorFlow-PyTorch-and-their-Distributed-Counterparts: the ResNet-50
training / inference step and the BERT-base (IMDB) fine-tuning step.

    import b200net
    model = b200net.models.resnet50(pretrained=False)     # instead of torchvision.models.resnet50
    model.fc = torch.nn.Sequential(...)                   # the reference's head, stock modules are fine
    optimizer = b200net.optim.Adam(model.fc.parameters(), lr=0.003)
    model = b200net.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)

    model = b200net.bert.BertForSequenceClassification.from_pretrained(...)   # instead of transformers'
    optimizer = b200net.optim.AdamW(model.parameters(), lr=2e-5, eps=1e-8)    # instead of transformers.AdamW

The rest of the reference's train / eval / inference loops runs unchanged.  All arithmetic on the path is done by
hand-written CUDA kernels in libb200net.so (C-ABI: include/b200net.h); importing this package without the built
library raises.
"""
from . import _C  # noqa: F401  (fails loudly when libb200net.so is missing)
from . import ops, functional, nn, models, optim, utils, parallel, bert, text  # noqa: F401
from .functional import set_precision, get_precision, set_cuda_graphs  # noqa: F401

__version__ = "0.1.0"
