"""Drop-in layer classes: same constructor signatures, `repr`, parameter names and state_dict layout as torch.nn
(they subclass it), but `forward` enqueues b200net kernels.  Inputs/outputs at this level are ordinary NCHW fp32
tensors, as everywhere in the reference's scripts.  Inside `models.ResNet` these modules are only parameter
containers: the backbone runs as one fused autograd node (functional.resnet_backbone).
"""
import torch
import torch.nn as _nn

from . import functional as F

Sequential = _nn.Sequential
Module = _nn.Module
Parameter = _nn.Parameter


class Conv2d(_nn.Conv2d):
    def forward(self, x):
        return F.conv2d_module(self, x)


class BatchNorm2d(_nn.BatchNorm2d):
    def forward(self, x):
        return F.batch_norm_module(self, x)


class ReLU(_nn.ReLU):
    def forward(self, x):
        return F.relu(x)


class MaxPool2d(_nn.MaxPool2d):
    def forward(self, x):
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        s = self.stride if isinstance(self.stride, int) else self.stride[0]
        p = self.padding if isinstance(self.padding, int) else self.padding[0]
        if (k, s, p) != (3, 2, 1) or self.ceil_mode or self.dilation not in (1, (1, 1)):
            raise NotImplementedError("b200net MaxPool2d: kernel 3, stride 2, padding 1 only (the ResNet stem pool)")
        return F.maxpool3x3s2(x)


class AdaptiveAvgPool2d(_nn.AdaptiveAvgPool2d):
    def forward(self, x):
        if self.output_size not in (1, (1, 1)):
            raise NotImplementedError("b200net AdaptiveAvgPool2d: output size (1,1) only")
        return F.global_avgpool(x)


class Linear(_nn.Linear):
    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


class Dropout(_nn.Dropout):
    def forward(self, x):
        return F.dropout(x, self.p, self.training)


class LogSoftmax(_nn.LogSoftmax):
    def forward(self, x):
        return F.log_softmax(x, self.dim if self.dim is not None else -1)


class NLLLoss(_nn.NLLLoss):
    def forward(self, logp, target):
        if self.weight is not None or self.reduction != "mean" or self.ignore_index != -100:
            raise NotImplementedError("b200net NLLLoss: default arguments only")
        return F.nll_loss(logp, target)


def run_head(fc, feat):
    """Runs the classifier head on b200net kernels even when the user assigned STOCK torch.nn modules, as the
    reference does: `model.fc = nn.Sequential(nn.Linear(2048,512), nn.ReLU(), nn.Dropout(0.2), nn.Linear(512,10),
    nn.LogSoftmax(dim=1))` (pytorch_training_inference_on_image.ipynb:439-443).  Module types outside that
    vocabulary are called as they are (user code, outside the hot path)."""
    mods = list(fc) if isinstance(fc, _nn.Sequential) else [fc]
    x = feat
    for m in mods:
        if isinstance(m, _nn.Linear):
            x = F.linear(x, m.weight, m.bias)
        elif isinstance(m, _nn.ReLU):
            x = F.relu(x)
        elif isinstance(m, _nn.Dropout):
            x = F.dropout(x, m.p, m.training)
        elif isinstance(m, _nn.LogSoftmax) and x.dim() == 2 and m.dim in (1, -1):
            x = F.log_softmax(x, 1)
        elif isinstance(m, _nn.Flatten):
            x = x.flatten(m.start_dim, m.end_dim)
        elif isinstance(m, _nn.Sequential):
            x = run_head(m, x)
        else:
            x = m(x)
    return x
