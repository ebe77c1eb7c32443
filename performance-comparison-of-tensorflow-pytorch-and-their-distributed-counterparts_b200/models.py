"""torchvision-shaped ResNet built on b200net kernels.

`resnet50(pretrained=...)` returns a module whose attribute names (conv1, bn1, relu, maxpool, layer1..4, avgpool,
fc), `repr`, parameter order and state_dict keys are those of torchvision.models.resnet50 -- the repr stored at
pytorch_training_inference_on_image.ipynb:454-635 -- so the reference's statements run unchanged:
freezing (`param.requires_grad = False`, :436-437), head replacement (`model.fc = nn.Sequential(...)`, :439-443),
`model.to(device)`, `model.forward(inputs)` (:669), `model.eval()` / `model(input)` (:853-860), `torch.save(model)`.
"""
from typing import List, Optional, Type

import torch
import torch.nn as _nn

from . import nn
from . import functional as F


def conv3x3(in_planes, out_planes, stride=1):
    return nn.Conv2d(in_planes, out_planes, kernel_size=3, stride=stride, padding=1, bias=False)


def conv1x1(in_planes, out_planes, stride=1):
    return nn.Conv2d(in_planes, out_planes, kernel_size=1, stride=stride, bias=False)


class Bottleneck(_nn.Module):
    """ResNet v1.5 bottleneck: 1x1 -> 3x3 (stride here) -> 1x1 (x4), residual add, ReLU."""
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        width = planes
        self.conv1 = conv1x1(inplanes, width)
        self.bn1 = nn.BatchNorm2d(width)
        self.conv2 = conv3x3(width, width, stride)
        self.bn2 = nn.BatchNorm2d(width)
        self.conv3 = conv1x1(width, planes * self.expansion)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        # stand-alone use (layer-by-layer, NCHW); the full model takes the fused path in ResNet.forward
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)


class ResNet(_nn.Module):
    def __init__(self, block: Type[Bottleneck], layers: List[int], num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, self.inplanes, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(self.inplanes)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        # torchvision's initialisation (kaiming-normal fan_out convs, BN gamma=1 / beta=0)
        for m in self.modules():
            if isinstance(m, _nn.Conv2d):
                _nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, _nn.BatchNorm2d):
                _nn.init.constant_(m.weight, 1)
                _nn.init.constant_(m.bias, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       nn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def features(self, x):
        """[N,3,H,W] -> [N,2048]: the whole backbone as one fused autograd node."""
        return F.resnet_backbone(self, x)

    def _pools_to_1x1(self):
        p = self.avgpool
        if not isinstance(p, _nn.AdaptiveAvgPool2d):
            return False
        size = p.output_size if isinstance(p.output_size, (tuple, list)) else (p.output_size, p.output_size)
        return tuple(size) == (1, 1)

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("b200net.models.ResNet runs on CUDA tensors only (no CPU fallback); "
                               "call model.to('cuda') and inputs.to('cuda') as the reference loop does")
        if self._pools_to_1x1():
            feat = self.features(x)
        else:
            # `model.avgpool = nn.Identity()`: the Keras recipe of the reference, ResNet50(include_top=False) ->
            # Flatten -> Dense(n_classes, softmax) on the whole 5x5x2048 map (resnet.py:17-23), written against the
            # torchvision graph.  The fused node hands out the layer4 map; torchvision's forward then applies
            # whatever the user put in `avgpool`, flattens, and calls `fc`.
            feat = torch.flatten(self.avgpool(F.resnet_backbone(self, x, pool=False)), 1)
        return nn.run_head(self.fc, feat)


def resnet50(pretrained: bool = False, progress: bool = True, weights=None, **kwargs) -> ResNet:
    """Same call as `models.resnet50(pretrained=True)` (pytorch_training_inference_on_image.ipynb:389).
    There is no network on the build / GPU boxes, so `pretrained=True` tries torchvision's local cache and
    otherwise raises; load a checkpoint with `model.load_state_dict(torch.load(path))` (keys are torchvision's)."""
    model = ResNet(Bottleneck, [3, 4, 6, 3], **kwargs)
    if pretrained or weights is not None:
        import os
        path = os.path.join(torch.hub.get_dir(), "checkpoints", "resnet50-0676ba61.pth")
        alt = os.path.join(torch.hub.get_dir(), "checkpoints", "resnet50-19c8e357.pth")
        for p in (path, alt):
            if os.path.exists(p):
                model.load_state_dict(torch.load(p, map_location="cpu"))
                return model
        raise RuntimeError("b200net.models.resnet50(pretrained=True): no cached torchvision checkpoint under %s and "
                           "no network; build with pretrained=False and load_state_dict() a local file" %
                           os.path.join(torch.hub.get_dir(), "checkpoints"))
    return model


from .vgg import VGG, vgg16  # noqa: E402,F401  (models.vgg16, another_neural_net.py:244)
