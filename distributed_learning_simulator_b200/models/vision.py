"""Vision models: LeNet5, densenet40 (depth 40, growth 12), CIFAR ResNet-18, Resnet50.

Layer inventories follow SURVEY.md appendix A: LeNet5 has 61 706 parameters in 10
tensors; densenet40 decomposes into 38 (BatchNorm2d, ReLU, Conv2d) runs, which is the
first block type FedOBD looks for
(/root/reference/simulation_lib/method/fed_obd/obd_algorithm.py:18).
"""
from typing import Any

import torch
from torch import nn

from ..ops.bn import bn_act, conv_bn_act
from .registry import register_model


def _in_channels(dc: Any) -> int:
    return int(dc.spec["shape"][0])


def _classes(dc: Any) -> int:
    return int(dc.spec["classes"])


class LeNet5(nn.Module):
    def __init__(self, in_channels: int = 1, num_classes: int = 10, input_size: int = 28) -> None:
        super().__init__()
        pad = 2 if input_size == 28 else 0
        self.conv1 = nn.Conv2d(in_channels, 6, kernel_size=5, padding=pad)
        self.act1 = nn.ReLU()
        self.pool1 = nn.MaxPool2d(2)
        self.conv2 = nn.Conv2d(6, 16, kernel_size=5)
        self.act2 = nn.ReLU()
        self.pool2 = nn.MaxPool2d(2)
        self.flatten = nn.Flatten()
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.act3 = nn.ReLU()
        self.fc2 = nn.Linear(120, 84)
        self.act4 = nn.ReLU()
        self.fc3 = nn.Linear(84, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.pool1(self.act1(self.conv1(x)))
        x = self.pool2(self.act2(self.conv2(x)))
        x = self.flatten(x)
        x = self.act3(self.fc1(x))
        x = self.act4(self.fc2(x))
        return self.fc3(x)


class _DenseLayer(nn.Module):
    def __init__(self, in_ch: int, growth: int) -> None:
        super().__init__()
        self.norm = nn.BatchNorm2d(in_ch)
        self.relu = nn.ReLU(inplace=True)
        self.conv = nn.Conv2d(in_ch, growth, kernel_size=3, padding=1, bias=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return torch.cat([x, self.conv(bn_act(self.norm, x, relu=True))], dim=1)


class _Transition(nn.Module):
    def __init__(self, in_ch: int, out_ch: int) -> None:
        super().__init__()
        self.norm = nn.BatchNorm2d(in_ch)
        self.relu = nn.ReLU(inplace=True)
        self.conv = nn.Conv2d(in_ch, out_ch, kernel_size=1, bias=False)
        self.pool = nn.AvgPool2d(2)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.pool(self.conv(bn_act(self.norm, x, relu=True)))


class DenseNet40(nn.Module):
    """DenseNet-BC-free basic variant: 3 dense stages x 12 layers, growth 12."""

    def __init__(self, in_channels: int = 3, num_classes: int = 10, growth: int = 12, layers: int = 12) -> None:
        super().__init__()
        ch = 2 * growth
        self.stem = nn.Conv2d(in_channels, ch, kernel_size=3, padding=1, bias=False)
        stages = []
        for stage in range(3):
            for _ in range(layers):
                stages.append(_DenseLayer(ch, growth))
                ch += growth
            if stage != 2:
                stages.append(_Transition(ch, ch))
        self.features = nn.Sequential(*stages)
        self.final_norm = nn.BatchNorm2d(ch)
        self.final_relu = nn.ReLU(inplace=True)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.flatten = nn.Flatten()
        self.classifier = nn.Linear(ch, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(self.stem(x))
        x = bn_act(self.final_norm, x, relu=True)
        return self.classifier(self.flatten(self.pool(x)))


class _BasicBlock(nn.Module):
    def __init__(self, in_ch: int, out_ch: int, stride: int) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(in_ch, out_ch, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(out_ch)
        self.relu1 = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(out_ch, out_ch, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(out_ch)
        self.downsample = None
        if stride != 1 or in_ch != out_ch:
            self.downsample = nn.Sequential(nn.Conv2d(in_ch, out_ch, 1, stride, bias=False), nn.BatchNorm2d(out_ch))
        self.relu2 = nn.ReLU(inplace=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        identity = x if self.downsample is None else conv_bn_act(self.downsample[0], self.downsample[1], x)
        out = conv_bn_act(self.conv1, self.bn1, x, relu=True)
        return conv_bn_act(self.conv2, self.bn2, out, relu=True, residual=identity)


class ResNet18(nn.Module):
    """ResNet-18 with the CIFAR stem (3x3 conv, no max-pool): 11 173 962 parameters for 10 classes."""

    def __init__(self, in_channels: int = 3, num_classes: int = 10) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, 64, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        cfg = [(64, 1), (64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1)]
        blocks, ch = [], 64
        for out_ch, stride in cfg:
            blocks.append(_BasicBlock(ch, out_ch, stride))
            ch = out_ch
        self.layers = nn.Sequential(*blocks)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.flatten = nn.Flatten()
        self.fc = nn.Linear(512, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = conv_bn_act(self.conv1, self.bn1, x, relu=True)
        x = self.layers(x)
        return self.fc(self.flatten(self.pool(x)))


def _resnet50(dataset_collection: Any, **kwargs: Any) -> nn.Module:
    import torchvision

    model = torchvision.models.resnet50(weights=None, num_classes=_classes(dataset_collection))
    if _in_channels(dataset_collection) != 3:
        model.conv1 = nn.Conv2d(_in_channels(dataset_collection), 64, 7, 2, 3, bias=False)
    return model


register_model(
    "LeNet5",
    lambda dataset_collection, **kw: LeNet5(
        _in_channels(dataset_collection), _classes(dataset_collection), int(dataset_collection.spec["shape"][1])
    ),
)
register_model(
    "densenet40",
    lambda dataset_collection, **kw: DenseNet40(_in_channels(dataset_collection), _classes(dataset_collection)),
)
register_model(
    "resnet18",
    lambda dataset_collection, **kw: ResNet18(_in_channels(dataset_collection), _classes(dataset_collection)),
)
register_model("Resnet50", _resnet50)
