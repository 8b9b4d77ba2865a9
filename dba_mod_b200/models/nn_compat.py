"""``torch.nn`` definitions of the four model families with the reference's ``state_dict`` layout.

Used for (a) default initialisation (same init scheme and, seeded alike, the same values as the reference),
(b) checkpoint interchange and (c) the single-client fp32 oracle in the tests.  The hot paths never run these
modules: training / evaluation go through the batched functional networks in ``batched.py`` (torch backend) or
the native sm_100a executor.

Key layouts: CIFAR = models/resnet_cifar.py:67-104 (narrow 32/64/128/256, 3x3 stem, ``shortcut.{0,1}``,
``linear``); Tiny = models/resnet_tinyimagenet.py:122-238 (torchvision layout, ``downsample.{0,1}``, ``fc`` 512->200,
Kaiming fan-out conv init); MNIST = models/MnistNet.py:7-31; LOAN = models/loan_model.py:10-27.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import config as C


class _ResBlock(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int, skip_name: str) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.skip_name = skip_name
        proj = stride != 1 or cin != cout
        if skip_name == "shortcut":           # CIFAR flavour always owns a (possibly empty) Sequential
            self.shortcut = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout)) \
                if proj else nn.Sequential()
        elif proj:                              # torchvision flavour only registers it when needed
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        skip = getattr(self, self.skip_name, None)
        y = y + (skip(x) if skip is not None and len(skip) > 0 else x)
        return F.relu(y)


class ResNet18Compat(nn.Module):
    """flavour='cifar' | 'tiny'."""

    def __init__(self, flavour: str) -> None:
        super().__init__()
        self.flavour = flavour
        if flavour == "cifar":
            widths, skip, ncls = (32, 64, 128, 256), "shortcut", 10
            self.conv1 = nn.Conv2d(3, 32, 3, 1, 1, bias=False)
        else:
            widths, skip, ncls = (64, 128, 256, 512), "downsample", 200
            self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(widths[0])
        cin = widths[0]
        for li, w in enumerate(widths):
            stride = 1 if li == 0 else 2
            setattr(self, f"layer{li + 1}", nn.Sequential(_ResBlock(cin, w, stride, skip), _ResBlock(w, w, 1, skip)))
            cin = w
        head = nn.Linear(widths[-1], 1000 if flavour == "tiny" else ncls)
        if flavour == "cifar":
            self.linear = head
        else:
            self.fc = head
            for m in self.modules():               # resnet_tinyimagenet.py:160-165
                if isinstance(m, nn.Conv2d):
                    nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                elif isinstance(m, nn.BatchNorm2d):
                    nn.init.constant_(m.weight, 1)
                    nn.init.constant_(m.bias, 0)
            self.fc = nn.Linear(widths[-1], ncls)  # resnet_tinyimagenet.py:176-178 re-creates the 200-way head

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        if self.flavour == "tiny":
            x = F.max_pool2d(x, 3, 2, 1)
        for li in range(4):
            x = getattr(self, f"layer{li + 1}")(x)
        if self.flavour == "cifar":
            x = F.avg_pool2d(x, 4).flatten(1)
            return self.linear(x)
        return self.fc(F.adaptive_avg_pool2d(x, 1).flatten(1))


class MnistNetCompat(nn.Module):
    def __init__(self) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(1, 20, 5, 1)
        self.conv2 = nn.Conv2d(20, 50, 5, 1)
        self.fc1 = nn.Linear(800, 500)
        self.fc2 = nn.Linear(500, 10)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2, 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2, 2)
        x = F.relu(self.fc1(x.flatten(1)))
        return F.log_softmax(self.fc2(x), dim=1)  # Q20: CE is applied on top of log_softmax


class LoanNetCompat(nn.Module):
    def __init__(self, in_dim=91, h1=46, h2=23, out_dim=9) -> None:
        super().__init__()
        self.layer1 = nn.Sequential(nn.Linear(in_dim, h1), nn.Dropout(0.5), nn.ReLU())  # Q19: dropout before ReLU
        self.layer2 = nn.Sequential(nn.Linear(h1, h2), nn.Dropout(0.5), nn.ReLU())
        self.layer3 = nn.Sequential(nn.Linear(h2, out_dim))

    def forward(self, x):
        return self.layer3(self.layer2(self.layer1(x)))


def build_module(kind: str) -> nn.Module:
    if kind == C.TYPE_CIFAR:
        return ResNet18Compat("cifar")
    if kind == C.TYPE_TINYIMAGENET:
        return ResNet18Compat("tiny")
    if kind == C.TYPE_MNIST:
        return MnistNetCompat()
    if kind == C.TYPE_LOAN:
        return LoanNetCompat()
    raise ValueError(f"no model for type {kind!r}")


def display_name(kind: str, role: str) -> str:
    """``model.name`` strings used in the reference's log lines (resnet_cifar.py:106-107, MnistNet.py:9)."""
    if kind in (C.TYPE_CIFAR, C.TYPE_TINYIMAGENET):
        return f"{role}_ResNet_18"
    return f"{role}_Simple"


# ---------------------------------------------------------------------------------------------------------------------
# Models the reference defines but never instantiates from main.py (ResNet34/50/101/152 + Bottleneck in
# models/resnet_cifar.py:39-64,106-116, SimpleMnist in models/simple.py:223-240).  Provided as plain modules with the
# same state_dict layout so user code written against them keeps working; they can be trained through
# ``ParamSpec`` + the torch backend's generic autograd path but have no native executor.
class _Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin: int, planes: int, stride: int) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.shortcut = nn.Sequential()
        if stride != 1 or cin != planes * 4:
            self.shortcut = nn.Sequential(nn.Conv2d(cin, planes * 4, 1, stride, bias=False), nn.BatchNorm2d(planes * 4))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = F.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return F.relu(y + self.shortcut(x))


class _BasicCifar(_ResBlock):
    expansion = 1

    def __init__(self, cin: int, planes: int, stride: int) -> None:
        super().__init__(cin, planes, stride, "shortcut")


class CifarResNetCompat(nn.Module):
    """Narrow (32/64/128/256) CIFAR ResNet of arbitrary depth."""

    def __init__(self, block, num_blocks, num_classes: int = 10) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(3, 32, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(32)
        cin = 32
        for li, (planes, n) in enumerate(zip((32, 64, 128, 256), num_blocks), start=1):
            layers = []
            for bi in range(n):
                layers.append(block(cin, planes, (1 if li == 1 else 2) if bi == 0 else 1))
                cin = planes * block.expansion
            setattr(self, f"layer{li}", nn.Sequential(*layers))
        self.linear = nn.Linear(256 * block.expansion, num_classes)

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        for li in range(1, 5):
            x = getattr(self, f"layer{li}")(x)
        return self.linear(F.avg_pool2d(x, 4).flatten(1))


def ResNet34():
    return CifarResNetCompat(_BasicCifar, [3, 4, 6, 3])


def ResNet50():
    return CifarResNetCompat(_Bottleneck, [3, 4, 6, 3])


def ResNet101():
    return CifarResNetCompat(_Bottleneck, [3, 4, 23, 3])


def ResNet152():
    return CifarResNetCompat(_Bottleneck, [3, 8, 36, 3])


class SimpleMnistCompat(nn.Module):
    def __init__(self) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        x = F.relu(F.max_pool2d(self.conv2_drop(self.conv2(x)), 2))
        x = F.relu(self.fc1(x.flatten(1)))
        x = F.dropout(x, training=self.training)
        return F.log_softmax(self.fc2(x), dim=1)
