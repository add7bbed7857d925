"""Backbones of the reference, on stock PyTorch / torchvision (not the product; SURVEY.md
section 2 rows 6-7).  Only what the head needs is kept: the trunk up to layer4.

* ``resnet18_2d(modal)``: torchvision ResNet-18 with layer4 stride 1, so a 224x224 frame
  gives a 14x14 map (reference ``models/base_models.py:149``); the audio variant reads a
  1-channel spectrogram through ``conv1_a`` (``base_models.py:135,197-198``).  Parameter
  names follow the reference so its checkpoints load (``imgnet.*`` / ``audnet.*``).
* ``resnet18_3d()``: 3-D ResNet-18 without temporal stride or max-pool, layer2-4 stride
  (1,2,2) (reference ``models/resnet3D.py:134-148``, ``model.py:20``): [b,3,T,224,224] ->
  [b,512,T,14,14].
"""
from __future__ import annotations

import torch
from torch import nn
from torchvision.models.resnet import BasicBlock, ResNet


class ResNet18Trunk2d(ResNet):
    def __init__(self, modal: str = "vision"):
        super().__init__(BasicBlock, [2, 2, 2, 2])
        assert modal in ("vision", "audio", "flow")
        self.modal = modal
        self.conv1_a = nn.Conv2d(1, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.conv1_flow = nn.Conv2d(6, 64, kernel_size=7, stride=2, padding=3, bias=False)
        # layer4 keeps the 14x14 resolution
        blk = self.layer4[0]
        blk.conv1.stride = (1, 1)
        blk.downsample[0].stride = (1, 1)

    def forward(self, x):
        stem = {"vision": self.conv1, "audio": self.conv1_a, "flow": self.conv1_flow}[self.modal]
        x = self.maxpool(self.relu(self.bn1(stem(x))))
        return self.layer4(self.layer3(self.layer2(self.layer1(x))))


def resnet18_2d(modal: str = "vision") -> ResNet18Trunk2d:
    return ResNet18Trunk2d(modal)


class _Block3d(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv3d(cin, cout, 3, stride=stride, padding=1, bias=False)
        self.bn1 = nn.BatchNorm3d(cout)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv3d(cout, cout, 3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm3d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv3d(cin, cout, 1, stride=stride, bias=False),
                                            nn.BatchNorm3d(cout))

    def forward(self, x):
        r = x if self.downsample is None else self.downsample(x)
        y = self.bn2(self.conv2(self.relu(self.bn1(self.conv1(x)))))
        return self.relu(y + r)


class ResNet18Trunk3d(nn.Module):
    def __init__(self, n_classes: int = 1039):
        super().__init__()
        self.conv1 = nn.Conv3d(3, 64, kernel_size=7, stride=(1, 2, 2), padding=3, bias=False)
        self.bn1 = nn.BatchNorm3d(64)
        self.relu = nn.ReLU(inplace=True)
        widths, cin, layers = [64, 128, 256, 512], 64, []
        for i, w in enumerate(widths):
            s = 1 if i == 0 else (1, 2, 2)
            layers.append(nn.Sequential(_Block3d(cin, w, s), _Block3d(w, w, 1)))
            cin = w
        self.layer1, self.layer2, self.layer3, self.layer4 = layers
        self.fc = nn.Linear(512, n_classes)          # kept for checkpoint compatibility only
        for m in self.modules():
            if isinstance(m, nn.Conv3d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = self.relu(self.bn1(self.conv1(x)))
        return self.layer4(self.layer3(self.layer2(self.layer1(x))))


def resnet18_3d(n_classes: int = 1039) -> ResNet18Trunk3d:
    return ResNet18Trunk3d(n_classes)
