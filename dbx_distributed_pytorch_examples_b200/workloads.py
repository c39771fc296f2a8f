"""The models and synthetic batches BASELINE.json's configs name.

* :func:`cifar_resnet18` — the CIFAR-style ResNet-18 of the reference
  (setup/resnet18.py:3-67: 3x3 stem, stride 1, max-pool, stages
  [64,128,256,512] x 2, 1x1-conv+BN projection on a shape change).  The module
  tree uses the same attribute names, so ``state_dict`` keys, parameter
  registration order and the RNG consumed by a seeded construction are those
  of the reference class (checked by
  tests/test_host_cpu.py::test_workload_matches_reference_model when the
  reference is mounted).
* :func:`resnet50` — torchvision's ResNet-50, the model of
  01_torch_distributor/03_tiny_imagenet_torch_distributor_resnet.py:125-143 and
  02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:121-139, random-initialised
  (there is no network for pretrained weights) with all parameters trainable.
"""
from __future__ import annotations

import torch
import torch.nn as nn


def _conv(cin, cout, k, stride):
    return nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)


class ResidualUnit(nn.Module):
    """conv3x3-BN-ReLU-conv3x3-BN, plus the (projected) input, then ReLU."""

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.conv1, self.bn1 = _conv(cin, cout, 3, stride), nn.BatchNorm2d(cout)
        self.relu = nn.ReLU(inplace=True)
        self.conv2, self.bn2 = _conv(cout, cout, 3, 1), nn.BatchNorm2d(cout)
        project = stride != 1 or cin != cout
        self.skip_connection = nn.Sequential(
            *([nn.Conv2d(cin, cout, kernel_size=1, stride=stride, bias=False), nn.BatchNorm2d(cout)]
              if project else []))

    def forward(self, x):
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        y = y + self.skip_connection(x)
        return self.relu(y)


class CifarResNet18(nn.Module):
    WIDTHS = (64, 128, 256, 512)

    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1, self.bn1 = _conv(3, 64, 3, 1), nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        cin = 64
        for idx, width in enumerate(self.WIDTHS, start=1):
            stride = 1 if idx == 1 else 2
            setattr(self, f"layer{idx}", nn.Sequential(ResidualUnit(cin, width, stride), ResidualUnit(width, width, 1)))
            cin = width
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        for idx in range(1, 5):
            x = getattr(self, f"layer{idx}")(x)
        return self.fc(torch.flatten(self.avgpool(x), 1))


def cifar_resnet18(num_classes: int = 10) -> nn.Module:
    return CifarResNet18(num_classes)


def resnet50(num_classes: int = 1000) -> nn.Module:
    import torchvision

    return torchvision.models.resnet50(weights=None, num_classes=num_classes)


def synthetic_batch(batch: int, hw: int, num_classes: int, seed: int, device="cpu", dtype=torch.float32):
    """Config-1 style synthetic tensors: randn images, uniform labels."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    y = torch.randint(0, num_classes, (batch,), generator=g)
    return x.to(device=device, dtype=dtype), y.to(device)


def synthetic_u8_images(batch: int, hw: int, seed: int, pin: bool = True) -> torch.Tensor:
    """uint8 NHWC source images for the GPU transform (config 4: 256x256 -> crop 224)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (batch, hw, hw, 3), dtype=torch.uint8, generator=g)
    return x.pin_memory() if pin and torch.cuda.is_available() else x
