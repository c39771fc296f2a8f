"""Rewire ResNet blocks so that BN -> ReLU and BN -> add -> ReLU run as one kernel.

The reference models spell these patterns as separate modules:
  setup/resnet18.py:23   output = relu(bn1(conv1(x)))
  setup/resnet18.py:24-26 output = bn2(conv2(output)); output += skip(x); relu
  setup/resnet18.py:57   stem relu(bn1(conv1(x)))
and torchvision's BasicBlock / Bottleneck (used by
01_torch_distributor/02_cifar_torch_distributor_resnet.py:146 and
03_tiny_imagenet_torch_distributor_resnet.py:130) do the same.  After
``convert_sync_batchnorm`` has put :class:`B200SyncBatchNorm` layers in place,
:func:`fuse_bn_relu` switches on their fused epilogues and replaces the block
``forward`` with one that hands the skip tensor to the last BN of the block.
Parameters, buffers and state_dict keys are untouched.
"""
from __future__ import annotations

import types

import torch.nn as nn

from . import pool
from .syncbn import B200SyncBatchNorm


def _is_ours(m) -> bool:
    return isinstance(m, B200SyncBatchNorm)


def _skip_of(block):
    # torchvision calls it `downsample` (None when absent); the reference model
    # calls it `skip_connection` (an empty nn.Sequential when absent)
    skip = getattr(block, "downsample", None)
    if skip is None:
        skip = getattr(block, "skip_connection", None)
    if isinstance(skip, nn.Sequential) and len(skip) == 0:
        skip = None
    return skip


def _basic_forward(self, x):
    identity = x
    out = self.bn1(self.conv1(x))                 # BN + ReLU
    skip = _skip_of(self)
    if skip is not None:
        identity = skip(x)
    return self.bn2(self.conv2(out), identity)    # BN + add + ReLU


def _bottleneck_forward(self, x):
    identity = x
    out = self.bn1(self.conv1(x))                 # BN + ReLU
    out = self.bn2(self.conv2(out))               # BN + ReLU
    skip = _skip_of(self)
    if skip is not None:
        identity = skip(x)
    return self.bn3(self.conv3(out), identity)    # BN + add + ReLU


def _looks_like_basic(m) -> bool:
    return (all(hasattr(m, a) for a in ("conv1", "bn1", "conv2", "bn2", "relu")) and not hasattr(m, "conv3")
            and (hasattr(m, "downsample") or hasattr(m, "skip_connection"))
            and _is_ours(m.bn1) and _is_ours(m.bn2))


def _looks_like_bottleneck(m) -> bool:
    return (all(hasattr(m, a) for a in ("conv1", "bn1", "conv2", "bn2", "conv3", "bn3", "relu"))
            and hasattr(m, "downsample") and _is_ours(m.bn1) and _is_ours(m.bn2) and _is_ours(m.bn3))


def _looks_like_resnet_stem(m) -> bool:
    return (all(hasattr(m, a) for a in ("conv1", "bn1", "relu", "maxpool", "layer1", "fc"))
            and _is_ours(m.bn1) and isinstance(m.relu, nn.ReLU))


def fuse_bn_relu(model: nn.Module) -> nn.Module:
    """Enable the fused epilogues on every ResNet-style block found in ``model``.

    Recognised by structure, not by class: the reference ``Block``
    (setup/resnet18.py:3-27), torchvision ``BasicBlock`` and ``Bottleneck``, and
    the ResNet stem.  Anything else is left exactly as it was.  Returns ``model``.
    """
    for m in model.modules():
        if _looks_like_bottleneck(m):
            m.bn1.fuse_relu = m.bn2.fuse_relu = m.bn3.fuse_relu = True
            m.forward = types.MethodType(_bottleneck_forward, m)
        elif _looks_like_basic(m):
            m.bn1.fuse_relu = m.bn2.fuse_relu = True
            m.forward = types.MethodType(_basic_forward, m)
        elif _looks_like_resnet_stem(m):
            # the stem's ReLU module is used nowhere else in these models
            m.bn1.fuse_relu = True
            m.relu = nn.Identity()
            # the pooling layer after it (setup/resnet18.py:37,58): same arithmetic, one kernel each way
            if pool.supported(m.maxpool):
                m.maxpool = pool.B200MaxPool2d.from_module(m.maxpool)
    return model


def unfuse_bn_relu(model: nn.Module) -> nn.Module:
    """Undo :func:`fuse_bn_relu` in place: the blocks get their class ``forward`` back, the BN layers stop
    applying ReLU / taking the skip tensor, the stem gets its ``nn.ReLU`` and ``nn.MaxPool2d`` again.
    ``B200DDP.__getstate__`` pickles such a copy: a module whose ``forward`` is an instance attribute bound
    to a function of this file cannot be unpickled, a plain torchvision / reference block can."""
    for m in model.modules():
        if "forward" in m.__dict__ and getattr(m.__dict__["forward"], "__func__", None) in (_basic_forward, _bottleneck_forward):
            del m.__dict__["forward"]
            for name in ("bn1", "bn2", "bn3"):
                bn = getattr(m, name, None)
                if _is_ours(bn):
                    bn.fuse_relu = False
        elif (all(hasattr(m, a) for a in ("conv1", "bn1", "relu", "maxpool", "layer1", "fc")) and _is_ours(m.bn1)
              and isinstance(m.relu, nn.Identity) and m.bn1.fuse_relu):   # a stem that fuse_bn_relu rewired
            m.bn1.fuse_relu = False
            m.relu = nn.ReLU(inplace=True)
            if isinstance(m.maxpool, pool.B200MaxPool2d):
                p = m.maxpool
                m.maxpool = nn.MaxPool2d(p.kernel_size, p.stride, p.padding, p.dilation, p.return_indices, p.ceil_mode)
    return model
