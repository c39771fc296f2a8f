"""b200dp — the B200-native data-parallel hot path.

Three drop-ins for a torch DDP ResNet train loop
(01_torch_distributor/01_basic_torch_distributor.py:222-237,291):

* :class:`B200DDP` — DistributedDataParallel-compatible wrapper whose
  gradient-bucket all-reduce is fused with the 1/N average and the optimizer
  step and runs over NVLink peer memory;
* :class:`B200SyncBatchNorm` / :func:`convert_sync_batchnorm` /
  :func:`fuse_bn_relu` — SyncBatchNorm with fused (+residual)+ReLU, NHWC;
* :class:`BatchTransform` — uint8 NHWC -> crop/flip/normalize -> bf16 in one kernel.
* :class:`ResizeBatchTransform` — the same with Pillow's bilinear resize (Resize / RandomResizedCrop).

All arithmetic runs in libb200dp.so (hand-written sm_100a CUDA behind the C ABI
of include/b200dp.h).  There is no CPU or eager fallback: without the built
library or without a GPU the entry points raise.
"""
from ._lib import B200DPError, build  # noqa: F401
from .comm import Communicator, all_gather_scalars  # noqa: F401
from .ddp import B200DDP, prepare_model  # noqa: F401
from .distributor import DeepspeedTorchDistributor, TorchDistributor  # noqa: F401
from .feed import DeviceFeeder, LossReader, PinnedImageDataset, ShardedBatchSampler  # noqa: F401
from .fusion import fuse_bn_relu, unfuse_bn_relu  # noqa: F401
from .syncbn import B200SyncBatchNorm, convert_sync_batchnorm  # noqa: F401
from .pool import B200MaxPool2d  # noqa: F401
from .transform import (BatchTransform, ResizeBatchTransform, draw_crop_flip_params,  # noqa: F401
                        draw_resized_crop_flip_params)

__all__ = [
    "B200DDP", "prepare_model", "B200SyncBatchNorm", "BatchTransform", "Communicator", "TorchDistributor",
    "DeepspeedTorchDistributor", "convert_sync_batchnorm", "fuse_bn_relu", "draw_crop_flip_params", "all_gather_scalars", "B200MaxPool2d", "ResizeBatchTransform", "draw_resized_crop_flip_params",
    "B200DPError", "build", "DeviceFeeder", "LossReader", "PinnedImageDataset", "ShardedBatchSampler",
]
