"""Batch transform: uint8 NHWC -> random crop / flip / normalize -> bf16 on the GPU.

Slots into the reference train loop at the line
``data, target = data.to(device), target.to(device)``
(01_torch_distributor/01_basic_torch_distributor.py:225,
01_torch_distributor/02_cifar_torch_distributor_resnet.py:232): the dataset
yields raw uint8 images (its ``transform=None`` path,
utils/hf_dataset_utilities.py:50-51) and one kernel launch per batch does what
``Compose([RandomCrop(size, padding), RandomHorizontalFlip(), ToTensor(),
Normalize(mean, std)])`` did per sample on the CPU
(04_accelerate/01_cifar_accelerate.ipynb:295-298, utils/hf_dataset_utilities.py:64-77).

The crop offsets and flip decisions are drawn on the host, one sample after
the other, in torchvision's order (``RandomCrop.get_params``: i then j,
TV/transforms/transforms.py:653-654; then ``torch.rand(1) < p``, :720) so that
the same generator state gives the same indices as the reference pipeline.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib, _mt19937

IMAGENET_MEAN = (0.485, 0.456, 0.406)  # utils/hf_dataset_utilities.py:75-78
IMAGENET_STD = (0.229, 0.224, 0.225)
CIFAR_MEAN = (0.4914, 0.4822, 0.4465)  # 04_accelerate/01_cifar_accelerate.ipynb:298
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def draw_crop_flip_params(n: int, src_hw, out_hw, padding: int = 0, flip_p: float = 0.5,
                          generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """int32 [n, 3] = (i, j, flip) per sample, torchvision's draws in torchvision's order.

    i, j index the zero-padded image, as ``RandomCrop.get_params`` returns them.
    With ``generator=None`` the global CPU generator is consumed, exactly as
    torchvision's transforms consume it.
    """
    h, w = src_hw[0] + 2 * padding, src_hw[1] + 2 * padding
    th, tw = out_hw
    if th > h or tw > w:
        raise ValueError(f"Required crop size {(th, tw)} is larger than input image size {(h, w)}")
    # Same stream as n x (randint, randint, rand) on this generator, produced in bulk by a
    # bit-exact replica of torch's mt19937 (the per-sample loop costs ~10 ms per 256 images).
    draws = _mt19937.crop_flip_draws(n, h - th + 1, w - tw + 1, flip_p, generator)
    return torch.from_numpy(draws)


def _draw_crop_flip_params_loop(n: int, src_hw, out_hw, padding: int = 0, flip_p: float = 0.5,
                                generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """The literal per-sample loop torchvision performs; kept as the check of the replica."""
    h, w = src_hw[0] + 2 * padding, src_hw[1] + 2 * padding
    th, tw = out_hw
    out = torch.empty((n, 3), dtype=torch.int32)
    acc = out.numpy()
    no_draw = (h == th and w == tw)  # get_params returns (0, 0) without touching the RNG
    for k in range(n):
        if no_draw:
            i = j = 0
        else:
            i = int(torch.randint(0, h - th + 1, size=(1,), generator=generator).item())
            j = int(torch.randint(0, w - tw + 1, size=(1,), generator=generator).item())
        flip = bool(torch.rand(1, generator=generator) < flip_p)
        acc[k, 0], acc[k, 1], acc[k, 2] = i, j, int(flip)
    return out


class BatchTransform:
    """``transform(batch_u8_nhwc, generator=...) -> bf16 [B, 3, H, W]`` (channels_last).

    batch_u8_nhwc: uint8 ``[B, Hs, Ws, 3]``, on the GPU or in pinned host memory
    (pageable host memory works too, through a synchronous copy).  The result
    lives on ``device`` and has channels_last strides, ready for a
    channels_last bf16 model.
    """

    def __init__(self, size, padding: int = 0, mean: Sequence[float] = IMAGENET_MEAN,
                 std: Sequence[float] = IMAGENET_STD, flip_p: float = 0.5,
                 device: Optional[torch.device] = None, zero_copy: bool = False):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.padding = int(padding)
        self.flip_p = float(flip_p)
        if len(mean) != 3 or len(std) != 3:
            raise ValueError("mean and std must have 3 entries (RGB)")
        # torchvision builds float32 tensors from the python floats (Normalize)
        self._mean = (ctypes.c_float * 3)(*[float(torch.tensor(m, dtype=torch.float32)) for m in mean])
        self._std = (ctypes.c_float * 3)(*[float(torch.tensor(s, dtype=torch.float32)) for s in std])
        if any(s == 0 for s in self._std):
            raise ValueError("std evaluated to zero after conversion to float32")
        self.device = torch.device(device) if device is not None else None
        self.zero_copy = bool(zero_copy)
        self._staging = None
        self._params_dev = None
        self.last_params = None  # int32 [B,3] on the host, for inspection / parity tests

    def _device(self, batch):
        if self.device is not None:
            return self.device
        if batch.is_cuda:
            return batch.device
        return torch.device("cuda", torch.cuda.current_device())

    def __call__(self, batch: torch.Tensor, *, generator: Optional[torch.Generator] = None,
                 params: Optional[torch.Tensor] = None) -> torch.Tensor:
        if batch.dtype != torch.uint8 or batch.dim() != 4 or batch.shape[-1] != 3:
            raise ValueError("expected a uint8 tensor of shape [B, H, W, 3]")
        if not batch.is_contiguous():
            raise ValueError("expected a contiguous NHWC batch")
        B, Hs, Ws, _ = batch.shape
        H, W = self.size
        if params is None:
            params = draw_crop_flip_params(B, (Hs, Ws), (H, W), self.padding, self.flip_p, generator)
        elif params.dtype != torch.int32 or tuple(params.shape) != (B, 3) or params.is_cuda:
            raise ValueError("params must be a host int32 tensor of shape [B, 3]")
        self.last_params = params
        dev = self._device(batch)
        lib = _lib.lib()  # raises if the CUDA library is not built: there is no CPU path
        with torch.cuda.device(dev):
            if batch.is_cuda:
                src = batch
            elif self.zero_copy and batch.is_pinned():
                src = batch  # the kernel reads the mapped host pages over PCIe
            else:
                if self._staging is None or self._staging.shape != batch.shape or self._staging.device != dev:
                    self._staging = torch.empty(batch.shape, dtype=torch.uint8, device=dev)
                self._staging.copy_(batch, non_blocking=batch.is_pinned())
                src = self._staging
            if self._params_dev is None or self._params_dev.shape[0] != B or self._params_dev.device != dev:
                self._params_dev = torch.empty((B, 3), dtype=torch.int32, device=dev)
            # a fresh pinned block per call: torch's host allocator keeps it alive
            # until the asynchronous copy below has run
            self._params_dev.copy_(params.pin_memory(), non_blocking=True)
            out = torch.empty((B, H, W, 3), dtype=torch.bfloat16, device=dev)
            _lib.check(lib.b200dp_augment_u8_bf16(
                src.data_ptr(), B, Hs, Ws, self._params_dev.data_ptr(), self.padding, H, W,
                self._mean, self._std, out.data_ptr(), _lib.current_stream()), "b200dp_augment_u8_bf16")
        return out.permute(0, 3, 1, 2)


def draw_resized_crop_flip_params(n: int, src_hw, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0),
                                  flip_p: float = 0.5, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """int32 [n, 5] = (i, j, h, w, flip) per sample: ``RandomResizedCrop.get_params``
    (TV/transforms/transforms.py:929-970: up to 10 tries of a uniform area fraction and a
    log-uniform aspect ratio, then two randint, else the central-crop fallback) followed by
    the flip draw (:720), in that order per sample, from one generator stream.  ``src_hw`` is
    ``(height, width)`` of every image, or an int array ``[n, 2]`` with one size per sample."""
    hw = np.asarray(src_hw)
    if hw.ndim == 2:
        if hw.shape != (n, 2):
            raise ValueError("per-sample sizes must have shape [n, 2]")
        draws = _mt19937.resized_crop_flip_draws(n, hw[:, 0], hw[:, 1], tuple(scale), tuple(ratio), flip_p, generator)
    else:
        draws = _mt19937.resized_crop_flip_draws(n, int(hw[0]), int(hw[1]), tuple(scale), tuple(ratio), flip_p, generator)
    return torch.from_numpy(draws)


def _draw_resized_crop_flip_params_loop(n: int, src_hw, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0),
                                        flip_p: float = 0.5, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """The literal per-sample torch calls torchvision makes; kept as the check of the replica."""
    import math

    hw = np.asarray(src_hw)
    log_ratio = torch.log(torch.tensor(ratio))
    out = torch.empty((n, 5), dtype=torch.int32)
    for k in range(n):
        height, width = (int(hw[k, 0]), int(hw[k, 1])) if hw.ndim == 2 else (int(hw[0]), int(hw[1]))
        area = height * width
        for _ in range(10):
            target_area = area * torch.empty(1).uniform_(scale[0], scale[1], generator=generator).item()
            aspect = torch.exp(torch.empty(1).uniform_(log_ratio[0], log_ratio[1], generator=generator)).item()
            w = int(round(math.sqrt(target_area * aspect)))
            h = int(round(math.sqrt(target_area / aspect)))
            if 0 < w <= width and 0 < h <= height:
                i = int(torch.randint(0, height - h + 1, size=(1,), generator=generator).item())
                j = int(torch.randint(0, width - w + 1, size=(1,), generator=generator).item())
                break
        else:
            in_ratio = float(width) / float(height)
            if in_ratio < min(ratio):
                w = width
                h = int(round(w / min(ratio)))
            elif in_ratio > max(ratio):
                h = height
                w = int(round(h * max(ratio)))
            else:
                w, h = width, height
            i, j = (height - h) // 2, (width - w) // 2
        flip = bool(torch.rand(1, generator=generator) < flip_p)
        out[k] = torch.tensor([i, j, h, w, int(flip)], dtype=torch.int32)
    return out


class ResizeBatchTransform:
    """``transform(batch_u8_nhwc, generator=...) -> bf16 [B, 3, H, W]`` (channels_last) with a resize.

    ``random_resized_crop=True``:  ``Compose([RandomResizedCrop(size, scale, ratio),
    RandomHorizontalFlip(flip_p), ToTensor(), Normalize(mean, std)])`` on PIL images
    (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53).
    ``random_resized_crop=False``: ``Compose([Resize(size), RandomHorizontalFlip(flip_p),
    ToTensor(), Normalize(mean, std)])`` (utils/hf_dataset_utilities.py:64-79; gray images are
    stored with their channel repeated, which is what the reference's Lambda does after ToTensor).
    The resize is Pillow's bilinear resampler, reproduced bit for bit on the uint8 level.
    """

    def __init__(self, size, random_resized_crop: bool = True, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0),
                 mean: Sequence[float] = IMAGENET_MEAN, std: Sequence[float] = IMAGENET_STD, flip_p: float = 0.5,
                 device: Optional[torch.device] = None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.random_resized_crop = bool(random_resized_crop)
        self.scale, self.ratio = tuple(scale), tuple(ratio)
        self.flip_p = float(flip_p)
        if len(mean) != 3 or len(std) != 3:
            raise ValueError("mean and std must have 3 entries (RGB)")
        self._mean = (ctypes.c_float * 3)(*[float(torch.tensor(m, dtype=torch.float32)) for m in mean])
        self._std = (ctypes.c_float * 3)(*[float(torch.tensor(s, dtype=torch.float32)) for s in std])
        if any(s == 0 for s in self._std):
            raise ValueError("std evaluated to zero after conversion to float32")
        self.device = torch.device(device) if device is not None else None
        self._staging = None
        self._params_dev = None
        self.last_params = None  # int32 [B,5] on the host

    def draw(self, B: int, src_hw, generator=None) -> torch.Tensor:
        """int32 [B, 5] boxes and flips.  ``src_hw``: ``(height, width)`` of every image, or ``[B, 2]`` sizes of
        ragged images (each stored in the top-left corner of its slab)."""
        if self.random_resized_crop:
            return draw_resized_crop_flip_params(B, src_hw, self.scale, self.ratio, self.flip_p, generator)
        flips = draw_crop_flip_params(B, (1, 1), (1, 1), 0, self.flip_p, generator)[:, 2]  # one rand(1) per sample
        out = torch.zeros((B, 5), dtype=torch.int32)
        hw = torch.as_tensor(np.asarray(src_hw), dtype=torch.int32)
        out[:, 2], out[:, 3], out[:, 4] = (hw[:, 0], hw[:, 1], flips) if hw.dim() == 2 else (int(hw[0]), int(hw[1]), flips)
        return out

    def __call__(self, batch: torch.Tensor, *, generator: Optional[torch.Generator] = None,
                 params: Optional[torch.Tensor] = None, sizes=None) -> torch.Tensor:
        """``sizes``: optional ``[B, 2]`` (height, width) of ragged images, each in the top-left corner of its
        ``[Hs, Ws]`` slab; the boxes are then drawn per image size, as torchvision does per PIL image."""
        if batch.dtype != torch.uint8 or batch.dim() != 4 or batch.shape[-1] != 3:
            raise ValueError("expected a uint8 tensor of shape [B, H, W, 3]")
        if not batch.is_contiguous():
            raise ValueError("expected a contiguous NHWC batch")
        B, Hs, Ws, _ = batch.shape
        H, W = self.size
        if params is None:
            if sizes is not None:
                hw = np.asarray(sizes)
                if hw.shape != (B, 2) or (hw < 1).any() or (hw[:, 0] > Hs).any() or (hw[:, 1] > Ws).any():
                    raise ValueError("sizes must be [B, 2] with 1 <= height <= Hs and 1 <= width <= Ws")
                params = self.draw(B, hw, generator)
            else:
                params = self.draw(B, (Hs, Ws), generator)
        elif params.dtype != torch.int32 or tuple(params.shape) != (B, 5) or params.is_cuda:
            raise ValueError("params must be a host int32 tensor of shape [B, 5]")
        self.last_params = params
        dev = self.device if self.device is not None else (
            batch.device if batch.is_cuda else torch.device("cuda", torch.cuda.current_device()))
        lib = _lib.lib()  # raises if the CUDA library is not built: there is no CPU path
        with torch.cuda.device(dev):
            if batch.is_cuda:
                src = batch
            else:
                if self._staging is None or self._staging.shape != batch.shape or self._staging.device != dev:
                    self._staging = torch.empty(batch.shape, dtype=torch.uint8, device=dev)
                self._staging.copy_(batch, non_blocking=batch.is_pinned())
                src = self._staging
            if self._params_dev is None or self._params_dev.shape[0] != B or self._params_dev.device != dev:
                self._params_dev = torch.empty((B, 5), dtype=torch.int32, device=dev)
            self._params_dev.copy_(params.pin_memory(), non_blocking=True)
            out = torch.empty((B, H, W, 3), dtype=torch.bfloat16, device=dev)
            _lib.check(lib.b200dp_resized_crop_u8_bf16(
                src.data_ptr(), B, Hs, Ws, self._params_dev.data_ptr(), H, W,
                self._mean, self._std, out.data_ptr(), _lib.current_stream()), "b200dp_resized_crop_u8_bf16")
        return out.permute(0, 3, 1, 2)
