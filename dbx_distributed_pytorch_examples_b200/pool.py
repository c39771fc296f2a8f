"""NHWC max pooling for the ResNet stem (``relu(bn1(conv1(x)))`` -> ``maxpool``).

Drop-in for ``nn.MaxPool2d(kernel_size=3, stride=2, padding=1)`` as the reference model uses it
(setup/resnet18.py:37,58) and as torchvision's resnet18/50 do.  Same arithmetic as torch's
``max_pool2d`` (first maximum of the window wins, gradient routed to it); one kernel forward,
one backward (csrc/pool.cu) instead of torch's NHWC kernels, which cost 2.5 ms of a 27 ms
ResNet-50 224^2 step on B200.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib


def supported(m: nn.Module) -> bool:
    """Can :class:`B200MaxPool2d` take over this ``nn.MaxPool2d``?"""
    if not isinstance(m, nn.MaxPool2d) or isinstance(m, B200MaxPool2d):
        return False

    def one(v):
        if isinstance(v, (tuple, list)):
            return v[0] if len(set(v)) == 1 else None
        return v

    k, s, p, d = one(m.kernel_size), one(m.stride if m.stride is not None else m.kernel_size), one(m.padding), one(m.dilation)
    return (k is not None and s is not None and p is not None and d == 1 and 1 <= k <= 3 and s >= 1
            and 0 <= 2 * p <= k and not m.ceil_mode and not m.return_indices)


class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        if x.dim() != 4:
            raise ValueError(f"expected a 4D input (got {x.dim()}D)")
        if not x.is_cuda:
            raise ValueError("B200MaxPool2d expected a CUDA tensor; there is no CPU path here")
        lib = _lib.lib()
        x = x if x.is_contiguous(memory_format=torch.channels_last) else x.contiguous(memory_format=torch.channels_last)
        N, C, H, W = x.shape
        Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
        dt = _lib.dtype_code(x.dtype)
        y = torch.empty((N, C, Ho, Wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((N, Ho, Wo, C), dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.b200dp_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), dt, N, H, W, C, k, stride, pad,
                                              _lib.current_stream()), "b200dp_maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.geom = (dt, N, H, W, C, k, stride, pad, x.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        dt, N, H, W, C, k, stride, pad, dtype = ctx.geom
        lib = _lib.lib()
        if dy.dtype != dtype:
            dy = dy.to(dtype)
        dy = dy if dy.is_contiguous(memory_format=torch.channels_last) else dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, C, H, W), dtype=dtype, device=dy.device, memory_format=torch.channels_last)
        with torch.cuda.device(dy.device):
            _lib.check(lib.b200dp_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), dt, N, H, W, C, k, stride, pad,
                                              _lib.current_stream()), "b200dp_maxpool_bwd")
        return dx, None, None, None


class B200MaxPool2d(nn.MaxPool2d):
    """``nn.MaxPool2d`` for square windows up to 3x3 (dilation 1, ceil_mode off) on channels_last CUDA tensors."""

    def __init__(self, kernel_size, stride=None, padding=0, dilation=1, return_indices=False, ceil_mode=False):
        super().__init__(kernel_size, stride, padding, dilation, return_indices, ceil_mode)
        probe = nn.MaxPool2d(kernel_size, stride, padding, dilation, return_indices, ceil_mode)
        if not supported(probe):
            raise NotImplementedError("B200MaxPool2d handles square windows <= 3, dilation 1, ceil_mode=False, "
                                      "return_indices=False, padding <= kernel_size / 2")

        def one(v):
            return v[0] if isinstance(v, (tuple, list)) else v

        self._k = int(one(self.kernel_size))
        self._s = int(one(self.stride if self.stride is not None else self.kernel_size))
        self._p = int(one(self.padding))

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        C = input.shape[1] if input.dim() == 4 else 0
        if not input.is_cuda:
            raise ValueError("B200MaxPool2d expected a CUDA tensor; there is no CPU path here")
        if (input.dim() != 4 or input.dtype not in (torch.float32, torch.bfloat16)
                or C % (8 if input.dtype == torch.bfloat16 else 4)):
            # no second implementation behind this module: say so instead of quietly running another kernel
            raise ValueError("B200MaxPool2d takes 4-D tensors of float32 (C % 4 == 0) or bfloat16 (C % 8 == 0); "
                             f"got {tuple(input.shape)} {input.dtype} on {input.device}. Keep nn.MaxPool2d for this layer "
                             "(fuse_bn_relu only swaps the ResNet stem's pooling layer).")
        return _MaxPool.apply(input, self._k, self._s, self._p)

    @classmethod
    def from_module(cls, m: nn.MaxPool2d) -> "B200MaxPool2d":
        return cls(m.kernel_size, m.stride, m.padding, m.dilation, m.return_indices, m.ceil_mode)
