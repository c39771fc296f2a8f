"""SyncBatchNorm replacement with fused (+residual)+ReLU epilogues, NHWC.

Drop-in for ``torch.nn.SyncBatchNorm`` (TORCH/nn/modules/batchnorm.py:615-905):
same constructor, ``convert_sync_batchnorm`` classmethod, state_dict keys
(``weight, bias, running_mean, running_var, num_batches_tracked``) and error
behaviour.  It replaces the ``nn.BatchNorm2d`` layers of the reference models
(setup/resnet18.py:8,12,19,35; torchvision resnet18/50 used at
01_torch_distributor/02_cifar_torch_distributor_resnet.py:146 and
03_tiny_imagenet_torch_distributor_resnet.py:130) and, with ``fuse_relu`` /
``forward(x, residual)``, also the ReLU and skip-add that follow them
(setup/resnet18.py:23-26,57).

Forward = two kernels (Welford statistics + exchange; merge + normalise),
backward = two kernels (reduce + exchange; sum + apply) — see csrc/bn.cu.  The
oracle of the arithmetic is TORCH/nn/modules/_functions.py:7-209.
"""
from __future__ import annotations

from typing import Optional

import os

import torch
import torch.distributed as dist
from torch.nn.modules.batchnorm import _BatchNorm

from . import _lib
from .comm import stats_communicator

_scratch = {}
# Residual layers (bf16): save a 1-bit-per-element ReLU mask (stored as one word per 16-byte vector) for the
# backward instead of re-reading y there.  It removes 6 % of the SyncBN bytes of a ResNet-50 step, but A/B runs
# of the whole step on B200 showed no difference (27.09 vs 27.13 ms, power-capped), and it costs 25 % of y in
# extra memory, so it is opt-in.
_USE_MASK_WORDS = os.environ.get("B200DP_BN_MASKW", "0") == "1"


def _bn_scratch(device_index: int, stream: int) -> int:
    """Address of the zero-filled reduction scratch for (device, stream): calls that share it are
    ordered by that stream."""
    key = (device_index, stream)
    buf = _scratch.get(key)
    if buf is None:
        nbytes = int(_lib.lib().b200dp_bn_scratch_bytes(0))
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=torch.device("cuda", device_index))
        _scratch[key] = buf
    return buf.data_ptr()


def _as_rows(x: torch.Tensor) -> torch.Tensor:
    """Return x laid out so that its memory is [M, C] row-major (channels last)."""
    d = x.dim()
    if d == 4:
        return x if x.is_contiguous(memory_format=torch.channels_last) else x.contiguous(memory_format=torch.channels_last)
    if d == 2:
        return x.contiguous()
    if d == 5:
        return x.contiguous(memory_format=torch.channels_last_3d)
    if d == 3:  # [N, C, L]: view as [N, C, L, 1]
        return x.unsqueeze(-1).contiguous(memory_format=torch.channels_last).squeeze(-1)
    raise ValueError(f"expected at least 2D input (got {x.dim()}D input)")


def _f32(t):
    """fp32 contiguous view of an affine parameter (no copy in the usual case)."""
    if t is None:
        return None
    t = t.detach()
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()


class _DeviceGuard:
    """`with torch.cuda.device(...)` only when the current device is not already the tensor's
    (the context manager costs ~10 us, a BN layer of a small model runs in about as much)."""

    __slots__ = ("idx", "prev")

    def __init__(self, idx):
        self.idx = idx

    def __enter__(self):
        self.prev = torch._C._cuda_getDevice()
        if self.prev != self.idx:
            torch._C._cuda_setDevice(self.idx)

    def __exit__(self, *a):
        if self.prev != self.idx:
            torch._C._cuda_setDevice(self.prev)


class _FusedSyncBN(torch.autograd.Function):
    """y = act(bn(x) [+ residual]) with batch statistics shared across ranks."""

    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, eps, momentum, relu, comm, count_ptr=0):
        lib = _lib.lib()
        x = _as_rows(x)
        C = x.shape[1]
        M = x.numel() // C
        world = comm.world if comm is not None else 1
        if M == 1 and world < 2:
            raise ValueError(f"Expected more than 1 value per channel when training, got input size {M}")
        dt = _lib.dtype_code(x.dtype)
        if residual is not None:
            if residual.shape != x.shape or residual.dtype != x.dtype:
                raise ValueError("residual must have the shape and dtype of the input")
            residual = _as_rows(residual)
        w32, b32 = _f32(weight), _f32(bias)
        dev = x.device
        idx = dev.index
        with _DeviceGuard(idx):
            stream = _lib.raw_stream(idx)
            # one fp32 allocation: [stats_local (2C+1) | pad | save = mean, invstd, count (2C+1)]
            off = (2 * C + 1 + 3) // 4 * 4
            aux = torch.empty(off + 2 * C + 1, dtype=torch.float32, device=dev)
            y = torch.empty_like(x)
            h = comm.handle if world > 1 else None
            seq = comm.next_seq() if world > 1 else 0
            aux_p = aux.data_ptr()
            sv = aux_p + 4 * off
            xp, rp, yp = x.data_ptr(), 0 if residual is None else residual.data_ptr(), y.data_ptr()
            wp, bp = 0 if w32 is None else w32.data_ptr(), 0 if b32 is None else b32.data_ptr()
            rmp = 0 if running_mean is None else running_mean.data_ptr()
            rvp = 0 if running_var is None else running_var.data_ptr()
            scr = _bn_scratch(idx, stream)
            # ReLU after a residual add: the backward needs the sign of y.  bf16: the kernel also writes
            # one mask word per 16-byte vector (0.5 B/element) and the backward reads that instead of y.
            maskw = None
            if _USE_MASK_WORDS and relu and residual is not None and dt == _lib.BF16 and torch.is_grad_enabled():
                maskw = torch.empty(x.numel() // 8, dtype=torch.int32, device=dev)
            mp = 0 if maskw is None else maskw.data_ptr()
            fast = _lib.fast()
            if _lib.PROFILER is None and fast is not None:
                _lib.CALLS["b200dp_bn_forward"] = _lib.CALLS.get("b200dp_bn_forward", 0) + 1
                rc = fast.bn_forward(xp, rp, yp, dt, M, C, scr, aux_p, wp, bp, float(eps), float(momentum), rmp, rvp, sv,
                                     int(bool(relu)), h.value if h else 0, seq, stream, mp, count_ptr)
                if rc:
                    _lib.check(rc, "b200dp_bn_forward")
            elif _lib.PROFILER is None:
                _lib.check(lib.b200dp_bn_forward(xp, rp, yp, dt, M, C, scr, aux_p, wp, bp, float(eps), float(momentum),
                                                 rmp, rvp, sv, int(bool(relu)), h, seq, stream, mp, count_ptr or None),
                           "b200dp_bn_forward")
            elif lib.b200dp_bn_one_kernel(dt, M, C, 0):  # timing pass, small layer: one cluster kernel does both
                _lib.check(lib.b200dp_bn_forward(xp, rp, yp, dt, M, C, scr, aux_p, wp, bp, float(eps), float(momentum),
                                                 rmp, rvp, sv, int(bool(relu)), h, seq, stream, mp, count_ptr or None),
                           "b200dp_bn_forward")
            else:  # per-kernel timing pass: the same two launches through their own entry points
                _lib.check(lib.b200dp_bn_fwd_stats(xp, dt, M, C, scr, aux_p, h, seq, stream), "b200dp_bn_fwd_stats")
                _lib.check(lib.b200dp_bn_fwd_apply(xp, rp, yp, dt, M, C, aux_p, wp, bp, float(eps), float(momentum),
                                                   rmp, rvp, sv, sv + 4 * C, sv + 8 * C, int(bool(relu)), h, seq,
                                                   stream, mp), "b200dp_bn_fwd_apply")
        # ReLU mask in the backward: recomputed from x when nothing was added,
        # otherwise taken from the saved output.
        ctx.relu_mode = 0 if not relu else (1 if residual is None else (3 if maskw is not None else 2))
        ctx.comm = comm
        ctx.dims = (M, C, dt, off)
        ctx.has_residual = residual is not None
        ctx.save_for_backward(x, y if ctx.relu_mode == 2 else maskw, weight, bias, w32, b32, aux)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.lib()
        x, y, weight, bias, w32, b32, aux = ctx.saved_tensors
        M, C, dt, off = ctx.dims
        comm = ctx.comm
        world = comm.world if comm is not None else 1
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dy = _as_rows(dy)
        dev = x.device
        idx = dev.index
        with _DeviceGuard(idx):
            stream = _lib.raw_stream(idx)
            # one fp32 allocation: [sums (2C) | grad_weight (C) | grad_bias (C)]
            out = torch.empty(4 * C, dtype=torch.float32, device=dev)
            dx = torch.empty_like(x)
            dres = torch.empty_like(x) if (ctx.has_residual and ctx.needs_input_grad[1]) else None
            h = comm.handle if world > 1 else None
            seq = comm.next_seq() if world > 1 else 0
            out_p = out.data_ptr()
            sv = aux.data_ptr() + 4 * off
            dyp, xp, yp = dy.data_ptr(), x.data_ptr(), 0 if y is None else y.data_ptr()
            dxp, drp = dx.data_ptr(), 0 if dres is None else dres.data_ptr()
            wp, bp = 0 if w32 is None else w32.data_ptr(), 0 if b32 is None else b32.data_ptr()
            gwp = out_p + 8 * C if weight is not None else 0
            gbp = out_p + 12 * C if bias is not None else 0
            scr = _bn_scratch(idx, stream)
            fast = _lib.fast()
            if _lib.PROFILER is None and fast is not None:
                _lib.CALLS["b200dp_bn_backward"] = _lib.CALLS.get("b200dp_bn_backward", 0) + 1
                rc = fast.bn_backward(dyp, xp, yp, dxp, drp, dt, M, C, sv, wp, bp, ctx.relu_mode, scr, out_p, gwp, gbp,
                                      h.value if h else 0, seq, stream)
                if rc:
                    _lib.check(rc, "b200dp_bn_backward")
            elif _lib.PROFILER is None:
                _lib.check(lib.b200dp_bn_backward(dyp, xp, yp, dxp, drp, dt, M, C, sv, wp, bp, ctx.relu_mode, scr,
                                                  out_p, gwp, gbp, h, seq, stream), "b200dp_bn_backward")
            elif lib.b200dp_bn_one_kernel(dt, M, C, 1):
                _lib.check(lib.b200dp_bn_backward(dyp, xp, yp, dxp, drp, dt, M, C, sv, wp, bp, ctx.relu_mode, scr,
                                                  out_p, gwp, gbp, h, seq, stream), "b200dp_bn_backward")
            else:
                _lib.check(lib.b200dp_bn_bwd_reduce(dyp, xp, yp, dt, M, C, sv, sv + 4 * C, wp, bp, ctx.relu_mode, scr,
                                                    out_p, gwp, gbp, h, seq, stream), "b200dp_bn_bwd_reduce")
                _lib.check(lib.b200dp_bn_bwd_apply(dyp, xp, yp, dxp, drp, dt, M, C, sv, sv + 4 * C, wp, bp, out_p,
                                                   sv + 8 * C, ctx.relu_mode, h, seq, stream), "b200dp_bn_bwd_apply")
        gw = out[2 * C:3 * C] if (weight is not None and ctx.needs_input_grad[2]) else None
        gb = out[3 * C:4 * C] if (bias is not None and ctx.needs_input_grad[3]) else None
        if gw is not None and weight.dtype != torch.float32:
            gw = gw.to(weight.dtype)
        if gb is not None and bias.dtype != torch.float32:
            gb = gb.to(bias.dtype)
        return dx, dres, gw, gb, None, None, None, None, None, None, None


class _FusedBNInference(torch.autograd.Function):
    """Eval-mode y = act(bn(x; running stats) [+ residual])."""

    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, eps, relu):
        lib = _lib.lib()
        x = _as_rows(x)
        C = x.shape[1]
        M = x.numel() // C
        dt = _lib.dtype_code(x.dtype)
        if residual is not None:
            residual = _as_rows(residual)
        w32 = None if weight is None else weight.detach().float().contiguous()
        b32 = None if bias is None else bias.detach().float().contiguous()
        rm, rv = running_mean.float().contiguous(), running_var.float().contiguous()
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(lib.b200dp_bn_inference(x.data_ptr(), _lib.ptr(residual), y.data_ptr(), dt, M, C,
                                               rm.data_ptr(), rv.data_ptr(), _lib.ptr(w32), _lib.ptr(b32),
                                               float(eps), int(bool(relu)), _lib.current_stream()),
                       "b200dp_bn_inference")
        ctx.relu = bool(relu)
        ctx.eps = float(eps)
        ctx.has_residual = residual is not None
        ctx.save_for_backward(x, y, w32, rm, rv)
        return y

    @staticmethod
    def backward(ctx, dy):
        # Frozen statistics: the layer is an affine map per channel (not on the
        # training hot path; plain tensor algebra on the GPU).
        x, y, w32, rm, rv = ctx.saved_tensors
        shape = [1, -1] + [1] * (x.dim() - 2)
        g = dy * (y > 0) if ctx.relu else dy
        invstd = torch.rsqrt(rv + ctx.eps)
        scale = invstd if w32 is None else invstd * w32
        dx = (g.float() * scale.view(shape)).to(x.dtype)
        red = [d for d in range(x.dim()) if d != 1]
        gw = (g.float() * (x.float() - rm.view(shape)) * invstd.view(shape)).sum(red) if w32 is not None else None
        gb = g.float().sum(red) if w32 is not None else None
        return dx, (g if ctx.has_residual else None), gw, gb, None, None, None, None


class B200SyncBatchNorm(_BatchNorm):
    """``nn.SyncBatchNorm`` with optional fused ReLU and residual add.

    ``forward(input, residual=None)`` returns ``act(bn(input) + residual)`` where
    ``act`` is ReLU when ``fuse_relu`` is set.  Parameters, running statistics
    and all arithmetic of the statistics stay float32 whatever the activation
    dtype (float32 or bfloat16).
    """

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: Optional[float] = 0.1,
                 affine: bool = True, track_running_stats: bool = True, process_group=None,
                 device=None, dtype=None, fuse_relu: bool = False) -> None:
        super().__init__(num_features, eps, momentum, affine, track_running_stats,
                         device=device, dtype=dtype)
        self.process_group = process_group
        self.fuse_relu = fuse_relu

    def _apply(self, fn, recurse=True):
        # keep statistics and affine parameters in float32 through .to(bf16)/.half()
        super()._apply(fn, recurse)
        for name, p in list(self._parameters.items()):
            if p is not None and p.is_floating_point() and p.dtype != torch.float32:
                p.data = p.data.float()
                if p.grad is not None:
                    p.grad.data = p.grad.data.float()
        for name, b in list(self._buffers.items()):
            if b is not None and b.is_floating_point() and b.dtype != torch.float32:
                self._buffers[name] = b.float()
        return self

    def _check_input_dim(self, input):
        if input.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {input.dim()}D input)")

    def _check_non_zero_input_channels(self, input):
        if input.size(1) == 0:
            raise ValueError("SyncBatchNorm number of input channels should be non-zero")

    def extra_repr(self):
        return super().extra_repr() + f", fuse_relu={self.fuse_relu}"

    def forward(self, input: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        self._check_input_dim(input)
        self._check_non_zero_input_channels(input)
        if not input.is_cuda:
            # same condition as TORCH/nn/modules/batchnorm.py:800-809; there is no CPU path here
            raise ValueError("SyncBatchNorm expected input tensor to be on GPU or "
                             f"{torch._C._get_privateuse1_backend_name()}")

        # exponential_average_factor exactly as _BatchNorm / SyncBatchNorm choose it
        if self.momentum is None:
            factor = 0.0
        else:
            factor = self.momentum
        count_ptr = 0
        if self.training and self.track_running_stats:
            if self.momentum is None or _lib.PROFILER is not None or not self.num_batches_tracked.is_cuda:
                self.num_batches_tracked.add_(1)
            else:
                # the forward kernel adds the 1 (b200dp_bn_forward's last argument): one tiny torch kernel
                # less per layer and step, and the statistics kernel stays directly behind the convolution
                count_ptr = self.num_batches_tracked.data_ptr()
            if self.momentum is None:
                factor = 1.0 / self.num_batches_tracked.item()
            else:
                factor = self.momentum

        bn_training = self.training or (self.running_mean is None and self.running_var is None)
        running_mean = self.running_mean if not self.training or self.track_running_stats else None
        running_var = self.running_var if not self.training or self.track_running_stats else None

        if not bn_training:
            return _FusedBNInference.apply(input, residual, self.weight, self.bias, running_mean,
                                           running_var, self.eps, self.fuse_relu)

        comm = None
        if self.training and dist.is_available() and dist.is_initialized():
            world = dist.get_world_size(self.process_group)
            if world > 1:
                comm = stats_communicator(self.process_group)
        nat = _lib.native() if _lib.PROFILER is None else None
        if nat is not None:
            # the same node in C++ (csrc/autograd_bn.cpp): same checks, same two C-ABI calls
            return nat.fused_syncbn(input, residual, self.weight, self.bias,
                                    0 if running_mean is None else running_mean.data_ptr(),
                                    0 if running_var is None else running_var.data_ptr(),
                                    self.eps, factor, self.fuse_relu,
                                    comm.handle.value if comm is not None else 0,
                                    comm.world if comm is not None else 1, _USE_MASK_WORDS, count_ptr)
        return _FusedSyncBN.apply(input, residual, self.weight, self.bias, running_mean, running_var,
                                  self.eps, factor, self.fuse_relu, comm, count_ptr)

    @classmethod
    def convert_sync_batchnorm(cls, module, process_group=None):
        """Replace every ``_BatchNorm`` (and ``nn.SyncBatchNorm``) under ``module``.

        Mirrors ``nn.SyncBatchNorm.convert_sync_batchnorm``
        (TORCH/nn/modules/batchnorm.py:844-905): the new layer shares the source
        layer's Parameters and buffers; children are converted recursively.
        """
        out = module
        if isinstance(module, _BatchNorm) and not isinstance(module, cls):
            out = cls(module.num_features, module.eps, module.momentum, module.affine,
                      module.track_running_stats,
                      process_group if process_group is not None else getattr(module, "process_group", None))
            if module.affine:
                with torch.no_grad():
                    out.weight = module.weight
                    out.bias = module.bias
            out.running_mean = module.running_mean
            out.running_var = module.running_var
            out.num_batches_tracked = module.num_batches_tracked
            out.training = module.training
            if hasattr(module, "qconfig"):
                out.qconfig = module.qconfig
            out._apply(lambda t: t)  # statistics / affine back to float32 if the source was cast
        for name, child in module.named_children():
            out.add_module(name, cls.convert_sync_batchnorm(child, process_group))
        del module
        return out


convert_sync_batchnorm = B200SyncBatchNorm.convert_sync_batchnorm
