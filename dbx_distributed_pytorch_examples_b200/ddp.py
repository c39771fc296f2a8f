"""DistributedDataParallel-compatible wrapper with a fused all-reduce + optimizer step.

Drop-in for the call at 01_torch_distributor/01_basic_torch_distributor.py:291
``DDP(model, device_ids=[local_rank], output_device=local_rank)`` and for the
wrappers Accelerate / Ray build (04_accelerate/01_cifar_accelerate.ipynb:503,
05_ray/02_cifar_resnet_pytorch_ray.ipynb:280).  What torch's Reducer + NCCL do
per gradient bucket (TORCH/nn/parallel/distributed.py:330-2472,
TORCH/include/torch/csrc/distributed/c10d/reducer.hpp:45-610) is done here by
two kernels of csrc/update.cu per bucket: pack (grad -> flat bucket, x 1/N) and
the fused reduce(+optimizer)(+parameter push) over NVLink peer memory.

Layout.  Parameters of one dtype form an *arena*: a flat parameter buffer and a
flat gradient buffer of the same length, both inside the communicator's
peer-visible heap (every ``param.data`` is re-pointed to a view of the flat
parameter buffer), plus — once an optimizer is fused — local fp32 state buffers
and, for bf16 parameters, fp32 master weights.  Buckets are contiguous element
ranges of an arena, filled in reverse registration order with torch's caps
(first 1 MiB, then ``bucket_cap_mb``; distributed.py:831-853).

Modes.
* plain (default): after ``backward()`` every ``param.grad`` holds the gradient
  averaged over the ranks, exactly what stock DDP leaves there, and the user's
  optimizer runs as usual.
* fused (``fuse_optimizer(optimizer)``): the bucket kernel also applies
  SGD-momentum / Adam to this rank's 1/N slice and pushes the new parameters to
  every rank; ``optimizer.step()`` only does the bookkeeping.  Precedent in
  torch: ``_register_fused_optim`` (distributed.py:2097-2167).
"""
from __future__ import annotations

import ctypes
import math
import os
from contextlib import contextmanager
from typing import List, Optional

import torch
import torch.distributed as dist
from torch.autograd import Variable
from torch.nn.parallel import DistributedDataParallel

from . import _lib
from .comm import Communicator

_ALIGN_ELEMS = 8
_FIRST_BUCKET_BYTES = 1024 * 1024  # distributed.py:853 (_DEFAULT_FIRST_BUCKET_BYTES)


def plan_buckets(numels: List[int], elem_size: int, world: int, cap_bytes: int,
                 first_cap_bytes: int = _FIRST_BUCKET_BYTES):
    """Lay parameters (already in bucket order) out in a flat arena.

    Returns ``(offsets, buckets, total)``: the element offset of every parameter
    (multiples of 8), ``buckets`` as ``(start, numel, first_param, last_param+1)``
    with ``numel`` a multiple of ``8 * world`` (so that every rank's slice is
    whole 16-byte vectors), and the arena length.  A bucket closes once it holds
    at least its cap, like ``_compute_bucket_assignment_by_size``.
    """
    quantum = _ALIGN_ELEMS * world
    offsets, buckets = [], []
    cur, start, first = 0, 0, 0
    cap = first_cap_bytes
    for k, n in enumerate(numels):
        offsets.append(cur)
        cur += (n + _ALIGN_ELEMS - 1) // _ALIGN_ELEMS * _ALIGN_ELEMS
        if (cur - start) * elem_size >= cap:
            cur = start + (cur - start + quantum - 1) // quantum * quantum
            buckets.append((start, cur - start, first, k + 1))
            start, first, cap = cur, k + 1, cap_bytes
    if first < len(numels):
        cur = start + (cur - start + quantum - 1) // quantum * quantum
        buckets.append((start, cur - start, first, len(numels)))
    return offsets, buckets, cur


class _Arena:
    """All trainable parameters of one dtype."""

    def __init__(self, dtype, params):
        self.dtype = dtype
        self.code = _lib.dtype_code(dtype)
        self.esize = 2 if dtype == torch.bfloat16 else 4
        self.params = params  # in bucket (reverse registration) order
        self.offsets = self.buckets = None
        self.total = 0
        self.grad_off = self.param_off = 0  # byte offsets in the heap
        self.flat_param = self.flat_grad = None
        self.master = self.state1 = self.state2 = None
        self.grad_views = []


class _BucketState:
    def __init__(self, arena, index, start, numel, p0, p1):
        self.arena, self.index = arena, index
        self.start, self.numel, self.p0, self.p1 = start, numel, p0, p1
        n = p1 - p0
        self.numels = (ctypes.c_int64 * n)(*[p.numel() for p in arena.params[p0:p1]])
        self.offsets = (ctypes.c_int64 * n)(*[o - start for o in arena.offsets[p0:p1]])
        self.srcs = (ctypes.c_void_p * n)()
        self.pending = n
        self.launched = False


class _HookBucket:
    """Duck-typed ``dist.GradBucket`` handed to hooks registered with
    ``register_comm_hook`` (TORCH/include/torch/csrc/distributed/c10d/comm.hpp:20-100)."""

    def __init__(self, b: _BucketState, last: bool):
        self._b, self._last = b, last
        a = b.arena
        self._buf = a.flat_grad[b.start:b.start + b.numel]

    def index(self):
        return self._b.index

    def buffer(self):
        return self._buf

    def set_buffer(self, t):
        self._buf = t

    def is_last(self):
        return self._last

    def parameters(self):
        return list(self._b.arena.params[self._b.p0:self._b.p1])

    def gradients(self):
        return list(self._b.arena.grad_views[self._b.p0:self._b.p1])


class B200DDP(DistributedDataParallel):
    """``DistributedDataParallel``-compatible module wrapper (see module docstring).

    Passes ``isinstance(m, torch.nn.parallel.DistributedDataParallel)`` (the Ray
    example checks it, 05_ray/02_cifar_resnet_pytorch_ray.ipynb:317) and exposes
    ``.module``; none of torch's Reducer machinery is constructed.
    """

    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 init_sync=True, process_group=None, bucket_cap_mb=None, find_unused_parameters=False,
                 check_reduction=False, gradient_as_bucket_view=False, static_graph=False,
                 delay_all_reduce_named_params=None, param_to_hook_all_reduce=None,
                 mixed_precision=None, device_mesh=None, skip_all_reduce_unused_params=False, *,
                 one_shot_bytes: int = 256 * 1024, max_ctas: int = 48, nvls=None):
        torch.nn.Module.__init__(self)
        if mixed_precision is not None or device_mesh is not None or delay_all_reduce_named_params is not None:
            raise NotImplementedError("B200DDP does not implement mixed_precision / device_mesh / delayed all-reduce")
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            # same condition and exception type as distributed.py:737-745
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any "
                               "parameter that requires a gradient.")
        if device_ids is not None and len(device_ids) > 1:
            raise ValueError("device_ids can only be None or contain a single element.")  # distributed.py:747-751
        devices = {p.device for p in params}
        if len(devices) != 1 or next(iter(devices)).type != "cuda":
            raise ValueError("B200DDP needs all parameters on one CUDA device; it has no CPU path "
                             f"(found {sorted(str(d) for d in devices)})")
        self.device = next(iter(devices))
        self.module = module
        self.dim = dim
        self.device_ids = [self.device.index]
        self.output_device = self.device.index if output_device is None else output_device
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = find_unused_parameters
        self.static_graph = static_graph
        self.gradient_as_bucket_view = True  # by construction
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self.process_group = process_group
        self.one_shot_bytes = int(one_shot_bytes)
        self.max_ctas = int(max_ctas)
        self.materialize_grads = True
        self._optimizer = None
        self._opt_step_count = 0
        self._comm_hook = None
        self._iteration_armed = False
        self._bcast_buffers = None
        self._iterations = 0
        self._step_counted = False
        self.error_poll_interval = 64
        self._err_word = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.bucket_bytes_cap = int((25 if bucket_cap_mb is None else bucket_cap_mb) * 1024 * 1024)

        if dist.is_available() and dist.is_initialized():
            self.world = dist.get_world_size(process_group)
            self.rank = dist.get_rank(process_group)
        else:
            self.world, self.rank = 1, 0

        # ---- arenas, buckets, heap -------------------------------------------------
        by_dtype = {}
        for p in reversed(params):  # reverse registration order ~ gradient-ready order
            by_dtype.setdefault(p.dtype, []).append(p)
        self._arenas = [_Arena(dt, ps) for dt, ps in by_dtype.items()]
        heap = 0
        for a in self._arenas:
            a.offsets, a.buckets, a.total = plan_buckets([p.numel() for p in a.params], a.esize, self.world,
                                                         self.bucket_bytes_cap)
            nbytes = (a.total * a.esize + 511) // 512 * 512
            a.grad_off, a.param_off = heap, heap + nbytes
            heap += 2 * nbytes
        with torch.cuda.device(self.device):
            # nvls: True = the heap must have a multicast (NVLS) mapping, "auto" = use one when the platform
            # offers it, False = plain IPC heap and the peer load/store kernel; None = what B200DP_NVLS says
            # in the environment ("1", "auto", anything else or unset = off)
            if nvls is None:
                nvls = {"1": True, "auto": "auto"}.get(os.environ.get("B200DP_NVLS", "0"), False)
            if nvls == "auto":   # the switch reduction pays from 4 ranks on (at 2 it moves more bytes than peer loads)
                multicast = None if self.world >= 4 else False
            else:
                multicast = bool(nvls)
            self.comm = Communicator(process_group, heap_bytes=heap, device=self.device, multicast=multicast)
            self._comm_stream = torch.cuda.Stream(device=self.device)
        self._param_index = {}
        self._buckets: List[_BucketState] = []
        with torch.no_grad():
            for a in self._arenas:
                a.flat_grad = self.comm.heap_view(a.grad_off, a.total, a.dtype)
                a.flat_param = self.comm.heap_view(a.param_off, a.total, a.dtype)
                for p, off in zip(a.params, a.offsets):
                    # keep the parameter's physical layout (channels_last conv weights stay
                    # KRSC): gradients then arrive in the same layout and pack as flat memory
                    view = _strided_like(a.flat_param[off:off + p.numel()], p)
                    view.copy_(p.data)
                    p.data = view
                    a.grad_views.append(_strided_like(a.flat_grad[off:off + p.numel()], p))
                for (start, numel, p0, p1) in a.buckets:
                    b = _BucketState(a, len(self._buckets), start, numel, p0, p1)
                    self._buckets.append(b)
                    for k in range(p0, p1):
                        self._param_index[a.params[k]] = (b, k - p0, a.grad_views[k])
        if self.world > 1 and init_sync:
            self._sync_module_states()
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad_ready)
        # checkpoints must not drag the whole heap along: hand out copies
        module._register_state_dict_hook(_clone_state_dict_hook)
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            self.comm.barrier()

    # ---- init-time synchronisation (distributed.py:879-890) -------------------------
    def _sync_module_states(self):
        for a in self._arenas:
            dist.broadcast(a.flat_param, src=dist.get_global_rank(self.process_group, 0)
                           if self.process_group is not None else 0, group=self.process_group)
        for buf in self.module.buffers():
            dist.broadcast(buf, src=dist.get_global_rank(self.process_group, 0)
                           if self.process_group is not None else 0, group=self.process_group)

    def _needs_buffer_broadcast(self):
        from .syncbn import B200SyncBatchNorm

        owned = set()
        for m in self.module.modules():
            if isinstance(m, B200SyncBatchNorm):
                owned.update(id(b) for b in m.buffers(recurse=False))
        return [b for b in self.module.buffers() if id(b) not in owned]

    # ---- forward ---------------------------------------------------------------------
    def forward(self, *inputs, **kwargs):
        if getattr(self, "_unpickled", False):
            if torch.is_grad_enabled() and any(p.requires_grad for p in self.module.parameters()) and self.training:
                raise RuntimeError("this B200DDP was restored from a pickle and has no communicator: use it for "
                                   "inference, or wrap .module in a new B200DDP to train")
            return self.module(*inputs, **kwargs)
        if self.world > 1 and self.broadcast_buffers and self.require_forward_param_sync:
            # buffers of B200SyncBatchNorm layers are identical on all ranks by construction
            if self._bcast_buffers is None:  # computed once: the module tree does not change between steps
                self._bcast_buffers = self._needs_buffer_broadcast()
            if self._bcast_buffers:
                # one coalesced broadcast for all of them, like DDP's _sync_buffers
                # (TORCH/nn/parallel/distributed.py:2214-2257 -> _distributed_broadcast_coalesced)
                pg = self.process_group if self.process_group is not None else dist.group.WORLD
                dist._broadcast_coalesced(pg, self._bcast_buffers, 250 * 1024 * 1024, 0)
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            self._arm_iteration()
        return self.module(*inputs, **kwargs)

    def _arm_iteration(self):
        for b in self._buckets:
            b.pending = b.p1 - b.p0
            b.launched = False
        self._next_bucket = 0
        self._callback_queued = False
        self._iteration_armed = True
        # one sequence number per bucket, reserved up front so that it does not
        # depend on the order in which gradients become ready
        self._seqs = [self.comm.next_seq() for _ in self._buckets]
        # the optimizer's step number advances when an update is actually issued (first bucket launch of
        # this iteration), not here: a grad-enabled forward that is never followed by backward() must not
        # shift Adam's bias correction or consume SGD's first step
        self._step_counted = False

    @contextmanager
    def no_sync(self):
        """Accumulate gradients locally (distributed.py:1474-1500)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    # ---- backward hooks ----------------------------------------------------------------
    def _on_grad_ready(self, param):
        if not (self.require_backward_grad_sync and self._iteration_armed):
            return
        if not self._callback_queued:
            Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
        b, _, _ = self._param_index[param]
        b.pending -= 1
        while self._next_bucket < len(self._buckets):
            nb = self._buckets[self._next_bucket]
            if nb.pending > 0:
                break
            self._launch_bucket(nb)
            self._next_bucket += 1

    def _finalize_backward(self):
        # buckets with parameters that received no gradient this iteration
        while self._next_bucket < len(self._buckets):
            self._launch_bucket(self._buckets[self._next_bucket])
            self._next_bucket += 1
        torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        self._iteration_armed = False
        self._iterations += 1
        if self.world > 1 and self._iterations % self.error_poll_interval == 0:
            self._poll_error()

    def _poll_error(self, blocking: bool = False):
        """A kernel that gave up waiting for a peer raises the rank's error word and skips its update; the
        replicas are then out of step and training must stop.  Every `error_poll_interval` iterations the
        word is copied to pinned host memory behind the bucket kernels (no host synchronisation) and the
        value the previous copy delivered is looked at; checkpoint paths read it synchronously."""
        if blocking:
            self.comm.check_error()
            return
        if int(self._err_word[0]) != 0:
            raise _lib.B200DPError("a libb200dp kernel gave up waiting for a peer rank (peer dead or more than "
                                   "B200DP_TIMEOUT_S seconds behind): the replicas are out of step, stop training")
        _lib.check(_lib.lib().b200dp_comm_error_async(self.comm.handle, self._err_word.data_ptr(),
                                                       self._comm_stream.cuda_stream), "b200dp_comm_error_async")

    def _launch_bucket(self, b: _BucketState):
        lib = _lib.lib()
        a = b.arena
        if self._optimizer is not None and self._comm_hook is None and not self._step_counted:
            self._opt_step_count += 1  # every bucket of this iteration uses the same step number
            self._step_counted = True
            self._state_stale = True
        cur = torch.cuda.current_stream(self.device)
        self._comm_stream.wait_stream(cur)
        with torch.cuda.stream(self._comm_stream):
            stream = self._comm_stream.cuda_stream
            n = 0
            keep_alive = []  # converted copies must outlive the pack launch below
            for k in range(b.p0, b.p1):
                p = a.params[k]
                g = p.grad
                if g is None:
                    a.grad_views[k].zero_()
                    continue
                if g.dtype != a.dtype or not _same_layout(g, p):
                    g = torch.empty_like(p).copy_(g)  # rare: layout or dtype differs from the parameter's
                    keep_alive.append(g)
                g.record_stream(self._comm_stream)
                b.srcs[n] = g.data_ptr()
                b.numels[n] = p.numel()
                b.offsets[n] = a.offsets[k] - b.start
                n += 1
            bucket_ptr = a.flat_grad.data_ptr() + b.start * a.esize
            _lib.check(lib.b200dp_bucket_pack(b.srcs, b.numels, b.offsets, n, bucket_ptr, a.code,
                                              1.0 / self.world, stream), "b200dp_bucket_pack")
            if self._comm_hook is not None:
                self._run_comm_hook(b)
            else:
                opt = self._opt_args(a)
                fused = opt.kind != _lib.OPT_NONE
                write_grads = 1 if (self.materialize_grads or not fused) else 0
                one_shot = int(self.world > 1 and not write_grads and b.numel * a.esize <= self.one_shot_bytes)
                off = b.start
                _lib.check(lib.b200dp_allreduce_update(
                    self.comm.handle, a.grad_off, a.code, a.param_off, a.code, off, b.numel,
                    None if a.master is None else a.master.data_ptr() + 4 * off,
                    None if a.state1 is None else a.state1.data_ptr() + 4 * off,
                    None if a.state2 is None else a.state2.data_ptr() + 4 * off,
                    ctypes.byref(opt), one_shot, write_grads, self.max_ctas,
                    self._seqs[b.index], stream), "b200dp_allreduce_update")
        keep = self.materialize_grads or self._optimizer is None or self._comm_hook is not None
        for k in range(b.p0, b.p1):
            a.params[k].grad = a.grad_views[k] if keep else None
        b.launched = True

    # ---- comm hook compatibility (distributed.py:1987) -----------------------------------
    def register_comm_hook(self, state, hook):
        if not callable(hook):
            raise TypeError("Communication hook must be callable.")
        if self._optimizer is not None:
            raise RuntimeError("a comm hook cannot be combined with a fused optimizer")
        self._comm_hook = (state, hook)

    def _run_comm_hook(self, b: _BucketState):
        state, hook = self._comm_hook
        hb = _HookBucket(b, b.index == len(self._buckets) - 1)
        # the pack kernel pre-divided by world like the Reducer does; hooks such as
        # default_hooks.allreduce_hook divide themselves, so undo it
        hb.buffer().mul_(self.world)
        fut = hook(state, hb)
        out = fut.wait() if hasattr(fut, "wait") else fut
        if isinstance(out, (list, tuple)):
            out = out[0]
        if out.data_ptr() != hb._b.arena.flat_grad.data_ptr() + b.start * b.arena.esize:
            b.arena.flat_grad[b.start:b.start + b.numel].copy_(out.view(-1)[:b.numel])

    # ---- fused optimizer -----------------------------------------------------------------
    def fuse_optimizer(self, optimizer, materialize_grads: bool = False):
        """Move ``optimizer``'s update into the bucket kernels.

        Supports ``torch.optim.SGD`` (momentum, dampening, nesterov, weight_decay;
        01_basic_torch_distributor.py:294) and ``torch.optim.Adam`` (coupled L2
        weight decay; 02_cifar_torch_distributor_resnet.py:213,
        04_accelerate/01_cifar_accelerate.ipynb:482-486), one set of
        hyper-parameters for all wrapped parameters.  Hyper-parameters are read
        from ``optimizer.param_groups`` at every step, so LR schedulers keep
        working.  ``optimizer.step()`` becomes bookkeeping only.
        """
        if self._comm_hook is not None:
            raise RuntimeError("a fused optimizer cannot be combined with a comm hook")
        if isinstance(optimizer, torch.optim.Adam) and not isinstance(optimizer, torch.optim.AdamW):
            kind = _lib.OPT_ADAM
            if any(g.get("amsgrad") or g.get("maximize") for g in optimizer.param_groups):
                raise NotImplementedError("amsgrad / maximize are not fused")
        elif isinstance(optimizer, torch.optim.SGD):
            kind = _lib.OPT_SGD
            if any(g.get("maximize") for g in optimizer.param_groups):
                raise NotImplementedError("maximize is not fused")
        else:
            raise NotImplementedError(f"cannot fuse {type(optimizer).__name__}; use SGD or Adam")
        ours = set(self._param_index)
        groups = [g for g in optimizer.param_groups if any(p in ours for p in g["params"])]
        covered = {p for g in groups for p in g["params"]}
        if not ours <= covered:
            raise ValueError("the optimizer does not hold all parameters of the wrapped module")
        keys = ("lr", "momentum", "dampening", "nesterov", "weight_decay", "betas", "eps")
        first = {k: groups[0].get(k) for k in keys}
        for g in groups[1:]:
            if {k: g.get(k) for k in keys} != first:
                raise NotImplementedError("parameter groups with different hyper-parameters are not fused")
        self._optimizer, self._opt_kind, self._opt_group = optimizer, kind, groups[0]
        self.materialize_grads = bool(materialize_grads)
        for a in self._arenas:
            a.state1 = torch.zeros(a.total, dtype=torch.float32, device=self.device)
            a.state2 = torch.zeros(a.total, dtype=torch.float32, device=self.device) if kind == _lib.OPT_ADAM else None
            a.master = a.flat_param.float() if a.dtype != torch.float32 else None
        # an optimizer that has already taken steps (or was restored from a checkpoint) keeps its state
        self._adopt_optimizer_state()
        self._state_stale = False
        optimizer.register_state_dict_pre_hook(self._check_state_consolidated)
        # optimizer.load_state_dict() replaces optimizer.state with fresh tensors: copy them into the arenas
        # the kernels use and restore the step number (resume path of 01_basic_torch_distributor.py:109-119)
        optimizer.register_load_state_dict_post_hook(lambda opt: self._adopt_optimizer_state())
        # module.load_state_dict() writes the (bf16) parameters in place: the fp32 master copy follows
        self._module_hook = self.module.register_load_state_dict_post_hook(lambda mod, keys: self._refresh_master())
        ddp = self

        def step(opt_self, closure=None):
            loss = closure() if closure is not None else None
            if ddp._opt_kind == _lib.OPT_ADAM:
                for a in ddp._arenas:
                    for p in a.params:
                        opt_self.state[p]["step"] += 1
            return loss

        # a bound method, like the one it replaces: LR schedulers wrap optimizer.step and look at
        # its __func__ / __self__ (torch/optim/lr_scheduler.py, patch_track_step_called)
        import types
        optimizer.step = types.MethodType(step, optimizer)
        return optimizer

    @torch.no_grad()
    def _adopt_optimizer_state(self):
        """Copy whatever per-parameter state the optimizer holds (momentum_buffer / exp_avg / exp_avg_sq /
        step — from earlier un-fused steps or from ``optimizer.load_state_dict``) into the flat arenas the
        kernels update, re-point the optimizer's entries at views of those arenas, and restore the step
        number (Adam's bias correction, SGD's "first step" rule)."""
        opt, kind = self._optimizer, self._opt_kind

        def adopt(st, key, arena_slice, p):
            old = st.get(key)
            if old is None:
                arena_slice.zero_()
            elif old.data_ptr() != arena_slice.data_ptr():
                arena_slice.copy_(_flat_like(old, p))
            st[key] = _strided_like(arena_slice, p)
            return old is not None

        count = 0
        for a in self._arenas:
            for p, off in zip(a.params, a.offsets):
                st = opt.state[p]
                v1 = a.state1[off:off + p.numel()]
                if kind == _lib.OPT_SGD:
                    if adopt(st, "momentum_buffer", v1, p):
                        count = max(count, 1, self._opt_step_count)  # a buffer exists: the first step is behind us
                else:
                    adopt(st, "exp_avg", v1, p)
                    adopt(st, "exp_avg_sq", a.state2[off:off + p.numel()], p)
                    t = st.get("step")
                    t = float(t) if t is not None else 0.0
                    st["step"] = torch.tensor(t)
                    count = max(count, int(t))
        self._opt_step_count = count
        self._state_stale = False  # written whole: every rank holds all of it until the next sharded update

    @torch.no_grad()
    def _refresh_master(self):
        for a in self._arenas:
            if a.master is not None:
                a.master.copy_(a.flat_param)

    def _opt_args(self, arena) -> _lib.OptArgs:
        o = _lib.OptArgs()
        if self._optimizer is None:
            o.kind = _lib.OPT_NONE
            return o
        g = self._opt_group
        o.kind = self._opt_kind
        o.lr = float(g["lr"])
        o.weight_decay = float(g.get("weight_decay", 0.0))
        t = self._opt_step_count
        if self._opt_kind == _lib.OPT_SGD:
            o.momentum = float(g.get("momentum", 0.0))
            o.dampening = float(g.get("dampening", 0.0))
            o.nesterov = int(bool(g.get("nesterov", False)))
            o.first_step = int(t == 1)
        else:
            b1, b2 = g["betas"]
            o.beta1, o.beta2, o.eps = float(b1), float(b2), float(g["eps"])
            # computed in double on the host, like _single_tensor_adam (adam.py)
            o.step_size = float(g["lr"]) / (1.0 - b1 ** t)
            o.bias_correction2_sqrt = math.sqrt(1.0 - b2 ** t)
        return o

    # ---- checkpoint support ----------------------------------------------------------------
    def consolidate_optimizer_state(self):
        """Make every rank's fused optimizer state complete.  COLLECTIVE: call it on all ranks
        (e.g. right before ``if rank == 0: save_checkpoint(...)``,
        01_torch_distributor/01_basic_torch_distributor.py:239-245,298-299).

        Under the two-shot kernel rank r only maintains the r-th 1/N slice of each bucket's
        momentum / Adam moments / fp32 master weights; this all-gathers the slices over the
        process group (epoch-rate, so torch.distributed is fine here), after which
        ``optimizer.state_dict()`` holds exactly what a replicated torch optimizer would."""
        if self._optimizer is None or self.world == 1:
            self._state_stale = False
            return
        torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        self._poll_error(blocking=True)  # never write a checkpoint from replicas that fell out of step
        for b in self._buckets:
            a = b.arena
            sl = b.numel // self.world
            for buf in (a.state1, a.state2, a.master):
                if buf is None:
                    continue
                whole = buf[b.start:b.start + b.numel]
                mine = whole[self.rank * sl:(self.rank + 1) * sl].clone()
                dist.all_gather_into_tensor(whole, mine, group=self.process_group)
        self._state_stale = False

    def _check_state_consolidated(self, *_):
        if self.world > 1 and getattr(self, "_state_stale", False):
            raise RuntimeError(
                "the fused optimizer's state is sharded across ranks (each rank updates 1/N of every bucket): "
                "call ddp.consolidate_optimizer_state() on ALL ranks before optimizer.state_dict()")

    # ---- housekeeping ---------------------------------------------------------------------
    def bucket_sizes_bytes(self):
        return [b.numel * b.arena.esize for b in self._buckets]

    def __getstate__(self):
        """Pickle support for ``mlflow.pytorch.log_model(ddp_model, ...)``
        (01_torch_distributor/01_basic_torch_distributor.py:303), which pickles the wrapper itself.  What is
        saved is the wrapped module with its parameters copied OUT of the peer-visible heap (a parameter here is
        a view of one large allocation, which must not travel) plus the constructor's settings; the
        communicator, streams and hooks are not."""
        import copy

        memo = {id(p): torch.nn.Parameter(p.detach().clone(memory_format=torch.preserve_format), requires_grad=p.requires_grad)
                for p in self.module.parameters()}
        from .fusion import unfuse_bn_relu

        module = copy.deepcopy(self.module, memo)
        hook = getattr(self, "_module_hook", None)
        if hook is not None:  # the master-weight refresh hook closes over this wrapper
            module._load_state_dict_post_hooks.pop(hook.id, None)
        unfuse_bn_relu(module)  # fused blocks carry an instance-bound forward that cannot be unpickled
        return {"module": module, "device_ids": self.device_ids,
                "output_device": self.output_device, "broadcast_buffers": self.broadcast_buffers,
                "bucket_cap_mb": self.bucket_bytes_cap / (1024 * 1024), "one_shot_bytes": self.one_shot_bytes,
                "max_ctas": self.max_ctas}

    def __setstate__(self, state):
        """An unpickled wrapper carries the module for inference / export (``.module``, ``state_dict()``,
        ``forward`` under ``torch.no_grad()``); to train again wrap ``.module`` in a new ``B200DDP`` inside an
        initialised process group (the peer mappings cannot be restored from a file)."""
        torch.nn.Module.__init__(self)
        self.module = state["module"]
        self.device_ids, self.output_device = state["device_ids"], state["output_device"]
        self.broadcast_buffers = state["broadcast_buffers"]
        self.bucket_bytes_cap = int(state["bucket_cap_mb"] * 1024 * 1024)
        self.one_shot_bytes, self.max_ctas = state["one_shot_bytes"], state["max_ctas"]
        self.world, self.rank = 1, 0
        self.process_group = None
        self.require_backward_grad_sync = self.require_forward_param_sync = True
        self._optimizer = self._comm_hook = None
        self._unpickled = True

    def _get_name(self):
        return "B200DDP"

    def _get_ddp_logging_data(self):
        return {"backend_name": "b200dp-nvlink-p2p", "world_size": self.world, "rank": self.rank}


def prepare_model(model: torch.nn.Module, device=None, *, sync_batchnorm: bool = True, fuse: bool = True,
                  channels_last: bool = True, dtype: Optional[torch.dtype] = None, **ddp_kwargs) -> "B200DDP":
    """One call for what Accelerate's ``accelerator.prepare(model)`` (04_accelerate/01_cifar_accelerate.ipynb:503)
    and Ray's ``ray.train.torch.prepare_model(model)`` (05_ray/02_cifar_resnet_pytorch_ray.ipynb:280) do for the
    reference — move the model to this rank's GPU and wrap it for data-parallel training — plus the
    conversions this library adds (SyncBatchNorm replacement, BN+ReLU fusion, channels_last, dtype).
    The result passes ``isinstance(m, DistributedDataParallel)`` and exposes ``.module`` as both expect."""
    import os

    from .fusion import fuse_bn_relu
    from .syncbn import convert_sync_batchnorm

    if device is None:
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", torch.cuda.current_device())))
    device = torch.device(device)
    if sync_batchnorm:
        model = convert_sync_batchnorm(model)
        if fuse:
            model = fuse_bn_relu(model)
    model = model.to(device)
    if dtype is not None:
        model = model.to(dtype)
    if channels_last:
        model = model.to(memory_format=torch.channels_last)
    return B200DDP(model, device_ids=[device.index], output_device=device.index, **ddp_kwargs)


def _flat_like(t: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """`t` (a state tensor shaped like p) as a 1-D fp32 tensor in p's physical element order."""
    t = t.to(device=p.device, dtype=torch.float32)
    if not _same_layout(t, p):
        t = torch.empty_like(p, dtype=torch.float32).copy_(t)
    return t.as_strided((p.numel(),), (1,), t.storage_offset()) if p.numel() else t.reshape(-1)


def _same_layout(g: torch.Tensor, p: torch.Tensor) -> bool:
    """Same physical element order: strides agree on every dimension longer than 1
    (a 1x1 conv weight is the same memory whether it is called channels_last or not)."""
    return all(sg == sp for sg, sp, n in zip(g.stride(), p.stride(), p.shape) if n > 1)


def _strided_like(flat: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """View of the 1-D slice `flat` with p's shape and, when p is a dense permutation
    (e.g. channels_last conv weights), p's strides."""
    if p.is_contiguous() or torch.empty_like(p).stride() != p.stride():
        return flat.view(p.shape)
    return flat.as_strided(p.shape, p.stride())


def _clone_state_dict_hook(module, state_dict, prefix, local_metadata):
    for k, v in list(state_dict.items()):
        if isinstance(v, torch.Tensor) and v.is_cuda:
            state_dict[k] = v.detach().clone()
    return state_dict
