"""GPU, one rank: bucket pack and the fused (all-reduce +) optimizer kernel through the
C ABI, against torch.optim golden trajectories and the numpy oracle.  Bar (north_star):
updated weights within rtol 1e-5 in fp32, 2e-2 in bf16."""
import ctypes

import numpy as np
import pytest
import torch

from dbx_distributed_pytorch_examples_b200 import _lib
from dbx_distributed_pytorch_examples_b200.comm import Communicator
from oracle import optim_ref

pytestmark = pytest.mark.gpu


def opt_args(kind, step, **kw):
    import math
    o = _lib.OptArgs()
    o.kind, o.lr, o.weight_decay = kind, kw.get("lr", 0.0), kw.get("weight_decay", 0.0)
    if kind == _lib.OPT_SGD:
        o.momentum, o.dampening = kw.get("momentum", 0.0), kw.get("dampening", 0.0)
        o.nesterov, o.first_step = int(kw.get("nesterov", False)), int(step == 1)
    elif kind == _lib.OPT_ADAM:
        b1, b2 = kw.get("betas", (0.9, 0.999))
        o.beta1, o.beta2, o.eps = b1, b2, kw.get("eps", 1e-8)
        o.step_size = kw["lr"] / (1 - b1 ** step)
        o.bias_correction2_sqrt = math.sqrt(1 - b2 ** step)
    return o


class Bucket:
    """One arena of n (padded) fp32/bf16 elements in a single-rank communicator's heap."""

    def __init__(self, n, dtype=torch.float32):
        self.n = (n + 7) // 8 * 8
        self.dtype, self.code = dtype, _lib.dtype_code(dtype)
        es = 2 if dtype == torch.bfloat16 else 4
        nbytes = (self.n * es + 511) // 512 * 512
        self.comm = Communicator(heap_bytes=2 * nbytes)
        self.grad_off, self.param_off = 0, nbytes
        self.grad = self.comm.heap_view(0, self.n, dtype)
        self.param = self.comm.heap_view(nbytes, self.n, dtype)
        self.s1 = torch.zeros(self.n, device="cuda")
        self.s2 = torch.zeros(self.n, device="cuda")
        self.master = None

    def step(self, opt, write_grads=0):
        lib = _lib.lib()
        _lib.check(lib.b200dp_allreduce_update(
            self.comm.handle, self.grad_off, self.code, self.param_off, self.code, 0, self.n,
            None if self.master is None else self.master.data_ptr(), self.s1.data_ptr(), self.s2.data_ptr(),
            ctypes.byref(opt), 0, write_grads, 0, self.comm.next_seq(), _lib.current_stream()))


CASES = {
    "sgd_m05": (_lib.OPT_SGD, dict(lr=1e-3, momentum=0.5)),
    "sgd_nesterov_wd": (_lib.OPT_SGD, dict(lr=1e-2, momentum=0.9, nesterov=True, weight_decay=1e-4)),
    "sgd_plain": (_lib.OPT_SGD, dict(lr=1e-2)),
    "adam": (_lib.OPT_ADAM, dict(lr=1e-3)),
    "adam_wd": (_lib.OPT_ADAM, dict(lr=1e-3, weight_decay=1e-4)),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_fused_update_matches_torch_optim_golden(golden, case):
    g = golden("optim.npz")
    kind, kw = CASES[case]
    n = len(g["p0"])
    bk = Bucket(n)
    bk.param[:n].copy_(torch.from_numpy(g["p0"]))
    for k, grad in enumerate(g["grads"]):
        bk.grad.zero_()
        bk.grad[:n].copy_(torch.from_numpy(grad))
        bk.step(opt_args(kind, k + 1, **kw))
        np.testing.assert_allclose(bk.param[:n].cpu().numpy(), g[case][k], rtol=1e-5, atol=1e-7)
    bk.comm.destroy()


def test_pack_scales_and_places_tensors():
    lib = _lib.lib()
    torch.manual_seed(0)
    for dtype in (torch.float32, torch.bfloat16):
        sizes = [1, 7, 8, 9, 1000, 4097, 64 * 3 * 3 * 3]
        srcs = [torch.randn(s, device="cuda").to(dtype) for s in sizes]
        srcs[3] = torch.randn(32, device="cuda").to(dtype)[1:10]  # misaligned view: scalar path
        offs, cur = [], 0
        for s in sizes:
            offs.append(cur)
            cur += (s + 7) // 8 * 8
        dst = torch.full((cur,), 123.0, device="cuda").to(dtype)
        n = len(sizes)
        _lib.check(lib.b200dp_bucket_pack((ctypes.c_void_p * n)(*[t.data_ptr() for t in srcs]),
                                          (ctypes.c_int64 * n)(*sizes), (ctypes.c_int64 * n)(*offs), n,
                                          dst.data_ptr(), _lib.dtype_code(dtype), 0.5, _lib.current_stream()))
        for t, o, s in zip(srcs, offs, sizes):
            want = (t.float() * 0.5).to(dtype)
            assert torch.equal(dst[o:o + s], want)
        assert float(dst[offs[1] + 7]) == 123.0  # gaps are left alone


def test_average_only_and_bf16_master():
    # OPT_NONE + write_grads on one rank is the identity on the bucket
    bk = Bucket(4096)
    bk.grad.normal_()
    before = bk.grad.clone()
    bk.step(opt_args(_lib.OPT_NONE, 1), write_grads=1)
    assert torch.equal(bk.grad, before)
    bk.comm.destroy()
    # bf16 parameters with fp32 master weights: the master follows the fp32 oracle to 1e-5,
    # the bf16 copy is its rounding
    n = 8192
    bk = Bucket(n, torch.bfloat16)
    p0 = torch.randn(n, device="cuda")
    bk.master = p0.clone()
    bk.param.copy_(p0)
    p, buf = p0.cpu().numpy(), None
    for k in range(3):
        gr = torch.randn(n, device="cuda").bfloat16()
        bk.grad.copy_(gr)
        bk.step(opt_args(_lib.OPT_SGD, k + 1, lr=0.1, momentum=0.9))
        p, buf = optim_ref.sgd_step(p, gr.float().cpu().numpy(), buf, 0.1, 0.9, first=(k == 0))
    np.testing.assert_allclose(bk.master.cpu().numpy(), p, rtol=1e-5, atol=1e-6)
    assert torch.equal(bk.param, bk.master.bfloat16())
    bk.comm.destroy()


def test_argument_contract():
    lib = _lib.lib()
    bk = Bucket(64)
    o = opt_args(_lib.OPT_SGD, 1, lr=0.1)
    h, s = bk.comm.handle, _lib.current_stream()
    assert lib.b200dp_allreduce_update(h, 0, 0, bk.param_off, 0, 0, 60, None, None, None, ctypes.byref(o), 0, 0, 0, 1, s) == -1  # numel % 8
    assert lib.b200dp_allreduce_update(h, 0, 0, bk.param_off, 1, 0, 64, None, None, None, ctypes.byref(o), 0, 0, 0, 1, s) == -2  # mixed dtypes
    assert lib.b200dp_allreduce_update(h, 0, 0, bk.param_off, 0, 0, 1 << 30, None, None, None, ctypes.byref(o), 0, 0, 0, 1, s) == -1  # beyond heap
    bk.comm.destroy()
