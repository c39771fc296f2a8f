"""GPU, one rank: fused BN(+residual)+ReLU forward/backward through the C ABI against
the numpy oracle and against torch's own F.batch_norm on the same device."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import dbx_distributed_pytorch_examples_b200 as b200
from oracle import syncbn_ref

pytestmark = pytest.mark.gpu

TOL = {torch.float32: dict(rtol=1e-4, atol=1e-5), torch.bfloat16: dict(rtol=2e-2, atol=2e-2)}


@pytest.fixture(autouse=True, params=["two_kernels", "one_kernel_cs16", "one_kernel_cs8"])
def bn_path(request):
    """Every test of this file runs through both code paths of b200dp_bn_forward / _backward: the statistics
    (reduce) kernel followed by the apply kernel, and the single cluster kernel that small layers take (with
    clusters of 16 CTAs and of 8).  The library's defaults come back afterwards."""
    from dbx_distributed_pytorch_examples_b200 import _lib
    lib = _lib.lib()
    if request.param == "two_kernels":
        lib.b200dp_bn_set_one_kernel_limits(0, 0, -1, -1)
    else:
        lib.b200dp_bn_set_one_kernel_limits(1 << 20, 1 << 20, 16 if request.param.endswith("16") else 8, 1)
    yield request.param
    lib.b200dp_bn_set_one_kernel_limits(56, 56, 16, 0)


def rows(t):
    C = t.shape[1]
    return t.detach().float().permute(0, 2, 3, 1).reshape(-1, C).cpu().numpy()


def run_ours(x, res, dy, w, b, rm, rv, relu, momentum=0.1, eps=1e-5):
    bn = b200.B200SyncBatchNorm(x.shape[1], eps=eps, momentum=momentum, fuse_relu=relu).cuda()
    with torch.no_grad():
        bn.weight.copy_(w); bn.bias.copy_(b); bn.running_mean.copy_(rm); bn.running_var.copy_(rv)
    x = x.clone().requires_grad_(True)
    res = None if res is None else res.clone().requires_grad_(True)
    y = bn(x, res)
    y.backward(dy)
    return bn, y, x.grad, None if res is None else res.grad


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(8, 64, 16, 16), (4, 256, 7, 7), (3, 72, 5, 5), (130, 64, 32, 32), (2, 2048, 3, 3)])
@pytest.mark.parametrize("mode", ["plain", "relu", "addrelu"])
def test_against_oracle(dtype, shape, mode):
    torch.manual_seed(sum(shape) + len(mode))
    N, C, H, W = shape
    relu, use_res = mode != "plain", mode == "addrelu"
    mk = lambda s=1.0, o=0.0: (torch.randn(shape, device="cuda") * s + o).to(dtype).contiguous(memory_format=torch.channels_last)  # noqa: E731
    x, res, dy = mk(2.0, 0.7), (mk() if use_res else None), mk()
    w, b = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    rm, rv = torch.randn(C, device="cuda") * 0.1, torch.rand(C, device="cuda") + 0.5
    bn, y, dx, dres = run_ours(x, res, dy, w, b, rm, rv, relu)
    ys, mean, invstd, _, rm_new, rv_new = syncbn_ref.forward(
        [rows(x)], w.cpu().numpy(), b.cpu().numpy(), rm.cpu().numpy(), rv.cpu().numpy(), 1e-5, 0.1, relu,
        [rows(res)] if use_res else None)
    tol = TOL[dtype]
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)
    np.testing.assert_allclose(rows(y), ys[0], **tol)
    np.testing.assert_allclose(bn.running_mean.cpu().numpy(), rm_new, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(bn.running_var.cpu().numpy(), rv_new, rtol=1e-4, atol=1e-6)
    assert int(bn.num_batches_tracked) == 1
    # The backward oracle takes the ReLU mask from ITS OWN y.  Where that y is within rounding of zero
    # the sign is not a property of the input, so those few elements are left out of the element-wise
    # comparison (they still enter the sums on both sides; the tolerance covers a flipped one).
    dxs, drs, gws, gbs = syncbn_ref.backward([rows(dy)], [rows(x)], [ys[0]], w.cpu().numpy(), mean, invstd, relu)
    sure = (np.abs(ys[0]) > (2.0 ** -7 if dtype == torch.bfloat16 else 1e-5) * np.maximum(1.0, np.abs(rows(x)))) | (not relu)
    assert sure.mean() > 0.4
    scale = max(1.0, float(np.abs(dxs[0]).max()))
    np.testing.assert_allclose(np.where(sure, rows(dx), 0), np.where(sure, dxs[0], 0), rtol=tol["rtol"], atol=tol["atol"] * scale)
    gtol = dict(rtol=2e-3, atol=2e-3 * N * H * W ** 0.5) if dtype == torch.float32 else dict(rtol=3e-2, atol=0.5)
    np.testing.assert_allclose(bn.weight.grad.cpu().numpy(), gws[0], **gtol)
    np.testing.assert_allclose(bn.bias.grad.cpu().numpy(), gbs[0], **gtol)
    if use_res:  # d_residual is the masked dy
        np.testing.assert_allclose(np.where(sure, rows(dres), 0), np.where(sure, drs[0], 0), **tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_against_torch_batch_norm(dtype):
    torch.manual_seed(0)
    shape = (32, 128, 14, 14)
    x = torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    C = shape[1]
    w, b = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    bn, y, dx, _ = run_ours(x, None, dy, w, b, rm, rv, True)
    xr = x.clone().float().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    rm2, rv2 = rm.clone(), rv.clone()
    yr = torch.relu(F.batch_norm(xr, rm2, rv2, wr, br, True, 0.1, 1e-5))
    yr.backward(dy.float())
    tol = TOL[dtype]
    torch.testing.assert_close(y.float(), yr, **tol)
    torch.testing.assert_close(bn.running_mean, rm2, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(bn.running_var, rv2, rtol=1e-4, atol=1e-6)
    if dtype == torch.float32:
        torch.testing.assert_close(dx, xr.grad, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(bn.weight.grad, wr.grad, rtol=1e-3, atol=1e-2)
        torch.testing.assert_close(bn.bias.grad, br.grad, rtol=1e-3, atol=1e-2)
    else:
        # bf16 against torch's OWN bf16 batch norm (cuDNN / ATen NHWC kernels, bf16 in and out) fed the same
        # bits: both round a float32 result to bf16 once, so they may differ by one bf16 ulp of the value
        xb = x.clone().requires_grad_(True)
        wb, bb = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
        yb = torch.relu(F.batch_norm(xb, rm.clone(), rv.clone(), wb, bb, True, 0.1, 1e-5))
        yb.backward(dy)
        torch.testing.assert_close(y, yb, rtol=2 ** -7, atol=2 ** -7)
        agree = (y > 0) == (yb > 0)        # an element whose ReLU fell on the other side has another gradient
        assert float(agree.float().mean()) > 0.999
        torch.testing.assert_close(torch.where(agree, dx, 0), torch.where(agree, xb.grad, 0), rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(bn.weight.grad, wb.grad.float(), rtol=2e-2, atol=0.5)
        torch.testing.assert_close(bn.bias.grad, bb.grad.float(), rtol=2e-2, atol=0.5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_extreme_statistics(dtype):
    """Channels a one-pass variance gets wrong: |mean| / std = 257 (in bf16 the values 256 and 258 are
    neighbours: nothing narrower can be stored around that mean), a constant channel (variance exactly 0:
    invstd = 1/sqrt(eps), y = bias, dx = 0) and a large constant one.  Against float64 on the same bits."""
    torch.manual_seed(11)
    N, C, H, W = 64, 64, 16, 16
    x = torch.randn(N, C, H, W, device="cuda")
    x[:, 0] = 256.0 + 2.0 * torch.randint(0, 2, (N, H, W), device="cuda")       # mean 257, std 1
    x[:, 1] = 3.0                                                              # constant
    x[:, 2] = -1024.0                                                          # constant, large
    x[:, 3] = 0.0
    x[:, 4] = 4096.0 + 32.0 * torch.randint(0, 2, (N, H, W), device="cuda")    # the same ratio, larger scale
    x = x.to(dtype).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, C, H, W, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)
    w, b = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    bn, y, dx, _ = run_ours(x, None, dy, w, b, torch.zeros(C, device="cuda"), torch.ones(C, device="cuda"), False)
    x64 = x.double()
    mean = x64.mean((0, 2, 3)); var = x64.var((0, 2, 3), unbiased=False)
    n = N * H * W
    torch.testing.assert_close(bn.running_mean.double(), 0.1 * mean, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(bn.running_var.double(), 0.9 + 0.1 * var * n / (n - 1), rtol=2e-3, atol=1e-6)
    assert float(bn.running_var[1]) == pytest.approx(0.9, abs=1e-6) and float(bn.running_var[2]) == pytest.approx(0.9, abs=1e-3)
    want = (x64 - mean.view(1, -1, 1, 1)) / torch.sqrt(var + 1e-5).view(1, -1, 1, 1) * w.double().view(1, -1, 1, 1) \
        + b.double().view(1, -1, 1, 1)
    tol = dict(rtol=1e-3, atol=2e-3) if dtype == torch.float32 else dict(rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(y.double()[:, [0, 4]], want[:, [0, 4]], **tol)   # the high-ratio channels
    torch.testing.assert_close(y.double()[:, 5:], want[:, 5:], **tol)
    for c in (1, 3):   # constant channels: y is the bias, the gradient vanishes (the sum term cancels dy's mean only)
        torch.testing.assert_close(y[:, c].float(), b[c].expand(N, H, W).to(dtype).float(), rtol=0, atol=1e-2)
    assert torch.isfinite(dx.float()).all() and torch.isfinite(y.float()).all()


def test_single_layer_gradient_error_against_float64():
    """The backward of ONE layer, ours and torch's float32 kernels, each against torch's float64 batch norm on
    the same (x, dy): the float32 implementations must be equally far from the truth.  (The whole-model
    version of this measurement is in test_gpu_ddp.py.)"""
    torch.manual_seed(12)
    for shape in [(128, 64, 32, 32), (128, 512, 4, 4), (32, 256, 14, 14)]:
        C = shape[1]
        x = (torch.randn(shape, device="cuda") * 1.5 + 0.3).contiguous(memory_format=torch.channels_last)
        dy = torch.randn(shape, device="cuda").contiguous(memory_format=torch.channels_last)
        w, b = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
        bn, y, dx, _ = run_ours(x, None, dy, w, b, torch.zeros(C, device="cuda"), torch.ones(C, device="cuda"), False)
        outs = {}
        for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
            xr = x.to(dt).requires_grad_(True)
            wr, br = w.to(dt).requires_grad_(True), b.to(dt).requires_grad_(True)
            yr = F.batch_norm(xr, None, None, wr, br, True, 0.1, 1e-5)
            yr.backward(dy.to(dt))
            outs[name] = (yr.detach().double(), xr.grad.double(), wr.grad.double(), br.grad.double())
        ours = (y.double(), dx.double(), bn.weight.grad.double(), bn.bias.grad.double())
        for k, what in enumerate(("y", "dx", "dweight", "dbias")):
            ref = outs["f64"][k]
            e_ours = float((ours[k] - ref).norm() / ref.norm())
            e_torch = float((outs["f32"][k] - ref).norm() / ref.norm())
            assert e_ours <= 2.0 * e_torch + 2e-7, (shape, what, e_ours, e_torch)


def test_mask_word_mode_equals_mask_from_y(monkeypatch):
    """Opt-in backward for residual layers that reads the forward's ReLU bit words instead of y: same bits out."""
    from dbx_distributed_pytorch_examples_b200 import syncbn
    torch.manual_seed(3)
    shape = (6, 128, 9, 9)
    mk = lambda: torch.randn(shape, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)  # noqa: E731
    x, res, dy = mk(), mk(), mk()
    w, b = torch.rand(128, device="cuda") + 0.5, torch.randn(128, device="cuda")
    rm, rv = torch.zeros(128, device="cuda"), torch.ones(128, device="cuda")
    outs = []
    for flag in (False, True):
        monkeypatch.setattr(syncbn, "_USE_MASK_WORDS", flag)
        bn, y, dx, dres = run_ours(x, res, dy, w, b, rm, rv, True)
        outs.append((y, dx, dres, bn.weight.grad.clone(), bn.bias.grad.clone()))
    for a, c in zip(*outs):
        assert torch.equal(a, c)


def test_eval_mode_and_nchw_input_and_2d():
    torch.manual_seed(1)
    ref = torch.nn.BatchNorm2d(64).cuda()
    with torch.no_grad():
        ref.running_mean.normal_(); ref.running_var.uniform_(0.5, 2); ref.weight.uniform_(0.5, 1.5); ref.bias.normal_()
    ours = b200.convert_sync_batchnorm(torch.nn.Sequential(ref))[0]
    assert isinstance(ours, b200.B200SyncBatchNorm) and ours.weight is ref.weight
    x = torch.randn(4, 64, 9, 9, device="cuda")  # NCHW-contiguous input is accepted
    ours.eval(); ref.eval()
    torch.testing.assert_close(ours(x), ref(x), rtol=1e-5, atol=1e-5)
    ours.train(); ref.train()
    y, yr = ours(x), ref(x)
    torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-5)
    # [N, C] input (BatchNorm1d shape)
    bn1 = b200.B200SyncBatchNorm(32).cuda()
    z = torch.randn(50, 32, device="cuda")
    torch.testing.assert_close(bn1(z), F.batch_norm(z, None, None, bn1.weight, bn1.bias, True), rtol=1e-4, atol=1e-5)
    with pytest.raises(ValueError, match="more than 1 value per channel"):
        bn1(torch.randn(1, 32, device="cuda"))


def test_momentum_none_and_no_running_stats():
    x = torch.randn(16, 16, 4, 4, device="cuda")
    ours = b200.B200SyncBatchNorm(16, momentum=None).cuda()
    ref = torch.nn.BatchNorm2d(16, momentum=None).cuda()
    for _ in range(3):
        ours(x); ref(x)
    torch.testing.assert_close(ours.running_mean, ref.running_mean, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ours.running_var, ref.running_var, rtol=1e-4, atol=1e-6)
    free = b200.B200SyncBatchNorm(16, track_running_stats=False, affine=False).cuda()
    torch.testing.assert_close(free(x), F.batch_norm(x, None, None, None, None, True), rtol=1e-4, atol=1e-5)


def test_linearity_and_large_layer_property():
    """Size-independent checks on the largest BN layer of ResNet-50 224^2 at a reduced batch:
    output statistics are (beta, gamma^2) and dx sums to zero per channel."""
    torch.manual_seed(2)
    x = (torch.randn(16, 64, 112, 112, device="cuda") * 3 + 1).bfloat16().contiguous(memory_format=torch.channels_last)
    bn = b200.B200SyncBatchNorm(64).cuda()
    with torch.no_grad():
        bn.weight.fill_(2.0); bn.bias.fill_(0.5)
    xr = x.clone().requires_grad_(True)
    y = bn(xr)
    m = y.float().mean((0, 2, 3)); v = y.float().var((0, 2, 3), unbiased=False)
    torch.testing.assert_close(m, torch.full_like(m, 0.5), rtol=0, atol=2e-2)
    torch.testing.assert_close(v, torch.full_like(v, 4.0), rtol=2e-2, atol=0)
    y.backward(torch.randn_like(y))
    s = xr.grad.float().sum((0, 2, 3))
    assert float(s.abs().max()) < 0.05 * float(xr.grad.float().abs().sum((0, 2, 3)).max())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("mode", ["plain", "relu", "addrelu"])
def test_cpp_autograd_node_equals_python_node(monkeypatch, dtype, mode):
    """csrc/autograd_bn.cpp is host plumbing over the same two C-ABI calls: identical bits, identical
    side effects (running statistics, counters), identical errors."""
    from dbx_distributed_pytorch_examples_b200 import _lib
    assert _lib.native() is not None, "the C++ autograd node was not built"
    torch.manual_seed(5)
    shape = (5, 96, 7, 9)
    mk = lambda: torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)  # noqa: E731
    x, dy = mk(), mk()
    res = mk() if mode == "addrelu" else None
    w, b = torch.rand(96, device="cuda") + 0.5, torch.randn(96, device="cuda")
    rm, rv = torch.randn(96, device="cuda"), torch.rand(96, device="cuda") + 0.5
    outs = []
    for use_cpp in (True, False):
        if not use_cpp:
            monkeypatch.setattr(_lib, "_native", False)
        before = _lib.total_calls().get("b200dp_bn_forward", 0)
        bn, y, dx, dres = run_ours(x, res, dy, w, b, rm, rv, mode != "plain")
        assert _lib.total_calls().get("b200dp_bn_forward", 0) == before + 1
        outs.append((y, dx, dres, bn.weight.grad.clone(), bn.bias.grad.clone(), bn.running_mean.clone(),
                     bn.running_var.clone(), bn.num_batches_tracked.clone()))
    for a, c in zip(*outs):
        assert (a is None and c is None) or torch.equal(a, c)


def test_cpp_autograd_node_errors_and_frozen_inputs():
    bn = b200.B200SyncBatchNorm(4).cuda()
    with pytest.raises(ValueError, match="more than 1 value per channel"):
        bn(torch.randn(1, 4, 1, 1, device="cuda"))
    with pytest.raises(ValueError, match="residual must have the shape"):
        bn(torch.randn(2, 4, 3, 3, device="cuda"), torch.randn(2, 4, 3, 2, device="cuda"))
    # input without grad, affine frozen: the backward still runs and returns nothing for them
    bn.weight.requires_grad_(False)
    x = torch.randn(2, 4, 3, 3, device="cuda")
    r = torch.randn(2, 4, 3, 3, device="cuda", requires_grad=True)
    bn(x, r).sum().backward()
    assert bn.weight.grad is None and bn.bias.grad is not None and r.grad is not None
    # NCHW-contiguous and 2D inputs go through the same node
    y = bn(torch.randn(2, 4, 3, 3, device="cuda").contiguous())
    assert y.shape == (2, 4, 3, 3)
    assert b200.B200SyncBatchNorm(4).cuda()(torch.randn(6, 4, device="cuda")).shape == (6, 4)


def test_num_batches_tracked_is_counted_by_the_kernel(monkeypatch):
    """The forward kernel increments the layer's counter (no separate torch kernel); momentum=None,
    eval mode and the python node count like nn.BatchNorm2d does."""
    from dbx_distributed_pytorch_examples_b200 import _lib
    x = torch.randn(4, 16, 5, 5, device="cuda")
    bn = b200.B200SyncBatchNorm(16).cuda()
    ref = torch.nn.BatchNorm2d(16).cuda()
    for _ in range(3):
        bn(x); ref(x)
    assert int(bn.num_batches_tracked) == int(ref.num_batches_tracked) == 3
    bn.eval(); bn(x)
    assert int(bn.num_batches_tracked) == 3
    bn.train()
    with torch.no_grad():
        bn(x)
    assert int(bn.num_batches_tracked) == 4
    monkeypatch.setattr(_lib, "_native", False)   # the python autograd node passes the same pointer
    bn(x)
    assert int(bn.num_batches_tracked) == 5
    cum = b200.B200SyncBatchNorm(16, momentum=None).cuda()
    cum(x); cum(x)
    assert int(cum.num_batches_tracked) == 2
