"""GPU: NHWC max pooling (b200dp_maxpool_fwd / _bwd) through the C ABI against the torch golden vectors,
the oracle, and torch's own CUDA max_pool2d on the same device.  Forward values are inputs, so bit-exact;
the winning tap follows torch's first-maximum rule (checked through the gradient routing on tie-heavy inputs)."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import dbx_distributed_pytorch_examples_b200 as b200
from dbx_distributed_pytorch_examples_b200 import _lib
from oracle import pool_ref

pytestmark = pytest.mark.gpu


def nhwc(t):
    return t.detach().permute(0, 2, 3, 1).contiguous().float().cpu().numpy()


@pytest.mark.parametrize("tag", ["stem", "odd", "k2", "k3p0"])
def test_golden_torch(golden, tag):
    g = golden("maxpool.npz")
    k, s, p = (int(v) for v in g[f"{tag}_geom"])
    x = torch.from_numpy(g[f"{tag}_x"]).permute(0, 3, 1, 2).cuda().requires_grad_(True)
    pool = b200.B200MaxPool2d(k, s, p)
    y = pool(x)
    assert y.is_contiguous(memory_format=torch.channels_last)
    y.backward(torch.from_numpy(g[f"{tag}_dy"]).permute(0, 3, 1, 2).cuda())
    assert np.array_equal(nhwc(y), g[f"{tag}_y"])
    np.testing.assert_allclose(nhwc(x.grad), g[f"{tag}_dx"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape,k,s,p", [((2, 64, 16, 16), 3, 2, 1), ((3, 8, 7, 9), 3, 2, 1), ((1, 16, 5, 5), 3, 1, 1),
                                         ((2, 8, 6, 6), 2, 2, 0), ((2, 8, 3, 3), 3, 3, 0), ((4, 24, 13, 10), 2, 1, 1),
                                         ((2, 8, 1, 1), 1, 1, 0)])
def test_against_oracle_and_torch_cuda(dtype, shape, k, s, p):
    torch.manual_seed(sum(shape) + k)
    x = torch.relu(torch.randn(shape, device="cuda")).mul(4).round().div(4).to(dtype)   # ties everywhere
    x = x.contiguous(memory_format=torch.channels_last)
    xo, xt = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    yo, yt = b200.B200MaxPool2d(k, s, p)(xo), F.max_pool2d(xt, k, s, p)
    dy = torch.randn_like(yt)
    yo.backward(dy); yt.backward(dy)
    assert torch.equal(yo, yt)
    assert torch.equal(xo.grad, xt.grad)        # same winner in every window, same (exact) sums
    y_ref, tap = pool_ref.maxpool_forward(nhwc(x), k, s, p)
    assert np.array_equal(nhwc(yo), y_ref)
    dx_ref = pool_ref.maxpool_backward(nhwc(dy), tap, shape[2:], k, s, p)
    np.testing.assert_allclose(nhwc(xo.grad), dx_ref, rtol=1e-2 if dtype == torch.bfloat16 else 1e-6, atol=1e-6)


def test_stem_size_properties_and_fusion():
    """ResNet-50 224^2 stem size at a reduced batch: equality with torch's kernel, gradient mass conserved;
    fuse_bn_relu swaps the module in and keeps the state_dict."""
    x = torch.relu(torch.randn(8, 64, 112, 112, device="cuda")).bfloat16().contiguous(memory_format=torch.channels_last)
    xo, xt = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    yo, yt = b200.B200MaxPool2d(3, 2, 1)(xo), F.max_pool2d(xt, 3, 2, 1)
    assert torch.equal(yo, yt)
    dy = torch.rand_like(yt)
    yo.backward(dy); yt.backward(dy)
    assert torch.equal(xo.grad, xt.grad)
    torch.testing.assert_close(xo.grad.float().sum(), dy.float().sum(), rtol=1e-2, atol=0)
    from dbx_distributed_pytorch_examples_b200 import workloads
    m = workloads.resnet50(10)
    keys = list(m.state_dict())
    m = b200.fuse_bn_relu(b200.convert_sync_batchnorm(m))
    assert isinstance(m.maxpool, b200.B200MaxPool2d) and list(m.state_dict()) == keys
    weird = nn.MaxPool2d(3, 2, 1, ceil_mode=True)
    assert not b200.pool.supported(weird)
    with pytest.raises(NotImplementedError):
        b200.B200MaxPool2d(5, 2, 2)


def test_argument_errors():
    lib = _lib.lib()
    x = torch.zeros((1, 8, 8, 8), device="cuda")
    y = torch.zeros((1, 4, 4, 8), device="cuda")
    i = torch.zeros((1, 4, 4, 8), dtype=torch.uint8, device="cuda")
    f = lib.b200dp_maxpool_fwd
    assert f(x.data_ptr(), y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 8, 3, 2, 1, None) == 0
    assert f(x.data_ptr(), y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 8, 5, 2, 1, None) == -2   # window > 3
    assert f(x.data_ptr(), y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 6, 3, 2, 1, None) == -2   # C % 4
    assert f(x.data_ptr(), y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 8, 3, 2, 2, None) == -1   # pad > k/2
    assert f(None, y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 8, 3, 2, 1, None) == -1
    assert f(x.data_ptr() + 4, y.data_ptr(), i.data_ptr(), 0, 1, 8, 8, 8, 3, 2, 1, None) == -3
    assert f(x.data_ptr(), y.data_ptr(), i.data_ptr(), 0, 0, 8, 8, 8, 3, 2, 1, None) == 0    # empty batch
    # channel counts the kernel does not take are refused (no second implementation behind the module)
    odd = torch.randn(2, 3, 8, 8, device="cuda")
    with pytest.raises(ValueError, match="Keep nn.MaxPool2d"):
        b200.B200MaxPool2d(3, 2, 1)(odd)
    # NaN propagates like torch's
    z = torch.randn(1, 8, 6, 6, device="cuda").contiguous(memory_format=torch.channels_last)
    z[0, 0, 2, 2] = float("nan")
    a, b = b200.B200MaxPool2d(3, 2, 1)(z), F.max_pool2d(z, 3, 2, 1)
    assert torch.equal(torch.isnan(a), torch.isnan(b)) and torch.equal(torch.nan_to_num(a), torch.nan_to_num(b))
    with pytest.raises(ValueError):
        b200.pool._MaxPool.apply(torch.zeros(1, 8, 4, 4), 3, 2, 1)
