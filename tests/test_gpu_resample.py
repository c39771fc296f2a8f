"""GPU: the resize transform kernel (b200dp_resized_crop_u8_bf16) through the C ABI against the
torchvision/Pillow golden vectors and the oracle.  Bar: the crop box and flip draws bit-exact,
the resampled uint8 pixels bit-exact with Pillow, the bf16 output bit-exact with
ToTensor + Normalize rounded to bf16."""
import ctypes

import numpy as np
import pytest
import torch

import dbx_distributed_pytorch_examples_b200 as b200
from dbx_distributed_pytorch_examples_b200 import _lib, transform as T
from oracle import resample_ref, transform_ref

pytestmark = pytest.mark.gpu


def bits(t):
    return t.permute(0, 2, 3, 1).contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


@pytest.mark.parametrize("name", ["resample_rrc56.npz", "resample_rrc224.npz", "resample_resize64.npz",
                                  "resample_down24.npz"])
def test_golden_torchvision(golden, name):
    g = golden(name)
    S = int(g["out_hw"])
    tf = b200.ResizeBatchTransform(S, random_resized_crop=bool(int(g["random_resized_crop"])),
                                   mean=g["mean"].tolist(), std=g["std"].tolist())
    torch.manual_seed(int(g["seed"]))  # the global generator state torchvision started from
    out = tf(torch.from_numpy(g["src"]).cuda())
    assert np.array_equal(tf.last_params.numpy(), g["params"])
    assert out.shape == (len(g["src"]), 3, S, S)
    assert out.dtype == torch.bfloat16 and out.is_contiguous(memory_format=torch.channels_last)
    assert np.array_equal(bits(out), g["out_bf16"])


def test_uint8_stage_is_pillows(golden):
    """mean 0 / std 1/255-free check of the integer stage: with mean = 0 and std = 1 the output is
    v/255, which identifies v uniquely in bf16 only up to rounding, so compare through the table."""
    g = golden("resample_rrc56.npz")
    S = int(g["out_hw"])
    tf = b200.ResizeBatchTransform(S, mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0))
    params = torch.from_numpy(g["params"].copy())
    params[:, 4] = 0  # no flip: resized_u8 in the fixture is the image before the flip
    out = tf(torch.from_numpy(g["src"]).cuda(), params=params)
    table = transform_ref.f32_to_bf16_bits(np.arange(256, dtype=np.float32) / np.float32(255.0))
    assert np.array_equal(bits(out), table[g["resized_u8"]])


@pytest.mark.parametrize("B,hs,ws,H,W", [
    (3, 64, 64, 64, 64),      # identity in both directions
    (4, 48, 80, 32, 32), (2, 32, 32, 64, 64), (3, 100, 37, 30, 30),   # W*3 % 8 != 0: scalar stores
    (2, 500, 375, 224, 224), (5, 17, 19, 8, 40), (1, 1, 1, 8, 8), (2, 900, 64, 16, 16),
    (70, 40, 40, 32, 32),
])
def test_against_oracle_random_boxes(B, hs, ws, H, W):
    rng = np.random.default_rng(B * 7 + hs)
    x = rng.integers(0, 256, size=(B, hs, ws, 3), dtype=np.uint8)
    params = np.zeros((B, 5), dtype=np.int32)
    for b in range(B):
        h, w = int(rng.integers(1, hs + 1)), int(rng.integers(1, ws + 1))
        if b == 0:
            h, w = hs, ws  # the whole image (plain Resize)
        params[b] = (rng.integers(0, hs - h + 1), rng.integers(0, ws - w + 1), h, w, rng.integers(0, 2))
    tf = b200.ResizeBatchTransform((H, W), mean=T.CIFAR_MEAN, std=T.CIFAR_STD)
    got = tf(torch.from_numpy(x).cuda(), params=torch.from_numpy(params))
    want = resample_ref.resized_crop_flip_normalize(x, params, (H, W), np.float32(T.CIFAR_MEAN), np.float32(T.CIFAR_STD))
    assert np.array_equal(bits(got), transform_ref.f32_to_bf16_bits(want))


def test_full_size_properties():
    """BASELINE config 4 sizes (256 images of 256x256 -> 224x224): flipping the box result equals the
    flipped output; a constant image stays constant; identical boxes give identical outputs."""
    B = 256
    x = torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device="cuda")
    x[1] = 77
    tf = b200.ResizeBatchTransform(224)
    p = T.draw_resized_crop_flip_params(B, (256, 256), generator=torch.Generator().manual_seed(5))
    p[2] = p[3]
    x[2] = x[3]
    a = tf(x, params=p)
    q = p.clone()
    q[:, 4] ^= 1
    b = tf(x, params=q)
    assert torch.equal(a, b.flip(3))
    assert torch.equal(a[2], a[3])
    const = a[1].float()
    assert all(float(const[c].min()) == float(const[c].max()) for c in range(3))
    # pinned host input takes the same path after one copy
    c = tf(x.cpu().pin_memory(), params=p)
    assert torch.equal(a, c)


def test_argument_errors():
    lib = _lib.lib()
    m = (ctypes.c_float * 3)(0.5, 0.5, 0.5)
    z = (ctypes.c_float * 3)(1.0, 0.0, 1.0)
    x = torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda")
    p = torch.zeros((1, 5), dtype=torch.int32, device="cuda")
    o = torch.empty((1, 8, 8, 3), dtype=torch.bfloat16, device="cuda")
    f = lib.b200dp_resized_crop_u8_bf16
    assert f(x.data_ptr(), -1, 8, 8, p.data_ptr(), 8, 8, m, m, o.data_ptr(), None) == -1
    assert f(None, 1, 8, 8, p.data_ptr(), 8, 8, m, m, o.data_ptr(), None) == -1
    assert f(x.data_ptr(), 1, 8, 8, p.data_ptr(), 8, 8, m, z, o.data_ptr(), None) == -1   # std == 0
    assert f(x.data_ptr(), 1, 8, 8, p.data_ptr(), 8, 8, m, m, o.data_ptr() + 2, None) == -3
    assert f(x.data_ptr(), 0, 8, 8, p.data_ptr(), 8, 8, m, m, o.data_ptr(), None) == 0    # empty batch
    assert f(x.data_ptr(), 1, 1 << 20, 8, p.data_ptr(), 1, 8, m, m, o.data_ptr(), None) == -2  # one row > smem
    with pytest.raises(ValueError):
        b200.ResizeBatchTransform(8)(torch.zeros((1, 8, 8, 3)))
    with pytest.raises(ValueError):
        b200.ResizeBatchTransform(8)(x, params=torch.zeros((1, 3), dtype=torch.int32))
    # a box outside the slab is forced inside instead of reading out of bounds
    p2 = torch.tensor([[5, 5, 100, 100, 0]], dtype=torch.int32)
    out = b200.ResizeBatchTransform(8, mean=(0, 0, 0), std=(1, 1, 1))(x, params=p2)
    assert torch.count_nonzero(out) == 0
