"""Oracle (TEST INFRASTRUCTURE ONLY): Pillow's bilinear resample and torchvision's
RandomResizedCrop parameter draw, restated in numpy / plain Python.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import
this file; the product path (dbx_distributed_pytorch_examples_b200) never does.

What it follows.  The reference resizes PIL images on the CPU:
``transforms.Resize((S, S))`` (utils/hf_dataset_utilities.py:64-66) and
``transforms.RandomResizedCrop(224)`` (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53).
For a PIL image torchvision calls ``img.crop(...)`` then ``img.resize(size, BILINEAR)``
(TV/transforms/functional.py ``resized_crop`` -> ``crop`` + ``resize`` ->
TV/transforms/_functional_pil.py ``resize``), i.e. Pillow's two-pass fixed-point
resampler (Pillow 12.2.0, src/libImaging/Resample.c: ``precompute_coeffs``,
``normalize_coeffs_8bpc``, ``ImagingResampleHorizontal_8bpc``,
``ImagingResampleVertical_8bpc``, ``ImagingResampleInner``).  Pillow's C source is not
under /root/reference; the algorithm is restated from its published source and pinned
against the installed Pillow itself (tests/golden/resample.npz, oracle/make_golden.py).

The parameter draw follows ``RandomResizedCrop.get_params``
(TV/transforms/transforms.py:929-970) and ``RandomHorizontalFlip.forward`` (:720).
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2  # Resample.c: coefficients are fixed point with 22 fractional bits


def precompute_coeffs(in_size: int, in0: float, in1: float, out_size: int):
    """(ksize, bounds int32 [out,2] = (xmin, xmax), kk int32 [out, ksize]) for the bilinear filter.

    Resample.c ``precompute_coeffs`` followed by ``normalize_coeffs_8bpc``.  ``in0``/``in1``
    are C floats in Pillow (the ``box``); every other quantity is a C double.
    """
    in0 = float(np.float32(in0))
    in1 = float(np.float32(in1))
    scale = float(np.float32(in1) - np.float32(in0)) / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale  # bilinear: support 1.0
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = in0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)  # C cast: truncation toward zero
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = np.zeros(ksize, dtype=np.float64)
        ww = 0.0
        for x in range(xmax):
            a = (x + xmin - center + 0.5) * ss
            if a < 0.0:
                a = -a
            w[x] = 1.0 - a if a < 1.0 else 0.0
            ww += w[x]
        for x in range(xmax):
            if ww != 0.0:
                w[x] /= ww
        for x in range(ksize):
            v = w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx, 0], bounds[xx, 1] = xmin, xmax
    return ksize, bounds, kk


def _clip8(acc: np.ndarray) -> np.ndarray:
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def _pass_1d(img: np.ndarray, bounds, kk, out_size: int) -> np.ndarray:
    """Resample axis 0 of ``img`` ([n, ...] uint8): out[o] = clip8(2^21 + sum_x img[xmin+x] * k[o, x])."""
    out = np.empty((out_size,) + img.shape[1:], dtype=np.uint8)
    wide = img.astype(np.int64)
    for o in range(out_size):
        xmin, xmax = int(bounds[o, 0]), int(bounds[o, 1])
        acc = np.full(img.shape[1:], 1 << (PRECISION_BITS - 1), dtype=np.int64)
        for x in range(xmax):
            acc += wide[xmin + x] * int(kk[o, x])
        out[o] = _clip8(acc)
    return out


def resize_bilinear_u8(img: np.ndarray, out_hw) -> np.ndarray:
    """uint8 [h, w, C] -> uint8 [H, W, C], as ``PIL.Image.resize((W, H), BILINEAR)`` on the whole image.

    ImagingResampleInner: horizontal pass over the rows the vertical pass will touch
    (into an 8-bit temporary), then the vertical pass; a pass whose size does not change
    is skipped, and an unchanged size is a copy (Image.resize, ``self.copy()``).
    """
    h, w = img.shape[:2]
    H, W = out_hw
    if (h, w) == (H, W):
        return img.copy()
    _, bh, kh = precompute_coeffs(w, 0.0, float(w), W)
    _, bv, kv = precompute_coeffs(h, 0.0, float(h), H)
    cur = img
    if W != w:
        first = int(bv[0, 0])
        last = int(bv[H - 1, 0] + bv[H - 1, 1])
        rows = cur[first:last]
        cur = np.swapaxes(_pass_1d(np.swapaxes(rows, 0, 1), bh, kh, W), 0, 1)
        bv = bv.copy()
        bv[:, 0] -= first
    if H != h:
        cur = _pass_1d(cur, bv, kv, H)
    return np.ascontiguousarray(cur)


def resized_crop_flip_normalize(src: np.ndarray, params: np.ndarray, out_hw, mean, std) -> np.ndarray:
    """float32 [B, H, W, 3] from uint8 [B, Hs, Ws, 3] and int32 params [B, 5] = (i, j, h, w, flip).

    Per sample: crop rows i..i+h, columns j..j+w; Pillow bilinear resize to (H, W);
    horizontal flip if flip != 0; ``.div(255)`` (TV/transforms/functional.py:174-176);
    ``sub_(mean).div_(std)`` in float32 (TV/transforms/_functional_tensor.py:921-928).
    """
    B = src.shape[0]
    H, W = out_hw
    mean32 = np.asarray(mean, dtype=np.float32)
    std32 = np.asarray(std, dtype=np.float32)
    out = np.empty((B, H, W, 3), dtype=np.float32)
    for b in range(B):
        i, j, h, w, flip = (int(v) for v in params[b])
        tile = resize_bilinear_u8(src[b, i:i + h, j:j + w, :], (H, W))
        if flip:
            tile = tile[:, ::-1, :]
        out[b] = (tile.astype(np.float32) / np.float32(255.0) - mean32) / std32
    return out
