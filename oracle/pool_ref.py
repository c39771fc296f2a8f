"""Oracle (TEST INFRASTRUCTURE ONLY): torch's max_pool2d forward / backward restated in numpy.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this file.

Follows ATen's max_pool2d_with_indices (the op behind nn.MaxPool2d, which the reference model
applies after the stem's BN + ReLU, setup/resnet18.py:37,58): for every output position the
window is scanned rows first, then columns, starting from -inf at the first in-range
position, and a value replaces the running maximum when ``val > max or isnan(val)`` — the
FIRST maximum wins a tie.  The backward adds each output gradient to the input position that
won (max_pool2d_with_indices_backward).  Pinned by tests/golden/maxpool.npz, made with torch
itself on inputs full of ties (oracle/make_golden.py).
"""
import numpy as np


def maxpool_forward(x: np.ndarray, k: int, stride: int, pad: int):
    """x [N, H, W, C] -> (y [N, Ho, Wo, C], tap uint8 [N, Ho, Wo, C]); tap = row * k + column of the winner
    relative to the unclipped window."""
    N, H, W, C = x.shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    y = np.empty((N, Ho, Wo, C), dtype=x.dtype)
    tap = np.empty((N, Ho, Wo, C), dtype=np.uint8)
    for oh in range(Ho):
        for ow in range(Wo):
            h0, w0 = oh * stride - pad, ow * stride - pad
            m = np.full((N, C), -np.inf, dtype=np.float64)
            first = (max(-h0, 0)) * k + max(-w0, 0)
            best = np.full((N, C), first, dtype=np.uint8)
            for i in range(k):
                h = h0 + i
                if h < 0 or h >= H:
                    continue
                for j in range(k):
                    w = w0 + j
                    if w < 0 or w >= W:
                        continue
                    v = x[:, h, w, :].astype(np.float64)
                    take = (v > m) | np.isnan(v)
                    m = np.where(take, v, m)
                    best = np.where(take, np.uint8(i * k + j), best)
            y[:, oh, ow, :] = m.astype(x.dtype)
            tap[:, oh, ow, :] = best
    return y, tap


def maxpool_backward(dy: np.ndarray, tap: np.ndarray, in_hw, k: int, stride: int, pad: int) -> np.ndarray:
    """dx [N, H, W, C] float32: every dy added at the position its window's winner came from."""
    N, Ho, Wo, C = dy.shape
    H, W = in_hw
    dx = np.zeros((N, H, W, C), dtype=np.float32)
    n_idx, c_idx = np.meshgrid(np.arange(N), np.arange(C), indexing="ij")
    for oh in range(Ho):
        for ow in range(Wo):
            t = tap[:, oh, ow, :].astype(np.int64)
            h = oh * stride - pad + t // k
            w = ow * stride - pad + t % k
            np.add.at(dx, (n_idx, h, w, c_idx), dy[:, oh, ow, :].astype(np.float32))
    return dx
