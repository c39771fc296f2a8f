"""Bit-exact replica of torch's CPU generator stream (at::mt19937), vectorised with numpy.

torchvision draws crop offsets and flip decisions one sample at a time from the CPU
generator (``torch.randint(0, n, (1,))`` twice, then ``torch.rand(1)``), and a
vectorised torch draw does NOT give the same stream.  Doing the three tiny torch calls per
sample costs milliseconds per batch on the host, so this module reads the generator's
state (``Generator.get_state()``: the mt19937 words plus the ``left`` / ``next`` cursors),
produces the same 32-bit outputs in bulk, applies torch's own scalar transformations

* ``randint(0, n)``  ->  ``random() % n``                      (ATen uniform_int_from_to, n < 2**32)
* ``rand(1)`` float32 ->  ``(random() & (2**24 - 1)) * 2**-24``  (ATen uniform_real<float>)

and writes the advanced state back, so that the generator continues exactly where the
per-sample loop would have left it.  Checked against the loop in tests/test_host_cpu.py.
"""
from __future__ import annotations

import math

import numpy as np
import torch

_N, _M = 624, 397
_STATE_BYTES = 5056           # sizeof(at::CPUGeneratorImplState)
_OFF_LEFT, _OFF_NEXT, _OFF_STATE = 8, 16, 24
_UPPER, _LOWER, _MATRIX_A = np.uint32(0x80000000), np.uint32(0x7FFFFFFF), np.uint32(0x9908B0DF)


def _twist(s: np.ndarray) -> None:
    """at::mt19937::next_state(), in place on the 624 uint32 words."""
    def tw(u, v):
        y = (u & _UPPER) | (v & _LOWER)
        return (y >> np.uint32(1)) ^ np.where(v & np.uint32(1), _MATRIX_A, np.uint32(0))
    k = _N - _M  # 227
    s[0:k] = s[_M:_N] ^ tw(s[0:k], s[1:k + 1])
    # the remaining words depend on freshly written ones at distance 227: do them in strides
    j = k
    while j < _N - 1:
        e = min(j + k, _N - 1)
        s[j:e] = s[j - k:e - k] ^ tw(s[j:e], s[j + 1:e + 1])
        j = e
    s[_N - 1] = s[_M - 1] ^ tw(s[_N - 1:_N], s[0:1])[0]


def _temper(y: np.ndarray) -> np.ndarray:
    y = y ^ (y >> np.uint32(11))
    y = y ^ ((y << np.uint32(7)) & np.uint32(0x9D2C5680))
    y = y ^ ((y << np.uint32(15)) & np.uint32(0xEFC60000))
    return y ^ (y >> np.uint32(18))


def take_uint32(generator: torch.Generator | None, count: int) -> np.ndarray:
    """The next ``count`` outputs of ``generator->random()``; the generator is advanced."""
    gen = generator if generator is not None else torch.default_generator
    if gen.device.type != "cpu":
        raise ValueError("the crop/flip parameters are drawn from a CPU generator, as torchvision does")
    raw = gen.get_state()
    if raw.numel() != _STATE_BYTES:
        raise RuntimeError(f"unexpected CPU generator state size {raw.numel()}")
    buf = raw.numpy().copy()
    left = int(buf[_OFF_LEFT:_OFF_LEFT + 4].view(np.int32)[0])
    nxt = int(buf[_OFF_NEXT:_OFF_NEXT + 8].view(np.uint64)[0])
    words64 = buf[_OFF_STATE:_OFF_STATE + 8 * _N].view(np.uint64)
    s = words64.astype(np.uint32)
    out = np.empty(count, dtype=np.uint32)
    got = 0
    while got < count:
        if left <= 1:  # the next call twists first, then returns word 0
            _twist(s)
            left, nxt = _N, 0
            out[got] = s[0]
            got, nxt = got + 1, 1
            continue
        m = min(count - got, left - 1)
        out[got:got + m] = s[nxt:nxt + m]
        got, nxt, left = got + m, nxt + m, left - m
    buf[_OFF_LEFT:_OFF_LEFT + 4].view(np.int32)[0] = left
    buf[_OFF_NEXT:_OFF_NEXT + 8].view(np.uint64)[0] = nxt
    words64[:] = s.astype(np.uint64)
    gen.set_state(torch.from_numpy(buf))
    return _temper(out)


def crop_flip_draws(n: int, range_i: int, range_j: int, flip_p: float, generator=None) -> np.ndarray:
    """int32 [n, 3] = (i, j, flip) as n sequential torchvision draws would give them.
    ``range_i == range_j == 1`` means RandomCrop.get_params draws nothing (equal sizes)."""
    out = np.zeros((n, 3), dtype=np.int32)
    if range_i == 1 and range_j == 1:
        r = take_uint32(generator, n)
        u = r
    else:
        r = take_uint32(generator, 3 * n).reshape(n, 3)
        out[:, 0] = (r[:, 0].astype(np.uint64) % np.uint64(range_i)).astype(np.int32)
        out[:, 1] = (r[:, 1].astype(np.uint64) % np.uint64(range_j)).astype(np.int32)
        u = r[:, 2]
    uniform = (u & np.uint32(0xFFFFFF)).astype(np.float32) * np.float32(2.0 ** -24)
    out[:, 2] = (uniform < np.float32(flip_p)).astype(np.int32)
    return out


def _peek_uint32(generator, count: int):
    """(saved state, next ``count`` outputs); the caller rewinds with ``set_state(saved)``."""
    gen = generator if generator is not None else torch.default_generator
    saved = gen.get_state()
    return gen, saved, take_uint32(gen, count)


def resized_crop_flip_draws(n: int, height, width, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0),
                            flip_p: float = 0.5, generator=None) -> np.ndarray:
    """int32 [n, 5] = (i, j, h, w, flip): n sequential ``RandomResizedCrop.get_params`` +
    ``RandomHorizontalFlip`` draws (TV/transforms/transforms.py:929-970, :720) from one stream.

    Per try torchvision consumes one word for ``empty(1).uniform_(scale)`` and one for
    ``exp(empty(1).uniform_(log ratio))``; an accepted try consumes two more (randint i, j);
    the flip consumes one.  ATen's float uniform is ``(word & (2**24-1)) * 2**-24`` in double,
    then ``* (to - from) + from`` (to/from rounded to float32), rounded to float32.  The number
    of words used depends on the data, so a block is read ahead, walked, and the generator
    is then advanced by exactly what the walk consumed.  ``height`` / ``width`` are the image size:
    one pair for the whole batch, or one value per sample (ragged images in a common slab).
    """
    heights = [int(height)] * n if np.ndim(height) == 0 else [int(v) for v in height]
    widths = [int(width)] * n if np.ndim(width) == 0 else [int(v) for v in width]
    if len(heights) != n or len(widths) != n:
        raise ValueError("per-sample heights / widths must have one entry per sample")
    log_ratio = torch.log(torch.tensor(ratio))  # float32, as torchvision builds it
    s0, s1 = np.float32(scale[0]), np.float32(scale[1])
    r0, r1 = np.float32(float(log_ratio[0])), np.float32(float(log_ratio[1]))
    out = np.zeros((n, 5), dtype=np.int32)
    block = 23 * n  # worst case per sample: 10 rejected tries (2 words each) + i, j + flip
    while True:
        gen, saved, raw = _peek_uint32(generator, block)
        u = (raw & np.uint32(0xFFFFFF)).astype(np.float64) * (2.0 ** -24)
        as_scale = (u * np.float64(s1 - s0) + np.float64(s0)).astype(np.float32).astype(np.float64).tolist()
        as_logr = (u * np.float64(r1 - r0) + np.float64(r0)).astype(np.float32)
        as_ratio = torch.exp(torch.from_numpy(as_logr)).numpy().astype(np.float64).tolist()  # torch's own float32 exp
        as_flip = (u.astype(np.float32) < np.float32(flip_p)).tolist()
        raw64 = raw.tolist()  # python ints: the walk below is scalar
        p, ok = 0, True
        for k in range(n):
            if p + 23 > block:  # worst case: 10 tries + flip
                ok = False
                break
            height, width = heights[k], widths[k]
            area = height * width
            done = False
            for _ in range(10):
                target_area = area * as_scale[p]
                aspect = as_ratio[p + 1]
                p += 2
                w = int(round(math.sqrt(target_area * aspect)))
                h = int(round(math.sqrt(target_area / aspect)))
                if 0 < w <= width and 0 < h <= height:
                    i = raw64[p] % (height - h + 1)
                    j = raw64[p + 1] % (width - w + 1)
                    p += 2
                    done = True
                    break
            if not done:  # central crop fallback
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
            out[k] = (i, j, h, w, int(as_flip[p]))
            p += 1
        gen.set_state(saved)
        if ok:
            take_uint32(gen, p)
            return out
        block *= 3
