#!/usr/bin/env python
"""How fast can this GPU READ? torch reductions over a large tensor vs a device copy (the roofline denominator)."""
import torch

dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def t(fn, n=10):
    ts = []
    for _ in range(n):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[0], ts[len(ts) // 2]


for nbytes in (411 << 20, 1 << 30, 4 << 30):
    a = torch.randn(nbytes // 2, device=dev, dtype=torch.bfloat16)
    b = torch.empty_like(a)
    af = a.view(torch.float32)
    for name, fn, moved in (("copy", lambda: b.copy_(a), 2 * nbytes), ("sum bf16", lambda: a.sum(), nbytes),
                            ("sum f32 view", lambda: af.sum(), nbytes), ("amax f32 view", lambda: af.amax(), nbytes),
                            ("fill (write only)", lambda: b.fill_(1.0), nbytes)):
        best, med = t(fn)
        print(f"{nbytes >> 20:5d} MiB {name:18s} best {moved / best / 1e6:7.0f} GB/s  median {moved / med / 1e6:7.0f} GB/s")
