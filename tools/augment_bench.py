#!/usr/bin/env python
"""Micro-benchmark of b200dp_augment_u8_bf16: device source, pinned-host source (copy, then kernel) and
zero-copy (kernel reads host memory over PCIe).  Algorithmic bytes = 3 per output element."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dbx_distributed_pytorch_examples_b200 as b200  # noqa: E402
from dbx_distributed_pytorch_examples_b200 import transform as T  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--once", action="store_true", help="single launch per config (for ncu)")
    a = ap.parse_args()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (B, src, out, pad, name) in [(256, 256, 224, 0, "config4 256x256->224"), (128, 32, 32, 4, "cifar 32 pad4"),
                                     (256, 64, 64, 0, "tinyimagenet 64"), (1024, 256, 224, 0, "4x config4")]:
        x = torch.randint(0, 256, (B, src, src, 3), dtype=torch.uint8)
        xp = x.pin_memory()
        xd = x.to(dev)
        params = T.draw_crop_flip_params(B, (src, src), (out, out), pad, generator=torch.Generator().manual_seed(0))
        nbytes = B * out * out * 3 * 3
        for label, tf, inp in (("device", b200.BatchTransform(out, padding=pad), xd),
                               ("pinned+copy", b200.BatchTransform(out, padding=pad), xp),
                               ("zero-copy", b200.BatchTransform(out, padding=pad, zero_copy=True), xp)):
            if a.once:
                tf(inp, params=params); torch.cuda.synchronize()
                continue
            for _ in range(3):
                tf(inp, params=params)
            ts = []
            for _ in range(a.iters):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); tf(inp, params=params); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            ms = ts[len(ts) // 2]
            print(f"{name:24s} {label:12s} {ms*1e3:8.1f} us  {nbytes/ms/1e6:7.0f} GB/s algorithmic ({nbytes/ms/1e6/peak*100:4.1f}% of {peak:.0f})"
                  f"  h2d {(x.numel()/ms/1e6) if label != 'device' else 0:6.1f} GB/s")


if __name__ == "__main__":
    main()
