#!/usr/bin/env python
"""Micro-benchmark of b200dp_resized_crop_u8_bf16 (the C-ABI call alone, CUDA events, L2 flushed between
launches).  Algorithmic bytes per image = h*w*3 read (the crop box) + H*W*3*2 written."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib, transform as T  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--once", action="store_true", help="single launch per config (for ncu)")
    a = ap.parse_args()
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
    dev = torch.device("cuda")
    lib = _lib.lib()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tf = T.ResizeBatchTransform(224)
    for (B, hs, ws, out, rrc, name) in [(256, 256, 256, 224, True, "rrc 256x256->224"),
                                        (256, 256, 256, 224, False, "resize 256x256->224"),
                                        (256, 500, 375, 224, True, "rrc 500x375->224"),
                                        (256, 64, 64, 64, True, "rrc 64x64->64"),
                                        (128, 32, 32, 64, False, "resize 32x32->64")]:
        x = torch.randint(0, 256, (B, hs, ws, 3), dtype=torch.uint8, device=dev)
        t = T.ResizeBatchTransform(out, random_resized_crop=rrc)
        p = t.draw(B, (hs, ws), torch.Generator().manual_seed(0))
        pd = p.to(dev)
        y = torch.empty((B, out, out, 3), dtype=torch.bfloat16, device=dev)
        nbytes = int((p[:, 2].long() * p[:, 3].long()).sum()) * 3 + B * out * out * 6
        s = _lib.current_stream()

        def call():
            _lib.check(lib.b200dp_resized_crop_u8_bf16(x.data_ptr(), B, hs, ws, pd.data_ptr(), out, out,
                                                       t._mean, t._std, y.data_ptr(), s))
        if a.once:
            call(); torch.cuda.synchronize()
            continue
        for _ in range(3):
            call()
        ts = []
        for _ in range(a.iters):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); call(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        ms = ts[len(ts) // 2]
        print(f"{name:22s} {ms*1e3:8.1f} us  {nbytes/1e6:7.1f} MB  {nbytes/ms/1e6:7.0f} GB/s algorithmic "
              f"({nbytes/ms/1e6/peak*100:4.1f}% of {peak:.0f})  {B/ms*1e3:9.0f} img/s")


if __name__ == "__main__":
    main()
