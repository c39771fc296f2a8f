#!/usr/bin/env python
"""Run the four SyncBN kernels once per listed layer shape (for `ncu --set full` captures).

usage: python tools/bn_one.py C:HW[:res] [C:HW[:res] ...] [--batch 256] [--reps 2]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("layers", nargs="+")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--calls", action="store_true",
                    help="use the one-call entry points b200dp_bn_forward / _backward (small layers then run as ONE cluster kernel)")
    ap.add_argument("--mask-mode", type=int, default=3, choices=[2, 3])
    a = ap.parse_args()
    lib = _lib.lib()
    dev = torch.device("cuda")
    scratch = torch.zeros(int(lib.b200dp_bn_scratch_bytes(0)), dtype=torch.uint8, device=dev)
    s = _lib.current_stream()
    for spec in a.layers:
        parts = spec.split(":")
        C, HW, res = int(parts[0]), int(parts[1]), len(parts) > 2
        M = a.batch * HW * HW
        dt = torch.bfloat16
        x = torch.randn(M, C, device=dev).to(dt)
        dy = torch.randn(M, C, device=dev).to(dt)
        r = torch.randn(M, C, device=dev).to(dt) if res else None
        y, dx = torch.empty_like(x), torch.empty_like(x)
        dr = torch.empty_like(x) if res else None
        w, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        stats, save = torch.empty(2 * C + 1, device=dev), torch.empty(2 * C + 1, device=dev)
        sums, gw, gb = torch.empty(2 * C, device=dev), torch.empty(C, device=dev), torch.empty(C, device=dev)
        P = lambda t: 0 if t is None else t.data_ptr()  # noqa: E731
        mode = a.mask_mode if res else 1
        mk = torch.empty(M * C // 8, dtype=torch.int32, device=dev) if (res and mode == 3) else None
        for _ in range(a.reps if a.calls else 0):
            _lib.check(lib.b200dp_bn_forward(P(x), P(r), P(y), 1, M, C, P(scratch), P(stats), P(w), P(b), 1e-5, 0.1, None, None,
                                             P(save), 1, None, 0, s, P(mk), None))
            _lib.check(lib.b200dp_bn_backward(P(dy), P(x), P(mk if mk is not None else y), P(dx), P(dr), 1, M, C, P(save), P(w), P(b),
                                              mode, P(scratch), P(sums), P(gw), P(gb), None, 0, s))
        for _ in range(0 if a.calls else a.reps):
            _lib.check(lib.b200dp_bn_fwd_stats(P(x), 1, M, C, P(scratch), P(stats), None, 0, s))
            _lib.check(lib.b200dp_bn_fwd_apply(P(x), P(r), P(y), 1, M, C, P(stats), P(w), P(b), 1e-5, 0.1, None, None,
                                               P(save), P(save) + 4 * C, P(save) + 8 * C, 1, None, 0, s, P(mk)))
            _lib.check(lib.b200dp_bn_bwd_reduce(P(dy), P(x), P(mk if mk is not None else y), 1, M, C, P(save), P(save) + 4 * C, P(w), P(b), mode,
                                                P(scratch), P(sums), P(gw), P(gb), None, 0, s))
            _lib.check(lib.b200dp_bn_bwd_apply(P(dy), P(x), P(mk if mk is not None else y), P(dx), P(dr), 1, M, C, P(save), P(save) + 4 * C,
                                               P(w), P(b), P(sums), P(save) + 8 * C, mode, None, 0, s))
        torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
