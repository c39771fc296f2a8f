#!/usr/bin/env python
"""Run the four SyncBN entry points one by one on a small tensor, synchronising after each, and say which
one the first CUDA error came from.  usage: python tools/bn_steps.py {f32|bf16} RELU_MODE [M C]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib  # noqa: E402


def main():
    dt = torch.float32 if sys.argv[1] == "f32" else torch.bfloat16
    mode = int(sys.argv[2])
    M, C = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (2048, 64)
    lib = _lib.lib()
    code = _lib.dtype_code(dt)
    dev = torch.device("cuda")
    x, dy = torch.randn(M, C, device=dev).to(dt), torch.randn(M, C, device=dev).to(dt)
    y, dx = torch.empty_like(x), torch.empty_like(x)
    w, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    stats, save = torch.empty(2 * C + 1, device=dev), torch.empty(2 * C + 1, device=dev)
    sums, gw, gb = torch.empty(2 * C, device=dev), torch.empty(C, device=dev), torch.empty(C, device=dev)
    scratch = torch.zeros(int(lib.b200dp_bn_scratch_bytes(0)), dtype=torch.uint8, device=dev)
    s = _lib.current_stream()
    P = lambda t: t.data_ptr()  # noqa: E731
    steps = [
        ("fwd_stats", lambda: lib.b200dp_bn_fwd_stats(P(x), code, M, C, P(scratch), P(stats), None, 0, s)),
        ("fwd_apply", lambda: lib.b200dp_bn_fwd_apply(P(x), 0, P(y), code, M, C, P(stats), P(w), P(b), 1e-5, 0.1, None, None,
                                                      P(save), P(save) + 4 * C, P(save) + 8 * C, int(mode > 0), None, 0, s, None)),
        ("bwd_reduce", lambda: lib.b200dp_bn_bwd_reduce(P(dy), P(x), P(y), code, M, C, P(save), P(save) + 4 * C, P(w), P(b), mode,
                                                        P(scratch), P(sums), P(gw), P(gb), None, 0, s)),
        ("bwd_apply", lambda: lib.b200dp_bn_bwd_apply(P(dy), P(x), P(y), P(dx), 0, code, M, C, P(save), P(save) + 4 * C, P(w), P(b),
                                                      P(sums), P(save) + 8 * C, mode, None, 0, s)),
        ("forward(1 call)", lambda: lib.b200dp_bn_forward(P(x), 0, P(y), code, M, C, P(scratch), P(stats), P(w), P(b), 1e-5, 0.1, None,
                                                          None, P(save), int(mode > 0), None, 0, s, None, None)),
        ("backward(1 call)", lambda: lib.b200dp_bn_backward(P(dy), P(x), P(y), P(dx), 0, code, M, C, P(save), P(w), P(b), mode,
                                                            P(scratch), P(sums), P(gw), P(gb), None, 0, s)),
    ]
    for name, fn in steps:
        rc = fn()
        try:
            torch.cuda.synchronize()
            err = None
        except Exception as e:  # noqa: BLE001
            err = str(e).splitlines()[0]
        print(f"{sys.argv[1]} mode {mode} M {M} C {C}: {name:18s} rc={rc} {'ok' if err is None else err}", flush=True)
        if err is not None or rc:
            return


if __name__ == "__main__":
    main()
