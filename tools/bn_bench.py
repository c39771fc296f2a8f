#!/usr/bin/env python
"""Per-layer micro-benchmark of the four SyncBN kernels on ResNet-50 224^2 shapes.

Each kernel is timed alone with CUDA events; a 256 MiB buffer is written between
iterations so that nothing is served from L2 (cold numbers), or not (--warm: the state the
kernels meet inside a training step, where the producer has just written the tensor).
Prints achieved GB/s of ALGORITHMIC bytes against MEASURED_PEAKS.json's copy bandwidth.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib  # noqa: E402

# (C, HW, count in ResNet-50, has_residual)
LAYERS = [
    (64, 112, 1, False), (64, 56, 6, False), (256, 56, 3, True), (256, 56, 1, False),
    (128, 56, 1, False), (128, 28, 7, False), (512, 28, 4, True), (512, 28, 1, False),
    (256, 28, 1, False), (256, 14, 11, False), (1024, 14, 6, True), (1024, 14, 1, False),
    (512, 14, 1, False), (512, 7, 5, False), (2048, 7, 3, True), (2048, 7, 1, False),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warm", action="store_true")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--mask-mode", type=int, default=2, choices=[2, 3],
                    help="backward of residual layers: 2 = ReLU mask from y (the default in training), 3 = from mask words")
    ap.add_argument("--one-kernel-mb", type=int, nargs=2, default=None, metavar=("FWD", "BWD"),
                    help="size limits of the one-kernel path for the '(1 call)' columns (0 0 = always two kernels)")
    ap.add_argument("--cluster", type=int, default=-1, help="largest cluster size of the one-kernel path (1..16)")
    ap.add_argument("--any-geometry", type=int, default=-1, help="1: every shape under the limits takes the one-kernel path")
    ap.add_argument("--chain", type=int, default=1,
                    help="time this many back-to-back calls per sample (amortises the ~5 us an idle stream adds to one launch; "
                         "closer to what a kernel costs inside a training step)")
    a = ap.parse_args()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    lib = _lib.lib()
    if a.one_kernel_mb is not None or a.cluster > 0 or a.any_geometry >= 0:
        f, b_ = a.one_kernel_mb if a.one_kernel_mb is not None else (-1, -1)
        lib.b200dp_bn_set_one_kernel_limits(f, b_, a.cluster, a.any_geometry)
    print(f"# one-kernel limits: {a.one_kernel_mb} cluster {a.cluster} mask mode {a.mask_mode} {'warm' if a.warm else 'cold'}")
    dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
    code, es = _lib.dtype_code(dt), (2 if dt == torch.bfloat16 else 4)
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    scratch = torch.zeros(int(lib.b200dp_bn_scratch_bytes(0)), dtype=torch.uint8, device=dev)
    s = _lib.current_stream()
    totals = {}
    print(f"{'C':>5} {'HW':>4} {'n':>2} {'res':>3} | " + " | ".join(f"{k:>22}" for k in ("stats", "fwd_apply", "bwd_reduce", "bwd_apply", "fwd(1 call)", "bwd(1 call)")))
    for C, HW, count, res in LAYERS:
        M = a.batch * HW * HW
        x = torch.randn(M, C, device=dev).to(dt)
        dy = torch.randn(M, C, device=dev).to(dt)
        r = torch.randn(M, C, device=dev).to(dt) if res else None
        y, dx = torch.empty_like(x), torch.empty_like(x)
        dr = torch.empty_like(x) if res else None
        w, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        stats = torch.empty(2 * C + 1, device=dev)
        save = torch.empty(2 * C + 1, device=dev)
        sums, gw, gb = torch.empty(2 * C, device=dev), torch.empty(C, device=dev), torch.empty(C, device=dev)
        mode = a.mask_mode if res else 1
        mk = torch.empty(M * C // 8, dtype=torch.int32, device=dev) if (res and mode == 3) else None
        mfrac = {2: 1.0, 3: 0.25}[a.mask_mode]
        P = lambda t: 0 if t is None else t.data_ptr()  # noqa: E731
        calls = {
            "stats": (lambda: lib.b200dp_bn_fwd_stats(P(x), code, M, C, P(scratch), P(stats), None, 0, s), M * C * es),
            "fwd_apply": (lambda: lib.b200dp_bn_fwd_apply(P(x), P(r), P(y), code, M, C, P(stats), P(w), P(b), 1e-5, 0.1,
                                                         None, None, P(save), P(save) + 4 * C, P(save) + 8 * C, 1, None, 0, s, P(mk)),
                          M * C * es * (3 + (0.25 if mk is not None else 0) if res else 2)),
            "bwd_reduce": (lambda: lib.b200dp_bn_bwd_reduce(P(dy), P(x), P(mk if mk is not None else y), code, M, C, P(save), P(save) + 4 * C, P(w), P(b),
                                                           mode, P(scratch), P(sums), P(gw), P(gb), None, 0, s),
                           M * C * es * (2 + mfrac if res else 2)),
            "bwd_apply": (lambda: lib.b200dp_bn_bwd_apply(P(dy), P(x), P(mk if mk is not None else y), P(dx), P(dr), code, M, C, P(save), P(save) + 4 * C,
                                                         P(w), P(b), P(sums), P(save) + 8 * C, mode, None, 0, s),
                          M * C * es * (4 + mfrac if res else 3)),
        }
        nb = lambda k: calls[k][1]  # noqa: E731
        calls["fwd(1 call)"] = (lambda: lib.b200dp_bn_forward(P(x), P(r), P(y), code, M, C, P(scratch), P(stats), P(w), P(b),
                                                             1e-5, 0.1, None, None, P(save), 1, None, 0, s, P(mk), None),
                                nb("stats") + nb("fwd_apply"))
        calls["bwd(1 call)"] = (lambda: lib.b200dp_bn_backward(P(dy), P(x), P(mk if mk is not None else y), P(dx), P(dr), code, M, C, P(save), P(w), P(b),
                                                              mode, P(scratch), P(sums), P(gw), P(gb), None, 0, s),
                                nb("bwd_reduce") + nb("bwd_apply"))
        cells = []
        for name, (fn, nbytes) in calls.items():
            for _ in range(2):
                _lib.check(fn())
            ts = []
            for _ in range(a.iters):
                if not a.warm:
                    flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.chain):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) / a.chain)
            ts.sort()
            ms = ts[len(ts) // 2]
            gbs = nbytes / ms / 1e6
            cells.append(f"{ms*1e3:8.1f}us {gbs:6.0f}GB/s {gbs/peak*100:3.0f}%")
            t = totals.setdefault(name, [0.0, 0])
            t[0] += ms * count
            t[1] += nbytes * count
        print(f"{C:5d} {HW:4d} {count:2d} {str(res)[0]:>3} | " + " | ".join(cells))
    print("per training step (all 53 layers):")
    for name, (ms, nbytes) in totals.items():
        print(f"  {name:10s} {ms:7.3f} ms  {nbytes/1e9:6.2f} GB  {nbytes/ms/1e6:6.0f} GB/s  {nbytes/ms/1e6/peak*100:4.1f}% of {peak:.0f} GB/s")
    four = ("stats", "fwd_apply", "bwd_reduce", "bwd_apply")
    tot_ms = sum(totals[k][0] for k in four); tot_b = sum(totals[k][1] for k in four)
    print(f"  {'all':10s} {tot_ms:7.3f} ms  {tot_b/1e9:6.2f} GB  {tot_b/tot_ms/1e6:6.0f} GB/s  {tot_b/tot_ms/1e6/peak*100:4.1f}%")


if __name__ == "__main__":
    main()
