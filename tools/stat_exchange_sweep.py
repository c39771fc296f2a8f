#!/usr/bin/env python
"""BASELINE config 5, second half: SyncBN statistics exchange, per layer, ours vs NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/stat_exchange_sweep.py

For C in {64 ... 2048} and a layer so small that the data passes cost nothing (M = 64 rows per rank), per
layer and direction, microseconds (CUDA events over back-to-back repetitions, max over ranks):

  ours_fwd / ours_bwd        b200dp_bn_forward / b200dp_bn_backward with the cross-rank exchange inside
  local_fwd / local_bwd      the same calls without a communicator (no exchange): the difference is what
                             the exchange adds to a layer
  nccl_fwd                   dist.all_gather_into_tensor of the (2C+1) vector, as nn.SyncBatchNorm.forward
                             issues it (TORCH/nn/modules/_functions.py:74-76)
  nccl_bwd                   dist.all_reduce of the 2C sums, as its backward does (:159-164)
  torch_syncbn_fwd_bwd       nn.SyncBatchNorm forward + backward on the same tensor (5 ATen kernels + the 2
                             collectives), next to ours_fwd + ours_bwd
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib  # noqa: E402
from dbx_distributed_pytorch_examples_b200.comm import stats_communicator  # noqa: E402


def timed(fn, iters, dev):
    for _ in range(10):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t), 2)


def main():
    rank = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    world = dist.get_world_size()
    lib = _lib.lib()
    comm = stats_communicator()
    iters = int(os.environ.get("SWEEP_ITERS", 200))
    scratch = torch.zeros(int(lib.b200dp_bn_scratch_bytes(0)), dtype=torch.uint8, device=dev)
    s = _lib.current_stream()
    P = lambda t: 0 if t is None else t.data_ptr()  # noqa: E731
    for dtype in (torch.bfloat16, torch.float32):
        code = _lib.dtype_code(dtype)
        for C in (64, 128, 256, 512, 1024, 2048):
            M = 64
            x = torch.randn(M, C, device=dev).to(dtype)
            dy = torch.randn(M, C, device=dev).to(dtype)
            y, dx = torch.empty_like(x), torch.empty_like(x)
            w, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
            stats, save = torch.empty(2 * C + 1, device=dev), torch.empty(2 * C + 1, device=dev)
            sums, gw, gb = torch.empty(2 * C, device=dev), torch.empty(C, device=dev), torch.empty(C, device=dev)

            def fwd(h):
                _lib.check(lib.b200dp_bn_forward(P(x), 0, P(y), code, M, C, P(scratch), P(stats), P(w), P(b), 1e-5, 0.1, None,
                                                 None, P(save), 1, h, comm.next_seq() if h else 0, s, None, None))

            def bwd(h):
                _lib.check(lib.b200dp_bn_backward(P(dy), P(x), 0, P(dx), 0, code, M, C, P(save), P(w), P(b), 1, P(scratch),
                                                  P(sums), P(gw), P(gb), h, comm.next_seq() if h else 0, s))

            row = {"world": world, "dtype": str(dtype).split(".")[-1], "C": C, "rows_per_rank": M}
            row["ours_fwd_us"] = timed(lambda: fwd(comm.handle), iters, dev)
            row["ours_bwd_us"] = timed(lambda: bwd(comm.handle), iters, dev)
            row["local_fwd_us"] = timed(lambda: fwd(None), iters, dev)
            row["local_bwd_us"] = timed(lambda: bwd(None), iters, dev)
            vec = torch.randn(2 * C + 1, device=dev)
            out = torch.empty(world * (2 * C + 1), device=dev)
            row["nccl_fwd_us"] = timed(lambda: dist.all_gather_into_tensor(out, vec), iters, dev)
            v2 = torch.randn(2 * C, device=dev)
            row["nccl_bwd_us"] = timed(lambda: dist.all_reduce(v2), iters, dev)
            ref = torch.nn.SyncBatchNorm(C).to(dev)
            x4 = x.view(M, C, 1, 1).contiguous(memory_format=torch.channels_last).requires_grad_(True)
            dy4 = dy.view(M, C, 1, 1).contiguous(memory_format=torch.channels_last)

            def torch_both():
                x4.grad = None
                ref(x4).backward(dy4)

            row["torch_syncbn_fwd_bwd_us"] = timed(torch_both, max(20, iters // 4), dev)
            row["exchange_added_fwd_us"] = round(row["ours_fwd_us"] - row["local_fwd_us"], 2)
            row["exchange_added_bwd_us"] = round(row["ours_bwd_us"] - row["local_bwd_us"], 2)
            comm.check_error()
            if rank == 0:
                print(json.dumps(row), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
