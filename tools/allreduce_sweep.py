#!/usr/bin/env python
"""BASELINE config 5: gradient-bucket all-reduce + update sweep, 1 KB - 256 MB, ours vs NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/allreduce_sweep.py

ours  : b200dp_allreduce_update (fused 1/N + SGD-momentum + parameter push over peer memory)
nccl  : dist.all_reduce on the same bucket followed by torch.optim.SGD(foreach) — the DDP path
busBW = S/t * 2(N-1)/N (nccl-tests convention), t = CUDA-event time, max over ranks.
"""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dbx_distributed_pytorch_examples_b200 import _lib  # noqa: E402
from dbx_distributed_pytorch_examples_b200.comm import Communicator  # noqa: E402


def timed(fn, iters, world, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    world = dist.get_world_size()
    lib = _lib.lib()
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", 256 << 20))
    nvls = os.environ.get("SWEEP_NVLS", "0") == "1"   # heap with a multicast mapping: the two-shot kernel reduces in the switch
    rows = []
    for dtype in (torch.float32, torch.bfloat16):
        code, es = _lib.dtype_code(dtype), (2 if dtype == torch.bfloat16 else 4)
        size = 1024
        while size <= max_bytes:
            n = max(size // es, 8 * world) // (8 * world) * (8 * world)
            nbytes = (n * es + 511) // 512 * 512
            comm = Communicator(heap_bytes=2 * nbytes, multicast=True if nvls else False)
            grad, param = comm.heap_view(0, n, dtype), comm.heap_view(nbytes, n, dtype)
            grad.normal_(); param.normal_()
            master = param.float() if dtype == torch.bfloat16 else None
            s1 = torch.zeros(n, device=dev)
            o = _lib.OptArgs()
            o.kind, o.lr, o.momentum = _lib.OPT_SGD, 1e-3, 0.5
            stream = _lib.current_stream()
            res = {}
            for label, one_shot, ctas in (("two_shot_32", 0, 32), ("two_shot", 0, 64), ("two_shot_148", 0, 148), ("one_shot", 1, 64)):
                if one_shot and size > (4 << 20):
                    continue

                def ours():
                    _lib.check(lib.b200dp_allreduce_update(
                        comm.handle, 0, code, nbytes, code, 0, n, None if master is None else master.data_ptr(),
                        s1.data_ptr(), None, ctypes.byref(o), one_shot, 0, ctas, comm.next_seq(), stream))

                res[label] = timed(ours, 20 if size <= (16 << 20) else 8, world, dev)
            p = torch.nn.Parameter(param.clone())
            p.grad = grad.clone()
            opt = torch.optim.SGD([p], lr=1e-3, momentum=0.5)

            def nccl():
                p.grad.div_(world)
                dist.all_reduce(p.grad)
                opt.step()

            res["nccl_allreduce_plus_sgd"] = timed(nccl, 20 if size <= (16 << 20) else 8, world, dev)

            def nccl_only():
                dist.all_reduce(p.grad)

            res["nccl_allreduce_only"] = timed(nccl_only, 20 if size <= (16 << 20) else 8, world, dev)
            comm.check_error()
            comm_multicast = comm.multicast
            comm.destroy()
            if rank == 0:
                bus = lambda ms: n * es / (ms / 1e3) * 2 * (world - 1) / world / 1e9  # noqa: E731
                row = {"dtype": str(dtype).split(".")[-1], "bytes": n * es, "world": world, "nvls": bool(comm_multicast),
                       **{k + "_us": round(v * 1e3, 2) for k, v in res.items()},
                       **{k + "_busGBs": round(bus(v), 1) for k, v in res.items()}}
                rows.append(row)
                print(json.dumps(row), flush=True)
            size *= 4
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
