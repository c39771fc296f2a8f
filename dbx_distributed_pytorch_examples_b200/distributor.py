"""Local stand-in for ``pyspark.ml.torch.distributor.TorchDistributor``.

The reference launches every train function through
``TorchDistributor(num_processes=N, local_mode=True, use_gpu=True).run(fn, *args, **kwargs)``
(01_torch_distributor/01_basic_torch_distributor.py:367,
02_cifar_torch_distributor_resnet.py:340-353) or the DeepSpeed flavour
``DeepspeedTorchDistributor(numGpus=N, nnodes=1, localMode=True).run(fn, **kw)``
(02_deepspeed/01_cifar_deepspeed_resnet.py:104-109,313).  pyspark is not part
of this image, so this module offers the same two classes with the same call
shape: N worker processes on this node, each with the torchrun environment
(RANK, LOCAL_RANK, WORLD_SIZE, LOCAL_WORLD_SIZE, MASTER_ADDR, MASTER_PORT), the
train function shipped with cloudpickle (they are notebook closures), and rank
0's return value handed back to the caller.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import socket
import time
import traceback

import cloudpickle


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, payload, conn, extra_env):
    try:
        os.environ.update({
            "RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world),
            "LOCAL_WORLD_SIZE": str(world), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
        })
        os.environ.update(extra_env)
        fn, args, kwargs = cloudpickle.loads(payload)
        out = fn(*args, **kwargs)
        conn.send(("ok", cloudpickle.dumps(out if rank == 0 else None)))
    except BaseException as e:  # noqa: BLE001 - reported to the parent, which re-raises
        conn.send(("error", f"rank {rank}: {type(e).__name__}: {e}\n{traceback.format_exc()}"))
    finally:
        conn.close()


class TorchDistributor:
    """``TorchDistributor(num_processes, local_mode, use_gpu).run(train_fn, *args, **kwargs)``."""

    def __init__(self, num_processes: int = 1, local_mode: bool = True, use_gpu: bool = True, **_ignored):
        if num_processes < 1:
            raise ValueError("num_processes has to be a positive integer")
        if not local_mode:
            raise NotImplementedError("only local_mode=True (one node) is available without a Spark cluster")
        self.num_processes = num_processes
        self.local_mode = local_mode
        self.use_gpu = use_gpu
        self.failure_grace_s = 10.0

    def run(self, train_object, *args, **kwargs):
        if not callable(train_object):
            raise TypeError("train_object must be a callable (a path to a training script is not supported here)")
        payload = cloudpickle.dumps((train_object, args, kwargs))
        ctx = mp.get_context("spawn")
        port = _free_port()
        extra = {} if self.use_gpu else {"CUDA_VISIBLE_DEVICES": ""}
        procs, conns = [], []
        for rank in range(self.num_processes):
            parent, child = ctx.Pipe(duplex=False)
            p = ctx.Process(target=_worker, args=(rank, self.num_processes, port, payload, child, extra))
            p.start()
            child.close()
            procs.append(p)
            conns.append(parent)
        # Collect the results as they come.  Once one rank has failed the others get a few seconds to report
        # and are then terminated (what torchrun's elastic agent does): a rank blocked in a collective with a
        # dead peer would otherwise sit there until the backend's own timeout (10 minutes for NCCL).
        from multiprocessing.connection import wait as _wait

        results, errors = [None] * self.num_processes, []
        pending = {c: rank for rank, c in enumerate(conns)}
        deadline = None
        while pending:
            timeout = None if deadline is None else max(0.0, deadline - time.monotonic())
            ready = _wait(list(pending), timeout=timeout)
            if not ready:  # grace period after a failure is over
                for c, rank in pending.items():
                    procs[rank].terminate()
                    errors.append(f"rank {rank}: terminated after another rank failed")
                break
            for c in ready:
                rank = pending.pop(c)
                try:
                    status, body = c.recv()
                except EOFError:
                    status, body = "error", f"rank {rank}: worker exited without a result"
                if status == "ok":
                    results[rank] = body
                else:
                    errors.append(body)
                    if deadline is None:
                        deadline = time.monotonic() + self.failure_grace_s
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
        if errors:
            raise RuntimeError("TorchDistributor worker failed:\n" + "\n".join(errors))
        return cloudpickle.loads(results[0])


class DeepspeedTorchDistributor(TorchDistributor):
    """``DeepspeedTorchDistributor(numGpus, nnodes, localMode, useGpu, deepspeedConfig)``.

    The reference never calls ``deepspeed.initialize`` (its train functions use
    plain ``optim.Adam``, 02_deepspeed/01_cifar_deepspeed_resnet.py:206), so the
    launcher part is all there is to reproduce.
    """

    def __init__(self, numGpus: int = 1, nnodes: int = 1, localMode: bool = True, useGpu: bool = True,
                 deepspeedConfig=None):
        if nnodes != 1:
            raise NotImplementedError("only nnodes=1 is available without a Spark cluster")
        super().__init__(num_processes=numGpus, local_mode=localMode, use_gpu=useGpu)
        self.deepspeed_config = deepspeedConfig
