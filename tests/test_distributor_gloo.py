"""CPU, world size 2 over gloo: the launcher shim and BASELINE config 1's plumbing
(01_torch_distributor/01_basic_torch_distributor.py:248-328 with synthetic tensors),
checked against the golden loss curve of the reference model under stock DDP."""
import os

import numpy as np
import pytest

from dbx_distributed_pytorch_examples_b200 import TorchDistributor, DeepspeedTorchDistributor


def _main_fn(steps, tag):
    """The reference main_fn's distributed skeleton (init -> DDP -> SGD -> loop -> teardown)."""
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from torch.nn.parallel import DistributedDataParallel as DDP

    from dbx_distributed_pytorch_examples_b200 import workloads
    from oracle import ddp_ref

    torch.set_num_threads(2)
    if not dist.is_initialized():
        dist.init_process_group("gloo")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert int(os.environ["LOCAL_RANK"]) == rank and dist.get_world_size() == world
    torch.manual_seed(0)
    ddp_model = DDP(workloads.cifar_resnet18(10))
    opt = torch.optim.SGD(ddp_model.parameters(), lr=1e-3, momentum=0.5)
    losses = []
    for step in range(steps):
        g = torch.Generator().manual_seed((100 + rank) * 100000 + step)
        x, y = torch.randn(128, 3, 32, 32, generator=g), torch.randint(0, 10, (128,), generator=g)
        opt.zero_grad()
        loss = F.cross_entropy(ddp_model(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    # DistributedSampler shards as the oracle says
    from torch.utils.data.distributed import DistributedSampler
    s = DistributedSampler(list(range(1003)), num_replicas=world, rank=rank, shuffle=True, seed=0)
    perm = torch.randperm(1003, generator=torch.Generator().manual_seed(0)).numpy()
    assert list(iter(s)) == ddp_ref.distributed_sampler_indices(1003, world, rank, perm).tolist()
    dist.destroy_process_group()
    return {"tag": tag, "losses": losses}


def test_config1_plumbing_world2(golden):
    out = TorchDistributor(num_processes=2, local_mode=True, use_gpu=False).run(_main_fn, 2, tag="cfg1")
    assert out["tag"] == "cfg1"
    path = os.path.join(os.path.dirname(__file__), "golden", "ddp_config1.npz")
    if os.path.exists(path):
        want = np.load(path)["loss_rank0"][:2]
        np.testing.assert_allclose(out["losses"], want, rtol=1e-4)


def test_worker_errors_propagate():
    def boom():
        raise ValueError("bad rank")
    with pytest.raises(RuntimeError, match="bad rank"):
        TorchDistributor(num_processes=2, use_gpu=False).run(boom)


def test_one_failing_rank_does_not_strand_the_others():
    """Rank 0 raises while rank 1 is stuck (as it would be inside a collective with a dead peer): the launcher
    reports rank 0's error and terminates rank 1 after a short grace period instead of waiting for it."""
    import time

    def fn():
        if int(os.environ["RANK"]) == 0:
            raise ValueError("rank 0 gave up")
        time.sleep(600)

    d = TorchDistributor(num_processes=2, use_gpu=False)
    d.failure_grace_s = 2.0
    t0 = time.monotonic()
    with pytest.raises(RuntimeError, match="rank 0 gave up") as info:
        d.run(fn)
    assert time.monotonic() - t0 < 120
    assert "rank 1: terminated after another rank failed" in str(info.value)


def test_deepspeed_flavour_signature():
    d = DeepspeedTorchDistributor(numGpus=1, nnodes=1, localMode=True, useGpu=False)
    assert d.run(lambda a, b=0: a + b, 1, b=2) == 3
    with pytest.raises(NotImplementedError):
        TorchDistributor(num_processes=2, local_mode=False)
