"""CPU: host-side logic of the drop-ins, and the C-ABI library's export table.
No compute entry point is called here (there is no GPU in the build container)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch
import torch.nn as nn

import dbx_distributed_pytorch_examples_b200 as b200
from dbx_distributed_pytorch_examples_b200 import _lib, ddp, fusion, workloads
from oracle import ddp_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    declared = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(handle, name), f"{name} is declared in include/b200dp.h but not exported"
    assert declared == set(_lib.exported_symbols())
    lib = _lib.lib()
    assert lib.b200dp_version() == 100
    assert b"invalid" in lib.b200dp_error_string(-1)
    assert lib.b200dp_bn_scratch_bytes(64) > 0


def test_argument_checks_need_no_gpu():
    lib = _lib.lib()
    # contract violations are rejected before any CUDA call
    assert lib.b200dp_augment_u8_bf16(None, -1, 8, 8, None, 0, 8, 8, None, None, None, None) == -1
    assert lib.b200dp_bn_fwd_stats(None, 7, 4, 64, None, None, None, 0, None) == -2
    assert lib.b200dp_bn_fwd_stats(None, _lib.BF16, 4, 60, None, None, None, 0, None) == -2  # C % 8
    assert lib.b200dp_bucket_pack(None, None, None, 2000, None, 0, 1.0, None) == -5


def test_draws_match_torchvision_golden(golden):
    for name, src in (("transform_cifar.npz", 32), ("transform_crop56.npz", 64)):
        g = golden(name)
        torch.manual_seed(int(g["seed"]))
        got = b200.draw_crop_flip_params(len(g["params"]), (src, src), (int(g["out_hw"]),) * 2, int(g["pad"]))
        assert np.array_equal(got.numpy(), g["params"])  # bit-exact indices


def test_draws_match_live_torchvision():
    from torchvision import transforms as T
    torch.manual_seed(99)
    want = []
    for _ in range(40):
        i, j, _, _ = T.RandomCrop.get_params(torch.zeros(3, 40, 40), (32, 32))
        want.append((i, j, int(torch.rand(1) < 0.5)))
    gen = torch.Generator().manual_seed(99)
    got = b200.draw_crop_flip_params(40, (32, 32), (32, 32), 4, generator=gen)
    assert got.tolist() == [list(w) for w in want]
    # equal sizes: get_params draws nothing, only the flip consumes the generator
    gen = torch.Generator().manual_seed(5)
    ref = torch.Generator().manual_seed(5)
    got = b200.draw_crop_flip_params(10, (32, 32), (32, 32), 0, generator=gen)
    assert got[:, :2].abs().sum() == 0
    assert got[:, 2].tolist() == [int(torch.rand(1, generator=ref) < 0.5) for _ in range(10)]
    with pytest.raises(ValueError):
        b200.draw_crop_flip_params(1, (8, 8), (16, 16))


def test_bulk_draws_replicate_the_per_sample_loop():
    from dbx_distributed_pytorch_examples_b200.transform import _draw_crop_flip_params_loop
    for seed in (0, 7, 2 ** 31 + 11):
        ga, gb = torch.Generator().manual_seed(seed), torch.Generator().manual_seed(seed)
        for n, src, out, pad in ((3, 32, 32, 4), (700, 256, 224, 0), (40, 64, 64, 0), (1, 8, 8, 1), (650, 40, 32, 0)):
            want = _draw_crop_flip_params_loop(n, (src, src + 1), (out, out), pad, generator=ga)
            got = b200.draw_crop_flip_params(n, (src, src + 1), (out, out), pad, generator=gb)
            assert torch.equal(got, want)
        # the generator is left exactly where the loop leaves it
        assert torch.equal(ga.get_state(), gb.get_state())
        assert torch.equal(torch.randn(7, generator=ga), torch.randn(7, generator=gb))


def test_sharded_batch_sampler_matches_distributed_sampler(golden):
    from torch.utils.data.distributed import DistributedSampler
    g = golden("sampler_buckets.npz")
    for world in (2, 4):
        for rank in range(world):
            for epoch in (0, 1):
                s = b200.ShardedBatchSampler(1003, 64, rank=rank, world=world, seed=0)
                s.set_epoch(epoch)
                flat = torch.cat(list(s)).numpy()
                assert np.array_equal(flat, g[f"w{world}_r{rank}_e{epoch}"])        # golden from torch
                assert np.array_equal(flat, ddp_ref.distributed_sampler_indices(1003, world, rank, g[f"perm_e{epoch}"]))
                assert len(s) == len(list(s)) == -(-len(flat) // 64)
    live = DistributedSampler(list(range(10)), num_replicas=4, rank=3, shuffle=False)
    ours = b200.ShardedBatchSampler(10, 2, rank=3, world=4, shuffle=False)
    assert torch.cat(list(ours)).tolist() == list(iter(live))
    ds = b200.PinnedImageDataset(torch.arange(10 * 2 * 2 * 3, dtype=torch.uint8).view(10, 2, 2, 3), torch.arange(10), pin=False)
    batches = [(im.clone(), lb.clone()) for im, lb in ds.batches(ours)]
    assert [lb.tolist() for _, lb in batches] == [[3, 7], [1]]
    assert torch.equal(batches[0][0][1], ds.images[7])
    with pytest.raises(ValueError):
        b200.ShardedBatchSampler(10, 2, rank=4, world=4)


def test_bucket_plan_matches_ddp_assignment():
    model = workloads.cifar_resnet18()
    params = list(reversed(list(model.parameters())))
    numels = [p.numel() for p in params]
    for world in (1, 2, 8):
        offsets, buckets, total = ddp.plan_buckets(numels, 4, world, 25 * 1024 * 1024)
        assert all(o % 8 == 0 for o in offsets)
        assert all(n % (8 * world) == 0 for _, n, _, _ in buckets)
        # contiguous, non-overlapping, every parameter inside its bucket
        assert buckets[0][0] == 0 and sum(n for _, n, _, _ in buckets) == total
        for (s, n, p0, p1) in buckets:
            assert offsets[p0] == s and offsets[p1 - 1] + numels[p1 - 1] <= s + n
    # same grouping as torch's _compute_bucket_assignment_by_size up to the 8-element padding
    want = ddp_ref.bucket_assignment([4 * n for n in numels])
    _, buckets, _ = ddp.plan_buckets(numels, 4, 1, 25 * 1024 * 1024)
    assert [list(range(p0, p1)) for _, _, p0, p1 in buckets] == want


def test_convert_and_fuse_keep_state_dict_layout():
    model = workloads.cifar_resnet18()
    keys = list(model.state_dict().keys())
    w = model.layer1[0].bn1.weight
    out = b200.convert_sync_batchnorm(model)
    assert list(out.state_dict().keys()) == keys
    assert out.layer1[0].bn1.weight is w  # Parameters are shared, not copied
    n_bn = sum(isinstance(m, b200.B200SyncBatchNorm) for m in out.modules())
    assert n_bn == 20  # setup/resnet18.py has 20 BatchNorm2d layers
    assert not any(isinstance(m, nn.BatchNorm2d) for m in out.modules())
    b200.fuse_bn_relu(out)
    assert list(out.state_dict().keys()) == keys
    assert out.bn1.fuse_relu and isinstance(out.relu, nn.Identity)
    assert out.layer2[0].bn1.fuse_relu and out.layer2[0].bn2.fuse_relu
    assert not out.layer2[0].skip_connection[1].fuse_relu  # projection BN has no activation
    # statistics and affine parameters stay fp32 through a bf16 cast
    out.to(torch.bfloat16)
    assert out.bn1.weight.dtype == torch.float32 and out.bn1.running_var.dtype == torch.float32
    assert out.conv1.weight.dtype == torch.bfloat16


def test_unfuse_restores_the_plain_modules():
    """fuse_bn_relu -> unfuse_bn_relu is the identity on the module structure (B200DDP pickles the unfused form:
    a block whose forward is an instance attribute cannot be unpickled)."""
    import pickle
    for make in (lambda: workloads.cifar_resnet18(10), lambda: workloads.resnet50(10)):
        m = b200.fuse_bn_relu(b200.convert_sync_batchnorm(make()))
        assert any("forward" in mod.__dict__ for mod in m.modules())
        keys = list(m.state_dict())
        m = b200.unfuse_bn_relu(m)
        assert not any("forward" in mod.__dict__ for mod in m.modules())
        assert not any(getattr(mod, "fuse_relu", False) for mod in m.modules())
        assert isinstance(m.relu, nn.ReLU) and type(m.maxpool) is nn.MaxPool2d
        assert list(m.state_dict()) == keys
        back = pickle.loads(pickle.dumps(m))
        assert list(back.state_dict()) == keys
        assert b200.fuse_bn_relu(back).bn1.fuse_relu          # and it fuses again


def test_torchvision_resnet50_is_recognised():
    model = b200.fuse_bn_relu(b200.convert_sync_batchnorm(workloads.resnet50(10)))
    bns = [m for m in model.modules() if isinstance(m, b200.B200SyncBatchNorm)]
    assert len(bns) == 53
    assert sum(m.fuse_relu for m in bns) == 53 - 4  # all but the 4 downsample BNs
    assert model.layer1[0].forward.__func__ is fusion._bottleneck_forward


def test_no_cpu_fallback():
    bn = b200.B200SyncBatchNorm(8)
    with pytest.raises(ValueError, match="expected input tensor to be on GPU"):
        bn(torch.randn(2, 8, 4, 4))
    with pytest.raises(ValueError, match="no CPU path"):
        b200.B200DDP(nn.Linear(4, 4))
    frozen = nn.Linear(4, 4).requires_grad_(False)
    with pytest.raises(RuntimeError, match="doesn't have any parameter that requires a gradient"):
        b200.B200DDP(frozen)
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            b200.BatchTransform(8)(torch.zeros(2, 8, 8, 3, dtype=torch.uint8))


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libb200dp.so")
    with pytest.raises(_lib.B200DPError, match="no CPU or eager fallback"):
        _lib.lib()


def test_workload_matches_reference_model():
    ref_path = "/root/reference/setup/resnet18.py"
    if not os.path.exists(ref_path):
        pytest.skip("reference tree not mounted (GPU box)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("_ref_resnet18", ref_path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(0)
    ours = workloads.cifar_resnet18(10)
    torch.manual_seed(0)
    ref = mod.ResNet18(10)
    sa, sb = ours.state_dict(), ref.state_dict()
    assert list(sa) == list(sb) and all(torch.equal(sa[k], sb[k]) for k in sa)
    x = torch.randn(4, 3, 32, 32)
    assert torch.equal(ours(x), ref(x))


def test_autograd_extension_loads_and_binds():
    """csrc/autograd_bn.cpp builds in-tree and resolves the SyncBN entry points of the C ABI."""
    nat = _lib.native()
    assert nat is not None, "build it with __graft_entry__.build()"
    assert nat.counters() == (0, 0) or all(v >= 0 for v in nat.counters())
    for name in ("bind", "fused_syncbn", "counters"):
        assert callable(getattr(nat, name))
    with pytest.raises(RuntimeError, match="cannot load"):
        nat.bind("/nonexistent/libb200dp.so")
    nat.bind(_lib.LIB_PATH)
    # argument checks happen before any CUDA call, so they can be exercised here; they must come out as
    # python exceptions (an extension built with a compiler that brings its own libstdc++ crashed instead)
    with pytest.raises(ValueError, match="more than 1 value per channel"):
        nat.fused_syncbn(torch.randn(1, 32), None, None, None, 0, 0, 1e-5, 0.1, False, 0, 1, False, 0)
    with pytest.raises(ValueError, match="residual must have the shape"):
        nat.fused_syncbn(torch.randn(4, 8, 3, 3), torch.randn(4, 8, 3, 2), None, None, 0, 0, 1e-5, 0.1, False, 0, 1, False, 0)
    with pytest.raises(RuntimeError, match="float32 and bfloat16"):
        nat.fused_syncbn(torch.randn(4, 8).double(), None, None, None, 0, 0, 1e-5, 0.1, False, 0, 1, False, 0)


def test_stem_pooling_module_swap():
    """fuse_bn_relu puts B200MaxPool2d behind the stem's fused BN + ReLU (setup/resnet18.py:37,58) for the geometry
    the kernel handles, leaves anything else alone, keeps the state_dict, and has no CPU path."""
    m = workloads.cifar_resnet18(10)
    keys = list(m.state_dict())
    m = b200.fuse_bn_relu(b200.convert_sync_batchnorm(m))
    assert isinstance(m.maxpool, b200.B200MaxPool2d) and isinstance(m.maxpool, nn.MaxPool2d)
    assert (m.maxpool.kernel_size, m.maxpool.stride, m.maxpool.padding) == (3, 2, 1)
    assert list(m.state_dict()) == keys
    assert b200.pool.supported(nn.MaxPool2d(2)) and b200.pool.supported(nn.MaxPool2d((3, 3), (2, 2), (1, 1)))
    for odd in (nn.MaxPool2d(3, 2, 1, ceil_mode=True), nn.MaxPool2d(3, 2, 1, dilation=2), nn.MaxPool2d(5, 2, 2),
                nn.MaxPool2d((3, 2)), nn.MaxPool2d(3, 2, 1, return_indices=True)):
        assert not b200.pool.supported(odd)
    with pytest.raises(NotImplementedError):
        b200.B200MaxPool2d(3, 2, 1, ceil_mode=True)
    with pytest.raises(ValueError, match="no CPU path"):
        m.maxpool(torch.zeros(1, 8, 8, 8))
    lib = _lib.lib()
    assert lib.b200dp_maxpool_fwd(None, None, None, 0, 1, 8, 8, 8, 4, 2, 1, None) == -2      # window > 3
    assert lib.b200dp_maxpool_bwd(None, None, None, 1, 1, 8, 8, 12, 3, 2, 1, None) == -2     # C % 8 (bf16)
    assert lib.b200dp_maxpool_fwd(None, None, None, 0, 1, 8, 8, 8, 3, 2, 1, None) == -1      # null pointers
    assert lib.b200dp_resized_crop_u8_bf16(None, -1, 8, 8, None, 8, 8, None, None, None, None) == -1
    assert lib.b200dp_comm_all_gather32(None, None, 1, None, None) == -1


def test_pinned_dataset_with_ragged_sizes():
    from dbx_distributed_pytorch_examples_b200.feed import PinnedImageDataset, ShardedBatchSampler
    imgs = torch.randint(0, 256, (10, 12, 16, 3), dtype=torch.uint8)
    labels = torch.arange(10)
    sizes = torch.stack([torch.arange(3, 13), torch.arange(7, 17)], dim=1)
    ds = PinnedImageDataset(imgs, labels, sizes=sizes)
    sampler = ShardedBatchSampler(len(ds), 4, rank=0, world=1, shuffle=False, drop_last=False)
    seen = 0
    for data, target, hw in ds.batches(sampler):
        idx = target.tolist()
        assert torch.equal(data, imgs[idx]) and hw.dtype == torch.int32 and torch.equal(hw, sizes[idx].int())
        seen += len(idx)
    assert seen == 10
    plain = PinnedImageDataset(imgs, labels)
    assert all(len(b) == 2 for b in plain.batches(sampler))
    with pytest.raises(ValueError):
        PinnedImageDataset(imgs, labels, sizes=sizes + 10)      # larger than the slab
    with pytest.raises(ValueError):
        PinnedImageDataset(imgs, labels, sizes=sizes[:3])


def test_bucket_plan_invariants_property():
    """Random parameter lists: the plan tiles the arena, keeps 16-byte alignment per parameter and whole vectors
    per rank slice, and groups parameters like torch's Reducer (up to the padding)."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(st.lists(st.integers(1, 3_000_000), min_size=1, max_size=40), st.sampled_from([1, 2, 4, 8]),
           st.sampled_from([2, 4]), st.sampled_from([1, 5, 25]))
    def check(numels, world, esize, cap_mb):
        offsets, buckets, total = ddp.plan_buckets(numels, esize, world, cap_mb * 1024 * 1024)
        assert len(offsets) == len(numels) and all(o % 8 == 0 for o in offsets)
        assert buckets[0][0] == 0 and sum(n for _, n, _, _ in buckets) == total
        nxt_start, nxt_param = 0, 0
        for (s, n, p0, p1) in buckets:
            assert s == nxt_start and p0 == nxt_param and p1 > p0 and n % (8 * world) == 0
            for k in range(p0, p1):
                assert s <= offsets[k] and offsets[k] + numels[k] <= s + n
            for k in range(p0, p1 - 1):
                assert offsets[k] + numels[k] <= offsets[k + 1]
            nxt_start, nxt_param = s + n, p1
        assert nxt_param == len(numels)
        if world == 1:
            # the padding to 8 elements can only make a bucket close earlier than torch's byte count would
            ours = [list(range(p0, p1)) for _, _, p0, p1 in buckets]
            padded = [(n + 7) // 8 * 8 * esize for n in numels]
            assert ours == ddp_ref.bucket_assignment(padded, cap=cap_mb * 1024 * 1024)

    check()


def test_reference_arm_of_the_bench_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver runs beside every GPU number): stock DDP over gloo with
    two ranks on the host cores, bounded steps, one JSON line with the contract's keys and no GPU launches."""
    import json
    import subprocess
    import sys
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, stdin=subprocess.DEVNULL, timeout=900, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "images/s" and line["value"] > 0
    assert line["gpu_launches"] == 0 and line["higher_is_better"] is True
    assert line["config"]["parallelism"] == "cpu-gloo-2" and line["config"]["workload"] == "resnet50_224_syncbn_ddp_augment"
    cb = line["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 2 and "2 ranks" in cb["sample"] and cb["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
