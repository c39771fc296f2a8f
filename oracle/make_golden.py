#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REAL dependencies the oracle restates.

Run in the build container, where /root/reference is mounted:
    python oracle/make_golden.py [--skip-ddp]
Sources of truth used here (never the oracle, never the product):
  * torchvision 0.26.0 transforms (RandomCrop, RandomHorizontalFlip, ToTensor, Normalize) on PIL images;
  * torch 2.11.0 F.batch_norm + autograd on the rank-concatenated batch (nn.SyncBatchNorm's own
    kernels are CUDA-only), torch.optim.SGD / Adam, DistributedSampler, _compute_bucket_assignment_by_size;
  * the reference's own model, /root/reference/setup/resnet18.py, trained by stock
    DistributedDataParallel over gloo with world size 2 (BASELINE.json config 1).
The GPU box has no /root/reference: tests only read the committed .npz files.
"""
import argparse
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)


def gen_transform(name, B, src_hw, out, pad, mean, std, seed):
    from PIL import Image
    from torchvision import transforms as T

    rng = np.random.default_rng(seed)
    src = rng.integers(0, 256, size=(B, src_hw, src_hw, 3), dtype=np.uint8)
    crop = T.RandomCrop(out, padding=pad) if pad else T.RandomCrop(out)
    pipeline = T.Compose([crop, T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize(mean, std)])
    # pass 1: the real pipeline, sample after sample, from the global generator
    torch.manual_seed(seed)
    outs = [pipeline(Image.fromarray(src[b])) for b in range(B)]
    out_chw = torch.stack(outs)                      # [B,3,H,W] float32
    out_nhwc = out_chw.permute(0, 2, 3, 1).contiguous()
    # pass 2: the same draws through torchvision's own get_params, to record them
    torch.manual_seed(seed)
    params = np.zeros((B, 3), dtype=np.int32)
    for b in range(B):
        img = torch.zeros(3, src_hw + 2 * pad, src_hw + 2 * pad)
        i, j, _, _ = T.RandomCrop.get_params(img, (out, out))
        flip = bool(torch.rand(1) < 0.5)
        params[b] = (i, j, int(flip))
    np.savez_compressed(os.path.join(GOLD, name), src=src, params=params, pad=pad, out_hw=out,
                        mean=np.asarray(mean, np.float64), std=np.asarray(std, np.float64),
                        out_f32=out_nhwc.numpy(), out_bf16=bf16_bits(out_nhwc), seed=seed)


def gen_resample(name, B, src_hw, out, random_resized_crop, mean, std, seed):
    """Resize / RandomResizedCrop on PIL images through torchvision itself
    (utils/hf_dataset_utilities.py:64-79; 02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53)."""
    from PIL import Image
    from torchvision import transforms as T
    from torchvision.transforms import functional as TF

    rng = np.random.default_rng(seed)
    hs, ws = src_hw
    # smooth-ish content plus noise, so that rounding cases of the resampler differ from pure noise
    base = rng.integers(0, 256, size=(B, (hs + 7) // 8, (ws + 7) // 8, 3), dtype=np.uint8)
    src = np.repeat(np.repeat(base, 8, axis=1), 8, axis=2)[:, :hs, :ws, :].astype(np.int32)
    src = np.clip(src + rng.integers(-40, 41, size=src.shape), 0, 255).astype(np.uint8)
    first = T.RandomResizedCrop(out) if random_resized_crop else T.Resize((out, out))
    pipeline = T.Compose([first, T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize(mean, std)])
    torch.manual_seed(seed)
    outs = torch.stack([pipeline(Image.fromarray(src[b])) for b in range(B)]).permute(0, 2, 3, 1).contiguous()
    # the same draws through torchvision's own get_params, to record them, and the uint8 stage
    torch.manual_seed(seed)
    params = np.zeros((B, 5), dtype=np.int32)
    resized = np.zeros((B, out, out, 3), dtype=np.uint8)
    for b in range(B):
        img = Image.fromarray(src[b])
        if random_resized_crop:
            i, j, h, w = T.RandomResizedCrop.get_params(img, [0.08, 1.0], [3.0 / 4.0, 4.0 / 3.0])
        else:
            i, j, h, w = 0, 0, hs, ws
        flip = bool(torch.rand(1) < 0.5)
        params[b] = (i, j, h, w, int(flip))
        resized[b] = np.asarray(TF.resized_crop(img, i, j, h, w, [out, out]))  # before the flip
    np.savez_compressed(os.path.join(GOLD, name), src=src, params=params, out_hw=out,
                        mean=np.asarray(mean, np.float64), std=np.asarray(std, np.float64),
                        resized_u8=resized, out_bf16=bf16_bits(outs), seed=seed,
                        random_resized_crop=int(random_resized_crop))


def gen_maxpool():
    """F.max_pool2d + autograd on inputs full of ties (ReLU output, coarse values): nn.MaxPool2d(3, 2, 1) as in
    setup/resnet18.py:37,58 plus a 2x2/2 and a 3x3/1 window on odd sizes."""
    g = torch.Generator().manual_seed(77)
    out = {}
    for tag, (N, C, H, W, k, s, p) in {"stem": (3, 16, 14, 14, 3, 2, 1), "odd": (2, 8, 9, 11, 3, 1, 1),
                                        "k2": (2, 8, 7, 10, 2, 2, 0), "k3p0": (2, 8, 8, 8, 3, 2, 0)}.items():
        x = torch.relu(torch.randn(N, C, H, W, generator=g)).mul(4).round().div(4)  # many exact ties, many zeros
        x = x.contiguous(memory_format=torch.channels_last).requires_grad_(True)
        y = F.max_pool2d(x, k, s, p)
        dy = torch.randn(y.shape, generator=g)
        y.backward(dy)
        out[f"{tag}_geom"] = np.array([k, s, p])
        out[f"{tag}_x"] = x.detach().permute(0, 2, 3, 1).contiguous().numpy()
        out[f"{tag}_y"] = y.detach().permute(0, 2, 3, 1).contiguous().numpy()
        out[f"{tag}_dy"] = dy.permute(0, 2, 3, 1).contiguous().numpy()
        out[f"{tag}_dx"] = x.grad.permute(0, 2, 3, 1).contiguous().numpy()
    np.savez_compressed(os.path.join(GOLD, "maxpool.npz"), **out)


def gen_syncbn():
    torch.manual_seed(7)
    C, eps, mom = 16, 1e-5, 0.1
    shapes = [(3, C, 4, 4), (5, C, 4, 4)]  # uneven per-rank batch sizes
    out = {}
    for tag, relu, use_res in [("plain", False, False), ("relu", True, False), ("addrelu", True, True)]:
        xs = [torch.randn(s) * 2 + 0.5 for s in shapes]
        rs = [torch.randn(s) for s in shapes]
        dys = [torch.randn(s) for s in shapes]
        w, b = torch.rand(C) + 0.5, torch.randn(C)
        rm, rv = torch.randn(C) * 0.1, torch.rand(C) + 0.5
        x = torch.cat(xs).requires_grad_(True)
        r = torch.cat(rs).requires_grad_(True)
        wp, bp = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
        rm2, rv2 = rm.clone(), rv.clone()
        y = F.batch_norm(x, rm2, rv2, wp, bp, True, mom, eps)
        if use_res:
            y = y + r
        if relu:
            y = torch.relu(y)
        y.backward(torch.cat(dys))
        rows = lambda t: t.permute(0, 2, 3, 1).reshape(-1, C).detach().numpy()  # noqa: E731
        n0 = shapes[0][0]
        out.update({
            f"{tag}_x0": rows(xs[0]), f"{tag}_x1": rows(xs[1]), f"{tag}_r0": rows(rs[0]), f"{tag}_r1": rows(rs[1]),
            f"{tag}_dy0": rows(dys[0]), f"{tag}_dy1": rows(dys[1]), f"{tag}_w": w.numpy(), f"{tag}_b": b.numpy(),
            f"{tag}_rm": rm.numpy(), f"{tag}_rv": rv.numpy(), f"{tag}_rm_new": rm2.numpy(), f"{tag}_rv_new": rv2.numpy(),
            f"{tag}_y0": rows(y[:n0]), f"{tag}_y1": rows(y[n0:]), f"{tag}_dx0": rows(x.grad[:n0]),
            f"{tag}_dx1": rows(x.grad[n0:]), f"{tag}_dw": wp.grad.numpy(), f"{tag}_db": bp.grad.numpy(),
        })
        if use_res:
            out.update({f"{tag}_dr0": rows(r.grad[:n0]), f"{tag}_dr1": rows(r.grad[n0:])})
    np.savez_compressed(os.path.join(GOLD, "syncbn.npz"), eps=eps, momentum=mom, **out)


def gen_optim():
    torch.manual_seed(11)
    n, steps = 1000, 3
    p0 = torch.randn(n)
    grads = [torch.randn(n) for _ in range(steps)]
    out = {"p0": p0.numpy(), "grads": torch.stack(grads).numpy()}
    cases = {
        "sgd_m05": (torch.optim.SGD, dict(lr=1e-3, momentum=0.5)),  # 01_basic_torch_distributor.py:294
        "sgd_nesterov_wd": (torch.optim.SGD, dict(lr=1e-2, momentum=0.9, nesterov=True, weight_decay=1e-4)),
        "sgd_plain": (torch.optim.SGD, dict(lr=1e-2)),
        "adam": (torch.optim.Adam, dict(lr=1e-3)),                  # 02_cifar_torch_distributor_resnet.py:213
        "adam_wd": (torch.optim.Adam, dict(lr=1e-3, weight_decay=1e-4)),  # 04_accelerate…ipynb:482-486
    }
    for name, (cls, kw) in cases.items():
        p = torch.nn.Parameter(p0.clone())
        opt = cls([p], **kw)
        traj = []
        for g in grads:
            p.grad = g.clone()
            opt.step()
            traj.append(p.detach().clone())
        out[name] = torch.stack(traj).numpy()
    np.savez_compressed(os.path.join(GOLD, "optim.npz"), **out)


def gen_sampler_and_buckets():
    from torch.utils.data.distributed import DistributedSampler
    import torch.distributed as dist

    out = {}
    ds = list(range(1003))
    for world in (2, 4):
        for rank in range(world):
            for epoch in (0, 1):
                s = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True, seed=0)
                s.set_epoch(epoch)
                out[f"w{world}_r{rank}_e{epoch}"] = np.asarray(list(iter(s)), dtype=np.int64)
    for epoch in (0, 1):
        g = torch.Generator().manual_seed(0 + epoch)
        out[f"perm_e{epoch}"] = torch.randperm(len(ds), generator=g).numpy()
    # bucket assignment of the reference ResNet18 in reverse registration order
    model = load_reference_resnet18()
    params = list(reversed([p for p in model.parameters()]))
    idx, _ = dist._compute_bucket_assignment_by_size(params, [1024 * 1024, 25 * 1024 * 1024])
    out["resnet18_param_bytes_rev"] = np.asarray([p.numel() * 4 for p in params], dtype=np.int64)
    out["resnet18_bucket_of_param"] = np.asarray(
        [next(b for b, members in enumerate(idx) if k in members) for k in range(len(params))], dtype=np.int64)
    np.savez_compressed(os.path.join(GOLD, "sampler_buckets.npz"), **out)


def load_reference_resnet18(num_classes=10):
    spec = importlib.util.spec_from_file_location("_ref_resnet18", os.path.join(REF, "setup", "resnet18.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(0)
    return mod.ResNet18(num_classes)


def config1_batch(rank, step, batch=128):
    g = torch.Generator().manual_seed((100 + rank) * 100000 + step)
    return torch.randn(batch, 3, 32, 32, generator=g), torch.randint(0, 10, (batch,), generator=g)


def _ddp_worker(rank, world, port, steps, q):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(max(1, os.cpu_count() // world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    model = load_reference_resnet18()
    ddp = DDP(model)
    opt = torch.optim.SGD(ddp.parameters(), lr=1e-3, momentum=0.5)  # 01_basic_torch_distributor.py:149,294
    losses, snap = [], {}
    names = ["fc.weight", "fc.bias", "bn1.weight", "layer1.0.conv1.weight", "layer4.1.bn2.bias"]
    for step in range(steps):
        x, y = config1_batch(rank, step)
        opt.zero_grad()
        loss = F.cross_entropy(ddp(x), y)
        loss.backward()
        if step == 0 and rank == 0:
            sd = dict(model.named_parameters())
            for n in names:
                snap["grad1_" + n] = sd[n].grad.detach().clone().numpy()
        opt.step()
        losses.append(loss.item())
        if step == 0 and rank == 0:
            sd = dict(model.named_parameters())
            for n in names:
                snap["w1_" + n] = sd[n].detach().clone().numpy()
    if rank == 0:
        sd = dict(model.named_parameters())
        for n in names:
            snap["wN_" + n] = sd[n].detach().clone().numpy()
        snap["wN_sum"] = np.asarray([float(p.double().sum()) for p in model.parameters()])
        snap["wN_abs"] = np.asarray([float(p.double().abs().sum()) for p in model.parameters()])
        snap["running_mean_bn1"] = model.bn1.running_mean.numpy().copy()
        snap["running_var_bn1"] = model.bn1.running_var.numpy().copy()
    q.put((rank, losses, snap))
    dist.barrier()
    dist.destroy_process_group()


def gen_ddp_config1(steps=200):
    import torch.multiprocessing as mp
    from dbx_distributed_pytorch_examples_b200.distributor import _free_port

    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, steps, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get() for _ in procs]
    for p in procs:
        p.join()
    got.sort(key=lambda t: t[0])
    snap = got[0][2]
    np.savez_compressed(os.path.join(GOLD, "ddp_config1.npz"), steps=steps,
                        loss_rank0=np.asarray(got[0][1]), loss_rank1=np.asarray(got[1][1]), **snap)


if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-ddp", action="store_true")
    ap.add_argument("--only-ddp", action="store_true")
    ap.add_argument("--only-resample", action="store_true")
    ap.add_argument("--steps", type=int, default=200)
    a = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    imagenet = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
    if not a.only_ddp:
        # RandomResizedCrop (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53) at fixture size and at
        # the full 256 -> 224 size; Resize((64, 64)) of 32x32 images (utils/hf_dataset_utilities.py:64-66)
        gen_resample("resample_rrc56.npz", 12, (96, 80), 56, True, *imagenet, 2024)
        gen_resample("resample_rrc224.npz", 2, (256, 256), 224, True, *imagenet, 7)
        gen_resample("resample_resize64.npz", 6, (32, 32), 64, False, *imagenet, 99)
        gen_resample("resample_down24.npz", 4, (75, 100), 24, False, *imagenet, 5)
        gen_maxpool()
    if a.only_resample:
        sys.exit(0)
    if not a.only_ddp:
        # CIFAR recipe (04_accelerate/01_cifar_accelerate.ipynb:295-298) and the crop-without-padding
        # ImageNet-style recipe (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53) at fixture size
        gen_transform("transform_cifar.npz", 32, 32, 32, 4, (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010), 1234)
        gen_transform("transform_crop56.npz", 8, 64, 56, 0, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225), 4321)
        gen_syncbn()
        gen_optim()
        gen_sampler_and_buckets()
    if not a.skip_ddp:
        gen_ddp_config1(a.steps)
    print("golden fixtures written to", GOLD)
