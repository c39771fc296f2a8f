#!/usr/bin/env python
"""Headline benchmark: ResNet-50 224^2 bf16 channels_last, SyncBN + DDP + GPU input transform.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the data-parallel hot path over one batch of 256 synthetic images per
GPU (BASELINE.json config 4): uint8 256x256 source images -> crop 224 / flip / normalize kernel
-> torchvision ResNet-50 (convolutions on cuDNN) with fused SyncBN(+add)+ReLU kernels ->
cross-entropy -> backward with the gradient buckets packed, all-reduced over NVLink peer
memory, averaged and applied by the fused SGD-momentum kernel.

`value`   images/s, whole job, source images already resident in HBM, CUDA-event timed.
`e2e`     same metric through the public API with the source images in pinned HOST memory
          (H2D copy inside the timed region) and loss.item() read back every step.
`roofline` the library kernel with the largest share of the step, timed per launch with CUDA
          events in a separate instrumented pass of the same step.
`cpu_baseline` / `--impl reference`: the reference's own software stack on the host cores —
          torchvision's per-sample CPU transform pipeline + stock DistributedDataParallel over
          gloo + torch.optim.SGD on a torchvision ResNet-50 (fp32, reduced batch, stated).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402

METRIC = "resnet50_224_bf16_ddp_images_per_sec"
UNIT = "images/s"
PER_GPU_BATCH = 256
SRC_HW, OUT_HW, CLASSES = 256, 224, 1000
# The headline (and default) workload is BASELINE config 4; configs 2 and 3 are selectable for
# completeness (their lines carry their own metric name and are not the contract's number).
WORKLOADS = {
    "r50_224": dict(metric=METRIC, model="resnet50", src=256, out=224, pad=0, classes=1000, batch=256,
                    name="resnet50_224_syncbn_ddp_augment"),
    "r50_64": dict(metric="resnet50_64_bf16_ddp_images_per_sec", model="resnet50", src=64, out=64, pad=0,
                   classes=200, batch=256, name="resnet50_tinyimagenet64_ddp"),
    "r18_cifar": dict(metric="resnet18_cifar32_bf16_syncbn_ddp_images_per_sec", model="resnet18", src=32, out=32,
                      pad=4, classes=10, batch=128, name="resnet18_cifar32_syncbn_ddp"),
}


# ------------------------------------------------------------------ reference / CPU arm
CPU_WORLD = 2  # torchrun --nproc_per_node=2, gloo, OMP_NUM_THREADS = cores / 2


def _cpu_rank(rank, world, port, steps, warmup, batch, threads, out_q):
    """One rank of the reference's stack on host cores: per-sample torchvision transforms
    (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53 minus the resize), stock DDP over
    gloo (01_torch_distributor/01_basic_torch_distributor.py:269,291 with backend gloo),
    SGD(lr=1e-3, momentum=0.5) (:294), torchvision ResNet-50 fp32."""
    import torchvision
    from PIL import Image
    from torch.nn.parallel import DistributedDataParallel as DDP
    from torchvision import transforms as T

    torch.set_num_threads(threads)
    store = dist.TCPStore("127.0.0.1", port, world, rank == 0)
    dist.init_process_group("gloo", store=store, rank=rank, world_size=world)
    torch.manual_seed(0)
    model = DDP(torchvision.models.resnet50(weights=None, num_classes=CLASSES))
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.5)
    pipeline = T.Compose([T.RandomCrop(OUT_HW), T.RandomHorizontalFlip(), T.ToTensor(),
                          T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])])
    g = torch.Generator().manual_seed(1 + rank)
    src = torch.randint(0, 256, (batch, SRC_HW, SRC_HW, 3), dtype=torch.uint8, generator=g).numpy()
    images = [Image.fromarray(a) for a in src]
    target = torch.randint(0, CLASSES, (batch,), generator=g)
    model.train()

    def step():
        x = torch.stack([pipeline(im) for im in images])
        opt.zero_grad()
        loss = F.cross_entropy(model(x), target)
        loss.backward()
        opt.step()
        return loss.item()

    for _ in range(warmup):
        step()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dist.barrier()
    dt = time.perf_counter() - t0
    dist.destroy_process_group()
    if rank == 0:
        out_q.put(dt)


def cpu_reference(steps, warmup, batch=16, world=CPU_WORLD):
    """The reference's own stack (its dependencies' code, unmodified) on the box's host cores: `world` gloo
    ranks as processes, each with cores/world threads and `batch` images per step."""
    import socket

    import torch.multiprocessing as mp

    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    env_keep = {k: os.environ.pop(k) for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT",
                                                  "CUDA_VISIBLE_DEVICES") if k in os.environ}
    os.environ["CUDA_VISIBLE_DEVICES"] = ""       # the CPU arm must not touch a GPU
    os.environ["OMP_NUM_THREADS"] = str(threads)
    try:
        procs = [ctx.Process(target=_cpu_rank, args=(r, world, port, steps, warmup, batch, threads, q)) for r in range(world)]
        for p in procs:
            p.start()
        dt = q.get(timeout=1800)
        for p in procs:
            p.join(timeout=120)
    finally:
        os.environ.pop("OMP_NUM_THREADS", None)
        os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        os.environ.update(env_keep)
    return {"value": batch * world * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": threads * world, "batch": batch,
            "steps": steps, "world": world}


def cpu_sample_text(r, warmup):
    return (f"stock torch DDP over gloo, {r['world']} ranks x {r['cores'] // r['world']} threads (the same for every --gpus N)"
            f" + torchvision CPU transforms + SGD, ResNet-50 224^2 fp32, batch {r['batch']} per rank, "
            f"{r['steps']} timed steps after {warmup} warm-up")


def run_reference_arm(args, rank):
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 4)), max(1, min(args.warmup, 1))
    r = cpu_reference(steps, warmup, batch=16)
    sample = cpu_sample_text(r, warmup) + f" (requested {args.steps}/{args.warmup}, bounded for CPU time)"
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "resnet50_224_syncbn_ddp_augment", "per_gpu_batch": r["batch"], "image": OUT_HW,
                   "parallelism": f"cpu-gloo-{r['world']}"},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": sample},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # NVML missing: report nulls rather than fail the bench
            self.nv = None
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.reasons.update(name for bit, name in self.REASONS.items() if mask & bit)
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        if self.nv:
            self.thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.nv:
            self.thread.join()

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------ per-kernel pass
def algorithmic_bytes(name, a, basis="survey-8d"):
    """Bytes of one call, from its arguments.  basis "survey-8d": SURVEY.md section 8(d)'s per-element figures
    (bf16: statistics 2 B, normalise 4 B (+2 residual), backward reduce 4 B, backward apply 6 B (+2 d_residual);
    a one-kernel layer = the sum of its two passes).  basis "as-moved": what the kernels here must actually
    touch — the same plus the forward output re-read for the ReLU mask of residual layers (2 B per backward
    pass, or 0.5 B with mask words)."""
    es = lambda code: 2 if code == 1 else 4  # noqa: E731
    moved = basis == "as-moved"
    mask = (lambda mode: {2: 1, 3: 0.25}.get(mode, 0)) if moved else (lambda mode: 0)
    if name == "b200dp_bn_fwd_stats":
        return a[2] * a[3] * es(a[1])
    if name == "b200dp_bn_fwd_apply":
        return a[4] * a[5] * es(a[3]) * (2 + (1 if a[1] else 0) + (0.25 if moved and len(a) > 20 and a[20] else 0))
    if name == "b200dp_bn_forward":      # one kernel: both passes
        return a[4] * a[5] * es(a[3]) * (3 + (1 if a[1] else 0) + (0.25 if moved and a[19] else 0))
    if name == "b200dp_bn_bwd_reduce":
        return a[4] * a[5] * es(a[3]) * (2 + mask(a[10]))
    if name == "b200dp_bn_bwd_apply":
        return a[6] * a[7] * es(a[5]) * (3 + mask(a[14]) + (1 if a[4] else 0))
    if name == "b200dp_bn_backward":     # one kernel: both passes
        return a[6] * a[7] * es(a[5]) * (5 + 2 * mask(a[11]) + (1 if a[4] else 0))
    if name == "b200dp_augment_u8_bf16":
        return a[1] * a[6] * a[7] * 3 * 3
    if name in ("b200dp_maxpool_fwd", "b200dp_maxpool_bwd"):  # x (or dx) once, y (or dy) + one tap byte per element
        N, H, W, C, k, st, pad = a[4], a[5], a[6], a[7], a[8], a[9], a[10]
        ho, wo = (H + 2 * pad - k) // st + 1, (W + 2 * pad - k) // st + 1
        return N * C * (H * W * es(a[3]) + ho * wo * (es(a[3]) + 1))
    if name == "b200dp_allreduce_update":  # HBM side on one rank: g + master r/w + momentum r/w + param copy
        e = es(a[2])
        return a[6] * (e + 8 + 8 + (e if a[7] else 0))
    return 0


def traffic_for(entry):
    """(DRAM bytes per launch, source) of that kernel from a committed `ncu --set full` capture of one
    training step (profiles/r*_traffic_*.json; NOT measured in this run), or (None, None)."""
    for rnd in ("r2", "r1"):
        path = os.path.join(ROOT, "profiles", f"{rnd}_traffic_" + entry.replace("b200dp_", "") + ".json")
        try:
            d = json.load(open(path))
            return d["traffic_bytes_per_launch"], f"profiles/{os.path.basename(path)} ({d.get('captured', 'earlier capture')})"
        except Exception:
            continue
    return None, None


class KernelProfiler:
    def __init__(self):
        self.records = []

    def __call__(self, name, fn, args):
        # every caller launches on torch's current stream (syncbn/transform: the ambient one,
        # ddp: inside `with torch.cuda.stream(comm_stream)`); record on exactly that stream
        s = torch.cuda.current_stream()
        assert s.cuda_stream == int(args[-1] or 0), "kernel was not launched on the current stream"
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        rc = fn(*args)
        e1.record(s)
        shape = None
        if name == "b200dp_bn_fwd_stats":
            shape = (args[2], args[3])
        elif name == "b200dp_bn_fwd_apply":
            shape = (args[4], args[5], "res" if args[1] else "-")
        elif name == "b200dp_bn_bwd_reduce":
            shape = (args[4], args[5], f"mask{args[10]}")
        elif name == "b200dp_bn_bwd_apply":
            shape = (args[6], args[7], f"mask{args[14]}", "dres" if args[4] else "-")
        elif name == "b200dp_bn_forward":
            shape = (args[4], args[5], "res" if args[1] else "-", "one-kernel")
        elif name == "b200dp_bn_backward":
            shape = (args[6], args[7], f"mask{args[11]}", "dres" if args[4] else "-", "one-kernel")
        elif name == "b200dp_allreduce_update":
            shape = ("bucket", int(args[6]) * (2 if args[2] == 1 else 4), int(args[5]))
        self.records.append((name, algorithmic_bytes(name, args), e0, e1, shape, algorithmic_bytes(name, args, "as-moved")))
        return rc

    def detail(self):
        """Per (kernel, layer shape): launches, mean microseconds, GB/s of algorithmic bytes."""
        torch.cuda.synchronize()
        out = {}
        for name, nbytes, e0, e1, shape, _ in self.records:
            if shape is None:
                continue
            t = out.setdefault((name, shape), [0, 0.0, 0])
            t[0] += 1
            t[1] += e0.elapsed_time(e1)
            t[2] += nbytes
        return out

    def table(self):
        torch.cuda.synchronize()
        out = {}
        for name, nbytes, e0, e1, _, moved in self.records:
            t = out.setdefault(name, {"launches": 0, "ms": 0.0, "bytes": 0, "bytes_moved": 0})
            t["launches"] += 1
            t["ms"] += e0.elapsed_time(e1)
            t["bytes"] += nbytes
            t["bytes_moved"] += moved
        return out

    def buckets(self, world):
        """Per gradient bucket (in launch order, averaged over the recorded steps): bytes, microseconds from
        launch to `done` on all ranks, and bus GB/s = S/t * 2(N-1)/N (the nccl-tests convention)."""
        torch.cuda.synchronize()
        per = {}
        order = []
        for name, _, e0, e1, shape, _ in self.records:
            if name != "b200dp_allreduce_update":
                continue
            if shape not in per:
                order.append(shape)
            t = per.setdefault(shape, [0, 0.0])
            t[0] += 1
            t[1] += e0.elapsed_time(e1)
        out = []
        for shape in order:
            n, ms = per[shape]
            us = ms / n * 1e3
            out.append({"bytes": shape[1], "us": round(us, 1),
                        "bus_gbs": round(shape[1] / (us * 1e-6) * 2 * (world - 1) / world / 1e9, 1) if world > 1 else None})
        return out


# ------------------------------------------------------------------ cross-rank parity probe (N > 1)
def parity_probe(ddp, dev, rank, world):
    """Outside the timed region, at the world size of this run: the two cross-rank exchanges against the NCCL
    path they replace, on the same inputs — a 4 MB gradient bucket through the fused all-reduce kernel vs
    dist.all_reduce (fp32 and bf16), one [32, 256, 14, 14] layer through B200SyncBatchNorm vs nn.SyncBatchNorm
    (a different batch on every rank) — and whether the replicas' parameters are still bit-identical after the
    timed steps.  Tolerances are north_star's: reduced gradients rtol 1e-5 (fp32) / 2e-2 (bf16)."""
    import ctypes

    import dbx_distributed_pytorch_examples_b200 as b200
    from dbx_distributed_pytorch_examples_b200 import _lib
    from dbx_distributed_pytorch_examples_b200.comm import Communicator

    out = {}
    lib = _lib.lib()
    for dtype, key in ((torch.float32, "allreduce_fp32_max_rel"), (torch.bfloat16, "allreduce_bf16_max_rel")):
        es = 2 if dtype == torch.bfloat16 else 4
        n = (4 << 20) // es
        comm = Communicator(heap_bytes=2 * n * es, device=dev)
        grad = comm.heap_view(0, n, dtype)
        torch.manual_seed(31 + rank)
        g = (torch.randn(n, device=dev) / world).to(dtype)
        grad.copy_(g)
        want = g.float().clone()
        dist.all_reduce(want)
        torch.cuda.synchronize(); dist.barrier()
        o = _lib.OptArgs()
        o.kind = _lib.OPT_NONE
        _lib.check(lib.b200dp_allreduce_update(comm.handle, 0, _lib.dtype_code(dtype), n * es, _lib.dtype_code(dtype), 0, n,
                                               None, None, None, ctypes.byref(o), 0, 1, 0, comm.next_seq(),
                                               _lib.current_stream()), "b200dp_allreduce_update")
        torch.cuda.synchronize()
        ref = want.to(dtype).float()
        rel = (grad.float() - ref).abs().max() / ref.pow(2).mean().sqrt()   # relative to the rms of the sums
        dist.all_reduce(rel, op=dist.ReduceOp.MAX)
        out[key] = float(rel)
        comm.check_error()
        dist.barrier()
        comm.destroy()
    for dtype, tag in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        torch.manual_seed(77 + rank)
        c = 256
        x = (torch.randn(32 - 2 * rank, c, 14, 14, device=dev) * 2 + 1).to(dtype).contiguous(memory_format=torch.channels_last)
        dy = torch.randn_like(x)
        ours, ref = b200.B200SyncBatchNorm(c).to(dev), torch.nn.SyncBatchNorm(c).to(dev)
        xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
        ya, yb = ours(xa), ref(xb)
        ya.backward(dy); yb.backward(dy)
        dy_max = torch.stack([(ya.float() - yb.float()).abs().max(), (xa.grad.float() - xb.grad.float()).abs().max(),
                              (ours.running_var - ref.running_var).abs().max()])
        dist.all_reduce(dy_max, op=dist.ReduceOp.MAX)
        out[f"syncbn_{tag}_y_max_abs"], out[f"syncbn_{tag}_dx_max_abs"] = float(dy_max[0]), float(dy_max[1])
        out[f"syncbn_{tag}_running_var_max_abs"] = float(dy_max[2])
    flat = torch.cat([p.detach().reshape(-1).view(torch.int16 if p.dtype == torch.bfloat16 else torch.int32).to(torch.int64)
                      for p in ddp.module.parameters()])
    digest = torch.stack([flat.sum(), (flat * torch.arange(1, flat.numel() + 1, device=dev) % 1000003).sum()])
    all_d = [torch.empty_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    out["params_identical_across_ranks"] = bool(all(torch.equal(all_d[0], d) for d in all_d))
    out["tolerance"] = ("north_star: reduced gradients rtol 1e-5 fp32 / 2e-2 bf16 (allreduce_*_max_rel = max |ours - NCCL| over the "
                        "rms of the sums); SyncBN vs nn.SyncBatchNorm+NCCL on the same inputs, max abs difference")
    return out


# ------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "torch-gpu"],
                    help="b200 = this library; reference = the reference stack on host cores; torch-gpu = stock torch "
                         "BatchNorm/DDP(NCCL)/optim on the same GPUs (context only, not part of the contract)")
    ap.add_argument("--workload", default="r50_224", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the workload's own, 256 for r50_224)")
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "adam"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-pass", action="store_true")
    ap.add_argument("--kernel-detail", default="", help="write per-layer-shape kernel timings of the kernel pass to this file")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run for N > 1")
    if args.warmup < 3:
        raise SystemExit("at least 3 warm-up steps are required by the timing rules")

    import dbx_distributed_pytorch_examples_b200 as b200
    from dbx_distributed_pytorch_examples_b200 import _lib, workloads

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True

    wl = WORKLOADS[args.workload]
    global METRIC, SRC_HW, OUT_HW, CLASSES
    METRIC, SRC_HW, OUT_HW, CLASSES = wl["metric"], wl["src"], wl["out"], wl["classes"]
    torch.manual_seed(0)
    model = workloads.resnet50(CLASSES) if wl["model"] == "resnet50" else workloads.cifar_resnet18(CLASSES)
    stock = args.impl == "torch-gpu"
    if stock:  # stock torch on the same GPU(s): nn.SyncBatchNorm, DistributedDataParallel over NCCL, torch.optim
        if world > 1:
            model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
        model = model.to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank]) if world > 1 else model
    else:
        model = b200.fuse_bn_relu(b200.convert_sync_batchnorm(model)).to(dev)
        model = model.to(torch.bfloat16).to(memory_format=torch.channels_last)
        ddp = b200.B200DDP(model, device_ids=[local_rank], output_device=local_rank)
    if args.optimizer == "sgd":  # 01_basic_torch_distributor.py:294
        opt = torch.optim.SGD(ddp.parameters(), lr=1e-3, momentum=0.5)
    else:                        # 02_cifar_torch_distributor_resnet.py:213
        opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)
    if not stock:
        ddp.fuse_optimizer(opt)
    cifar = wl["pad"] > 0
    tf = b200.BatchTransform(OUT_HW, padding=wl["pad"], device=dev,
                             mean=b200.transform.CIFAR_MEAN if cifar else b200.transform.IMAGENET_MEAN,
                             std=b200.transform.CIFAR_STD if cifar else b200.transform.IMAGENET_STD)

    B = args.batch or wl["batch"]
    host_u8 = workloads.synthetic_u8_images(B, SRC_HW, seed=100 + rank, pin=True)
    host_y = torch.randint(0, CLASSES, (B,), generator=torch.Generator().manual_seed(100 + rank)).pin_memory()
    dev_u8, dev_y = host_u8.to(dev), host_y.to(dev)
    gen = torch.Generator().manual_seed(1234 + rank)
    ddp.train()

    def step(images, labels):
        x = tf(images, generator=gen)
        y = labels if labels.is_cuda else labels.to(dev, non_blocking=True)
        opt.zero_grad()
        loss = F.cross_entropy(ddp(x).float(), y)
        loss.backward()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(args.warmup):
        step(dev_u8, dev_y)
    launches0 = int(_lib.lib().b200dp_kernel_launches())
    with ClockSampler(local_rank) as clocks:
        ms_dev = timed(lambda: step(dev_u8, dev_y), args.steps)
    launches = int(_lib.lib().b200dp_kernel_launches()) - launches0   # counted by the library at each launch
    loss_dev = float(step(dev_u8, dev_y).detach())

    # End to end through the public API: every step's source batch starts in pinned HOST memory
    # (DeviceFeeder copies batch k+1 while step k computes) and every step's loss is read on the
    # host (LossReader, one step late so the GPU queue never drains); both inside the timed region.
    def e2e_epoch(n):
        reader = b200.LossReader(dev)
        feeder = b200.DeviceFeeder(((host_u8, host_y) for _ in range(n)), dev)
        for images, labels in feeder:
            reader.push(step(images, labels))
        return reader.flush()

    e2e_epoch(3)
    e2e_losses = []
    ms_e2e = timed(lambda: e2e_losses.extend(e2e_epoch(args.steps)), 1)
    assert len(e2e_losses) == args.steps
    if not stock:
        ddp.comm.check_error()

    images = B * world * args.steps
    value = images / (ms_dev / 1e3)
    e2e = images / (ms_e2e / 1e3)

    line = {
        **({"impl": "torch-gpu"} if stock else {}),
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": wl["name"], "global_batch": B * world, "per_gpu_batch": B,
                   "source_image": SRC_HW, "image": OUT_HW, "classes": CLASSES, "optimizer": args.optimizer,
                   "parallelism": f"dp{world}", "l2": ("inputs_exceed_l2 (>=10 GB of activations per step)" if args.workload == "r50_224"
                          else "working set of a step exceeds L2; no explicit flush")},
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": host_u8.numel() + host_y.numel() * 8 + B * 12, "d2h_bytes_per_step": 4},
        "gpu_launches": launches,
        "e2e_note": "H2D of each uint8 source batch prefetched one step ahead on a copy stream; loss read back one step late",
        "clocks": clocks.summary(),
        "loss": loss_dev,
    }

    if world > 1 and not stock:
        line["parity"] = parity_probe(ddp, dev, rank, world)

    peak, peaks = 6650.0, {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak = peaks.get("hbm_gbs", peak)
    except Exception:
        pass
    if not args.no_kernel_pass and not stock:
        line["roofline_peak_source"] = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        # every rank runs the instrumented steps (they contain collectives); rank 0 reports
        prof = KernelProfiler()
        _lib.PROFILER = prof
        for _ in range(2):
            step(dev_u8, dev_y)
        _lib.PROFILER = None
        table = prof.table()
        if rank == 0 and args.kernel_detail:
            with open(args.kernel_detail, "w") as f:
                for (name, shape), (n, ms, nb) in sorted(prof.detail().items(), key=lambda kv: (kv[0][0], -kv[1][1])):
                    f.write(f"{name:24s} {str(shape):44s} n={n:3d} {ms / n * 1e3:8.1f} us  {nb / ms / 1e6:7.0f} GB/s\n")
        if rank == 0:
            bn = [k for k in table if k.startswith("b200dp_bn_") and table[k]["bytes"] > 0]
            top = max((k for k in table if table[k]["bytes"] > 0 and k != "b200dp_allreduce_update"), key=lambda k: table[k]["ms"])
            t = table[top]
            achieved = t["bytes"] / (t["ms"] / 1e3) / 1e9
            moved = t["bytes_moved"] / (t["ms"] / 1e3) / 1e9
            step_ms = ms_dev / args.steps
            traffic, traffic_source = traffic_for(top)
            line["roofline"] = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                                "frac": achieved / peak, "bytes_basis": "survey-8d",
                                "achieved_as_moved": moved, "frac_as_moved": moved / peak,
                                "traffic": traffic, "traffic_source": traffic_source,
                                "algorithmic_bytes_per_launch": t["bytes"] / max(t["launches"], 1),
                                "launches_per_step": t["launches"] // 2, "ms_per_step": t["ms"] / 2,
                                "share_of_step": t["ms"] / 2 / step_ms}
            line["kernels"] = {k: {"launches_per_step": v["launches"] // 2, "ms_per_step": round(v["ms"] / 2, 4),
                                   "gbps": round(v["bytes"] / (v["ms"] / 1e3) / 1e9, 1) if v["ms"] > 0 and k != "b200dp_allreduce_update" else None,
                                   "frac": round(v["bytes"] / (v["ms"] / 1e3) / 1e9 / peak, 3) if v["ms"] > 0 and k.startswith("b200dp_bn_") else None,
                                   "frac_as_moved": round(v["bytes_moved"] / (v["ms"] / 1e3) / 1e9 / peak, 3) if v["ms"] > 0 and k.startswith("b200dp_bn_") else None}
                               for k, v in sorted(table.items())}
            if bn:   # the SyncBN family as a whole: section 8(d) bytes over the summed kernel time
                tb, tm = sum(table[k]["bytes"] for k in bn), sum(table[k]["ms"] for k in bn)
                line["syncbn_family"] = {"ms_per_step": round(tm / 2, 3), "gb_per_step": round(tb / 2 / 1e9, 2),
                                         "gbps": round(tb / (tm / 1e3) / 1e9, 1), "frac": round(tb / (tm / 1e3) / 1e9 / peak, 3)}
            bk = prof.buckets(world)
            if world > 1:   # BASELINE.json's second metric: grad-allreduce bus GB/s, per bucket, in-step
                line["allreduce_buckets"] = bk
                tot_b, tot_us = sum(b["bytes"] for b in bk), sum(b["us"] for b in bk)
                line["allreduce_bus_gbs"] = round(tot_b / (tot_us * 1e-6) * 2 * (world - 1) / world / 1e9, 1)
                line["allreduce_note"] = ("in-step: the bucket kernel runs on a side stream with a bounded grid while the "
                                          "backward continues; t = launch to `done` on all ranks incl. rank skew. "
                                          "Stand-alone sweep: tools/allreduce_sweep.py")
            else:
                line["allreduce_buckets"] = bk

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference(steps=3, warmup=1, batch=16)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "reference",
                                "sample": cpu_sample_text(r, 1)}
    if world > 1:
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
