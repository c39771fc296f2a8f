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
def cpu_reference(steps, warmup, batch=16, threads=None):
    """The reference's stack on host cores: per-sample torchvision transforms
    (02_deepspeed/03_1k_imagenet_deepspeed_resnet.py:45-53 minus the resize), stock DDP over
    gloo (01_torch_distributor/01_basic_torch_distributor.py:269,291 with backend gloo),
    SGD(lr=1e-3, momentum=0.5) (:294), torchvision ResNet-50 fp32."""
    import torchvision
    from PIL import Image
    from torch.nn.parallel import DistributedDataParallel as DDP
    from torchvision import transforms as T

    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    own_group = not dist.is_initialized()
    if own_group:
        store = dist.TCPStore("127.0.0.1", 0, 1, True)
        dist.init_process_group("gloo", store=store, rank=0, world_size=1)
    torch.manual_seed(0)
    model = DDP(torchvision.models.resnet50(weights=None, num_classes=CLASSES))
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.5)
    pipeline = T.Compose([T.RandomCrop(OUT_HW), T.RandomHorizontalFlip(), T.ToTensor(),
                          T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])])
    g = torch.Generator().manual_seed(1)
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
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    if own_group:
        dist.destroy_process_group()
    return {"value": batch * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": threads, "batch": batch,
            "steps": steps}


def run_reference_arm(args, rank):
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 4)), max(1, min(args.warmup, 1))
    r = cpu_reference(steps, warmup, batch=16)
    sample = (f"stock torch DDP(gloo, world 1)+torchvision CPU transforms+SGD, ResNet-50 fp32, batch {r['batch']}, "
              f"{steps} timed steps after {warmup} warm-up (requested {args.steps}/{args.warmup}, bounded for CPU time)")
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "resnet50_224_syncbn_ddp_augment", "per_gpu_batch": r["batch"], "image": OUT_HW,
                   "parallelism": "cpu-gloo-1"},
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
def algorithmic_bytes(name, a):
    """Bytes each entry point must move at minimum, from its arguments."""
    es = lambda code: 2 if code == 1 else 4  # noqa: E731
    if name == "b200dp_bn_fwd_stats":
        return a[2] * a[3] * es(a[1])
    if name == "b200dp_bn_fwd_apply":
        return a[4] * a[5] * es(a[3]) * (2 + (1 if a[1] else 0) + (0.25 if len(a) > 20 and a[20] else 0))
    if name == "b200dp_bn_bwd_reduce":
        return a[4] * a[5] * es(a[3]) * (2 + {2: 1, 3: 0.25}.get(a[10], 0))
    if name == "b200dp_bn_bwd_apply":
        return a[6] * a[7] * es(a[5]) * (3 + {2: 1, 3: 0.25}.get(a[14], 0) + (1 if a[4] else 0))
    if name == "b200dp_augment_u8_bf16":
        return a[1] * a[6] * a[7] * 3 * 3
    if name in ("b200dp_maxpool_fwd", "b200dp_maxpool_bwd"):  # x (or dx) once, y (or dy) + one tap byte per element
        N, H, W, C, k, st, pad = a[4], a[5], a[6], a[7], a[8], a[9], a[10]
        ho, wo = (H + 2 * pad - k) // st + 1, (W + 2 * pad - k) // st + 1
        return N * C * (H * W * es(a[3]) + ho * wo * (es(a[3]) + 1))
    if name == "b200dp_allreduce_update":  # world 1: g + master r/w + momentum r/w + param copy
        e = es(a[2])
        return a[6] * (e + 8 + 8 + (e if a[7] else 0))
    return 0


def traffic_for(entry):
    """DRAM bytes (read + write) per launch of that kernel from the committed `ncu --set full` capture of
    one training step (profiles/r1_traffic_*.json), or None when no capture exists for it."""
    path = os.path.join(ROOT, "profiles", "r1_traffic_" + entry.replace("b200dp_", "") + ".json")
    try:
        return json.load(open(path))["traffic_bytes_per_launch"]
    except Exception:
        return None


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
        self.records.append((name, algorithmic_bytes(name, args), e0, e1, shape))
        return rc

    def detail(self):
        """Per (kernel, layer shape): launches, mean microseconds, GB/s of algorithmic bytes."""
        torch.cuda.synchronize()
        out = {}
        for name, nbytes, e0, e1, shape in self.records:
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
        for name, nbytes, e0, e1, _ in self.records:
            t = out.setdefault(name, {"launches": 0, "ms": 0.0, "bytes": 0})
            t["launches"] += 1
            t["ms"] += e0.elapsed_time(e1)
            t["bytes"] += nbytes
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
    _lib.reset_calls()
    with ClockSampler(local_rank) as clocks:
        ms_dev = timed(lambda: step(dev_u8, dev_y), args.steps)
    launches = sum(n * _lib.LAUNCHES_PER_CALL.get(k, 1) for k, n in _lib.total_calls().items())
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

    if rank == 0 and not args.no_kernel_pass and not stock:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = peaks.get("hbm_gbs", 6650.0)
        line["roofline_peak_source"] = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    if not args.no_kernel_pass and not stock:
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
                    f.write(f"{name:24s} {str(shape):32s} n={n:3d} {ms / n * 1e3:8.1f} us  {nb / ms / 1e6:7.0f} GB/s\n")
        if rank == 0:
            top = max((k for k in table if table[k]["bytes"] > 0), key=lambda k: table[k]["ms"])
            t = table[top]
            achieved = t["bytes"] / (t["ms"] / 1e3) / 1e9
            step_ms = ms_dev / args.steps
            line["roofline"] = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                                "frac": achieved / peak, "traffic": traffic_for(top),
                                "algorithmic_bytes_per_launch": t["bytes"] / max(t["launches"], 1),
                                "launches_per_step": t["launches"] // 2, "ms_per_step": t["ms"] / 2,
                                "share_of_step": t["ms"] / 2 / step_ms}
            line["kernels"] = {k: {"launches_per_step": v["launches"] // 2, "ms_per_step": round(v["ms"] / 2, 4),
                                   "gbps": round(v["bytes"] / (v["ms"] / 1e3) / 1e9, 1) if v["ms"] > 0 else None}
                               for k, v in sorted(table.items())}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference(steps=3, warmup=1, batch=16)
        line["cpu_baseline"] = {
            "value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "reference",
            "sample": "stock torch DDP(gloo, world 1)+torchvision CPU transforms+SGD, ResNet-50 224^2 fp32, "
                      "batch 16, 3 timed steps after 1 warm-up"}
    if world > 1:
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
