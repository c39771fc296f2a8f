#!/usr/bin/env python
"""Where does the HOST time of a training step go?  cProfile over a few steps of a bench workload, plus the
wall time the host needs to enqueue a step (no synchronisation) next to the GPU time of the step."""
import cProfile
import os
import pstats
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dbx_distributed_pytorch_examples_b200 as b200  # noqa: E402
from dbx_distributed_pytorch_examples_b200 import workloads  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "r50_64"
stock = len(sys.argv) > 2 and sys.argv[2] == "stock"
cfg = {"r50_64": (workloads.resnet50, 200, 64, 256), "r50_224": (workloads.resnet50, 1000, 224, 256),
       "r18_cifar": (workloads.cifar_resnet18, 10, 32, 128)}[which]
dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
model = cfg[0](cfg[1])
if stock:
    ddp = model.to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(ddp.parameters(), lr=1e-3, momentum=0.5)
else:
    ddp = b200.prepare_model(model, dev, dtype=torch.bfloat16)
    opt = ddp.fuse_optimizer(torch.optim.SGD(ddp.parameters(), lr=1e-3, momentum=0.5))
x = torch.randn(cfg[3], 3, cfg[2], cfg[2], device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, cfg[1], (cfg[3],), device=dev)


def step():
    opt.zero_grad()
    loss = F.cross_entropy(ddp(x).float(), y)
    loss.backward()
    opt.step()


for _ in range(10):
    step()
torch.cuda.synchronize()
n = 20
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(n):
    step()
t_host = time.perf_counter() - t0
e1.record(); torch.cuda.synchronize()
print(f"{which} {'stock' if stock else 'b200'}: host enqueue {t_host/n*1e3:.2f} ms/step, GPU {e0.elapsed_time(e1)/n:.2f} ms/step")
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(18)
