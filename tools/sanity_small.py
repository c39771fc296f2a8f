#!/usr/bin/env python
"""One small call of every kernel family (for `compute-sanitizer --tool memcheck`)."""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dbx_distributed_pytorch_examples_b200 as b200  # noqa: E402
from dbx_distributed_pytorch_examples_b200 import transform as T, workloads  # noqa: E402

dev = torch.device("cuda")
g = torch.Generator().manual_seed(0)
for (B, src, out, pad) in ((5, 40, 32, 0), (3, 32, 32, 4), (2, 30, 30, 3)):
    x = torch.randint(0, 256, (B, src, src + 1, 3), dtype=torch.uint8)
    for env in ({}, {"B200DP_AUGMENT_FORCE_STAGED": "1"}, {"B200DP_AUGMENT_FORCE_TABLE": "1"}):
        os.environ.update(env)
        b200.BatchTransform(out, padding=pad, mean=T.CIFAR_MEAN, std=T.CIFAR_STD)(x.to(dev), generator=g)
        for k in env:
            del os.environ[k]
torch.manual_seed(0)
for dtype in (torch.bfloat16, torch.float32):
    model = b200.prepare_model(workloads.cifar_resnet18(10), dev, dtype=dtype)
    opt = model.fuse_optimizer(torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.5))
    xb = torch.randn(8, 3, 32, 32, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    yb = torch.randint(0, 10, (8,), device=dev)
    for _ in range(2):
        opt.zero_grad()
        F.cross_entropy(model(xb).float(), yb).backward()
        opt.step()
    model.eval()
    with torch.no_grad():
        model(xb)
torch.cuda.synchronize()
print("sanity ok")
