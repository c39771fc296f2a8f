#!/usr/bin/env python
"""How much does cuDNN's stem convolution (7x7/2, 3 -> 64 channels, bf16 channels_last, batch 256, 224^2) gain
from an input padded to 4 or 8 channels with zeros (same result, tensor-core friendly layout)?"""
import torch
import torch.nn.functional as F

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")
B = 256
for cin in (3, 4, 8):
    x = torch.randn(B, cin, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    w = torch.randn(64, cin, 7, 7, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    dy = torch.randn(B, 64, 112, 112, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    for _ in range(3):
        y = F.conv2d(x, w, None, 2, 3); y.backward(dy); w.grad = None
    torch.cuda.synchronize()
    tf, tb = [], []
    for _ in range(10):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(); y = F.conv2d(x, w, None, 2, 3); e1.record(); y.backward(dy); e2.record(); torch.cuda.synchronize()
        tf.append(e0.elapsed_time(e1)); tb.append(e1.elapsed_time(e2)); w.grad = None
    tf.sort(); tb.sort()
    print(f"Cin={cin}: forward {tf[5]*1e3:7.1f} us   weight-grad {tb[5]*1e3:7.1f} us")
# the cost of padding: one pass over the bf16 input batch
x3 = torch.randn(B, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    x4 = torch.zeros((B, 4, 224, 224), dtype=torch.bfloat16, device=dev).contiguous(memory_format=torch.channels_last)
    x4[:, :3] = x3
    e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ts.sort(); print(f"padding 3 -> 4 channels with torch ops: {ts[5]*1e3:.1f} us")
