"""Oracle: torch.optim.SGD / Adam single steps and DDP gradient averaging (numpy fp32)."""
import math

import numpy as np

f32 = np.float32


def ddp_average(grads):
    """What DDP leaves in .grad: every rank's bucket is divided by the world size
    BEFORE the sum (TORCH/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-32;
    Reducer div_factor_), so the result is sum_r(g_r / N), not (sum_r g_r) / N."""
    n = f32(len(grads))
    acc = np.zeros_like(grads[0], dtype=np.float32)
    for g in grads:
        acc = acc + g.astype(np.float32) / n
    return acc


def sgd_step(p, g, buf, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, first=False):
    """TORCH/optim/sgd.py:322-380 (_single_tensor_sgd).  Returns (p, buf)."""
    p, g = p.astype(np.float32), g.astype(np.float32)
    if weight_decay != 0:
        g = g + f32(weight_decay) * p
    if momentum != 0:
        if first or buf is None:
            buf = g.copy()
        else:
            buf = buf.astype(np.float32) * f32(momentum) + f32(1 - dampening) * g
        g = g + f32(momentum) * buf if nesterov else buf
    return p - f32(lr) * g, buf


def adam_step(p, g, m, v, step, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
    """TORCH/optim/adam.py:347+ (_single_tensor_adam, non-capturable, no amsgrad).
    `step` is the 1-based step count; bias corrections are python doubles."""
    b1, b2 = betas
    p, g, m, v = (a.astype(np.float32) for a in (p, g, m, v))
    if weight_decay != 0:
        g = g + f32(weight_decay) * p
    m = m + f32(1 - b1) * (g - m)                 # exp_avg.lerp_(grad, 1 - beta1)
    v = v * f32(b2) + f32(1 - b2) * g * g         # exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    step_size = lr / bc1
    denom = np.sqrt(v) / f32(math.sqrt(bc2)) + f32(eps)
    p = p - f32(step_size) * (m / denom)          # param.addcdiv_(exp_avg, denom, value=-step_size)
    return p, m, v
