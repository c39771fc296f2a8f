"""Oracle: SyncBatchNorm forward/backward over R ranks, restated in numpy.

Follows TORCH/nn/modules/_functions.py:7-209 step by step.  Inputs are per-rank
[M_r, C] float arrays (channels last, rows = N*H*W); all statistics are computed
in float64 and returned as float32, which is at least as accurate as the fp32
kernels on either side.
"""
import numpy as np


def forward(xs, weight, bias, running_mean, running_var, eps, momentum, relu=False, residuals=None):
    """Returns (ys, mean, invstd, counts, new_running_mean, new_running_var)."""
    C = xs[0].shape[1]
    means, invstds, counts = [], [], []
    for x in xs:  # batch_norm_stats (:39): per-rank mean and biased-variance invstd
        n = x.shape[0]
        if n == 0:  # (:50-57) empty ranks still take part, with zero count
            continue
        x64 = x.astype(np.float64)
        m = x64.mean(0)
        v = x64.var(0)
        means.append(m)
        invstds.append(1.0 / np.sqrt(v + eps))
        counts.append(float(n))
    # batch_norm_gather_stats_with_counts (:96-105): variance recovered as
    # 1/invstd^2 - eps, count-weighted merge (Chan), unbiased variance for running_var
    total = sum(counts)
    mean = sum(m * c for m, c in zip(means, counts)) / total
    var = sum(((1.0 / (s * s) - eps) + (m - mean) ** 2) * c for m, s, c in zip(means, invstds, counts)) / total
    invstd = 1.0 / np.sqrt(var + eps)
    new_rm = new_rv = None
    if running_mean is not None:
        new_rm = ((1 - momentum) * running_mean.astype(np.float64) + momentum * mean).astype(np.float32)
        unbiased = var * total / (total - 1)
        new_rv = ((1 - momentum) * running_var.astype(np.float64) + momentum * unbiased).astype(np.float32)
    w = np.ones(C) if weight is None else weight.astype(np.float64)
    b = np.zeros(C) if bias is None else bias.astype(np.float64)
    ys = []
    for r, x in enumerate(xs):  # batch_norm_elemt (:119-121) + the epilogues of setup/resnet18.py:23-26
        y = (x.astype(np.float64) - mean) * invstd * w + b
        if residuals is not None:
            y = y + residuals[r].astype(np.float64)
        if relu:
            y = np.maximum(y, 0.0)
        ys.append(y.astype(np.float32))
    return ys, mean.astype(np.float32), invstd.astype(np.float32), counts, new_rm, new_rv


def backward(dys, xs, ys, weight, mean, invstd, relu=False):
    """Returns (dxs, dresiduals, grad_weights, grad_biases): per-rank lists.

    grad_weight / grad_bias are the LOCAL sums, as SyncBatchNorm returns them
    (_functions.py:176-178: DDP averages them afterwards); dx uses the sums
    all-reduced over the ranks (:159-164) and the global count (:166-175).
    """
    C = xs[0].shape[1]
    w = np.ones(C) if weight is None else weight.astype(np.float64)
    mean64, invstd64 = mean.astype(np.float64), invstd.astype(np.float64)
    gs, sum_dy, sum_dy_xmu = [], [], []
    for dy, x, y in zip(dys, xs, ys):  # batch_norm_backward_reduce (:139-152)
        g = dy.astype(np.float64)
        if relu:
            g = g * (y > 0)
        gs.append(g)
        sum_dy.append(g.sum(0))
        sum_dy_xmu.append((g * (x.astype(np.float64) - mean64)).sum(0))
    total = float(sum(x.shape[0] for x in xs))
    S1, S2 = sum(sum_dy), sum(sum_dy_xmu)  # all_reduce SUM
    dxs, dres, gws, gbs = [], [], [], []
    for g, x, s1, s2 in zip(gs, xs, sum_dy, sum_dy_xmu):  # batch_norm_backward_elemt (:170-179)
        xm = x.astype(np.float64) - mean64
        dx = (g - S1 / total - xm * invstd64 * invstd64 * S2 / total) * invstd64 * w
        dxs.append(dx.astype(np.float32))
        dres.append(g.astype(np.float32))
        gws.append((s2 * invstd64).astype(np.float32))
        gbs.append(s1.astype(np.float32))
    return dxs, dres, gws, gbs
