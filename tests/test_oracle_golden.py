"""CPU: the oracle against the golden vectors made from torch / torchvision / the
reference model (oracle/make_golden.py).  No product code involved."""
import numpy as np

from oracle import ddp_ref, optim_ref, syncbn_ref, transform_ref


def test_bf16_rounding_matches_torch():
    import torch
    x = np.random.default_rng(0).standard_normal(4096).astype(np.float32) * 3
    x[:4] = [0.0, -0.0, 1.0, 65504.0]
    want = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(transform_ref.f32_to_bf16_bits(x), want)


def _check_transform(g):
    out = transform_ref.crop_flip_normalize(g["src"], g["params"], int(g["pad"]), (int(g["out_hw"]),) * 2,
                                            g["mean"], g["std"])
    # float32 arithmetic restated op for op: bit-exact, and so is the bf16 rounding
    assert np.array_equal(out, g["out_f32"])
    assert np.array_equal(transform_ref.f32_to_bf16_bits(out), g["out_bf16"])


def test_transform_cifar_recipe(golden):
    _check_transform(golden("transform_cifar.npz"))


def test_transform_crop_no_padding(golden):
    _check_transform(golden("transform_crop56.npz"))


def test_syncbn_forward_backward(golden):
    g = golden("syncbn.npz")
    eps, mom = float(g["eps"]), float(g["momentum"])
    for tag, relu, res in [("plain", False, False), ("relu", True, False), ("addrelu", True, True)]:
        xs = [g[f"{tag}_x0"], g[f"{tag}_x1"]]
        rs = [g[f"{tag}_r0"], g[f"{tag}_r1"]] if res else None
        ys, mean, invstd, counts, rm, rv = syncbn_ref.forward(xs, g[f"{tag}_w"], g[f"{tag}_b"], g[f"{tag}_rm"],
                                                              g[f"{tag}_rv"], eps, mom, relu, rs)
        np.testing.assert_allclose(ys[0], g[f"{tag}_y0"], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(ys[1], g[f"{tag}_y1"], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(rm, g[f"{tag}_rm_new"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(rv, g[f"{tag}_rv_new"], rtol=1e-5, atol=1e-7)
        dxs, drs, gws, gbs = syncbn_ref.backward([g[f"{tag}_dy0"], g[f"{tag}_dy1"]], xs, ys, g[f"{tag}_w"],
                                                 mean, invstd, relu)
        np.testing.assert_allclose(dxs[0], g[f"{tag}_dx0"], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(dxs[1], g[f"{tag}_dx1"], rtol=1e-4, atol=1e-5)
        # SyncBatchNorm returns rank-local weight/bias grads; DDP's sum over ranks is autograd's total
        np.testing.assert_allclose(gws[0] + gws[1], g[f"{tag}_dw"], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(gbs[0] + gbs[1], g[f"{tag}_db"], rtol=1e-4, atol=1e-5)
        if res:
            np.testing.assert_allclose(drs[0], g[f"{tag}_dr0"], rtol=1e-6, atol=0)


def test_optimizer_steps(golden):
    g = golden("optim.npz")
    p0, grads = g["p0"], g["grads"]

    def run_sgd(**kw):
        p, buf, out = p0.copy(), None, []
        for k, gr in enumerate(grads):
            p, buf = optim_ref.sgd_step(p, gr, buf, first=(k == 0), **kw)
            out.append(p)
        return np.stack(out)

    def run_adam(**kw):
        p, m, v, out = p0.copy(), np.zeros_like(p0), np.zeros_like(p0), []
        for k, gr in enumerate(grads):
            p, m, v = optim_ref.adam_step(p, gr, m, v, k + 1, **kw)
            out.append(p)
        return np.stack(out)

    np.testing.assert_allclose(run_sgd(lr=1e-3, momentum=0.5), g["sgd_m05"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(run_sgd(lr=1e-2, momentum=0.9, nesterov=True, weight_decay=1e-4),
                               g["sgd_nesterov_wd"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(run_sgd(lr=1e-2), g["sgd_plain"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(run_adam(lr=1e-3), g["adam"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(run_adam(lr=1e-3, weight_decay=1e-4), g["adam_wd"], rtol=1e-5, atol=1e-7)


def test_ddp_average_divides_before_summing():
    rng = np.random.default_rng(3)
    gs = [rng.standard_normal(1000).astype(np.float32) for _ in range(3)]
    want = (gs[0] / np.float32(3) + gs[1] / np.float32(3)) + gs[2] / np.float32(3)
    assert np.array_equal(optim_ref.ddp_average(gs), want)


def test_sampler_and_bucket_assignment(golden):
    g = golden("sampler_buckets.npz")
    for world in (2, 4):
        for rank in range(world):
            for epoch in (0, 1):
                got = ddp_ref.distributed_sampler_indices(1003, world, rank, g[f"perm_e{epoch}"])
                assert np.array_equal(got, g[f"w{world}_r{rank}_e{epoch}"])
    buckets = ddp_ref.bucket_assignment(g["resnet18_param_bytes_rev"].tolist())
    bucket_of = np.empty(len(g["resnet18_param_bytes_rev"]), dtype=np.int64)
    for b, members in enumerate(buckets):
        bucket_of[members] = b
    assert np.array_equal(bucket_of, g["resnet18_bucket_of_param"])


def test_maxpool_forward_backward(golden):
    """oracle/pool_ref.py against torch's own max_pool2d + autograd on tie-heavy inputs."""
    from oracle import pool_ref
    g = golden("maxpool.npz")
    for tag in ("stem", "odd", "k2", "k3p0"):
        k, s, p = (int(v) for v in g[f"{tag}_geom"])
        x = g[f"{tag}_x"]
        y, tap = pool_ref.maxpool_forward(x, k, s, p)
        assert np.array_equal(y, g[f"{tag}_y"])
        dx = pool_ref.maxpool_backward(g[f"{tag}_dy"], tap, x.shape[1:3], k, s, p)
        np.testing.assert_allclose(dx, g[f"{tag}_dx"], rtol=1e-6, atol=1e-6)
