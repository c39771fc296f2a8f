"""CPU: the resample oracle against the torchvision / Pillow golden vectors, the live Pillow
in this image, and the RandomResizedCrop parameter draw against the literal torch calls."""
import numpy as np
import pytest
import torch

from dbx_distributed_pytorch_examples_b200 import transform as T
from oracle import resample_ref, transform_ref

GOLDENS = ["resample_rrc56.npz", "resample_rrc224.npz", "resample_resize64.npz", "resample_down24.npz"]


@pytest.mark.parametrize("name", GOLDENS)
def test_oracle_matches_torchvision_golden(golden, name):
    g = golden(name)
    out_hw = (int(g["out_hw"]),) * 2
    for b in range(len(g["src"])):
        i, j, h, w, _ = (int(v) for v in g["params"][b])
        got = resample_ref.resize_bilinear_u8(g["src"][b, i:i + h, j:j + w], out_hw)
        assert np.array_equal(got, g["resized_u8"][b])  # the uint8 stage, bit for bit
    out = resample_ref.resized_crop_flip_normalize(g["src"], g["params"], out_hw, g["mean"], g["std"])
    assert np.array_equal(transform_ref.f32_to_bf16_bits(out), g["out_bf16"])


def test_oracle_matches_installed_pillow():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(3)
    cases = [((256, 256), (224, 224)), ((32, 32), (64, 64)), ((64, 64), (64, 64)), ((20, 300), (56, 56)),
             ((300, 20), (56, 56)), ((224, 100), (224, 224)), ((100, 224), (224, 224)), ((1, 1), (8, 8)),
             ((3, 7), (5, 5)), ((700, 33), (16, 16))]
    cases += [((int(rng.integers(1, 200)), int(rng.integers(1, 200))),
               (int(rng.integers(1, 120)), int(rng.integers(1, 120)))) for _ in range(12)]
    for (h, w), (H, W) in cases:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        want = np.asarray(Image.fromarray(img).resize((W, H), Image.BILINEAR))
        assert np.array_equal(resample_ref.resize_bilinear_u8(img, (H, W)), want), ((h, w), (H, W))


@pytest.mark.parametrize("hw,scale,ratio", [
    ((256, 256), (0.08, 1.0), (3 / 4, 4 / 3)), ((64, 48), (0.08, 1.0), (3 / 4, 4 / 3)),
    ((40, 400), (0.5, 1.0), (3 / 4, 4 / 3)),      # mostly the central-crop fallback, in_ratio > max(ratio)
    ((400, 30), (0.9, 1.0), (0.9, 1.1)),          # fallback, in_ratio < min(ratio)
    ((100, 100), (0.35, 1.0), (0.5, 2.0)),
])
def test_resized_crop_draws_match_the_literal_loop(hw, scale, ratio):
    n = 97
    g1 = torch.Generator().manual_seed(11)
    g2 = torch.Generator().manual_seed(11)
    a = T._draw_resized_crop_flip_params_loop(n, hw, scale, ratio, 0.5, g1)
    b = T.draw_resized_crop_flip_params(n, hw, scale, ratio, 0.5, g2)
    assert torch.equal(a, b)
    # the generator is left exactly where the per-sample calls leave it
    assert torch.equal(torch.rand(5, generator=g1), torch.rand(5, generator=g2))


def test_resized_crop_draws_match_torchvision_golden(golden):
    g = golden("resample_rrc56.npz")
    torch.manual_seed(int(g["seed"]))
    got = T.draw_resized_crop_flip_params(len(g["src"]), g["src"].shape[1:3])
    assert np.array_equal(got.numpy(), g["params"])


def test_ragged_image_sizes_draw_like_torchvision_per_image():
    """Per-sample image sizes (ragged ImageNet-style images in a common slab): the boxes are what torchvision
    draws when it is handed the images one by one."""
    from PIL import Image
    import torchvision.transforms as TV
    rng = np.random.default_rng(9)
    sizes = np.stack([rng.integers(20, 300, 40), rng.integers(20, 300, 40)], axis=1)
    torch.manual_seed(21)
    want = []
    for h, w in sizes:
        i, j, hh, ww = TV.RandomResizedCrop.get_params(Image.new("RGB", (int(w), int(h))), [0.08, 1.0], [3 / 4, 4 / 3])
        want.append((i, j, hh, ww, int(bool(torch.rand(1) < 0.5))))
    torch.manual_seed(21)
    got = T.draw_resized_crop_flip_params(len(sizes), sizes)
    assert np.array_equal(got.numpy(), np.array(want, dtype=np.int32))
    assert torch.equal(T._draw_resized_crop_flip_params_loop(5, sizes[:5], generator=torch.Generator().manual_seed(3)),
                       T.draw_resized_crop_flip_params(5, sizes[:5], generator=torch.Generator().manual_seed(3)))
    # plain Resize of ragged images: the whole image is the box, one flip draw per sample
    rt = T.ResizeBatchTransform(32, random_resized_crop=False)
    g1, g2 = torch.Generator().manual_seed(4), torch.Generator().manual_seed(4)
    p = rt.draw(len(sizes), sizes, g1)
    flips = [int(bool(torch.rand(1, generator=g2) < 0.5)) for _ in sizes]
    assert np.array_equal(p[:, :4].numpy(), np.concatenate([np.zeros_like(sizes), sizes], axis=1))
    assert p[:, 4].tolist() == flips
    same = rt.draw(3, (48, 64), torch.Generator().manual_seed(1))
    assert same[:, 2].tolist() == [48] * 3 and same[:, 3].tolist() == [64] * 3
    with pytest.raises(ValueError):
        T.draw_resized_crop_flip_params(3, sizes[:2])
