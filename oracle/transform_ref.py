"""Oracle: per-sample torchvision pipeline restated for a uint8 NHWC batch (numpy)."""
import numpy as np


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even float32 -> bfloat16, returned as uint16 bit patterns
    (what ``tensor.to(torch.bfloat16)`` does for finite values)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    rounded = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return rounded.astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def crop_flip_normalize(src: np.ndarray, params: np.ndarray, pad: int, out_hw, mean, std) -> np.ndarray:
    """float32 [B, H, W, 3].

    Follows, per sample: zero padding + crop at (i, j) — RandomCrop.forward,
    TV/transforms/transforms.py:676-677,690 with get_params' (i, j) of :653-654;
    horizontal flip when params[:, 2] != 0 — RandomHorizontalFlip.forward :720-722;
    ToTensor's ``.div(255)`` in float32 — TV/transforms/functional.py:174-176;
    Normalize's ``sub_(mean).div_(std)`` in float32 — TV/transforms/_functional_tensor.py:921-928.
    (The reference applies these in utils/hf_dataset_utilities.py:46-53,64-77 and
    04_accelerate/01_cifar_accelerate.ipynb:295-298.)
    """
    B = src.shape[0]
    H, W = out_hw
    mean32 = np.asarray(mean, dtype=np.float32)
    std32 = np.asarray(std, dtype=np.float32)
    out = np.empty((B, H, W, 3), dtype=np.float32)
    padded = np.pad(src, ((0, 0), (pad, pad), (pad, pad), (0, 0))) if pad else src
    for b in range(B):
        i, j, flip = (int(v) for v in params[b])
        tile = padded[b, i:i + H, j:j + W, :]
        if flip:
            tile = tile[:, ::-1, :]
        x = tile.astype(np.float32) / np.float32(255.0)
        out[b] = (x - mean32) / std32
    return out
