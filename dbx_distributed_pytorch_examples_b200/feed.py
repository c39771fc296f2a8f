"""Host-side feed for the batch transform: pinned uint8 batches, sharded like
``DistributedSampler``, copied to the GPU one step ahead of the compute.

Replaces, around the GPU transform, what the reference builds from
``CustomDataset`` + ``DataLoader(..., sampler=DistributedSampler(dataset))``
(utils/hf_dataset_utilities.py:31-55, 01_torch_distributor/01_basic_torch_distributor.py:285-286;
``pin_memory=True`` in 04_accelerate/01_cifar_accelerate.ipynb:409-414): the dataset keeps raw
uint8 HWC images (its ``transform=None`` path), a batch is gathered into pinned memory, and
its host->device copy runs on a side stream while the previous step computes.  The kernel of
:class:`BatchTransform` then reads it from HBM.

* :class:`ShardedBatchSampler` — the index lists ``DistributedSampler`` + ``BatchSampler`` would
  give (TORCH/utils/data/distributed.py:105-134): ``randperm(len, seed + epoch)``, padded by
  wrapping to a multiple of the world size, ``indices[rank::world]``, cut into batches.
* :class:`PinnedImageDataset` — uint8 ``[N, H, W, 3]`` images + int64 labels in pinned memory.
* :class:`DeviceFeeder` — iterates ``(images_u8_cuda, labels_cuda)`` with the next batch's copy
  already in flight.
* :class:`LossReader` — reads each step's loss on the host one step late, so that the host never
  stalls the GPU queue (the reference calls ``loss.item()`` every step,
  01_torch_distributor/02_cifar_torch_distributor_resnet.py:246).
"""
from __future__ import annotations

from typing import Iterable, Iterator, Optional, Tuple

import torch


class ShardedBatchSampler:
    """Batches of dataset indices for one rank, as DistributedSampler + BatchSampler yield them."""

    def __init__(self, length: int, batch_size: int, rank: int = 0, world: int = 1, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if not 0 <= rank < world:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {world - 1}]")
        self.length, self.batch_size, self.rank, self.world = length, batch_size, rank, world
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        self.num_samples = -(-length // world)  # ceil: the sampler pads, it does not drop
        self.total_size = self.num_samples * world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.length, generator=g)
        else:
            idx = torch.arange(self.length)
        pad = self.total_size - self.length
        if pad > 0:
            reps = -(-pad // self.length)
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        return idx[self.rank:self.total_size:self.world]

    def __iter__(self) -> Iterator[torch.Tensor]:
        idx = self.indices()
        for s in range(0, len(idx), self.batch_size):
            b = idx[s:s + self.batch_size]
            if len(b) < self.batch_size and self.drop_last:
                return
            yield b

    def __len__(self) -> int:
        n = self.num_samples
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)


class PinnedImageDataset:
    """Raw uint8 NHWC images and labels held in page-locked host memory.

    ``sizes`` (optional, int ``[N, 2]`` = height, width): images of different sizes, each stored in the
    top-left corner of its ``[H, W]`` slab; ``batches`` then yields ``(images, labels, sizes)`` and the sizes go
    to ``ResizeBatchTransform(...)(images, sizes=sizes)``."""

    def __init__(self, images_u8: torch.Tensor, labels: torch.Tensor, pin: bool = True, sizes=None):
        if images_u8.dtype != torch.uint8 or images_u8.dim() != 4 or images_u8.shape[-1] != 3:
            raise ValueError("expected uint8 images of shape [N, H, W, 3]")
        if len(labels) != len(images_u8):
            raise ValueError("images and labels differ in length")
        self.sizes = None
        if sizes is not None:
            sz = torch.as_tensor(sizes).to(torch.int32).contiguous()
            if tuple(sz.shape) != (len(images_u8), 2) or bool((sz < 1).any()) or \
                    bool((sz[:, 0] > images_u8.shape[1]).any()) or bool((sz[:, 1] > images_u8.shape[2]).any()):
                raise ValueError("sizes must be [N, 2] with 1 <= height <= H and 1 <= width <= W")
            self.sizes = sz
        can_pin = pin and torch.cuda.is_available()
        self.images = images_u8.contiguous().pin_memory() if can_pin else images_u8.contiguous()
        self.labels = labels.to(torch.int64).contiguous()
        self._pin = can_pin

    def __len__(self) -> int:
        return len(self.images)

    def batches(self, sampler: Iterable[torch.Tensor]) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        """Gather each index batch into (reused) pinned staging tensors."""
        stage = [None, None, None]  # three deep: one being filled, one in flight, one being consumed
        k = 0
        for idx in sampler:
            slot = stage[k % 3]
            n = len(idx)
            if slot is None or slot[0].shape[0] != n:
                img = torch.empty((n,) + tuple(self.images.shape[1:]), dtype=torch.uint8)
                lab = torch.empty((n,), dtype=torch.int64)
                if self._pin:
                    img, lab = img.pin_memory(), lab.pin_memory()
                slot = stage[k % 3] = (img, lab)
            torch.index_select(self.images, 0, idx, out=slot[0])
            torch.index_select(self.labels, 0, idx, out=slot[1])
            k += 1
            if self.sizes is None:
                yield slot
            else:
                yield slot[0], slot[1], torch.index_select(self.sizes, 0, idx)  # host metadata, not copied to the GPU


class DeviceFeeder:
    """Iterate device batches with the next host->device copy already running on a side stream."""

    def __init__(self, host_batches: Iterable[Tuple[torch.Tensor, torch.Tensor]], device: Optional[torch.device] = None):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.host_batches = host_batches
        self.copy_stream = torch.cuda.Stream(device=self.device)

    def _start_copy(self, batch):
        images, labels, *rest = batch  # anything after the labels (image sizes) stays on the host
        with torch.cuda.stream(self.copy_stream):
            d_img = images.to(self.device, non_blocking=True)
            d_lab = labels.to(self.device, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.copy_stream)
        return d_img, d_lab, done, tuple(rest)

    def __iter__(self):
        it = iter(self.host_batches)
        try:
            pending = self._start_copy(next(it))
        except StopIteration:
            return
        while pending is not None:
            d_img, d_lab, done, rest = pending
            # This copy was issued a whole step ago, so the wait is free; it guarantees that the
            # pinned staging slot of this (and every older) batch may be refilled by next(it).
            done.synchronize()
            try:
                pending = self._start_copy(next(it))  # overlaps the step that consumes d_img
            except StopIteration:
                pending = None
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(done)
            # the copy stream allocated these; tell the allocator the compute stream uses them
            d_img.record_stream(cur)
            d_lab.record_stream(cur)
            yield (d_img, d_lab) + rest


class LossReader:
    """``push(loss)`` every step; the value comes back from ``push`` one step later (and the last
    one from ``flush``), read through pinned memory without draining the GPU queue."""

    def __init__(self, device: Optional[torch.device] = None, depth: int = 2):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.slots = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(depth)]
        self.events = [None] * depth
        self.k = 0
        self.values = []

    def push(self, loss: torch.Tensor) -> Optional[float]:
        """Queue this step's loss for read-back; return the previous step's value (None at first)."""
        d = len(self.slots)
        slot = self.k % d
        self.slots[slot].copy_(loss.detach().float(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.events[slot] = ev
        self.k += 1
        prev = (self.k - 2) % d
        out = None
        if self.k >= 2 and self.events[prev] is not None:
            self.events[prev].synchronize()  # step k-1 is done; step k is already queued behind it
            out = float(self.slots[prev])
            self.values.append(out)
            self.events[prev] = None
        return out

    def flush(self):
        """Read what is still in flight (the last step's loss); returns all values in order."""
        d = len(self.slots)
        for j in range(max(self.k - d, 0), self.k):
            slot = j % d
            if self.events[slot] is not None:
                self.events[slot].synchronize()
                self.values.append(float(self.slots[slot]))
                self.events[slot] = None
        return self.values
