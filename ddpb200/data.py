"""Data pipeline: CIFAR-10-shaped batches for the four training modes.

Reference: ``create_data_loaders`` builds torchvision CIFAR10 datasets with
RandomCrop(32, padding=4) / RandomHorizontalFlip / ToTensor / Normalize, a DataLoader with
2 worker processes, batch 256 (or 256/world), ``shuffle=False``, and -- in the distributed
parts -- ``DistributedSampler(shuffle=False, drop_last=False)`` on the train split only
(/root/reference/part1/main.py:19-50, part2/part2b/main.py:61-94).

Here the dataset stays as raw uint8 NHWC pixels in (pinned) host memory; a loader step is
an index gather into a pinned staging buffer.  Crop, flip, normalisation and the layout
change happen on the GPU inside the first-layer operand kernel (csrc/data_prep.cu), so
the host only draws the per-image crop offsets and flip bits.

The box has no network, so ``download=True`` becomes: use the CIFAR-10 python batches if
they already exist under ``root``; otherwise fall back to a deterministic synthetic set of
the same shape.
"""
from __future__ import annotations

import math
import os
import pickle
from pathlib import Path

import numpy as np
import torch

NUM_CLASSES = 10
IMAGE_HW = 32
TRAIN_SIZE = 50_000
TEST_SIZE = 10_000


class ImageDataset:
    """In-memory uint8 NHWC images + int64 labels."""

    def __init__(self, images: torch.Tensor, labels: torch.Tensor, train: bool, source: str):
        assert images.dtype == torch.uint8 and images.dim() == 4 and images.shape[-1] == 3
        self.images, self.labels = images.contiguous(), labels.to(torch.int64).contiguous()
        self.train, self.source = train, source

    def __len__(self) -> int:
        return self.images.shape[0]

    def pin(self) -> "ImageDataset":
        if torch.cuda.is_available() and not self.images.is_pinned():
            self.images, self.labels = self.images.pin_memory(), self.labels.pin_memory()
        return self


def synthetic_cifar10(train: bool, size: int | None = None, seed: int = 89395) -> ImageDataset:
    """Deterministic CIFAR-shaped data with learnable structure: every class has a smooth
    colour template, each image is its class template plus noise.  The same (seed, split,
    size) always yields the same bytes on every rank."""
    n = size if size is not None else (TRAIN_SIZE if train else TEST_SIZE)
    g = torch.Generator().manual_seed(seed)                    # templates are shared by both splits
    coarse = torch.rand(NUM_CLASSES, 3, 4, 4, generator=g)
    templates = torch.nn.functional.interpolate(coarse, size=(IMAGE_HW, IMAGE_HW), mode="bilinear",
                                                align_corners=False)  # [10,3,32,32] in [0,1]
    g2 = torch.Generator().manual_seed(seed + (1 if train else 2))
    labels = torch.randint(0, NUM_CLASSES, (n,), generator=g2)
    images = torch.empty(n, IMAGE_HW, IMAGE_HW, 3, dtype=torch.uint8)
    chunk = 5000
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        base = templates[labels[s:e]].permute(0, 2, 3, 1)      # NHWC
        noise = torch.randn(e - s, IMAGE_HW, IMAGE_HW, 3, generator=g2) * 0.25
        images[s:e] = ((base + noise).clamp_(0, 1) * 255).to(torch.uint8)
    return ImageDataset(images, labels, train, "synthetic")


def load_cifar10(root: str | os.PathLike, train: bool) -> ImageDataset | None:
    """Reads torchvision's on-disk layout (root/cifar-10-batches-py/*) if present."""
    d = Path(root) / "cifar-10-batches-py"
    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    if not all((d / n).exists() for n in names):
        return None
    xs, ys = [], []
    for n in names:
        with open(d / n, "rb") as f:
            entry = pickle.load(f, encoding="latin1")
        xs.append(np.asarray(entry["data"], dtype=np.uint8))
        ys.extend(entry.get("labels", entry.get("fine_labels")))
    x = np.concatenate(xs).reshape(-1, 3, IMAGE_HW, IMAGE_HW).transpose(0, 2, 3, 1)  # -> NHWC
    return ImageDataset(torch.from_numpy(np.ascontiguousarray(x)), torch.tensor(ys), train, str(d))


def open_dataset(root: str | os.PathLike | None, train: bool, synthetic: bool | None = None,
                 size: int | None = None) -> ImageDataset:
    if not synthetic and root is not None:
        ds = load_cifar10(root, train)
        if ds is not None:
            return ds
        if synthetic is False:
            raise FileNotFoundError(f"CIFAR-10 not found under {root} and there is no network to download it")
    return synthetic_cifar10(train, size)


# ---------------------------------------------------------------------------------------
# sampler (DistributedSampler(shuffle=False, drop_last=False) semantics)
# ---------------------------------------------------------------------------------------
class ShardSampler:
    """Rank r sees indices r, r+W, r+2W, ... of the (wrap-padded) index list, exactly like
    ``DistributedSampler(num_replicas=W, rank=r, shuffle=False, drop_last=False)``
    (part2/part2b/main.py:78-79).  With W ranks and per-rank batch 256/W the union of the
    ranks' i-th batches is the single-process i-th batch."""

    def __init__(self, n: int, num_replicas: int = 1, rank: int = 0, shuffle: bool = False, seed: int = 0):
        self.n, self.num_replicas, self.rank, self.shuffle, self.seed = n, num_replicas, rank, shuffle, seed
        self.epoch = 0
        self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        """Kept for API parity (part2b/main.py:189); only matters when shuffle=True."""
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        pad = self.total_size - self.n
        if pad > 0:
            reps = math.ceil(pad / self.n)
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        return idx[self.rank:self.total_size:self.num_replicas]

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self):
        return iter(self.indices().tolist())


class BatchLoader:
    """Iterating it gives what the reference's DataLoader gives: ``(data, target)`` pairs with
    ``data`` a normalised (and, for the train split, randomly cropped / flipped) fp32
    ``[N, 3, 32, 32]`` tensor and ``target`` int64 labels, on the host -- a loop written against
    /root/reference/part1/main.py:65-75 consumes it unchanged.

    ``raw()`` is the fast path ddpb200's own ``train_model`` / ``test_model`` use: it yields
    ``(uint8 NHWC batch, int64 labels, int8 aug-or-None)`` from pinned staging buffers and leaves
    crop / flip / normalise to the input kernel on the GPU (csrc/data_prep.cu).  Both views draw
    the same crop offsets and flip bits for the same (seed, epoch, rank).

    Two staging slots alternate so the async H2D copy of batch i can overlap the host
    gather of batch i+1.  ``len()`` and ``.dataset`` / ``.sampler`` / ``.batch_size`` mirror
    DataLoader, which is what the reference's loops rely on (part1/main.py:65,108-111).
    """

    def __init__(self, dataset: ImageDataset, batch_size: int, sampler: ShardSampler | None = None,
                 augment: bool = False, drop_last: bool = False, seed: int = 89395, pin: bool = True,
                 pad: int = 4):
        self.dataset, self.batch_size, self.augment, self.drop_last, self.pad = dataset, batch_size, augment, drop_last, pad
        self.sampler = sampler if sampler is not None else ShardSampler(len(dataset))
        self.seed = seed
        pin = pin and torch.cuda.is_available()
        hw = dataset.images.shape[1]
        self._img = [torch.empty(batch_size, hw, hw, 3, dtype=torch.uint8, pin_memory=pin) for _ in range(2)]
        self._lab = [torch.empty(batch_size, dtype=torch.int64, pin_memory=pin) for _ in range(2)]
        self._aug = [torch.zeros(batch_size, 4, dtype=torch.int8, pin_memory=pin) for _ in range(2)]

    def __len__(self) -> int:
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def __iter__(self):
        from .ops.reference import augment_normalize_u8

        for img, lab, aug in self.raw():
            yield augment_normalize_u8(img, aug, self.pad), lab.clone()

    def raw(self):
        idx = self.sampler.indices()
        g = torch.Generator().manual_seed(self.seed + 7919 * self.sampler.epoch + self.sampler.rank)
        slot = 0
        for b in range(len(self)):
            sel = idx[b * self.batch_size:(b + 1) * self.batch_size]
            n = sel.numel()
            img, lab, aug = self._img[slot][:n], self._lab[slot][:n], None
            torch.index_select(self.dataset.images, 0, sel, out=img)
            torch.index_select(self.dataset.labels, 0, sel, out=lab)
            if self.augment:
                aug = self._aug[slot][:n]
                aug[:, 0] = torch.randint(0, 2 * self.pad + 1, (n,), generator=g).to(torch.int8)
                aug[:, 1] = torch.randint(0, 2 * self.pad + 1, (n,), generator=g).to(torch.int8)
                aug[:, 2] = torch.randint(0, 2, (n,), generator=g).to(torch.int8)
            slot ^= 1
            yield img, lab, aug


def create_data_loaders(rank: int = 0, world_size: int = 1, batch_size: int = 256, root: str | None = "./../data",
                        synthetic: bool | None = None, train_size: int | None = None, test_size: int | None = None,
                        augment: bool = True):
    """Same roles as the reference's function (part1/main.py:19-50; part2b/main.py:61-94):
    returns ``(train_loader, test_loader)``; the train split is sharded over ranks without
    shuffling, the test split is NOT sharded (every rank evaluates all of it)."""
    train = open_dataset(root, True, synthetic, train_size).pin()
    test = open_dataset(root, False, synthetic, test_size).pin()
    sampler = ShardSampler(len(train), world_size, rank, shuffle=False)
    train_loader = BatchLoader(train, batch_size, sampler, augment=augment)
    test_loader = BatchLoader(test, batch_size, None, augment=False)
    return train_loader, test_loader
