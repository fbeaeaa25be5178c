"""Data path of the SwAV trainer (SURVEY 2.4 `vissl/data`): data sources (`synthetic`, `disk_folder`), the SSL
transforms the recipe names (ImgPilToMultiCrop, ImgPilColorDistortion, ImgPilGaussianBlur + the torchvision ones
referenced by class name), `multicrop_collator`, and an async host->device prefetch wrapper
(DataloaderAsyncGPUWrapper: side-stream copies overlapped with compute)."""
from __future__ import annotations

import random
from typing import Any, Callable, Dict, Iterator, List, Sequence

import torch
from torch.utils.data import DataLoader, Dataset

TRANSFORM_REGISTRY: Dict[str, Callable[..., Any]] = {}


def register_transform(name):
    def deco(cls):
        TRANSFORM_REGISTRY[name] = cls
        return cls

    return deco


class SyntheticImageDataset(Dataset):
    """Mean-gray images (vissl/data/synthetic_dataset.py:7-52): lets every recipe run without a dataset."""

    def __init__(self, size: int = 224, length: int = 100_000):
        self.size, self.length = size, length

    def __len__(self):
        return self.length

    def __getitem__(self, idx):
        from PIL import Image

        return Image.new("RGB", (self.size, self.size), color=(124, 116, 104)), True


class DiskImageDataset(Dataset):
    """`disk_folder`: images under <root>/<class>/<file> (labels unused: LABEL_TYPE = sample_index)."""

    def __init__(self, root: str):
        from torchvision.datasets import ImageFolder

        self.folder = ImageFolder(root)

    def __len__(self):
        return len(self.folder)

    def __getitem__(self, idx):
        try:
            return self.folder[idx][0], True
        except Exception:
            from PIL import Image

            return Image.new("RGB", (224, 224), color=(124, 116, 104)), False


@register_transform("ImgPilToMultiCrop")
class ImgPilToMultiCrop:
    """image -> list of `total_num_crops` random-resized crops: num_crops[i] crops of size_crops[i] with scale range
    crop_scales[i] (2 x 224 @ [0.14, 1] + 6 x 96 @ [0.05, 0.14] in the recipe)."""

    def __init__(self, total_num_crops, num_crops, size_crops, crop_scales):
        import torchvision.transforms as T

        assert sum(num_crops) == total_num_crops and len(num_crops) == len(size_crops) == len(crop_scales)
        self.transforms = []
        for n, size, scale in zip(num_crops, size_crops, crop_scales):
            self.transforms += [T.RandomResizedCrop(size, scale=tuple(scale))] * n

    def __call__(self, image):
        return [t(image) for t in self.transforms]


@register_transform("ImgPilColorDistortion")
class ImgPilColorDistortion:
    """SimCLR colour distortion: jitter (0.8s, 0.8s, 0.8s, 0.2s) with p = 0.8, then grayscale with p = 0.2."""

    def __init__(self, strength: float = 1.0):
        import torchvision.transforms as T

        s = strength
        self.t = T.Compose([T.RandomApply([T.ColorJitter(0.8 * s, 0.8 * s, 0.8 * s, 0.2 * s)], p=0.8),
                            T.RandomGrayscale(p=0.2)])

    def __call__(self, image):
        return self.t(image)


@register_transform("ImgPilGaussianBlur")
class ImgPilGaussianBlur:
    def __init__(self, p: float = 0.5, radius_min: float = 0.1, radius_max: float = 2.0):
        self.p, self.rmin, self.rmax = p, radius_min, radius_max

    def __call__(self, image):
        from PIL import ImageFilter

        if random.random() > self.p:
            return image
        return image.filter(ImageFilter.GaussianBlur(radius=random.uniform(self.rmin, self.rmax)))


class _PerCrop:
    """Applies a single-image transform to every crop once the image has become a list of crops."""

    def __init__(self, t):
        self.t = t

    def __call__(self, x):
        return [self.t(c) for c in x] if isinstance(x, list) else self.t(x)


def build_transforms(specs: Sequence[Dict[str, Any]]) -> Callable:
    import torchvision.transforms as T

    steps = []
    for spec in specs:
        spec = dict(spec)
        name = spec.pop("name")
        if name in TRANSFORM_REGISTRY:
            t = TRANSFORM_REGISTRY[name](**spec)
            steps.append(t if name == "ImgPilToMultiCrop" else _PerCrop(t))
        else:
            steps.append(_PerCrop(getattr(T, name)(**spec)))   # torchvision transform by class name
    return T.Compose(steps)


def multicrop_collator(batch: List[Dict[str, Any]]) -> Dict[str, Any]:
    """[{data: [crop_0..crop_7], ...}] -> {"input": [tensor[B,3,224,224] x2, tensor[B,3,96,96] x6], ...}
    (vissl/data/collators/multicrop_collator.py:7-55): crop-major so that equal resolutions can share a pass."""
    n_crops = len(batch[0]["data"])
    crops = [torch.stack([s["data"][c] for s in batch]) for c in range(n_crops)]
    return {"input": crops, "label": torch.tensor([s["label"] for s in batch]),
            "data_valid": torch.tensor([s["data_valid"] for s in batch]), "data_idx": torch.tensor([s["data_idx"] for s in batch])}


class GenericSSLDataset(Dataset):
    def __init__(self, cfg):
        tr = cfg.DATA.TRAIN
        src = tr.DATA_SOURCES[0]
        if src == "synthetic":
            self.base = SyntheticImageDataset()
        elif src == "disk_folder":
            assert tr.DATA_PATHS, "disk_folder needs +config.DATA.TRAIN.DATA_PATHS=[...]"
            self.base = DiskImageDataset(tr.DATA_PATHS[0])
        else:
            raise ValueError(f"unsupported data source {src}")
        self.transform = build_transforms(tr.TRANSFORMS)
        self.limit = tr.get("DATA_LIMIT", -1)

    def __len__(self):
        return len(self.base) if self.limit < 0 else min(self.limit, len(self.base))

    def __getitem__(self, idx):
        img, valid = self.base[idx]
        return {"data": self.transform(img), "label": idx, "data_valid": valid, "data_idx": idx}


class StatefulPeerSampler(torch.utils.data.Sampler):
    """Shuffled order that is a pure function of (seed, epoch) and can start in the middle of an epoch after a resume
    (vissl/data/data_helper.py:88-148 StatefulDistributedSampler). Every peer of a collaboration is a world of one,
    so the rank / world split of the original is replaced by a PER-PEER seed: peers that ran the same config with the
    same SEED_VALUE would otherwise walk through the dataset in the same order and average duplicate gradients
    (the ALBERT trainer seeds its shuffle with the hash of the peer's public key for the same reason)."""

    def __init__(self, num_samples: int, batch_size: int, seed: int):
        self.num_samples, self.batch_size, self.seed = int(num_samples), int(batch_size), int(seed)
        self.epoch, self.start_iter = 0, 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def set_start_iter(self, start_iter: int) -> None:
        """Skip the first `start_iter` batches of the epoch (they were consumed before the checkpoint)."""
        self.start_iter = int(start_iter)

    def __iter__(self):
        g = torch.Generator().manual_seed(self.seed + self.epoch)
        order = torch.randperm(self.num_samples, generator=g).tolist()
        skip = min(self.start_iter * self.batch_size, self.num_samples)
        self.start_iter = 0          # only the first epoch after a resume is partial
        return iter(order[skip:])

    def __len__(self):
        return self.num_samples


def peer_shuffle_seed(cfg) -> int:
    """SEED_VALUE mixed with something only this peer has: DATA.TRAIN.PEER_SHUFFLE_SALT when given (reproducible
    multi-peer runs), else the process id and the clock."""
    import hashlib
    import os
    import time

    salt = cfg.DATA.TRAIN.get("PEER_SHUFFLE_SALT") if hasattr(cfg.DATA.TRAIN, "get") else None
    if salt in (None, ""):
        salt = f"{os.getpid()}-{time.time_ns()}-{os.urandom(4).hex()}"
    digest = hashlib.sha256(f"{cfg.get('SEED_VALUE', 0)}|{salt}".encode()).digest()
    return int.from_bytes(digest[:4], "little")


class DataloaderAsyncGPUWrapper:
    """Prefetches the next batch to the GPU on a side stream (classy_vision DataloaderAsyncGPUWrapper)."""

    def __init__(self, loader, device):
        self.loader, self.device = loader, device
        self.sampler = getattr(loader, "sampler", None)
        self.stream = torch.cuda.Stream(device) if device.type == "cuda" else None

    def _to_device(self, batch):
        move = lambda t: t.to(self.device, non_blocking=True)
        return {k: ([move(c) for c in v] if isinstance(v, list) else move(v)) for k, v in batch.items()}

    def __iter__(self):
        it = iter(self.loader)
        if self.stream is None:
            yield from it
            return
        nxt = None
        for batch in it:
            with torch.cuda.stream(self.stream):
                cur = self._to_device(batch)
            if nxt is not None:
                yield nxt
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            nxt = cur
        if nxt is not None:
            yield nxt

    def __len__(self):
        return len(self.loader)


def build_dataloader(cfg, device) -> DataloaderAsyncGPUWrapper:
    ds = GenericSSLDataset(cfg)
    tr = cfg.DATA.TRAIN
    collate = multicrop_collator if tr.COLLATE_FUNCTION == "multicrop_collator" else None
    sampler = StatefulPeerSampler(len(ds), tr.BATCHSIZE_PER_REPLICA, peer_shuffle_seed(cfg))
    loader = DataLoader(ds, batch_size=tr.BATCHSIZE_PER_REPLICA, sampler=sampler, drop_last=tr.DROP_LAST,
                        num_workers=cfg.DATA.NUM_DATALOADER_WORKERS, collate_fn=collate,
                        pin_memory=device.type == "cuda", persistent_workers=cfg.DATA.NUM_DATALOADER_WORKERS > 0)
    return DataloaderAsyncGPUWrapper(loader, device)


def synthetic_multicrop_batch(batch_size: int, device, size_crops=(224, 96), num_crops=(2, 6), dtype=torch.float32,
                              pinned: bool = False) -> List[torch.Tensor]:
    """Random multi-crop batch with the recipe's shapes (for benchmarks; `pinned` = host memory for the e2e arm)."""
    out = []
    for n, s in zip(num_crops, size_crops):
        for _ in range(n):
            t = torch.randn(batch_size, 3, s, s, dtype=dtype)
            out.append(t.pin_memory() if pinned else t.to(device))
    return out
