"""Data path: device-placing loader, samplers, and the datasets the scripts need.

Parity targets
  * ``stoke_model.DataLoader(dataset=..., sampler=..., num_workers=...,
    multiprocessing_context='spawn')`` (Stoke-DDP.py:286-298): batch size is injected
    by the facade, a ``DistributedSampler`` is required when distributed, and every
    yielded batch arrives on the local GPU (the training loop never moves tensors,
    Stoke-DDP.py:70-74).
  * ``DistributedSampler(dataset, num_replicas, rank)`` (Fairscale-DDP.py:45-55,
    Stoke-DDP.py:272-283).
  * ``CustomDataset(input_path, target_path)`` of paired low/high-resolution image
    patches (Fairscale-DDP.py:37, Stoke-DDP.py:264) — the module is absent from the
    reference, so :class:`PairedImageDataset` implements the evident contract.

B200-first: host batches are pinned and copied with ``non_blocking=True`` on a
dedicated copy stream one batch ahead of the consumer (double buffering), so the
H2D transfer overlaps the previous step instead of serialising in front of it.
"""
from __future__ import annotations

import math
import os
from typing import Any, Callable, Iterator, List, Optional, Sequence, Tuple

import torch
from torch.utils.data import DataLoader, Dataset, Sampler
from torch.utils.data.distributed import DistributedSampler as _TorchDistributedSampler

__all__ = ["StokeDataLoader", "DevicePrefetcher", "pin_channels_last", "DistributedSampler", "BucketedDistributedSampler", "SyntheticImageDataset",
           "SyntheticPairedDataset", "PairedImageDataset", "CustomDataset", "to_device"]


def to_device(batch: Any, device: torch.device, non_blocking: bool = True, channels_last: bool = False) -> Any:
    """Move tensors inside arbitrarily nested tuples / lists / dicts."""
    if isinstance(batch, torch.Tensor):
        out = batch.to(device, non_blocking=non_blocking)
        if channels_last and out.dim() == 4:
            out = out.contiguous(memory_format=torch.channels_last)
        return out
    if isinstance(batch, dict):
        return {k: to_device(v, device, non_blocking, channels_last) for k, v in batch.items()}
    if isinstance(batch, tuple) and hasattr(batch, "_fields"):
        return type(batch)(*(to_device(v, device, non_blocking, channels_last) for v in batch))
    if isinstance(batch, (list, tuple)):
        return type(batch)(to_device(v, device, non_blocking, channels_last) for v in batch)
    return batch


def pin_channels_last(x: torch.Tensor) -> torch.Tensor:
    """Pinned host copy of an NCHW-shaped batch stored NHWC (the order decoded images come in): the H2D copy
    then lands directly in the channels-last layout the kernels read, with no permute pass on the device."""
    if x.dim() != 4:
        return x.pin_memory() if torch.cuda.is_available() else x
    n, c, h, w = x.shape
    buf = torch.empty((n, h, w, c), dtype=x.dtype, pin_memory=torch.cuda.is_available()).permute(0, 3, 1, 2)
    buf.copy_(x)
    return buf


def _record_stream(batch: Any, stream: "torch.cuda.Stream") -> None:
    if isinstance(batch, torch.Tensor):
        if batch.is_cuda:
            batch.record_stream(stream)
    elif isinstance(batch, dict):
        for v in batch.values():
            _record_stream(v, stream)
    elif isinstance(batch, (list, tuple)):
        for v in batch:
            _record_stream(v, stream)


class DistributedSampler(_TorchDistributedSampler):
    """Rank-strided index partition with padding.  Same contract as torch's; the
    epoch is advanced automatically at each ``__iter__`` unless ``set_epoch`` is used
    (the reference never calls ``set_epoch`` — SURVEY §7 — which silently reuses one
    shuffle; this default fixes that without changing the call sites)."""

    def __init__(self, dataset: Dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False, auto_epoch: bool = True):
        if num_replicas is None or rank is None:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
                rank = dist.get_rank() if rank is None else rank
            else:
                num_replicas = 1 if num_replicas is None else num_replicas
                rank = 0 if rank is None else rank
        super().__init__(dataset, num_replicas=num_replicas, rank=rank, shuffle=shuffle, seed=seed,
                         drop_last=drop_last)
        self._auto_epoch = auto_epoch
        self._explicit_epoch = False

    def set_epoch(self, epoch: int) -> None:
        self._explicit_epoch = True
        super().set_epoch(epoch)

    def __iter__(self) -> Iterator[int]:
        it = super().__iter__()
        if self._auto_epoch and not self._explicit_epoch:
            self.epoch += 1
        return it


class BucketedDistributedSampler(Sampler[int]):
    """Distributed sampler that groups indices of similar length into buckets so a
    batch has little padding (Stoke ships one; SURVEY §2.2 E12 names the loader
    contract).  ``sorted_idx`` is the dataset order sorted by length."""

    def __init__(self, dataset: Dataset, buckets: int, batch_size: int, sorted_idx: Sequence[int],
                 num_replicas: int = 1, rank: int = 0, shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        self.num_replicas, self.rank, self.shuffle, self.seed, self.drop_last = num_replicas, rank, shuffle, seed, drop_last
        self.batch_size = batch_size
        self.epoch = 0
        n = len(sorted_idx)
        per = math.ceil(n / buckets)
        self.buckets = [list(sorted_idx[i * per:(i + 1) * per]) for i in range(buckets)]
        self.buckets = [b for b in self.buckets if b]
        self.slice = batch_size * num_replicas
        self.num_samples = sum((len(b) // self.slice if drop_last else math.ceil(len(b) / self.slice)) * batch_size
                               for b in self.buckets)

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self) -> Iterator[int]:
        g = torch.Generator().manual_seed(self.seed + self.epoch)
        batches: List[List[int]] = []
        for b in self.buckets:
            idx = [b[i] for i in torch.randperm(len(b), generator=g).tolist()] if self.shuffle else list(b)
            n_full = len(idx) // self.slice
            rem = len(idx) - n_full * self.slice
            if rem and not self.drop_last:
                idx = idx + idx[: self.slice - rem]
                n_full += 1
            for k in range(n_full):
                chunk = idx[k * self.slice:(k + 1) * self.slice]
                batches.append(chunk[self.rank * self.batch_size:(self.rank + 1) * self.batch_size])
        if self.shuffle:
            order = torch.randperm(len(batches), generator=g).tolist()
            batches = [batches[i] for i in order]
        return iter([i for b in batches for i in b])


class _ChannelsLastCollate:
    """Collate, then store image batches NHWC on the host (picklable, so it works with worker processes).
    ``pin_memory`` keeps the strides, so the H2D copy lands in the channels-last layout the kernels read and
    the device never runs a permute pass over the input."""

    def __init__(self, base: Any = None):
        from torch.utils.data import default_collate
        self.base = base if base is not None else default_collate

    def _fix(self, b: Any) -> Any:
        if isinstance(b, torch.Tensor):
            return b.contiguous(memory_format=torch.channels_last) if b.dim() == 4 and b.is_floating_point() else b
        if isinstance(b, dict):
            return {k: self._fix(v) for k, v in b.items()}
        if isinstance(b, tuple) and hasattr(b, "_fields"):
            return type(b)(*(self._fix(v) for v in b))
        if isinstance(b, (list, tuple)):
            return type(b)(self._fix(v) for v in b)
        return b

    def __call__(self, samples: Any) -> Any:
        return self._fix(self.base(samples))


class StokeDataLoader(DataLoader):
    """``DataLoader`` that forces the facade's per-device batch size and yields batches
    already resident on the training device."""

    def __init__(self, dataset: Dataset, batch_size: int, gpu: bool = False, device: Optional[torch.device] = None,
                 distributed: bool = False, channels_last: bool = False, prefetch_to_device: bool = True,
                 **kwargs: Any):
        sampler = kwargs.get("sampler")
        if distributed and kwargs.get("batch_sampler") is None and not isinstance(
                sampler, (_TorchDistributedSampler, BucketedDistributedSampler)):
            raise ValueError(
                "a distributed run needs sampler=DistributedSampler(...) (or BucketedDistributedSampler) so each "
                "rank reads a disjoint slice of the dataset")
        if gpu:
            kwargs.setdefault("pin_memory", True)
        if channels_last and kwargs.get("batch_sampler") is None:
            kwargs["collate_fn"] = _ChannelsLastCollate(kwargs.get("collate_fn"))
        if kwargs.get("num_workers", 0) == 0:
            kwargs.pop("multiprocessing_context", None)
            kwargs.pop("prefetch_factor", None)
            kwargs.pop("persistent_workers", None)
        super().__init__(dataset, batch_size=batch_size, **kwargs)
        self._gpu = gpu
        self._device = device
        self._channels_last = channels_last
        self._prefetch = prefetch_to_device

    def __iter__(self) -> Iterator[Any]:  # type: ignore[override]
        host_iter = super().__iter__()
        if not self._gpu or self._device is None or self._device.type != "cuda":
            for b in host_iter:
                yield b
            return
        if not self._prefetch:
            for b in host_iter:
                yield to_device(b, self._device, True, self._channels_last)
            return
        yield from DevicePrefetcher(host_iter, self._device, self._channels_last)


class DevicePrefetcher:
    """Wrap an iterator of (ideally pinned) host batches: batch *i+1* is copied on a side
    stream while batch *i* is being consumed, so the H2D transfer hides behind compute."""

    def __init__(self, host_iter: Any, device: torch.device, channels_last: bool = False):
        self._it = iter(host_iter)
        self._device = device
        self._channels_last = channels_last
        self._stream = torch.cuda.Stream(device=device)
        self._nxt = self._fetch()

    def _fetch(self) -> Any:
        try:
            host = next(self._it)
        except StopIteration:
            return None
        with torch.cuda.stream(self._stream):
            return to_device(host, self._device, True, self._channels_last)

    def __iter__(self) -> "DevicePrefetcher":
        return self

    def __next__(self) -> Any:
        if self._nxt is None:
            raise StopIteration
        cur = self._nxt
        main = torch.cuda.current_stream(self._device)
        main.wait_stream(self._stream)
        _record_stream(cur, main)
        self._nxt = self._fetch()
        return cur


# --------------------------------------------------------------------------- datasets
class SyntheticImageDataset(Dataset):
    """Deterministic random images + integer labels of a named shape (no network, no
    datasets on the box: BASELINE.json asks for synthetic data of the real shape)."""

    def __init__(self, length: int, shape: Tuple[int, int, int] = (3, 224, 224), num_classes: int = 1000,
                 seed: int = 0, dtype: torch.dtype = torch.float32, cache: int = 256):
        self.length, self.shape, self.num_classes, self.seed, self.dtype = length, shape, num_classes, seed, dtype
        g = torch.Generator().manual_seed(seed)
        n = min(length, cache)
        self._x = torch.randn((n, *shape), generator=g).to(dtype)
        self._y = torch.randint(0, num_classes, (n,), generator=g)

    def __len__(self) -> int:
        return self.length

    def __getitem__(self, i: int):
        k = i % self._x.shape[0]
        return self._x[k], self._y[k]


class SyntheticPairedDataset(Dataset):
    """Random low-res / high-res pairs for the super-resolution scripts."""

    def __init__(self, length: int, lr_size: int = 64, scale: int = 2, channels: int = 3, seed: int = 0, cache: int = 64):
        g = torch.Generator().manual_seed(seed)
        n = min(length, cache)
        self.length = length
        hr = torch.rand((n, channels, lr_size * scale, lr_size * scale), generator=g)
        self._hr = hr
        self._lr = torch.nn.functional.avg_pool2d(hr, scale)

    def __len__(self) -> int:
        return self.length

    def __getitem__(self, i: int):
        k = i % self._hr.shape[0]
        return self._lr[k], self._hr[k]


_IMG_EXT = (".png", ".jpg", ".jpeg", ".bmp", ".webp", ".tif", ".tiff")


class PairedImageDataset(Dataset):
    """Paired image folders: file *i* (sorted) of ``input_path`` is the network input
    and file *i* of ``target_path`` is the target, both returned as float CHW tensors in
    [0, 1] (``CustomDataset(input_path, target_path)``, Fairscale-DDP.py:37)."""

    def __init__(self, input_path: str, target_path: str, transform: Optional[Callable] = None,
                 target_transform: Optional[Callable] = None):
        self.input_files = self._list(input_path)
        self.target_files = self._list(target_path)
        if len(self.input_files) != len(self.target_files):
            raise ValueError(f"{len(self.input_files)} inputs vs {len(self.target_files)} targets")
        self.transform, self.target_transform = transform, target_transform

    @staticmethod
    def _list(path: str) -> List[str]:
        return sorted(os.path.join(path, f) for f in os.listdir(path) if f.lower().endswith(_IMG_EXT))

    @staticmethod
    def _load(path: str) -> torch.Tensor:
        from PIL import Image, ImageFile
        import numpy as np
        ImageFile.LOAD_TRUNCATED_IMAGES = True  # Fairscale-DDP.py:12
        with Image.open(path) as im:
            arr = np.asarray(im.convert("RGB"), dtype=np.uint8)
        return torch.from_numpy(arr.copy()).permute(2, 0, 1).float().div_(255.0)

    def __len__(self) -> int:
        return len(self.input_files)

    def __getitem__(self, i: int):
        x, y = self._load(self.input_files[i]), self._load(self.target_files[i])
        if self.transform is not None:
            x = self.transform(x)
        if self.target_transform is not None:
            y = self.target_transform(y)
        return x, y


CustomDataset = PairedImageDataset
