"""Batch loader.

The reference's ``DataLoader(batch_size, sampler, num_workers=2)`` (cifar10_mpi_mobilenet_224.py:126-133)
collates pre-transformed fp32 224x224 images in worker processes and ships 77 MB per step through
pageable memory with a blocking ``.to(device)`` (:174).  This loader moves *raw* samples:

* a batch is one ``index_select`` of the uint8 dataset by the sampler's index tensor;
* host datasets are staged in double-buffered pinned memory by a background thread and copied
  on a dedicated stream (393 KB + labels + 12 floats/sample of augmentation parameters);
* device-resident datasets are gathered on the GPU, no PCIe traffic at all;
* the ragged last batch is kept (reference behaviour) and reported as ``Batch.n``.

Resize / augmentation / normalisation happen downstream on the device (``ops.preprocess``) or,
in CPU mode, in ``transforms.apply_pipeline_torch``.
"""
from __future__ import annotations

import queue
import threading
from dataclasses import dataclass
from typing import Iterator, List, Optional

import torch

from ..parallel.sampler import PartitionSampler
from .datasets import ArrayDataset
from .transforms import identity_params, sample_aug_params


@dataclass
class Batch:
    images: torch.Tensor      # uint8 [n,32,32,3] on the target device
    labels: torch.Tensor      # int64 [n]
    params: torch.Tensor      # float32 [n,12] augmentation parameters
    n: int
    ready: Optional["torch.cuda.Event"] = None   # recorded on the copy stream; wait before use
    x: Optional[torch.Tensor] = None             # CPU mode: the transformed fp32 [n,3,S,S] input, made by the loader thread

    def wait(self, stream: Optional["torch.cuda.Stream"] = None) -> "Batch":
        """Make ``stream`` (default: current) wait for the H2D copies, and tell the caching allocator that the
        tensors are used there.  Without ``record_stream`` the memory -- allocated on the copy stream -- is handed
        to a later batch's copy as soon as Python drops this one, while the compute stream (which runs many steps
        behind the host) has not read it yet: the GPU would then train on whichever batches were copied last."""
        if self.ready is not None:
            st = stream or torch.cuda.current_stream()
            st.wait_event(self.ready)
            for t in (self.images, self.labels, self.params):
                if t.is_cuda:
                    t.record_stream(st)
        return self


class BatchLoader:
    def __init__(self, dataset: ArrayDataset, sampler: PartitionSampler, batch_size: int,
                 device: torch.device, train: bool, augment: bool = True, drop_last: bool = False,
                 seed: int = 0, prefetch: int = 2, max_batches: int = 0, cpu_pipeline: int = 0,
                 img_size: int = 224) -> None:
        self.ds = dataset
        self.sampler = sampler
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.train = train
        self.augment = augment and train
        self.drop_last = drop_last
        self.seed = seed
        self.prefetch = max(1, prefetch)
        self.max_batches = max_batches
        # CPU mode: image size for which a background thread runs the transform pipeline one batch ahead of the
        # training step -- the role of the reference's DataLoader workers (ref :126-133); 0 = leave it to the engine
        self.cpu_pipeline = int(cpu_pipeline) if self.device.type == "cpu" else 0
        self.img_size = int(cpu_pipeline) if cpu_pipeline else int(img_size)     # crop boxes are drawn in output-pixel units
        self.resident = dataset.images.device.type == "cuda"
        self._copy_stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self._staging: List[dict] = []

    def __len__(self) -> int:
        n = self.sampler.num_batches(self.batch_size, self.drop_last)
        return min(n, self.max_batches) if self.max_batches else n

    def set_epoch(self, epoch: int) -> None:
        self.sampler.set_epoch(epoch)

    # -- parameter rows ----------------------------------------------------------------------
    def _params(self, n: int, gen: torch.Generator) -> torch.Tensor:
        return sample_aug_params(n, gen, size=self.img_size) if self.augment else identity_params(n)

    def _gen(self) -> torch.Generator:
        g = torch.Generator()
        g.manual_seed(self.seed * 1_000_003 + self.sampler.epoch * 977 + self.sampler.rank)
        return g

    # -- iteration ---------------------------------------------------------------------------
    def __iter__(self) -> Iterator[Batch]:
        batches = self.sampler.batches(self.batch_size, self.drop_last)
        if self.max_batches:
            batches = batches[: self.max_batches]
        gen = self._gen()
        if self.device.type != "cuda":
            if self.cpu_pipeline:
                yield from self._iter_cpu_ahead(batches, gen)
                return
            for idx in batches:
                yield Batch(self.ds.images.index_select(0, idx), self.ds.labels.index_select(0, idx),
                            self._params(idx.numel(), gen), idx.numel())
            return
        if self.resident:
            for idx in batches:
                di = idx.to(self.device, non_blocking=True)
                yield Batch(self.ds.images.index_select(0, di), self.ds.labels.index_select(0, di),
                            self._params(idx.numel(), gen).to(self.device, non_blocking=True), idx.numel())
            return
        yield from self._iter_pinned(batches, gen)

    def _iter_cpu_ahead(self, batches: List[torch.Tensor], gen: torch.Generator) -> Iterator[Batch]:
        """CPU mode: a producer thread gathers AND transforms the next batches while the model trains on this one."""
        from .transforms import apply_pipeline_torch
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def produce():
            try:
                for idx in batches:
                    if stop.is_set():
                        return
                    n = idx.numel()
                    img, lab = self.ds.images.index_select(0, idx), self.ds.labels.index_select(0, idx)
                    par = self._params(n, gen)
                    x = apply_pipeline_torch(img, par if self.train else None, self.cpu_pipeline)
                    q.put(Batch(img, lab, par, n, None, x))
                q.put(None)
            except BaseException as e:
                q.put(e)

        th = threading.Thread(target=produce, name="b200dp-cpu-loader", daemon=True)
        th.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    break
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
            while th.is_alive():                   # unblock a producer that is waiting on the full queue, then join it
                try:
                    q.get(timeout=0.05)
                except queue.Empty:
                    pass
            th.join()

    def _ensure_staging(self) -> None:
        want = self.prefetch + 1
        while len(self._staging) < want:
            B = self.batch_size
            self._staging.append({
                "img": torch.empty((B, 32, 32, 3), dtype=torch.uint8).pin_memory(),
                "lab": torch.empty((B,), dtype=torch.int64).pin_memory(),
                "par": torch.empty((B, 12), dtype=torch.float32).pin_memory(),
                "free": None,   # event: device copy out of this buffer finished
            })

    def _iter_pinned(self, batches: List[torch.Tensor], gen: torch.Generator) -> Iterator[Batch]:
        """Producer thread fills pinned buffers; this thread issues the H2D copies on the copy stream."""
        self._ensure_staging()
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        free: "queue.Queue" = queue.Queue()
        for s in self._staging:
            free.put(s)
        stop = threading.Event()

        def produce():
            try:
                for idx in batches:
                    if stop.is_set():
                        return
                    s = free.get()
                    if s is None or stop.is_set():     # the consumer went away (generator closed mid-epoch)
                        return
                    if s["free"] is not None:
                        s["free"].synchronize()
                    n = idx.numel()
                    torch.index_select(self.ds.images, 0, idx, out=s["img"][:n])
                    torch.index_select(self.ds.labels, 0, idx, out=s["lab"][:n])
                    s["par"][:n].copy_(self._params(n, gen))
                    q.put((s, n))
                q.put(None)
            except BaseException as e:  # surface loader failures in the consumer
                q.put(e)

        th = threading.Thread(target=produce, name="b200dp-loader", daemon=True)
        th.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    break
                if isinstance(item, BaseException):
                    raise item
                s, n = item
                with torch.cuda.stream(self._copy_stream):
                    img = s["img"][:n].to(self.device, non_blocking=True)
                    lab = s["lab"][:n].to(self.device, non_blocking=True)
                    par = s["par"][:n].to(self.device, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(self._copy_stream)
                s["free"] = ev
                free.put(s)
                yield Batch(img, lab, par, n, ev)
        finally:
            # Stop the producer for good: it may be blocked putting into the full queue or waiting for a free staging
            # buffer.  A daemon thread that is still inside native code when the interpreter exits aborts the process.
            stop.set()
            free.put(None)
            while th.is_alive():
                try:
                    q.get(timeout=0.05)
                except queue.Empty:
                    pass
            th.join()

    def h2d_bytes_per_batch(self, n: Optional[int] = None) -> int:
        n = self.batch_size if n is None else n
        return 0 if self.resident else n * (32 * 32 * 3 + 8 + 12 * 4)
