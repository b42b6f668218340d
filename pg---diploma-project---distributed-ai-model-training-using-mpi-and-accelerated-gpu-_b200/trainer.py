"""Epoch loop, evaluation, metric collection, best-checkpoint tracking.

One trainer replaces the reference's three scripts (cifar10_serial_mobilenet_224.py,
cifar10_128batch.py, cifar10_mpi_mobilenet_224.py): ``mode`` picks serial CPU, single GPU or one
rank per GPU.  Behaviour kept from the reference: 20 epochs, StepLR per epoch, ``set_epoch`` on the
train sampler (:165), evaluation every epoch inside the epoch time (:161-227), rank-0 logging with
the same lines, best-accuracy tracking and the same checkpoint file names and format.  Deliberate
differences: accuracy and best-model choice are global, not rank-local; the best file
is written when it improves; no per-step host synchronisation.
"""
from __future__ import annotations

import os
import time
from typing import Any, Dict, Optional

import torch

from .config import TrainConfig
from .data.datasets import ArrayDataset, get_datasets
from .data.loader import BatchLoader
from .optim import step_lr
from .parallel.bootstrap import DistContext, barrier, cleanup, rank_zero_first, setup_distributed
from .parallel.sampler import PartitionSampler
from .utils.checkpoint import load_resume, save_model, save_resume
from .utils.logging import RankLogger, format_epoch_line
from .utils.timers import PhaseTimer, max_over_ranks, sum_over_ranks
from .utils.watchdog import Watchdog


def make_engine(cfg: TrainConfig, ctx: DistContext):
    engine = cfg.engine
    if engine == "auto":
        engine = "torch"
        if ctx.device.type == "cuda":
            from . import engine_native
            engine = "native" if engine_native.available() else "torch"
    if engine == "native":
        if ctx.device.type != "cuda":
            raise RuntimeError("engine=native needs a CUDA device (mode gpu or ddp)")
        from .engine_native import NativeEngine
        return NativeEngine(cfg, ctx)
    from .engine_torch import TorchEngine
    return TorchEngine(cfg, ctx)


class _Range:
    """NVTX range around a phase when B200DP_NVTX=1 (shows up in Nsight timelines; free otherwise)."""

    on = os.environ.get("B200DP_NVTX", "0") == "1"

    def __init__(self, name: str) -> None:
        self.name = name

    def __enter__(self):
        if _Range.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        if _Range.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()
        return False


class Trainer:
    def __init__(self, cfg: TrainConfig, ctx: Optional[DistContext] = None,
                 train_set: Optional[ArrayDataset] = None, test_set: Optional[ArrayDataset] = None,
                 logger: Optional[RankLogger] = None) -> None:
        self.cfg = cfg.validate()
        self.own_ctx = ctx is None
        self.ctx = ctx or self._bootstrap(cfg)
        for d in (cfg.out_dir, os.path.dirname(cfg.metrics_jsonl), os.path.dirname(cfg.resume)):
            if d:
                os.makedirs(d, exist_ok=True)        # every rank: harmless, and rank 0 must not race the others to it
        self.log = logger or RankLogger(self.ctx.is_main, jsonl_path=cfg.metrics_jsonl)
        self.batch_size = cfg.resolved_batch_size()
        if train_set is None or test_set is None:
            with rank_zero_first(self.ctx):          # ref :93-102: rank 0 materialises the data first
                train_set, test_set = get_datasets(cfg.data_root, cfg.dataset, cfg.train_samples,
                                                   cfg.test_samples)
        self.train_set, self.test_set = train_set, test_set
        if self.ctx.device.type == "cuda" and train_set.images.device.type == "cpu":
            train_set.pin()
            test_set.pin()
        W, r = self.ctx.world_size, self.ctx.rank
        self.train_sampler = PartitionSampler(len(train_set), W, r, shuffle=True)    # ref :119-121
        self.test_sampler = PartitionSampler(len(test_set), W, r, shuffle=False)     # ref :122-124
        self.train_loader = BatchLoader(train_set, self.train_sampler, self.batch_size, self.ctx.device,
                                        train=True, augment=cfg.augment, drop_last=cfg.drop_last,
                                        seed=cfg.seed, max_batches=cfg.max_steps_per_epoch, cpu_pipeline=cfg.img_size,
                                        img_size=cfg.img_size)
        self.test_loader = BatchLoader(test_set, self.test_sampler, self.batch_size, self.ctx.device,
                                       train=False, augment=False, seed=cfg.seed,
                                       max_batches=cfg.max_eval_steps, cpu_pipeline=cfg.img_size)
        self.engine = make_engine(cfg, self.ctx)
        self.best_acc = 0.0
        self.best_state = self.engine.state_dict()        # ref :159-160
        self.start_epoch = 0
        self.history: list = []
        self.watchdog = Watchdog(timeout_s=0.0, rank=self.ctx.rank)
        self.timer = PhaseTimer(enabled=True, cuda=self.ctx.device.type == "cuda")
        if cfg.resume and os.path.exists(cfg.resume):      # a first run with --resume creates the file
            self._resume(cfg.resume)

    @staticmethod
    def _bootstrap(cfg: TrainConfig) -> DistContext:
        if cfg.mode == "ddp":
            return setup_distributed(cfg.master_addr, cfg.master_port, cfg.backend)
        from .parallel.bootstrap import RankInfo
        use_cuda = cfg.mode == "gpu" and torch.cuda.is_available()      # ref 128batch:19 falls back to CPU
        return setup_distributed(cfg.master_addr, cfg.master_port, cfg.backend,
                                 info=RankInfo(0, 1, 0, "single"), use_cuda=use_cuda)

    # ---- resume ----------------------------------------------------------------------------
    def _resume(self, path: str) -> None:
        st = load_resume(path)
        self.engine.load_state_dict(st["model"])
        self.engine.load_optimizer_state(st["optimizer"])
        self.best_acc = st["best_acc"]
        if st.get("best_model") is not None:
            self.best_state = st["best_model"]
        self.start_epoch = st["epoch"]
        self.log.print(f"Resumed from {path} at epoch {self.start_epoch} (best acc {self.best_acc:.4f})")

    def _ckpt_path(self) -> str:
        return os.path.join(self.cfg.out_dir, self.cfg.resolved_checkpoint_name())

    # ---- one epoch ---------------------------------------------------------------------------
    def train_epoch(self, epoch: int) -> Dict[str, float]:
        cfg = self.cfg
        lr = step_lr(cfg.lr, epoch, cfg.lr_step_size, cfg.lr_gamma)
        self.train_loader.set_epoch(epoch)                                # ref :165
        t0 = time.perf_counter()
        self.timer.reset()
        it = iter(self.train_loader)
        i = -1
        while True:                                                       # ref :173 hot loop
            i += 1
            sampled = (i & 63) == 0          # device-time one step in 64: events are cheap, but not free
            try:
                if sampled:
                    with self.timer.phase("data_wait"):
                        batch = next(it)
                        batch.wait()
                else:
                    batch = next(it)
            except StopIteration:
                break
            if sampled:
                with self.timer.phase("step"):
                    self.engine.train_step(batch, lr)
            else:
                self.engine.train_step(batch, lr)
            self.watchdog.kick()
            if cfg.check_replicas_every and (i + 1) % cfg.check_replicas_every == 0:
                self.check_replicas()
            if cfg.log_every and self.ctx.is_main and (i + 1) % cfg.log_every == 0:
                self.log.print(f"  step {i + 1}/{len(self.train_loader)}")
        loss_sum, correct, count = self.engine.read_metrics()
        g = self._sum([loss_sum, correct, count])                         # K-C6 (ref :188-195), plus accuracy
        dt = time.perf_counter() - t0
        ph = self.timer.collect()
        nsamp = max((i + 63) // 64, 1)
        return {"loss": g[0] / max(g[2], 1), "acc": g[1] / max(g[2], 1), "images": g[2], "seconds": dt, "lr": lr,
                "step_ms": ph.get("step", 0.0) / nsamp, "data_wait_ms": ph.get("data_wait", 0.0) / nsamp}

    # ---- small cross-rank reductions: on the engine's peer slot tables when it has them, else torch.distributed
    def _sum(self, values):
        if self.ctx.initialized and hasattr(self.engine, "reduce"):
            return self.engine.reduce(values, "sum")
        return sum_over_ranks(values, self.ctx.device)

    def _max(self, value: float) -> float:
        if self.ctx.initialized and hasattr(self.engine, "reduce"):
            return self.engine.reduce([value], "max")[0]
        return max_over_ranks(value, self.ctx.device)

    def evaluate(self) -> Dict[str, float]:
        self.engine.sync_buffers()                                        # K-C4 semantics (see engine)
        for batch in self.test_loader:                                    # ref :205-213
            self.engine.eval_step(batch)
            self.watchdog.kick()
        loss_sum, correct, count = self.engine.read_metrics()
        g = self._sum([loss_sum, correct, count])                         # K-C7 (ref :216-223)
        return {"loss": g[0] / max(g[2], 1), "acc": g[1] / max(g[2], 1), "images": g[2]}

    def check_replicas(self) -> None:
        """Replicas must hold identical parameters after every fused step (hard part 4)."""
        if not self.ctx.initialized:
            return
        h = self.engine.param_hash()
        if hasattr(self.engine, "reduce"):
            hi, neg_lo = self.engine.reduce([h, -h], "max")
            lo = -neg_lo
        else:
            import torch.distributed as dist
            t = torch.tensor([h], dtype=torch.float64, device=self.ctx.device)
            lo_t, hi_t = t.clone(), t.clone()
            dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
            lo, hi = lo_t.item(), hi_t.item()
        if lo != hi:
            raise RuntimeError(f"replica divergence: parameter hash spans [{lo!r}, {hi!r}]")

    # ---- full run ----------------------------------------------------------------------------
    def fit(self) -> Dict[str, Any]:
        cfg, ctx, log = self.cfg, self.ctx, self.log
        title = {"cpu": "Serial CPU MobileNetV2 CIFAR-10 (224x224)",
                 "gpu": "Single-GPU MobileNetV2 CIFAR-10 (224x224)",
                 "ddp": "Distributed (one rank per GPU) MobileNetV2 CIFAR-10 (224x224)"}[cfg.mode]
        log.banner(title)
        log.print(f"Backend: {ctx.backend}, World size: {ctx.world_size}, Device(rank0): {ctx.device}, "
                  f"Engine: {self.engine.name}, Rank source: {ctx.info.source}")
        log.print("Train samples:", len(self.train_set))
        log.print("Test samples:", len(self.test_set))
        log.print(f"Total parameters: {self.engine.parameter_count()}")       # ref :151-153
        log.print("Starting training...\n")
        self.watchdog = Watchdog(timeout_s=900.0, rank=ctx.rank).start()
        total_start = time.perf_counter()
        for epoch in range(self.start_epoch, cfg.epochs):
            epoch_start = time.perf_counter()
            with _Range(f"epoch {epoch + 1} train"):
                tr = self.train_epoch(epoch)
            with _Range(f"epoch {epoch + 1} eval"):
                ev = self.evaluate()
            if ctx.device.type == "cuda":
                torch.cuda.synchronize()
            epoch_time = self._max(time.perf_counter() - epoch_start)
            train_time = self._max(tr["seconds"])
            log.print(format_epoch_line(cfg.mode, epoch + 1, cfg.epochs, epoch_time, tr["loss"], tr["acc"],
                                        ev["loss"], ev["acc"]))
            rec = {"epoch": epoch + 1, "epoch_seconds": epoch_time, "train_loss": tr["loss"],
                   "train_acc": tr["acc"], "test_loss": ev["loss"], "test_acc": ev["acc"], "lr": tr["lr"],
                   "train_images_per_sec": tr["images"] / max(train_time, 1e-9), "world_size": ctx.world_size,
                   "step_ms": tr["step_ms"], "data_wait_ms": tr["data_wait_ms"],
                   "engine": self.engine.name, "comm": getattr(self.engine, "comm_algo", cfg.comm),
                   "comm_memory": getattr(getattr(getattr(self.engine, "fused", None), "arena", None), "kind", None),
                   # fraction of the measured HBM copy bandwidth the step sustains on its ~11.5 GB of unavoidable
                   # activation traffic (profiles/step_roofline.md)
                   "hbm_roofline_fraction": (11.5e9 * self.batch_size / 128 / 6.55e12) / max(tr["step_ms"] * 1e-3, 1e-9)
                   if tr["step_ms"] > 0 and ctx.device.type == "cuda" else None}
            self.history.append(rec)
            log.json(rec)
            if ev["acc"] > self.best_acc:                                  # ref :238-240
                self.best_acc = ev["acc"]
                self.best_state = self.engine.state_dict()
                if cfg.save_on_improve and ctx.is_main:
                    save_model(self.best_state, self._ckpt_path())
            if cfg.save_on_improve and cfg.resume:
                opt_state = self.engine.optimizer_state()          # collective: sharded moments are summed over ranks
                if ctx.is_main:
                    save_resume(cfg.resume, epoch=epoch + 1, best_acc=self.best_acc, optimizer=opt_state,
                                model=self.engine.state_dict(), best_model=self.best_state)
        total = time.perf_counter() - total_start
        self.watchdog.stop()
        log.print(f"\nBest test accuracy: {self.best_acc:.4f}")                # ref :244
        log.print(f"Total training time: {total:.2f}s ({total / 60:.2f} min)")  # ref :245-248
        if ctx.is_main:
            save_model(self.best_state, self._ckpt_path())                  # ref :249
            log.print(f"Saved {self._ckpt_path()}")
        barrier(ctx)
        return {"best_acc": self.best_acc, "total_seconds": total, "history": self.history,
                "checkpoint": self._ckpt_path()}

    def close(self) -> None:
        self.engine.close()
        self.log.close()
        if self.own_ctx:
            cleanup(self.ctx)                                               # ref :252
