"""One config object for all three training modes.

The reference has no config system: each hyper-parameter is a literal repeated in three
near-identical scripts.  Here the reference's literals are the defaults and
``--mode {cpu,gpu,ddp}`` replaces the three copies:

* batch size 128 (GPU/DDP) / 64 (CPU)  — cifar10_mpi_mobilenet_224.py:117, cifar10_serial_mobilenet_224.py:59
* lr 1e-4, StepLR(10, 0.1), 20 epochs  — cifar10_mpi_mobilenet_224.py:148-149, 158
* seed 42, port 29500, data root ./data — cifar10_mpi_mobilenet_224.py:58, 30, 96
* checkpoint names                      — cifar10_mpi_mobilenet_224.py:249, cifar10_128batch.py:152

Every field can be overridden by a CLI flag (``--lr 3e-4``) or an environment variable
(``B200DP_LR=3e-4``); CLI wins over env, env wins over defaults.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Sequence

MODES = ("cpu", "gpu", "ddp")


@dataclass
class TrainConfig:
    # ---- mode / placement -------------------------------------------------
    mode: str = "gpu"                 # cpu = serial CPU (mode 1), gpu = single GPU (mode 2), ddp = one rank per GPU (mode 3)
    engine: str = "auto"              # auto | native (sm_100a kernels) | torch (ATen ops; CPU mode and the comparison baseline)
    comm: str = "fused"               # fused (peer-memory all-reduce + Adam kernel) | nccl (baseline: NCCL all-reduce + separate Adam)
    comm_algo: str = "auto"           # fused kernel: one_shot (every rank pulls all peers' gradients) | two_shot (reduce-scatter +
                                      # all-gather over peer memory, 2(W-1)/W instead of W-1 gradient volumes per GPU) | nvls (two-shot
                                      # with multimem.ld_reduce / multimem.st through the NVSwitch) | auto = nvls / two_shot for W >= 4
    comm_overlap: bool = True         # run the fused kernel per gradient bucket from a branch of the step graph, overlapped with backward
    comm_buckets: str = ""            # block indices (0..16, descending) at which a new gradient bucket starts; "" = one bucket,
                                      # launched at the end of backward.  Measured on 8x B200 (ms/step): one bucket 5.167, buckets
                                      # "7" (96 % of the parameters reduced after ~30 % of backward) 5.224 -- a bucket launch next to
                                      # the persistent compute kernels costs more than the 60 us of exposed communication it hides
    comm_blocks: int = 148            # cap on the CTAs of an in-graph bucket launch (wide and short; the waiting is a separate one-CTA launch)
    # ---- data --------------------------------------------------------------
    data_root: str = "./data"
    dataset: str = "auto"             # auto = real CIFAR-10 if present under data_root, else synthetic
    train_samples: int = 50_000
    test_samples: int = 10_000
    img_size: int = 224               # cifar10_mpi_mobilenet_224.py:70
    num_classes: int = 10
    augment: bool = True              # train-time RandomResizedCrop/Flip/ColorJitter/Rotation (ref :72-82)
    drop_last: bool = False           # reference keeps the ragged last batch
    # ---- optimisation -------------------------------------------------------
    batch_size: int = 0               # 0 -> 128 on GPU, 64 on CPU (reference values)
    epochs: int = 20
    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.0
    lr_step_size: int = 10
    lr_gamma: float = 0.1
    dropout: float = 0.2
    seed: int = 42
    pretrained: str = ""              # path to a torchvision-format state_dict to fine-tune from ("" = random init)
    # ---- distributed --------------------------------------------------------
    master_addr: str = "127.0.0.1"
    master_port: int = 29500
    backend: str = "auto"             # auto -> nccl if CUDA else gloo (ref :34)
    check_replicas_every: int = 0     # >0: hash parameters across ranks every N steps (debug "race detector")
    # ---- checkpoint / logging ----------------------------------------------
    out_dir: str = "."
    checkpoint_name: str = ""         # "" -> reference file names depending on mode
    save_on_improve: bool = True      # reference only saves after the last epoch (a crash loses everything)
    resume: str = ""                  # path to a resume sidecar (optimizer state + epoch)
    metrics_jsonl: str = ""           # machine-readable per-epoch stream
    max_steps_per_epoch: int = 0      # 0 = full epoch (smoke tests set this)
    max_eval_steps: int = 0
    log_every: int = 0
    # ---- inference ----------------------------------------------------------
    serve_port: int = 7861            # GROUP03.pdf p.27
    serve_host: str = "0.0.0.0"

    def resolved_batch_size(self) -> int:
        if self.batch_size > 0:
            return self.batch_size
        return 64 if self.mode == "cpu" else 128

    def resolved_checkpoint_name(self) -> str:
        if self.checkpoint_name:
            return self.checkpoint_name
        if self.mode == "ddp":
            return "best_mobilenetv2_cifar10_224_mpi.pth"   # ref mpi:249
        return "best_mobilenetv2_cifar10_224.pth"           # ref 128batch:152

    def validate(self) -> "TrainConfig":
        if self.mode not in MODES:
            raise ValueError(f"mode must be one of {MODES}, got {self.mode!r}")
        if self.engine not in ("auto", "native", "torch"):
            raise ValueError(f"engine must be auto|native|torch, got {self.engine!r}")
        if self.comm not in ("fused", "nccl"):
            raise ValueError(f"comm must be fused|nccl, got {self.comm!r}")
        if self.comm_algo not in ("auto", "one_shot", "two_shot", "nvls"):
            raise ValueError(f"comm_algo must be auto|one_shot|two_shot|nvls, got {self.comm_algo!r}")
        try:
            starts = [int(x) for x in self.comm_buckets.split(",") if x.strip()]
        except ValueError:
            raise ValueError(f"comm_buckets must be a comma-separated list of block indices, got {self.comm_buckets!r}")
        if any(not (0 < b < 17) for b in starts) or any(a <= b for a, b in zip(starts, starts[1:])):
            raise ValueError(f"comm_buckets must descend within 1..16, got {self.comm_buckets!r}")
        if self.epochs < 0 or self.lr <= 0:
            raise ValueError("epochs must be >= 0 and lr > 0")
        if not (0.0 <= self.dropout < 1.0):
            raise ValueError("dropout must be in [0, 1)")
        return self

    def to_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)


_ENV_PREFIX = "B200DP_"


def _coerce(value: str, typ: type) -> Any:
    if typ is bool:
        return value.strip().lower() in ("1", "true", "yes", "on")
    return typ(value)


def add_config_args(parser: argparse.ArgumentParser) -> None:
    for f in dataclasses.fields(TrainConfig):
        flag = "--" + f.name.replace("_", "-")
        if f.type in ("bool", bool):
            parser.add_argument(flag, dest=f.name, default=None,
                                type=lambda s: _coerce(s, bool), nargs="?", const=True)
        else:
            typ = {"int": int, "float": float, "str": str}.get(str(f.type), str)
            parser.add_argument(flag, dest=f.name, default=None, type=typ)


def config_from_args(argv: Optional[Sequence[str]] = None,
                     env: Optional[Dict[str, str]] = None,
                     **overrides: Any) -> TrainConfig:
    """defaults <- environment (B200DP_*) <- CLI flags <- explicit overrides."""
    env = os.environ if env is None else env
    cfg = TrainConfig()
    for f in dataclasses.fields(TrainConfig):
        key = _ENV_PREFIX + f.name.upper()
        if key in env:
            typ = type(getattr(cfg, f.name))
            setattr(cfg, f.name, _coerce(env[key], typ))
    parser = argparse.ArgumentParser(prog="b200dp.train", description=__doc__,
                                     formatter_class=argparse.RawDescriptionHelpFormatter)
    add_config_args(parser)
    ns = parser.parse_args(list(argv) if argv is not None else [])
    for k, v in vars(ns).items():
        if v is not None:
            setattr(cfg, k, v)
    for k, v in overrides.items():
        if not hasattr(cfg, k):
            raise AttributeError(f"unknown config field {k!r}")
        setattr(cfg, k, v)
    return cfg.validate()
