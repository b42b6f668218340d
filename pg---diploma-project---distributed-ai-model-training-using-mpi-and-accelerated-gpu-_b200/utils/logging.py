"""Rank-0 logging with the reference's user-visible lines.

The per-epoch line is the contract users of the reference read from their SLURM ``.out`` files:

* mode 3 (cifar10_mpi_mobilenet_224.py:229-236):
  ``Epoch 3/20 Time: 253.76s Train Loss: 0.1234 Test Loss: 0.2345 Test Acc(local): 0.9500``
* modes 1/2 (cifar10_128batch.py:134-141):
  ``Epoch 3/20 Time: 528.43s Train Loss: 0.1234 Train Acc: 0.9600 Test Loss: 0.2345 Test Acc: 0.9500``

Both formats are reproduced; in mode 3 the accuracy printed here is the *global* one (the
reference prints rank 0's local share) and the label says so.  A JSONL stream
carries the same numbers plus images/sec for machines.
"""
from __future__ import annotations

import json
import sys
import time
from typing import Any, Dict, Optional, TextIO


class RankLogger:
    def __init__(self, is_main: bool = True, stream: Optional[TextIO] = None,
                 jsonl_path: str = "") -> None:
        self.is_main = is_main
        self.stream = stream or sys.stdout
        self._jsonl = open(jsonl_path, "a") if (jsonl_path and is_main) else None

    def print(self, *args: Any, **kw: Any) -> None:
        if self.is_main:
            print(*args, file=self.stream, flush=True, **kw)

    def banner(self, title: str) -> None:
        self.print("=" * 70)
        self.print(title)
        self.print("=" * 70)

    def json(self, record: Dict[str, Any]) -> None:
        if self._jsonl is not None:
            record = dict(record, ts=time.time())
            self._jsonl.write(json.dumps(record) + "\n")
            self._jsonl.flush()

    def close(self) -> None:
        if self._jsonl is not None:
            self._jsonl.close()
            self._jsonl = None


def format_epoch_line(mode: str, epoch: int, epochs: int, seconds: float, train_loss: float,
                      train_acc: float, test_loss: float, test_acc: float) -> str:
    if mode == "ddp":
        return (f"Epoch {epoch}/{epochs} Time: {seconds:.2f}s Train Loss: {train_loss:.4f} "
                f"Test Loss: {test_loss:.4f} Test Acc(global): {test_acc:.4f}")
    return (f"Epoch {epoch}/{epochs} Time: {seconds:.2f}s Train Loss: {train_loss:.4f} "
            f"Train Acc: {train_acc:.4f} Test Loss: {test_loss:.4f} Test Acc: {test_acc:.4f}")
