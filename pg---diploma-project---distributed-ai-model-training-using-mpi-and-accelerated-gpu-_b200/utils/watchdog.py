"""Failure detection.

The reference has none: a dead rank leaves the others spinning in NCCL until the process-group
timeout.  Two pieces here:

* ``Watchdog`` -- a host thread that aborts the process with a clear message if the training loop
  has not called ``kick()`` for ``timeout_s`` (hung collective, dead peer, stuck kernel);
  the launcher then tears the other ranks down (parallel/launcher.py).
* the device side: spin-waits inside the comm kernel are bounded and raise a host-visible error
  flag (csrc/comm_adam.cu, ``b200dp_comm_error``), which ``check_device_flag`` turns into an exception.
"""
from __future__ import annotations

import os
import sys
import threading
import time
from typing import Callable, Optional


class Watchdog:
    def __init__(self, timeout_s: float = 600.0, rank: int = 0,
                 on_timeout: Optional[Callable[[], None]] = None) -> None:
        self.timeout_s = timeout_s
        self.rank = rank
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._on_timeout = on_timeout
        self._thread: Optional[threading.Thread] = None
        self.fired = False

    def start(self) -> "Watchdog":
        if self.timeout_s <= 0:
            return self
        self._thread = threading.Thread(target=self._run, name="b200dp-watchdog", daemon=True)
        self._thread.start()
        return self

    def kick(self) -> None:
        self._last = time.monotonic()

    def stop(self) -> None:
        self._stop.set()

    def _run(self) -> None:
        while not self._stop.wait(min(1.0, self.timeout_s / 4)):
            if time.monotonic() - self._last > self.timeout_s:
                self.fired = True
                sys.stderr.write(f"[b200dp.watchdog] rank {self.rank}: no progress for "
                                 f"{self.timeout_s:.0f}s -- aborting so the launcher can stop the job\n")
                sys.stderr.flush()
                if self._on_timeout is not None:
                    self._on_timeout()
                else:
                    os._exit(75)
                return


class CommError(RuntimeError):
    pass


def check_device_flag(flag_value: int, rank: int = 0) -> None:
    if flag_value != 0:
        raise CommError(f"rank {rank}: device-side barrier timed out (flag={flag_value}); a peer rank is "
                        "dead or stalled")
