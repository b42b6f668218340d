"""Reference arm of bench.py: run the UNMODIFIED reference scripts (baseline/_ref) and time K of their steps.

The scripts are executed byte for byte with `runpy`.  What cannot exist offline is supplied from outside, without
touching them:
  * `mpi4py`                      -> baseline/shims/mpi4py (rank/size from the launcher's environment);
  * CIFAR-10 download             -> `torchvision.datasets.CIFAR10` replaced by a synthetic dataset of the same
                                     interface (PIL 32x32 RGB image + label, `transform` applied) so the reference's
                                     own PIL transform chain and `DataLoader(num_workers=2)` run as written;
  * ImageNet weights download     -> `models.mobilenet_v2(pretrained=True)` builds the same architecture with
                                     random weights;
  * `MASTER_ADDR=localhost:29500` -> `init_process_group` is pointed at the launcher's rendezvous (127.0.0.1).
N = 1 runs `cifar10_128batch.py` (mode 2), N > 1 runs `cifar10_mpi_mobilenet_224.py` (mode 3) under torchrun.
Timing: `torch.optim.Adam.step` is wrapped to count optimiser steps; CUDA events are recorded after step W and
after step W+K (device synchronised + barrier), then the script is stopped.  A reference "step" is its whole
loop body: DataLoader fetch (CPU transforms), H2D copy, forward, loss, backward (+DDP all-reduce), Adam, and its
two `.item()` host reads -- that IS its public code path, so `value` and `e2e` are the same number here.
"""
from __future__ import annotations

import contextlib
import os
import runpy
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


class _Done(BaseException):
    pass


def main(ns) -> dict:
    # baseline/_ref is git-ignored; on a box where build() has not run, materialise the same bytes from the
    # mount or the bundled archive into a private directory (one per process: ranks must not race on files).
    sys.path.insert(0, HERE)
    import install_reference
    ref = REF
    if not install_reference.installed(ref):
        ref = tempfile.mkdtemp(prefix="b200dp_refsrc_")
        try:
            if install_reference.main(ref, quiet=True) != 0:
                return {"impl": "reference", "unavailable": "no reference scripts: baseline/_ref, /root/reference and baseline/reference_scripts.tar.gz all absent"}
        except Exception as e:
            return {"impl": "reference", "unavailable": f"reference install failed: {e}"}
    cpu_mode = bool(getattr(ns, "cpu", False))       # BASELINE.json config 1: the serial CPU script, batch 64
    script = os.path.join(ref, "cifar10_serial_mobilenet_224.py" if cpu_mode else
                          ("cifar10_128batch.py" if ns.gpus == 1 else "cifar10_mpi_mobilenet_224.py"))
    try:
        import torch
        import torchvision
        from PIL import Image
        import numpy as np
    except Exception as e:
        return {"impl": "reference", "unavailable": f"missing dependency: {e}"}
    if not cpu_mode and not torch.cuda.is_available():
        return {"impl": "reference", "unavailable": "no CUDA device"}
    import torch.distributed as dist

    W, K, B = max(ns.warmup, 3), ns.steps, (64 if cpu_mode else 128)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world != ns.gpus:
        return {"impl": "reference", "unavailable": f"--gpus {ns.gpus} needs torchrun with {ns.gpus} ranks (WORLD_SIZE={world})"}
    sys.path.insert(0, os.path.join(HERE, "shims"))

    # ---- stand-ins for what needs the network ------------------------------------------------
    n_train = (W + K + 2) * B * world

    class SyntheticCIFAR10(torch.utils.data.Dataset):
        """Same interface as torchvision.datasets.CIFAR10: __getitem__ -> (transform(PIL image), label)."""

        def __init__(self, root=None, train=True, transform=None, target_transform=None, download=False):
            rng = np.random.RandomState(1 if train else 2)
            n = n_train if train else 2 * B * world
            self.data = rng.randint(0, 256, size=(n, 32, 32, 3), dtype=np.uint8)
            self.targets = rng.randint(0, 10, size=n).tolist()
            self.transform = transform
            self.classes = ['airplane', 'automobile', 'bird', 'cat', 'deer', 'dog', 'frog', 'horse', 'ship', 'truck']

        def __len__(self):
            return len(self.targets)

        def __getitem__(self, i):
            img = Image.fromarray(self.data[i])
            if self.transform is not None:
                img = self.transform(img)
            return img, self.targets[i]

    torchvision.datasets.CIFAR10 = SyntheticCIFAR10
    _orig_mv2 = torchvision.models.mobilenet_v2

    def mobilenet_v2(pretrained=False, **kw):
        kw.pop("weights", None)
        return _orig_mv2(weights=None, **kw)

    torchvision.models.mobilenet_v2 = mobilenet_v2

    _orig_init = dist.init_process_group
    launcher_addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    launcher_port = os.environ.get("MASTER_PORT", "29500")

    def init_process_group(*a, **kw):
        os.environ["MASTER_ADDR"] = "127.0.0.1" if launcher_addr == "localhost" else launcher_addr
        os.environ["MASTER_PORT"] = launcher_port
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        return _orig_init(*a, **kw)

    dist.init_process_group = init_process_group

    # ---- step counter / timer on the optimiser ---------------------------------------------------
    state = {"n": 0, "e0": None, "e1": None}
    _orig_step = torch.optim.Adam.step

    import time as _time

    def sync_all():
        if cpu_mode:
            return
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier()
            torch.cuda.synchronize()

    def step(self, *a, **kw):
        out = _orig_step(self, *a, **kw)
        state["n"] += 1
        if state["n"] == W:
            sync_all()
            if cpu_mode:
                state["e0"] = _time.perf_counter()
            else:
                state["e0"] = torch.cuda.Event(enable_timing=True)
                state["e0"].record()
        elif state["n"] == W + K:
            if cpu_mode:
                state["e1"] = _time.perf_counter()
            else:
                state["e1"] = torch.cuda.Event(enable_timing=True)
                state["e1"].record()
            sync_all()
            raise _Done()
        return out

    torch.optim.Adam.step = step

    cwd = os.getcwd()
    work = tempfile.mkdtemp(prefix="b200dp_ref_")
    os.chdir(work)
    try:
        with contextlib.redirect_stdout(sys.stderr):
            try:
                runpy.run_path(script, run_name="__main__")
            except _Done:
                pass
    finally:
        os.chdir(cwd)
        torch.optim.Adam.step = _orig_step
    if state["e1"] is None:
        return {"impl": "reference", "unavailable": f"reference stopped after {state['n']} optimiser steps"}
    ms = (state["e1"] - state["e0"]) * 1e3 if cpu_mode else state["e0"].elapsed_time(state["e1"])
    if dist.is_initialized():
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
        dist.destroy_process_group()
    value = world * K * B / (ms / 1e3)
    if rank != 0:
        return {}
    return {
        "impl": "reference", "metric": "MobileNetV2@224 train images/sec (max over ranks)", "value": value,
        "unit": "images/sec", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "weak", "dtype": "fp32 (reference default; TF32 convolutions as torch defaults)",
        "data": "synthetic CIFAR-10-shaped uint8 32x32x3 through the reference's PIL transforms and DataLoader(num_workers=2), random-init weights",
        "config": {"model": "MobileNetV2 (10-class head) @224", "global_batch": B * world, "per_rank_batch": B,
                   "parallelism": f"dp{world}", "script": os.path.basename(script), "optimizer": "Adam lr=1e-4"},
        "e2e": {"value": value, "unit": "images/sec", "note": "the reference's own loop body is end to end (CPU transforms + pageable H2D + .item() reads)"},
        "vs_baseline": None, "device": "cpu" if cpu_mode else "cuda",
    }
