"""CPU-side tests of the round-2 additions: reference installer, descriptor exchange, loader-thread transform, config."""
import hashlib
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_installer_from_archive_is_byte_identical(tmp_path, monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import install_reference as ir
    monkeypatch.setattr(ir, "SRC", str(tmp_path / "no-such-mount"))      # force the archive path
    dst = str(tmp_path / "_ref")
    assert not ir.installed(dst)
    assert ir.main(dst, quiet=True) == 0 and ir.installed(dst)
    for fn, h in ir.SHA256.items():
        assert hashlib.sha256(open(os.path.join(dst, fn), "rb").read()).hexdigest() == h
    # a tampered copy is detected
    with open(os.path.join(dst, "cifar10_128batch.py"), "a") as f:
        f.write("# edited\n")
    assert not ir.installed(dst)
    if os.path.isdir("/root/reference"):                                  # and the archive equals the mount
        for fn in ir.SHA256:
            assert hashlib.sha256(open(os.path.join("/root/reference", fn), "rb").read()).hexdigest() == ir.SHA256[fn]


def test_reference_arm_reports_cleanly_without_gpu():
    """`bench.py --impl reference` must print one JSON line and exit 0 even where it cannot run (no CUDA here)."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference"
    if not torch.cuda.is_available():
        assert "unavailable" in d and "CUDA" in d["unavailable"]


def test_fd_exchange_two_ranks_gloo():
    env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-m", "b200dp.launch", "-np", "2", "--", os.path.join(ROOT, "tests", "mp_fd_exchange.py")],
                       env=env, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "FD EXCHANGE OK" in r.stdout


def test_cpu_loader_thread_runs_the_same_pipeline():
    from b200dp.data.datasets import SyntheticCIFAR10
    from b200dp.data.loader import BatchLoader
    from b200dp.data.transforms import apply_pipeline_torch
    from b200dp.parallel.sampler import PartitionSampler
    ds = SyntheticCIFAR10(40, True)
    for train in (True, False):
        mk = lambda cp: BatchLoader(ds, PartitionSampler(len(ds), 1, 0, shuffle=False), 16, torch.device("cpu"), train=train,
                                    augment=True, seed=3, cpu_pipeline=cp, img_size=64)
        ahead, plain = list(mk(64)), list(mk(0))
        assert [b.n for b in ahead] == [16, 16, 8] == [b.n for b in plain]          # ragged tail kept
        for a, p in zip(ahead, plain):
            assert p.x is None and a.x is not None and a.x.shape == (a.n, 3, 64, 64)
            assert torch.equal(a.images, p.images) and torch.equal(a.params, p.params)
            assert torch.equal(a.x, apply_pipeline_torch(p.images, p.params if train else None, 64))


def test_color_jitter_subset_form_matches_select_form():
    """The per-op subset evaluation of ColorJitter against evaluating all four ops on everything and selecting."""
    from b200dp.data import transforms as T
    g = torch.Generator().manual_seed(9)
    x = torch.randint(0, 256, (12, 3, 20, 20), generator=g).float()
    p = T.sample_aug_params(12, g, size=20)
    got = T.color_jitter(x, p)
    order = torch.tensor(T._PERMS)[p[:, T.P_ORDER].long()]
    ref = x.clone()
    fb, fc, fs = (p[:, k].view(-1, 1, 1, 1) for k in (T.P_BRIGHT, T.P_CONTRAST, T.P_SAT))
    fh = p[:, T.P_HUE].view(-1, 1, 1)
    for pos in range(4):
        op = order[:, pos].view(-1, 1, 1, 1)
        xb = (ref * fb).clamp(0, 255)
        mean = torch.floor(T._gray(ref).mean(dim=(1, 2, 3), keepdim=True) + 0.5)
        xc = T._blend(mean.expand_as(ref), ref, fc)
        xs = T._blend(T._gray(ref).expand_as(ref), ref, fs)
        h, s, v = T._rgb_to_hsv(ref)
        xh = T._hsv_to_rgb(h + fh, s, v)
        ref = torch.where(op == 0, xb, torch.where(op == 1, xc, torch.where(op == 2, xs, xh)))
    assert torch.allclose(got, ref, atol=1e-4)


def test_config_comm_fields():
    from b200dp.config import TrainConfig, config_from_args
    assert TrainConfig().comm_buckets == "" and TrainConfig().comm_overlap
    cfg = config_from_args(["--comm-algo", "nvls", "--comm-buckets", "14,7,3", "--comm-blocks", "32"])
    assert (cfg.comm_algo, cfg.comm_buckets, cfg.comm_blocks) == ("nvls", "14,7,3", 32)
    for bad in ("3,7", "17", "7,7", "x"):
        with pytest.raises(ValueError):
            TrainConfig(comm_buckets=bad).validate()
    with pytest.raises(ValueError):
        TrainConfig(comm_algo="ring").validate()


def test_resume_file_is_created_on_first_run_and_loads_weights_only(tmp_path):
    """`--resume <new path>` starts fresh and creates the sidecar (round 1 died with FileNotFoundError); the sidecar holds
    tensors and primitives only, so it loads with weights_only=True."""
    from b200dp.config import TrainConfig
    from b200dp.trainer import Trainer
    from b200dp.utils.checkpoint import load_resume
    path = str(tmp_path / "resume.pt")
    cfg = TrainConfig(mode="cpu", img_size=32, batch_size=8, epochs=1, train_samples=16, test_samples=8, dataset="synthetic",
                      out_dir=str(tmp_path), augment=False, resume=path)
    t = Trainer(cfg)
    t.fit()
    t.close()
    st = load_resume(path)
    assert st["epoch"] == 1 and "m" in st["optimizer"] and st["optimizer"]["step"] == 2
    t2 = Trainer(TrainConfig(**{**cfg.to_dict(), "epochs": 2}))
    assert t2.start_epoch == 1
    t2.close()


def test_bench_cpu_arm_prints_one_json_line():
    """`bench.py --cpu` (BASELINE.json config 1 on this framework's mode 1) runs without a GPU and reports images/sec."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--cpu", "--steps", "1", "--warmup", "3"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["device"] == "cpu" and d["config"]["per_rank_batch"] == 64 and d["value"] > 0 and d["steps"] == 1
