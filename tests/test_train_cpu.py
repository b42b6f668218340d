"""End-to-end on CPU: mode 1 trains, loss falls, checkpoint is loadable the reference's way."""
import os
import subprocess
import sys

import pytest
import torch

from b200dp.config import TrainConfig
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.inference import Predictor, predict_cifar10_image
from b200dp.trainer import Trainer
from b200dp.utils.checkpoint import load_model, load_resume

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(tmp_path, **kw):
    base = dict(mode="cpu", img_size=64, batch_size=16, epochs=2, lr=2e-3, train_samples=64, test_samples=32,
                out_dir=str(tmp_path), augment=False, dataset="synthetic")
    base.update(kw)
    return TrainConfig(**base)


def test_mode1_cpu_trains_and_saves(tmp_path, capsys):
    cfg = _cfg(tmp_path, metrics_jsonl=str(tmp_path / "m.jsonl"))
    tr = Trainer(cfg, train_set=SyntheticCIFAR10(64, True), test_set=SyntheticCIFAR10(32, False))
    out = tr.fit()
    tr.close()
    text = capsys.readouterr().out
    assert "Total parameters: 2236682" in text            # logs_cifar10_cpu_27299.out:28
    assert "Epoch 1/2 Time:" in text and "Train Acc:" in text and "Test Acc:" in text
    h = out["history"]
    assert h[1]["train_loss"] < h[0]["train_loss"]
    path = out["checkpoint"]
    assert os.path.basename(path) == "best_mobilenetv2_cifar10_224.pth"
    sd = load_model(path)
    assert len(sd) == 314 and not any(k.startswith("module.") for k in sd)
    tv = pytest.importorskip("torchvision")
    ref = tv.models.mobilenet_v2(weights=None)
    ref.classifier[1] = torch.nn.Linear(ref.last_channel, 10)       # GROUP03.pdf p.26
    ref.load_state_dict(sd, strict=True)
    assert os.path.getsize(tmp_path / "m.jsonl") > 0


def test_resume_sidecar_roundtrip(tmp_path):
    side = str(tmp_path / "run.resume")
    cfg = _cfg(tmp_path, epochs=1, resume="")
    tr = Trainer(cfg, train_set=SyntheticCIFAR10(32, True), test_set=SyntheticCIFAR10(16, False))
    tr.cfg.resume = side
    tr.fit(); tr.close()
    st = load_resume(side)
    assert st["epoch"] == 1 and st["optimizer"]["step"] == 2 and len(st["model"]) == 314
    cfg2 = _cfg(tmp_path, epochs=2, resume=side)
    tr2 = Trainer(cfg2, train_set=SyntheticCIFAR10(32, True), test_set=SyntheticCIFAR10(16, False))
    assert tr2.start_epoch == 1 and tr2.engine.adam.step == 2
    tr2.fit(); tr2.close()
    assert load_resume(side)["epoch"] == 2


def test_predict_helper_and_threshold(tmp_path, capsys):
    from PIL import Image
    ds = SyntheticCIFAR10(4, False)
    img_path = str(tmp_path / "x.png")
    Image.fromarray(ds.images[0].numpy()).resize((96, 96)).save(img_path)
    pred = Predictor(None, device="cpu")
    res = predict_cifar10_image(img_path, topk=3, conf_threshold=0.99, predictor=pred)
    assert len(res) == 3 and all(isinstance(l, str) and 0 <= c <= 1 for l, c in res)
    assert res[0][1] >= res[1][1] >= res[2][1]
    assert "uncertain" in capsys.readouterr().out          # 128batch.py:181-183
    d = pred.predict(Image.open(img_path))
    assert len(d) == 3 and abs(sum(pred.probabilities(Image.open(img_path)).tolist()) - 1) < 1e-5


def test_ddp_gloo_two_ranks_via_launcher(tmp_path):
    """Mode 3 on the CPU fallback path (gloo), launched by our mpirun replacement."""
    env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "b200dp.launch", "-np", "2", "--", "-m", "b200dp.train", "--mode", "ddp",
           "--img-size", "32", "--batch-size", "8", "--epochs", "1", "--train-samples", "32",
           "--test-samples", "16", "--dataset", "synthetic", "--out-dir", str(tmp_path), "--augment", "false",
           "--check-replicas-every", "1"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "World size: 2" in r.stdout and "Rank source: openmpi" in r.stdout
    assert r.stdout.count("Epoch 1/1") == 1                 # rank-0-gated printing
    assert os.path.exists(tmp_path / "best_mobilenetv2_cifar10_224_mpi.pth")


def test_launcher_propagates_failure():
    env = dict(os.environ, PYTHONPATH=ROOT)
    code = "import os,sys,time; r=int(os.environ['OMPI_COMM_WORLD_RANK']); sys.exit(3) if r==1 else time.sleep(60)"
    r = subprocess.run([sys.executable, "-m", "b200dp.launch", "-np", "2", "--", "-c", code], env=env,
                       capture_output=True, text=True, timeout=60, cwd=ROOT)
    assert r.returncode == 3 and "rank 1 exited with code 3" in r.stderr
