"""Multi-GPU tests of the whole native engine (SURVEY §4 "(b)": bit-identical replicas after N real steps,
agreement with the NCCL path) and of its failure contract (a stalled peer is an error, never a diverged model).
Skipped when the box has fewer GPUs than the world size."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(world, *args, timeout=900):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, "-m", "b200dp.launch", "-np", str(world), "--",
                           os.path.join(ROOT, "tests", "mp_engine_check.py"), *args], env=env, capture_output=True,
                          text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_engine_replicas_identical_and_match_nccl(world):
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    r = _launch(world, "train")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "ENGINE CHECK OK" in r.stdout


def test_stalled_peer_is_a_comm_error():
    if torch.cuda.device_count() < 2:
        pytest.skip("not enough GPUs")
    r = _launch(2, "fault", timeout=600)
    assert r.returncode != 0, r.stdout[-2000:]
    assert "FAULT NOT DETECTED" not in r.stdout
    assert "did not reach a device barrier" in r.stderr or "CommError" in r.stderr, r.stderr[-3000:]


def test_sharded_optimiser_state_resumes_like_an_uninterrupted_run():
    if torch.cuda.device_count() < 2:
        pytest.skip("not enough GPUs")
    r = _launch(2, "resume", timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESUME CHECK OK" in r.stdout
