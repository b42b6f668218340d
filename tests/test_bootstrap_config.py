"""Rank discovery from each launcher's environment, config layering, logging contract."""
import pytest

from b200dp.config import TrainConfig, config_from_args
from b200dp.parallel.bootstrap import discover_rank_world
from b200dp.utils.logging import format_epoch_line
from b200dp.optim import step_lr, adam_scalars


@pytest.mark.parametrize("env,expect", [
    ({"OMPI_COMM_WORLD_RANK": "3", "OMPI_COMM_WORLD_SIZE": "8", "OMPI_COMM_WORLD_LOCAL_RANK": "3"}, (3, 8, 3, "openmpi")),
    ({"PMI_RANK": "1", "PMI_SIZE": "2"}, (1, 2, 1, "pmi")),
    ({"PMIX_RANK": "2", "SLURM_NTASKS": "4", "SLURM_LOCALID": "0", "SLURM_PROCID": "2"}, (2, 4, 0, "pmix")),
    ({"SLURM_PROCID": "5", "SLURM_NTASKS": "8", "SLURM_LOCALID": "1"}, (5, 8, 1, "slurm")),
    ({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, (1, 2, 1, "torchrun")),
    ({}, (0, 1, 0, "single")),
])
def test_discover(env, expect):
    info = discover_rank_world(env)
    assert (info.rank, info.world_size, info.local_rank, info.source) == expect


def test_mpi_wins_over_torchrun_vars():
    info = discover_rank_world({"OMPI_COMM_WORLD_RANK": "1", "OMPI_COMM_WORLD_SIZE": "2", "RANK": "0", "WORLD_SIZE": "1"})
    assert (info.rank, info.world_size) == (1, 2)


def test_bad_rank_rejected():
    with pytest.raises(RuntimeError):
        discover_rank_world({"RANK": "4", "WORLD_SIZE": "2"})


def test_config_defaults_are_reference_values():
    c = TrainConfig()
    assert (c.lr, c.epochs, c.lr_step_size, c.lr_gamma, c.seed, c.master_port) == (1e-4, 20, 10, 0.1, 42, 29500)
    assert TrainConfig(mode="cpu").resolved_batch_size() == 64       # serial:59
    assert TrainConfig(mode="gpu").resolved_batch_size() == 128      # 128batch:59
    assert TrainConfig(mode="ddp").resolved_checkpoint_name() == "best_mobilenetv2_cifar10_224_mpi.pth"
    assert TrainConfig(mode="gpu").resolved_checkpoint_name() == "best_mobilenetv2_cifar10_224.pth"


def test_config_layering():
    c = config_from_args(["--mode", "cpu", "--lr", "3e-4", "--augment", "false"], env={"B200DP_EPOCHS": "3", "B200DP_LR": "1"})
    assert c.mode == "cpu" and c.lr == 3e-4 and c.epochs == 3 and c.augment is False
    with pytest.raises(ValueError):
        config_from_args(["--mode", "tpu"], env={})


def test_step_lr_and_adam_scalars():
    assert step_lr(1e-4, 0) == 1e-4 and step_lr(1e-4, 9) == 1e-4
    assert abs(step_lr(1e-4, 10) - 1e-5) < 1e-12 and abs(step_lr(1e-4, 19) - 1e-5) < 1e-12
    ss, isb2 = adam_scalars(1, 1e-4, 0.9, 0.999)
    assert abs(ss - 1e-3) < 1e-9 and abs(isb2 - (1 - 0.999) ** -0.5) < 1e-6


def test_epoch_line_formats():
    a = format_epoch_line("gpu", 3, 20, 528.434, 0.12345, 0.96, 0.2345, 0.95)
    assert a == "Epoch 3/20 Time: 528.43s Train Loss: 0.1235 Train Acc: 0.9600 Test Loss: 0.2345 Test Acc: 0.9500"
    b = format_epoch_line("ddp", 3, 20, 253.76, 0.1, 0.9, 0.2, 0.95)
    assert b.startswith("Epoch 3/20 Time: 253.76s Train Loss: 0.1000 Test Loss: 0.2000 Test Acc(")
