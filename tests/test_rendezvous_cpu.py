"""Rank / world-size discovery of the command-line trainer (imagenet.py:224-244): the reference's SLURM
variables + `scontrol show hostnames`, or torchrun's variables, or a single process."""
import os

import imagent_distributed_training_pytorch_with_slurm_b200.imagenet as cli


def _clean(monkeypatch):
    for k in list(os.environ):
        if k.startswith("SLURM_") or k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT",
                                           "GROUP_RANK", "GROUP_WORLD_SIZE"):
            monkeypatch.delenv(k, raising=False)


def test_slurm_variables_as_in_the_reference(monkeypatch):
    _clean(monkeypatch)
    for k, v in dict(SLURM_JOB_NUM_NODES="8", SLURM_NODEID="3", SLURM_LOCALID="1", SLURM_PROCID="7", SLURM_NTASKS="16",
                     SLURM_JOB_NODELIST="ener[021-028]").items():
        monkeypatch.setenv(k, v)
    monkeypatch.setattr(cli.subprocess, "check_output", lambda cmd: b"ener021\nener022\n")
    assert cli._rendezvous() == (8, 3, 1, 7, 16)
    assert os.environ["MASTER_ADDR"] == "ener021" and os.environ["MASTER_PORT"] == "29500"   # imagenet.py:237-242
    assert os.environ["WORLD_SIZE"] == "16" and os.environ["RANK"] == "7"


def test_torchrun_variables_and_single_process(monkeypatch):
    _clean(monkeypatch)
    for k, v in dict(RANK="5", LOCAL_RANK="5", WORLD_SIZE="8", MASTER_ADDR="127.0.0.1", MASTER_PORT="29511").items():
        monkeypatch.setenv(k, v)
    assert cli._rendezvous() == (1, 0, 5, 5, 8)
    assert os.environ["MASTER_PORT"] == "29511"
    _clean(monkeypatch)
    assert cli._rendezvous() == (1, 0, 0, 0, 1)
    assert os.environ["MASTER_ADDR"] == "127.0.0.1"


def test_flags_and_defaults_match_the_reference():
    a = cli.build_parser().parse_args([])
    # imagenet.py:437-449
    assert (a.seed, a.backend, a.batch_size, a.epochs, a.lr, a.save_model) == (0, "nccl", 128, 100, 0.1, False)
    assert a.arch == "resnet18" and a.image_size == 448      # additions default to the reference's behaviour
    a = cli.build_parser().parse_args(["--batch-size", "64", "--save-model", "--backend", "gloo"])
    assert a.batch_size == 64 and a.save_model and a.backend == "gloo"
