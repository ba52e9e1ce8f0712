"""The command-line trainer end to end on one GPU (tiny synthetic run): the reference's stdout
format (imagenet.py:252-262,398-403,422-429), checkpoint name/keys (:392) and TensorBoard layout
(:363,405-421)."""
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cli_two_epochs_single_process(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29577", RANK="0", WORLD_SIZE="1",
               LOCAL_RANK="0")
    for k in list(env):
        if k.startswith("SLURM_"):
            del env[k]
    cmd = [sys.executable, "-m", "imagent_distributed_training_pytorch_with_slurm_b200.imagenet", "--synthetic", "32",
           "--batch-size", "8", "--epochs", "2", "--image-size", "64", "--source-size", "80", "--save-model",
           "--lr", "0.05"]
    r = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    out = r.stdout
    for line in ("0 - Number of nodes: 1", "0 - World size     : 1", "0 - Master         : True", "Backend: nccl",
                 "Initialize Model...", "Epoch 1 Summary: ", "Epoch 2 Summary: ", "\tLearning rate: 0.05",
                 "Training Summary:"):
        assert line in out, (line, out[-1500:])
    assert re.search(r"\tTrain loss: [0-9.]+ ; Test loss: [0-9.]+", out)
    assert re.search(r"\tTrain time: [0-9.e-]+ seconds; Test time:[0-9.e-]+ seconds", out)
    assert re.search(r"\tTop1 best epoch: [01]", out) and re.search(r"\tTraining time: [0-9.e-]+ minutes", out)
    # checkpoint: written only when validation top-1 improves (imagenet.py:388-392) - with 1000 random
    # classes and 16 validation images that usually never happens; when it does, it must be the reference's
    # file name and a DDP-prefixed float32/int64 state dict that stock torchvision loads
    ckpt = tmp_path / "imagenet_FR_resnet18.pt"
    if ckpt.exists():
        sd = torch.load(ckpt, map_location="cpu")
        assert len(sd) == 122 and next(iter(sd)) == "module.conv1.weight"
        import torchvision
        ref = torch.nn.Sequential()
        ref.add_module("module", torchvision.models.resnet18(weights=None))
        ref.load_state_dict(sd)
    else:
        assert "\t Best Test top1 accuracy: 0" in out
    # TensorBoard: add_scalars creates one sub-run per (tag, series)
    tb = tmp_path / "imagenet_FR"
    names = sorted(p.name for p in tb.iterdir())
    for d in ("Loss_train", "Loss_val", "Top1 accuracy_train", "Top5 accuracy_val"):
        assert d in names, names
    assert any(n.startswith("events.out.tfevents") for n in names)
