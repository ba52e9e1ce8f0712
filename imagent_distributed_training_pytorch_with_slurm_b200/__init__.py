"""B200-native data-parallel training step of the reference's ``imagenet.py`` (train(), :97-151).

Python host code mirroring the reference's interface over a C-ABI library of hand-written sm_100a
kernels (``include/b200_imagenet.h``).  There is no CPU, Triton or PyTorch-eager fallback for those
kernels: importing this package on a machine without the built library still works (so tooling can
inspect it), but any kernel call raises.

Reference name            here
------------------------  -----------------------------------------------
models.resnet18(...)      build_model(arch, num_classes)          (resnet.py)
DistributedDataParallel   FlatDDP / DistributedDataParallel        (ddp.py)
nn.CrossEntropyLoss       CrossEntropyLoss                         (loss.py)
accuracy                  accuracy                                 (loss.py)
SGD                       FlatSGD / SGD                            (optim.py)
DistributedSampler        DistributedSampler                       (sampler.py)
transforms + DataLoader   DeviceInputStage, SyntheticImageNet      (data.py)
datasets.ImageNet + DataLoader  ImageFolderIndex, ImageFolderLoader  (folder.py)
train / validate / ...    engine.py (same names and signatures; GraphedStep / SegmentedGraphStep replay
                          an iteration as CUDA graph(s))
run / CLI                 imagenet.py
"""
from . import _cabi  # noqa: F401
from ._cabi import B200Error  # noqa: F401
from .data import DeviceInputStage, SyntheticImageNet, draw_batch_params, rrc_params  # noqa: F401
from .ddp import DistributedDataParallel, FlatDDP  # noqa: F401
from .folder import ImageFolderIndex, ImageFolderLoader  # noqa: F401
from .engine import (AverageMeter, GraphedStep, SegmentedGraphStep, adjust_learning_rate, reduce_tensor, to_python_float, train,  # noqa: F401
                     validate)
from .loss import CrossEntropyLoss, accuracy  # noqa: F401
from .optim import SGD, FlatSGD  # noqa: F401
from .resnet import B200ResNet, ParamArena, build_model  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401

__all__ = ["B200Error", "B200ResNet", "ParamArena", "build_model", "FlatDDP", "DistributedDataParallel",
           "CrossEntropyLoss", "accuracy", "FlatSGD", "SGD", "DistributedSampler", "DeviceInputStage",
           "SyntheticImageNet", "ImageFolderIndex", "ImageFolderLoader", "draw_batch_params", "rrc_params", "AverageMeter", "reduce_tensor",
           "to_python_float", "adjust_learning_rate", "train", "validate", "GraphedStep", "SegmentedGraphStep"]
