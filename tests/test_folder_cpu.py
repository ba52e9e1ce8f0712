"""Host side of the on-disk front end (no GPU): the index must list what torchvision's ImageFolder lists, in its
order, and the sampler over it must give torch's DistributedSampler stream (imagenet.py:293-296,346-347,375)."""
import os

import torch
import torchvision

import imagent_distributed_training_pytorch_with_slurm_b200 as b200
from imagent_distributed_training_pytorch_with_slurm_b200 import folder
from tests.folder_fixture import make_dataset


def test_index_equals_torchvision_imagefolder(tmp_path):
    root = make_dataset(str(tmp_path))
    for split in ("train", "val"):
        ours = b200.ImageFolderIndex(root, split)
        ref = torchvision.datasets.ImageFolder(os.path.join(root, split))
        assert ours.classes == ref.classes and ours.class_to_idx == ref.class_to_idx
        assert ours.samples == ref.samples and ours.targets == ref.targets
    assert len(b200.ImageFolderIndex(root, "train")) == 15
    try:
        b200.ImageFolderIndex(root, "test")
        raise AssertionError("a missing split must raise")
    except FileNotFoundError:
        pass


def test_sampler_over_the_index_equals_torch_sampler(tmp_path):
    idx = b200.ImageFolderIndex(make_dataset(str(tmp_path)), "train")
    ref_ds = list(range(len(idx)))
    for world, rank, epoch in ((1, 0, 0), (2, 1, 0), (2, 0, 3), (4, 3, 1)):
        ref = torch.utils.data.distributed.DistributedSampler(ref_ds, num_replicas=world, rank=rank)
        ref.set_epoch(epoch)
        ours = b200.DistributedSampler(idx, num_replicas=world, rank=rank)
        ours.set_epoch(epoch)
        assert ours.indices().tolist() == list(ref) and len(ours) == len(ref)


def test_slab_layout_and_probe(tmp_path):
    root = make_dataset(str(tmp_path))
    idx = b200.ImageFolderIndex(root, "train")
    offs, total = folder.ImageFolderLoader._layout([(3, 5), (7, 2), (1, 1)])
    assert offs == [0, 48, 96] and total == 112 + folder._SLAB_TAIL and all(o % 16 == 0 for o in offs)
    ld = folder.ImageFolderLoader.__new__(folder.ImageFolderLoader)
    ld.limit = (8 * 64, 8 * 64)
    kinds = {}
    for path, _ in idx.samples:
        h, w, arr = ld._probe(open(path, "rb").read())
        kinds[os.path.relpath(path, root)] = arr is None
        if arr is not None:
            assert arr.shape == (h, w, 3)
    host = sorted(k for k, v in kinds.items() if not v)
    assert host == ["train/n02/img_02.JPEG", "train/n03/img_00.png"]   # the CMYK file and the PNG: decoded on the host
    ld.limit = (32, 32)                            # everything is now beyond the 8x limit: shrunk by an integer factor
    h, w, arr = ld._probe(open(idx.samples[0][0], "rb").read())
    assert arr is not None and h <= 32 and w <= 32
