"""Command-line trainer with the reference's flags, outputs and epoch loop (imagenet.py:213-452).

    srun python -m imagent_distributed_training_pytorch_with_slurm_b200.imagenet --backend=nccl        # Slurm
    python -m torch.distributed.run --nproc-per-node 8 -m ...imagenet --synthetic 40960 --arch resnet50  # one box

Kept byte for byte: the six flags and their defaults (imagenet.py:437-449), the rank summary
(:252-262), SGD hyper-parameters (:325), LR schedule (:154-162), the epoch summary lines (:398-403),
the TensorBoard directory / tags / step offsets (:363,405-421: ``lr`` is logged at ``epoch``, the rest
at ``epoch+1``), the best-epoch bookkeeping and the ``module.``-prefixed float32 checkpoint
``imagenet_FR_resnet18.pt`` (:388-392; ``imagenet_FR_<arch>.pt`` for other archs).

Added flags default to the reference's behaviour where it has one: ``--arch resnet18``,
``--image-size 448``, float32 arithmetic (``--bf16`` selects the bf16 path), cuDNN autotuning off
(``--cudnn-benchmark`` turns it on).  ``--data DIR`` points at the dataset (default: the reference's ``../data/imagenet``; layout
``DIR/{train,val}/<class>/<image>``, what ``datasets.ImageNet`` reads, imagenet.py:287-296); ``--synthetic N`` (images
per epoch) selects seeded synthetic uint8 data instead.  Either way the epoch walks this package's ``DistributedSampler``
with ``set_epoch`` (imagenet.py:346-347,375).
Rendezvous: the reference's SLURM variables (:225-244) when present, else torchrun's
``RANK/LOCAL_RANK/WORLD_SIZE``, else a single process.
"""
import argparse
import os
import socket
import subprocess

import torch
import torch.distributed as dist

from . import data as b200_data
from .ddp import FlatDDP
from .engine import adjust_learning_rate, train, validate
from .folder import ImageFolderIndex, ImageFolderLoader
from .loss import CrossEntropyLoss
from .optim import FlatSGD
from .resnet import build_model
from .sampler import DistributedSampler


def _rendezvous():
    env = os.environ
    if "SLURM_PROCID" in env and "SLURM_NTASKS" in env:
        n_nodes = int(env["SLURM_JOB_NUM_NODES"])
        node_id = int(env["SLURM_NODEID"])
        local_rank = int(env["SLURM_LOCALID"])
        global_rank = int(env["SLURM_PROCID"])
        world_size = int(env["SLURM_NTASKS"])
        hostnames = subprocess.check_output(["scontrol", "show", "hostnames", env["SLURM_JOB_NODELIST"]])
        env["MASTER_ADDR"] = hostnames.split()[0].decode("utf-8")
        env["MASTER_PORT"] = str(29500)
    else:
        n_nodes, node_id = int(env.get("GROUP_WORLD_SIZE", 1)), int(env.get("GROUP_RANK", 0))
        local_rank = int(env.get("LOCAL_RANK", 0))
        global_rank = int(env.get("RANK", 0))
        world_size = int(env.get("WORLD_SIZE", 1))
        env.setdefault("MASTER_ADDR", "127.0.0.1")
        env.setdefault("MASTER_PORT", str(29500))
    env["WORLD_SIZE"] = str(world_size)
    env["RANK"] = str(global_rank)
    return n_nodes, node_id, local_rank, global_rank, world_size


def make_loaders(args, world_size, global_rank, dtype):
    """Datasets, samplers and loaders as run() builds them (imagenet.py:280-296,346-359).  Returns
    (train_loader, val_loader, train_sampler, val_sampler, num_classes, n_train, n_val)."""
    if not args.synthetic and not args.data:
        # the reference's hard-wired location (imagenet.py:285-289): ../data/imagenet relative to the working directory
        args.data = os.path.join(os.path.abspath(os.path.join(".", os.pardir)), "data/imagenet")
    ref_mode = not args.random_crop
    mean, std = (b200_data.REFERENCE_MEAN, b200_data.REFERENCE_STD) if ref_mode else \
        (b200_data.IMAGENET_MEAN, b200_data.IMAGENET_STD)
    if args.synthetic:
        # a virtual dataset of --synthetic images (validation: 1/25 of that), walked in DistributedSampler order
        num_classes = 1000
        n_train, n_val = args.synthetic, max(1, args.synthetic // 25)
        train_sampler = DistributedSampler(n_train, num_replicas=world_size, rank=global_rank)
        val_sampler = DistributedSampler(n_val, num_replicas=world_size, rank=global_rank)
        common = dict(src_size=args.source_size, out_size=args.image_size, num_classes=num_classes, rank=global_rank,
                      mean=mean, std=std, dtype=dtype, host=True)
        train_loader = b200_data.SyntheticImageNet(args.batch_size, 0, random_crop=not ref_mode, sampler=train_sampler,
                                                   **common)
        val_loader = b200_data.SyntheticImageNet(args.batch_size, 0, random_crop=False, seed=4321, sampler=val_sampler,
                                                 **common)
    else:
        # datasets.ImageNet(root, split) x2, DistributedSampler x2, DataLoader x2 (imagenet.py:293-296,346-359)
        trainset = ImageFolderIndex(args.data, "train")
        valset = ImageFolderIndex(args.data, "val")
        num_classes = len(trainset.classes)
        n_train, n_val = len(trainset), len(valset)
        train_sampler = DistributedSampler(trainset, num_replicas=world_size, rank=global_rank)
        val_sampler = DistributedSampler(valset, num_replicas=world_size, rank=global_rank)
        common = dict(out_size=args.image_size, mean=mean, std=std, dtype=dtype, decoder=args.decoder,
                      workers=args.workers, rank=global_rank)
        train_loader = ImageFolderLoader(trainset, train_sampler, args.batch_size, random_crop=not ref_mode, **common)
        val_loader = ImageFolderLoader(valset, val_sampler, args.batch_size, random_crop=False, **common)
    return train_loader, val_loader, train_sampler, val_sampler, num_classes, n_train, n_val


def run(args):
    torch.manual_seed(args.seed)
    n_nodes, node_id, local_rank, global_rank, world_size = _rendezvous()
    n_gpu_per_node = world_size // n_nodes
    is_master = node_id == 0 and local_rank == 0

    prefix = "%i - " % global_rank
    print(prefix + "Number of nodes: %i" % n_nodes)
    print(prefix + "Node ID        : %i" % node_id)
    print(prefix + "Local rank     : %i" % local_rank)
    print(prefix + "Global rank    : %i" % global_rank)
    print(prefix + "World size     : %i" % world_size)
    print(prefix + "GPUs per node  : %i" % n_gpu_per_node)
    print(prefix + "Master         : %s" % str(is_master))
    print(prefix + "Multi-node     : %s" % str(n_nodes > 1))
    print(prefix + "Multi-GPU      : %s" % str(world_size > 1))
    print(prefix + "Hostname       : %s" % socket.gethostname())

    torch.cuda.set_device(local_rank)
    if args.cudnn_benchmark:          # the reference leaves cuDNN's autotuner off (the PyTorch default)
        torch.backends.cudnn.benchmark = True
    print("Initializing PyTorch distributed ...")
    dist.init_process_group(init_method="env://", backend=args.backend)
    print(f"Backend: {dist.get_backend()}")

    print("Initialize Dataloaders...")
    dtype = torch.bfloat16 if args.bf16 else torch.float32      # the reference trains in float32 (no autocast anywhere)
    train_loader, val_loader, train_sampler, val_sampler, num_classes, n_train, n_val = make_loaders(
        args, world_size, global_rank, dtype)
    print("Training samples: {} images ".format(n_train))
    print("Test samples: {} images ".format(n_val))
    print("number of classes: {}".format(num_classes))

    print("Initialize Model...")
    model = build_model(args.arch, num_classes=num_classes, dtype=dtype).cuda()
    model = FlatDDP(model, device_ids=[local_rank], output_device=local_rank)
    criterion = CrossEntropyLoss().cuda()
    optimizer = FlatSGD(model.parameters(), args.lr, momentum=0.9, weight_decay=1e-4)

    writer = None
    if is_master:
        from torch.utils.tensorboard import SummaryWriter

        writer = SummaryWriter("imagenet_FR")

    best_top5 = best_top1 = 0
    total_time = 0
    best_epoch_top1 = best_epoch_top5 = 0
    ckpt = "imagenet_FR_resnet18.pt" if args.arch == "resnet18" else f"imagenet_FR_{args.arch}.pt"
    for epoch in range(args.epochs):
        train_sampler.set_epoch(epoch)            # imagenet.py:375 (the validation sampler stays at epoch 0, as there)
        lr = adjust_learning_rate(args.lr, optimizer, epoch)
        train_loss, train_prec1, train_prec5, epoch_t1 = train(train_loader, model, criterion, optimizer, epoch,
                                                               world_size)
        val_loss, val_prec1, val_prec5, epoch_t2 = validate(val_loader, model, criterion, world_size)
        total_time += epoch_t1 + epoch_t2
        if val_prec1 > best_top1:
            best_top1, best_epoch_top1 = val_prec1, epoch
            if is_master and args.save_model:
                torch.save(model.state_dict(), ckpt)
        if val_prec5 > best_top5:
            best_top5, best_epoch_top5 = val_prec5, epoch
        if is_master:
            print(f"Epoch {epoch+1} Summary: ")
            print(f"\tLearning rate: {lr}")
            print(f"\tTrain loss: {train_loss} ; Test loss: {val_loss}")
            print(f"\tTrain top1 accuracy: {train_prec1} ; Test top1 accuracy: {val_prec1}")
            print(f"\tTrain top5 accuracy: {train_prec5} ; Test top5 accuracy: {val_prec5}")
            print(f"\tTrain time: {epoch_t1} seconds; Test time:{epoch_t2} seconds")
            writer.add_scalars("Loss", {"train": train_loss, "val": val_loss}, epoch + 1)
            writer.add_scalars("Top1 accuracy", {"train": train_prec1, "val": val_prec1}, epoch + 1)
            writer.add_scalars("Top5 accuracy", {"train": train_prec5, "val": val_prec5}, epoch + 1)
            writer.add_scalar("lr", lr, epoch)
    if is_master:
        print("\n")
        print("Training Summary:")
        print(f"\tTop1 best epoch: {best_epoch_top1}")
        print(f"\t Best Test top1 accuracy: {best_top1}")
        print(f"\tTop5 best epoch: {best_epoch_top5}")
        print(f"\t Best Test top5 accuracy: {best_top5}")
        print(f"\tTraining time: {total_time/60} minutes")
        writer.close()
    dist.destroy_process_group()


def build_parser():
    parser = argparse.ArgumentParser(description="Imagenet Pytorch")
    # the reference's flags (imagenet.py:437-449), defaults unchanged
    parser.add_argument("--seed", default=0, type=int, help="seed for initializing training. ")
    parser.add_argument("--backend", type=str, default="nccl")
    parser.add_argument("--batch-size", type=int, default=128, metavar="N", help="input batch size per GPU")
    parser.add_argument("--epochs", type=int, default=100, metavar="N", help="number of epochs to train")
    parser.add_argument("--lr", type=float, default=0.1, metavar="LR", help="learning rate (default: .1)")
    parser.add_argument("--save-model", action="store_true", default=False, help="For Saving the current Model")
    # additions
    parser.add_argument("--arch", type=str, default="resnet18", help="torchvision ResNet constructor name")
    parser.add_argument("--image-size", type=int, default=448, help="network input size (reference: 448)")
    parser.add_argument("--source-size", type=int, default=500, help="synthetic source image size")
    parser.add_argument("--synthetic", type=int, default=0, metavar="IMAGES", help="synthetic images per epoch")
    parser.add_argument("--data", type=str, default=None, metavar="DIR",
                        help="dataset root holding train/<class>/*.JPEG and val/<class>/*.JPEG, i.e. what "
                             "datasets.ImageNet reads after extraction (default: the reference's ../data/imagenet)")
    parser.add_argument("--decoder", type=str, default="pil", choices=["pil", "nvjpeg", "torchvision"],
                        help="JPEG decoding: Pillow in host threads (the reference's pixels), nvJPEG's batched API on the "
                             "GPU (JPEG engines where the files allow), or nvJPEG through torchvision")
    parser.add_argument("--workers", type=int, default=10, help="host threads reading / decoding files (reference: 10 "
                                                                "DataLoader workers)")
    parser.add_argument("--random-crop", action="store_true", default=False,
                        help="RandomResizedCrop + horizontal flip + ImageNet mean/std instead of the reference's "
                             "plain resize and 0.5/0.5 normalisation")
    parser.add_argument("--bf16", action="store_true", default=False,
                        help="bf16 activations and bf16 tensor-core convolutions over float32 master weights (the fast "
                             "path); default is float32 like the reference")
    parser.add_argument("--fp32", action="store_true", default=False, help="float32 activations (the default; kept so "
                                                                           "that older command lines still parse)")
    parser.add_argument("--cudnn-benchmark", action="store_true", default=False,
                        help="let cuDNN autotune its convolution algorithms (the reference does not)")
    return parser


def main(argv=None):
    run(build_parser().parse_args(argv))


if __name__ == "__main__":
    main()
