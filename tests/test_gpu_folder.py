"""The on-disk front end on the GPU: class folders -> DistributedSampler order -> decode -> ragged
input kernel, against the oracle (Pillow decode = torchvision's pil_loader, then oracle/input_pipeline.py, which is
pinned bit-exact on torchvision-on-PIL transforms by tests/golden/input_case.npz).

Tolerances: with ``decoder="pil"`` the whole path is BIT-EXACT (decode is the reference's own; K1 is bit-exact).
With ``decoder="nvjpeg"`` the decoded pixels are nvJPEG's, which differ from libjpeg-turbo's in IDCT rounding and
chroma up-sampling: the decoded slab must agree with Pillow's within 4 grey levels on 4:4:4 / grayscale files (measured:
1 and 3) and within mean 3.5 / max 24 grey levels on 4:2:0 files (measured: mean 1.1-2.3, max 6-11;
profiles/r02_nvjpeg_vs_pillow.json); everything AFTER the
decode is again bit-exact given the decoded pixels."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
from PIL import Image

from oracle import input_pipeline as oin
from oracle import sampler as osamp

import imagent_distributed_training_pytorch_with_slurm_b200 as b200
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi
from tests.folder_fixture import make_dataset

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_MEAN = REF_STD = (0.5, 0.5, 0.5)


def _pil(path):
    return np.asarray(Image.open(path).convert("RGB"))        # torchvision.datasets.folder.pil_loader


def _nhwc(x):
    return x.float().cpu().permute(0, 2, 3, 1).numpy()


@pytest.mark.parametrize("world,rank", [(1, 0), (2, 1)])
def test_pil_path_reference_mode_is_bit_exact_in_sampler_order(tmp_path, world, rank):
    idx = b200.ImageFolderIndex(make_dataset(str(tmp_path)), "train")
    sampler = b200.DistributedSampler(idx, num_replicas=world, rank=rank)
    sampler.set_epoch(2)
    want_ids = osamp.sampler_indices_oracle(len(idx), world, rank, 2).tolist()
    assert sampler.indices().tolist() == want_ids
    loader = b200.ImageFolderLoader(idx, sampler, batch_size=4, out_size=56, mean=REF_MEAN, std=REF_STD, random_crop=False,
                                    dtype=torch.float32, decoder="pil", workers=3, layout=_cabi.LAYOUT_NHWC)
    assert len(loader) == -(-len(want_ids) // 4)
    seen = 0
    for x, t in loader:
        ids = want_ids[seen:seen + 4]                              # the last batch is short (no drop_last)
        assert x.shape == (len(ids), 3, 56, 56) and t.tolist() == [idx.samples[i][1] for i in ids]
        got = _nhwc(x)
        for k, i in enumerate(ids):
            img = _pil(idx.samples[i][0])
            h, w = img.shape[:2]
            want = oin.input_stage_oracle(img[None], np.array([[0, 0, h, w]]), np.array([False]), 56, 56, REF_MEAN, REF_STD)
            assert np.array_equal(got[k], want[0]), (i, idx.samples[i][0])
        seen += len(ids)
    assert seen == len(want_ids)


def test_pil_path_random_crop_draws_and_stem_layout(tmp_path):
    """RandomResizedCrop + flip on images of different sizes: the boxes are torchvision's get_params draws for each
    image's own size, in loader order, from the per-rank generator; the stem-layout output equals the dense kernel's
    output for every image taken alone (bit for bit)."""
    idx = b200.ImageFolderIndex(make_dataset(str(tmp_path)), "train")
    sampler = b200.DistributedSampler(idx, num_replicas=1, rank=0)
    loader = b200.ImageFolderLoader(idx, sampler, batch_size=6, out_size=48, random_crop=True, dtype=torch.float32,
                                    decoder="pil", seed=77, rank=0, layout=_cabi.LAYOUT_NHWC)
    stem = b200.ImageFolderLoader(idx, sampler, batch_size=6, out_size=48, random_crop=True, dtype=torch.bfloat16,
                                  decoder="pil", seed=77, rank=0)
    g = torch.Generator().manual_seed(77)
    ids = sampler.indices().tolist()
    mean, std = b200.data.IMAGENET_MEAN, b200.data.IMAGENET_STD
    seen = 0
    for (x, t), (xs, ts) in zip(loader, stem):
        got = _nhwc(x)
        assert torch.equal(t, ts) and getattr(xs, "_b200_s2d_stem", None) == (48, 48)
        boxes, flips = loader.last_batch["boxes"].cpu(), loader.last_batch["flips"].cpu()
        for k in range(x.shape[0]):
            img = _pil(idx.samples[ids[seen + k]][0])
            box = oin.rrc_params_oracle(img.shape[0], img.shape[1], generator=g)
            fl = oin.flip_draw_oracle(generator=g)
            assert boxes[k].tolist() == list(box) and bool(flips[k]) == bool(fl)
            want = oin.input_stage_oracle(img[None], np.array([box]), np.array([fl]), 48, 48, mean, std)
            assert np.array_equal(got[k], want[0])
            dense = _cabi.input_stage(torch.from_numpy(img[None].copy()).cuda(), boxes[k:k + 1].cuda(), flips[k:k + 1].cuda(),
                                      48, 48, mean, std, torch.bfloat16, _cabi.LAYOUT_S2D_STEM)
            assert torch.equal(dense[0], xs[k])
        seen += x.shape[0]
    assert seen == len(ids)


@pytest.mark.parametrize("decoder", ["nvjpeg", "torchvision"])
def test_nvjpeg_path_agrees_with_pillow_within_decoder_tolerance(tmp_path, decoder):
    """decoder="nvjpeg": nvJPEG's batched API driven by libb200_jpeg.so (hardware engine or GPU-assisted backend) writing
    interleaved RGB into the slab; decoder="torchvision": nvJPEG through torchvision + the planar re-pack kernel."""
    idx = b200.ImageFolderIndex(make_dataset(str(tmp_path)), "train")
    sampler = b200.DistributedSampler(idx, num_replicas=1, rank=0, shuffle=False)
    loader = b200.ImageFolderLoader(idx, sampler, batch_size=8, out_size=56, mean=REF_MEAN, std=REF_STD, random_crop=False,
                                    dtype=torch.float32, decoder=decoder, layout=_cabi.LAYOUT_NHWC)
    seen, report = 0, []
    for x, t in loader:
        lb = loader.last_batch
        slab, table = lb["slab"].cpu().numpy(), lb["table"].cpu().numpy()
        got = _nhwc(x)
        for k in range(x.shape[0]):
            path = idx.samples[seen + k][0]
            ref = _pil(path)
            off, h, w, pitch = (int(v) for v in table[k])
            assert (h, w) == ref.shape[:2] and pitch == 3 * w
            dec = slab[off:off + h * w * 3].reshape(h, w, 3)
            d = np.abs(dec.astype(np.int32) - ref.astype(np.int32))
            im = Image.open(path)
            exact_decoder = im.format != "JPEG" or im.mode == "CMYK"          # decoded by Pillow on the host
            full_chroma = im.mode == "L" or "img_03" in path and "n03" in path  # grayscale / 4:4:4
            kind = "pillow" if exact_decoder else ("full-chroma" if full_chroma else "4:2:0")
            report.append({"file": os.path.relpath(path, str(tmp_path)), "kind": kind, "size": [h, w],
                           "max": int(d.max()), "mean": float(d.mean())})
            # everything after the decode is bit-exact on the decoded pixels
            want = oin.input_stage_oracle(dec[None], np.array([[0, 0, h, w]]), np.array([False]), 56, 56, REF_MEAN, REF_STD)
            assert np.array_equal(got[k], want[0]), path
        assert t.tolist() == [idx.samples[seen + k][1] for k in range(x.shape[0])]
        seen += x.shape[0]
    assert seen == len(idx)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        if decoder == "nvjpeg":
            report.append({"hardware_engine_images_in_last_batch": getattr(loader, "last_hw_count", None)})
        json.dump(report, open(os.path.join(out, f"{decoder}_vs_pillow.json"), "w"), indent=1)
        report = [r for r in report if "kind" in r]
    for r in report:
        if r["kind"] == "pillow":
            assert r["max"] == 0, r
        elif r["kind"] == "full-chroma":
            assert r["max"] <= 4, r                              # measured on B200: 1 (grayscale), 3 (4:4:4)
        else:
            assert r["mean"] <= 3.5 and r["max"] <= 24, r        # measured: mean 1.1-2.3, max 6-11


def test_interleave_planar_matches_permute():
    g = torch.Generator().manual_seed(1)
    sizes = [(5, 7), (33, 64), (1, 1), (64, 129), (17, 3)]
    planes = [torch.randint(0, 256, (3, h, w), dtype=torch.uint8, generator=g).cuda() for h, w in sizes]
    offs, total = b200.folder.ImageFolderLoader._layout(sizes)
    slab = torch.full((total,), 255, dtype=torch.uint8, device="cuda")
    tab = torch.tensor([[p.data_ptr(), h, w, o] for p, (h, w), o in zip(planes, sizes, offs)], dtype=torch.int64).cuda()
    _cabi.interleave_planar(tab, slab, max(h * w for h, w in sizes))
    for p, (h, w), o in zip(planes, sizes, offs):
        assert torch.equal(slab[o:o + 3 * h * w].view(h, w, 3), p.permute(1, 2, 0))
    used = torch.zeros(total, dtype=torch.bool)
    for (h, w), o in zip(sizes, offs):
        used[o:o + 3 * h * w] = True
    assert bool((slab.cpu()[~used] == 255).all())          # nothing written outside the images


@pytest.mark.parametrize("decoder", ["pil", "nvjpeg"])
def test_cli_trains_from_a_dataset_folder(tmp_path, decoder):
    data = make_dataset(str(tmp_path / "data"))
    env = dict(os.environ, PYTHONPATH=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29579", RANK="0", WORLD_SIZE="1",
               LOCAL_RANK="0")
    for k in list(env):
        if k.startswith("SLURM_"):
            del env[k]
    cmd = [sys.executable, "-m", "imagent_distributed_training_pytorch_with_slurm_b200.imagenet", "--data", data,
           "--batch-size", "4", "--epochs", "2", "--image-size", "64", "--lr", "0.01", "--decoder", decoder, "--workers", "2"]
    r = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for line in ("Training samples: 15 images ", "Test samples: 6 images ", "number of classes: 3", "Epoch 2 Summary: ",
                 "Training Summary:"):
        assert line in r.stdout, (line, r.stdout[-1500:])


@pytest.mark.parametrize("world,rank", [(1, 0), (4, 3)])
def test_cli_loaders_walk_the_sampler_in_both_modes(tmp_path, world, rank):
    """The loaders exactly as the command line builds them (imagenet.make_loaders = the data part of run()): with a
    dataset folder and with --synthetic, an epoch must serve this rank's DistributedSampler indices of THAT epoch -
    checked against the oracle sampler, which the goldens pin bit-exact on torch's (imagenet.py:346-347,375)."""
    from imagent_distributed_training_pytorch_with_slurm_b200 import imagenet as cli

    data = make_dataset(str(tmp_path / "data"))
    # --data: targets served == class of the sampler's indices, per epoch
    args = cli.build_parser().parse_args(["--data", data, "--batch-size", "4", "--image-size", "32", "--workers", "2"])
    tl, vl, ts, vs, ncls, n_train, n_val = cli.make_loaders(args, world, rank, torch.float32)
    assert (ncls, n_train, n_val) == (3, 15, 6)
    for epoch in (0, 1, 5):
        ts.set_epoch(epoch)
        want = osamp.sampler_indices_oracle(n_train, world, rank, epoch).tolist()
        got = torch.cat([t for _, t in tl]).tolist()
        assert got == [tl.index.samples[i][1] for i in want]
    want_val = osamp.sampler_indices_oracle(n_val, world, rank, 0).tolist()          # the reference never moves it
    assert torch.cat([t for _, t in vl]).tolist() == [vl.index.samples[i][1] for i in want_val]
    # --synthetic N: label of virtual image i is i % classes, so the targets reveal the index stream
    args = cli.build_parser().parse_args(["--synthetic", "1003", "--batch-size", "64", "--image-size", "32",
                                          "--source-size", "40"])
    tl, vl, ts, vs, ncls, n_train, n_val = cli.make_loaders(args, world, rank, torch.bfloat16)
    assert (ncls, n_train, n_val) == (1000, 1003, 40) and len(tl) == -(-len(ts) // 64)
    for epoch in (0, 2):
        ts.set_epoch(epoch)
        want = osamp.sampler_indices_oracle(1003, world, rank, epoch)
        batches = [(x.shape[0], t) for x, t in tl]
        got = torch.cat([t for _, t in batches]).cpu()
        assert torch.equal(got, torch.from_numpy(want) % 1000)
        assert [n for n, _ in batches][-1] == len(want) - 64 * (len(batches) - 1)      # short last batch, no drop_last
