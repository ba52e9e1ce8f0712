import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device here")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture
def cpu_kernels(monkeypatch):
    """Swap the C-ABI kernel wrappers for the ORACLE's CPU restatements so the host-side logic
    (arena layout, tape, hand-written backward, bucket all-reduce, optimizer plumbing) can be
    exercised without a GPU.  Test-only: the product never does this."""
    from tests import cpu_kernel_standins as s
    import imagent_distributed_training_pytorch_with_slurm_b200._cabi as cabi

    for name in ("bn_fwd_train", "bn_fwd_eval", "bn_bwd", "ce_topk_fwd", "ce_bwd", "sgd_flat", "input_stage", "maxpool_fwd",
                 "maxpool_bwd", "bn_relu_maxpool_fwd", "pack_grads", "bn_dual_supported", "bn_stats", "bn_fwd_train_dual", "bn_bwd_dual"):
        monkeypatch.setattr(cabi, name, getattr(s, name))
    return s
