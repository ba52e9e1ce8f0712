"""The C-ABI library loads and exports every symbol include/b200_imagenet.h declares (no compute)."""
import ctypes
import os
import re

import imagent_distributed_training_pytorch_with_slurm_b200._cabi as cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200_imagenet.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 14
    lib = ctypes.CDLL(cabi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(cabi.SIGNATURES), set(names) ^ set(cabi.SIGNATURES)


def test_jpeg_library_exports_every_declared_symbol():
    """include/b200_jpeg.h <-> libb200_jpeg.so <-> the ctypes table (loading it needs libnvjpeg, no GPU)."""
    import imagent_distributed_training_pytorch_with_slurm_b200._jpeg as jpeg

    src = open(os.path.join(ROOT, "include", "b200_jpeg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = sorted(set(re.findall(r"\b(b200_jpeg_[a-z0-9_]+)\s*\(", src)))
    assert len(names) == 7 and set(names) == set(jpeg.SIGNATURES), set(names) ^ set(jpeg.SIGNATURES)
    lib = ctypes.CDLL(jpeg.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    assert jpeg.load().b200_jpeg_error_string(1004).decode().endswith("not supported")
    assert jpeg.load().b200_jpeg_error_string(0) == b"success"


def test_binding_loads_and_reports_version_and_sizes():
    lib = cabi.load()
    assert lib.b200_abi_version() == cabi.ABI_VERSION
    assert lib.b200_error_string(1).decode().lower().startswith("invalid")
    assert lib.b200_ce_topk_workspace_bytes(256) == 16 + 12 * 256
    assert lib.b200_bn_workspace_bytes(256 * 56 * 56, 64, cabi.BF16) > 0
    assert lib.b200_input_workspace_bytes(4, 256, 256, 224, 224) > 0
    assert lib.b200_bn_vec(cabi.F32) == 4 and lib.b200_bn_vec(cabi.BF16) in (4, 8)
    assert lib.b200_tune_set(b"no_such_key", 1) != 0


def test_argument_errors_are_reported_not_crashed():
    lib = cabi.load()
    # null pointers / bad sizes are rejected before any launch
    assert lib.b200_sgd_flat_f32(None, None, None, None, 16, 0.1, 0.0, 0.9, 1.0, 0, None) == 1
    assert lib.b200_ce_topk_fwd(None, None, None, None, None, 0, 4, 10, 0, None) == 1
    assert lib.b200_bn_fwd_train(None, None, None, None, None, None, None, None, None, None, None, 0, 8, 6, 1e-5, 0.1,
                                 0, 1, None) == 1  # C=6 not a multiple of the bf16 vector width


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    """No fallback: without the built CUDA library every kernel call raises (nothing routes to the oracle)."""
    import pytest
    import torch

    monkeypatch.setattr(cabi, "_lib", None)
    monkeypatch.setattr(cabi, "LIB_PATH", str(tmp_path / "libb200_imagenet.so"))
    with pytest.raises(cabi.B200Error, match="no CPU or PyTorch fallback"):
        cabi.load()
    with pytest.raises(cabi.B200Error):
        cabi.tune_set("bn_pdl", 1)
    # and with the library present, CPU tensors are refused rather than computed some other way
    monkeypatch.undo()
    with pytest.raises(cabi.B200Error, match="CUDA tensors"):
        cabi.sgd_flat(torch.zeros(8), torch.zeros(8), torch.zeros(8), None, 0.1, 0.0, 0.9, 1.0, True)


def test_product_package_never_imports_the_oracle():
    import pathlib
    pkg = pathlib.Path(cabi.__file__).parent
    for f in pkg.rglob("*.py"):
        src = f.read_text()
        assert "import oracle" not in src and "from oracle" not in src, f
