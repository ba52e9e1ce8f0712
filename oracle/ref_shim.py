"""Import the unmodified reference ``imagenet.py`` (build container only).

``imagenet.py:36`` imports ``custom_optimizers`` (FR, Nadam), a module that is
not in the reference repository, so a bare ``import imagenet`` raises
ModuleNotFoundError.  The two names are only used in commented-out lines
(``imagenet.py:338,340``), so an empty stand-in module is enough.

``/root/reference`` does not exist on the GPU box: callers must treat a ``None``
return as "reference unavailable" and fall back to the committed golden vectors.
"""
import importlib
import os
import sys
import types

REFERENCE_DIR = os.environ.get("B200_REFERENCE_DIR", "/root/reference")


def load_reference():
    """Return the reference ``imagenet`` module, or None when it is not on disk."""
    path = os.path.join(REFERENCE_DIR, "imagenet.py")
    if not os.path.isfile(path):
        return None
    if "custom_optimizers" not in sys.modules:
        stub = types.ModuleType("custom_optimizers")
        stub.FR = type("FR", (), {})
        stub.Nadam = type("Nadam", (), {})
        sys.modules["custom_optimizers"] = stub
    spec = importlib.util.spec_from_file_location("_reference_imagenet", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
