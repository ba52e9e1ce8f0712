"""NVTX ranges around the phases of an iteration (input / forward / loss / backward / exchange / step), so a
timeline (Nsight Systems, or ncu's --nvtx filters) shows the step's structure.  The reference has no tracing beyond
wall-clock meters (imagenet.py:103-104,116,148); SURVEY section 5 adds these.  A push/pop pair costs well under a
microsecond when no profiler is attached; ``B200_NVTX=0`` turns them into no-ops."""
import os

import torch

_ON = os.environ.get("B200_NVTX", "1") != "0"
try:
    _push, _pop = torch.cuda.nvtx.range_push, torch.cuda.nvtx.range_pop
    if _ON:
        _push("b200.probe")
        _pop()
except Exception:  # noqa: BLE001  (a torch build without NVTX)
    _ON = False


def push(name):
    if _ON:
        _push(name)


def pop():
    if _ON:
        _pop()


class range:  # noqa: A001  (mirrors torch.cuda.nvtx.range)
    __slots__ = ("name",)

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        push(self.name)

    def __exit__(self, *exc):
        pop()
        return False
