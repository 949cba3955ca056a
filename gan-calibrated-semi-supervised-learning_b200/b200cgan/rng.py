"""Counter-based dropout-mask stream (replaces torch's bernoulli_ behind nn.Dropout, cgan/models.py:62,77).

Masks are a pure function of (seed, running offset), so a data-parallel run that sets the same seed and advances the
offset by GLOBAL element counts draws the same masks as the single-GPU run."""
from __future__ import annotations

import torch

from . import _lib

_state = {"seed": 42, "offset": 0, "base": None}


def manual_seed(seed: int, offset: int = 0):
    _state["seed"], _state["offset"] = int(seed), int(offset)


def set_device_base(base):
    """base: int64 device scalar added to the host offset inside the kernel (None = host offset only).  A captured
    CUDA graph records host offsets relative to the start of the step and advances `base` between replays."""
    _state["base"] = base


def get_state():
    return dict(_state)


def next_mask(numel: int, device):
    m = torch.empty(numel, dtype=torch.uint8, device=device)
    _lib.call("b2c_dropout_mask", m, numel, _state["seed"], _state["offset"], _state["base"])
    _state["offset"] += numel
    return m
