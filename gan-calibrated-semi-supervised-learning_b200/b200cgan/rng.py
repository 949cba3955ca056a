"""Counter-based random stream for dropout masks (replaces torch's bernoulli_ behind nn.Dropout, cgan/models.py:62,77)
and the gradient-penalty interpolation coefficients (torch.rand at cgan/losses.py:199).

Every draw is a pure function of (seed, global stream position), and the position of a per-sample quantity is keyed by
the GLOBAL sample index: a data-parallel rank that holds samples [rank*B, (rank+1)*B) of the global batch reads exactly
the slice of the stream the single-GPU run would use for those samples, and every rank advances the stream by the
GLOBAL element count.  So a 1 x (N*B) run and an N x B run draw identical masks / alphas (SURVEY.md App. E;
tests/test_dp2_gpu.py).  set_shard(rank, world) selects the slice; the default (0, 1) is the single-GPU stream."""
from __future__ import annotations

import torch

from . import _lib

_state = {"seed": 42, "offset": 0, "base": None, "rank": 0, "world": 1}


def manual_seed(seed: int, offset: int = 0):
    _state["seed"], _state["offset"] = int(seed), int(offset)


def set_shard(rank: int, world: int):
    """This process holds the rank-th of `world` equal contiguous slices of the global batch."""
    _state["rank"], _state["world"] = int(rank), int(world)


def set_device_base(base):
    """base: int64 device scalar added to the host offset inside the kernel (None = host offset only).  A captured
    CUDA graph records host offsets relative to the start of the step and advances `base` between replays."""
    _state["base"] = base


def get_state():
    return dict(_state)


def _advance(local_count: int) -> int:
    """Position of this rank's slice for a draw of `local_count` local elements; advances by the global count."""
    pos = _state["offset"] + _state["rank"] * local_count
    _state["offset"] += (_state["world"] * local_count + 7) // 8 * 8  # masks are drawn 8 elements at a time
    return pos


def skip(local_count: int):
    """Advance the stream as if `local_count` local elements had been drawn (a forward that is not executed)."""
    _advance(local_count)


def next_mask(numel: int, device):
    m = torch.empty(numel, dtype=torch.uint8, device=device)
    _lib.call("b2c_dropout_mask", m, numel, _state["seed"], _advance(numel), _state["base"])
    return m


def next_uniform(n: int, device):
    """n per-sample U[0,1) draws (fp32)."""
    out = torch.empty(n, dtype=torch.float32, device=device)
    _lib.call("b2c_uniform", out, n, _state["seed"], _advance(n), _state["base"])
    return out
