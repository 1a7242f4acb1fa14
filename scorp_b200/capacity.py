"""Opt-in: forward passes without the host round trip (include/gsr_b200.h: gsr_rasterize_forward_nosync).

The reference synchronises with the GPU in the middle of every forward to read the number of (Gaussian, tile)
instances (cuda_rasterizer/rasterizer_impl.cu:278-286); the stock entry points of this package keep that. A training
loop that synchronises once per iteration anyway (`loss.item()`, train_3dgs.py:158) can move the check there:

    from scorp_b200 import capacity
    capacity.enable()                      # GaussianRasterizer.forward of both packages stops synchronising
    for it in range(...):
        loss = step()                      # render(), loss, backward, optimizer: all enqueued, nothing waits
        value = loss.item()                # the loop's own synchronisation point
        if capacity.overflowed():          # rare: this view had more instances than any view before, plus the margin
            redo the iteration             # (the capacity has been raised; the step's outputs held the background only)

How the size is chosen: per (rasterizer, P, W, H) the largest instance count seen so far is kept; the binning buffer of a
no-sync forward holds `margin` x that plus `slack` instances. The first forward of a shape runs through the stock
(synchronising) path to learn the count. Every no-sync forward leaves its {num_rendered, longest list, overflow} in a
pinned slot; `overflowed()` / `check()` read the slots of the forwards issued since the last call (after synchronising
the current stream, which costs nothing when the caller just did) and raise the high-water marks.

CUDA graphs: a forward captured while this is enabled is recorded with the capacity in force at capture time; its
status slot is re-read by every `check()` / `overflowed()` after a replay. On overflow: re-capture.
"""
from __future__ import annotations

import threading
from typing import Dict, List, Optional, Tuple

import torch


class InstanceCapacityExceeded(RuntimeError):
    """A no-sync forward met more (Gaussian, tile) instances than its binning buffer holds; its outputs are the
    background only. The capacity has been raised: repeat the step."""


class _State(threading.local):
    def __init__(self):
        self.enabled = False
        self.margin = 1.25
        self.slack = 1 << 16
        self.high: Dict[Tuple, int] = {}
        self.pending: List[Tuple[Tuple, torch.Tensor, int, bool]] = []     # (key, status slot, capacity, captured)
        self.ring: Optional[torch.Tensor] = None
        self.ring_next = 0


_S = _State()
_RING = 256


def enable(margin: float = 1.25, slack: int = 1 << 16) -> None:
    if margin < 1.0 or slack < 0:
        raise ValueError("margin >= 1 and slack >= 0 expected")
    _S.enabled, _S.margin, _S.slack = True, float(margin), int(slack)


def disable() -> None:
    _S.enabled = False


def enabled() -> bool:
    return _S.enabled


def reset() -> None:
    """Forget every high-water mark and pending status (tests)."""
    _S.high.clear()
    _S.pending.clear()


def observe(key: Tuple, num_rendered: int) -> None:
    """A synchronising forward reports its instance count."""
    if num_rendered > _S.high.get(key, 0):
        _S.high[key] = int(num_rendered)


def capacity_for(key: Tuple) -> Optional[int]:
    """Instances the binning buffer of the next no-sync forward of this shape holds; None: not known yet (run the
    stock path once)."""
    hw = _S.high.get(key)
    if hw is None:
        return None
    return min(int(hw * _S.margin) + _S.slack, 0x7fffffff)


def status_slot() -> torch.Tensor:
    """A pinned int32[4] slot for one forward. Eager calls share a ring (a slot is reused after _RING forwards, long
    after check() has read it); a call under stream capture gets a slot of its own, since every replay writes to it."""
    if torch.cuda.is_current_stream_capturing():
        return torch.zeros(4, dtype=torch.int32).pin_memory()
    if _S.ring is None:
        _S.ring = torch.zeros((_RING, 4), dtype=torch.int32).pin_memory()
    slot = _S.ring[_S.ring_next]
    _S.ring_next = (_S.ring_next + 1) % _RING
    return slot


def note(key: Tuple, slot: torch.Tensor, capacity: int) -> None:
    _S.pending.append((key, slot, int(capacity), torch.cuda.is_current_stream_capturing()))
    if len(_S.pending) > _RING // 2:          # nobody is checking: keep the newest ones (their slots are still theirs)
        del _S.pending[: len(_S.pending) - _RING // 2]


def overflowed(synchronize: bool = True) -> bool:
    """True if a no-sync forward issued since the last call (or a captured one, since its last replay) exceeded its
    capacity. Raises the high-water marks either way. Call after the step's synchronisation point."""
    if not _S.pending:
        return False
    if synchronize and torch.cuda.is_available():
        torch.cuda.current_stream().synchronize()
    bad = False
    keep = []
    for key, slot, cap, captured in _S.pending:
        n, _, over, _ = (int(v) for v in slot.tolist())
        observe(key, n)
        bad = bad or over != 0 or n > cap
        if captured:
            keep.append((key, slot, cap, captured))
    _S.pending[:] = keep
    return bad


def forget_captured() -> None:
    """Drop the status slots of captured forwards (call before capturing a graph again)."""
    _S.pending[:] = [e for e in _S.pending if not e[3]]


def check() -> None:
    """`overflowed()` as an exception."""
    if overflowed():
        raise InstanceCapacityExceeded(
            "a forward exceeded the capacity of its binning buffer; the capacity has been raised, repeat the step "
            "(a captured CUDA graph must be captured again)")
