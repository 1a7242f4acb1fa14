"""View-sharded multi-GPU use of the rasterizer: one process per GPU, Gaussians replicated,
every rank renders its own cameras, per-Gaussian parameter gradients are summed with ONE collective.

The reference is single-GPU (gs3dgs/utils/general_utils.py:148; one random camera per iteration,
train_3dgs.py:89); this module is the harness SURVEY.md 8(e) describes, built on two facts of this library:
  * every gradient of one backward lives in one arena whose head is the parameter gradients
    (diff_gaussian_rasterization/_C.py: grad_arena, param_grad_floats), so the exchange is one all_reduce
    over P*(11+3M) floats (P*(10+3M) for surfels) instead of five;
  * the densification statistics are per-view quantities (a norm and a max, gaussian_model.py:603-605,
    train_3dgs.py:180) and reduce separately: sum, sum, max.
Works with any torch.distributed backend (NCCL over NVLink on the GPUs; gloo in the CPU tests of the
host-side logic).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def views_for_rank(n_views: int, rank: int, world: int) -> List[int]:
    """Camera indices rank `rank` renders: v = batch*world + rank. The union over ranks is
    range(n_views), the sets are disjoint, and sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_views, world))


def param_grad_floats(M: int, surfel: bool = False) -> int:
    """floats per Gaussian at the head of a gradient arena: means3D 3, sh 3M, opacity 1, scales 3 (2), rotations 4."""
    return 3 + 3 * M + 1 + (2 if surfel else 3) + 4


def arena_of(grad: torch.Tensor) -> torch.Tensor:
    """The flat allocation a gradient returned by this library is a view of."""
    base = grad._base
    while base is not None and base._base is not None:
        base = base._base
    if base is None:
        raise ValueError("tensor is not a view of a gradient arena")
    return base


def accumulate_param_grads(total: Optional[torch.Tensor], grad_means3D: torch.Tensor, P: int, M: int,
                           surfel: bool = False) -> torch.Tensor:
    """Local sum over this rank's views before the collective (ranks with several views)."""
    head = arena_of(grad_means3D)[: P * param_grad_floats(M, surfel)]
    if total is None:
        return head.clone()
    total.add_(head)
    return total


def allreduce_param_grads(head: torch.Tensor, group=None, average: bool = False) -> torch.Tensor:
    """ONE all_reduce(sum) over the contiguous parameter-gradient head; in place."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(head, op=dist.ReduceOp.SUM, group=group)
        if average:
            head.div_(dist.get_world_size(group))
    return head


def split_param_grads(head: torch.Tensor, P: int, M: int, surfel: bool = False):
    """(means3D[P,3], sh[P,M,3], opacity[P,1], scales[P,3|2], rotations[P,4]) views of a reduced head."""
    ns = 2 if surfel else 3
    widths = (3, 3 * M, 1, ns, 4)
    outs, o = [], 0
    for w in widths:
        outs.append(head[o:o + P * w]); o += P * w
    return (outs[0].view(P, 3), outs[1].view(P, M, 3), outs[2].view(P, 1), outs[3].view(P, ns), outs[4].view(P, 4))


def reduce_densification_stats(xyz_gradient_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor,
                               group=None) -> None:
    """Per-view statistics of add_densification_stats (gaussian_model.py:603-605) and the running max radius
    (train_3dgs.py:180): every rank accumulates its own views, then sum / sum / max across ranks; in place."""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return
    dist.all_reduce(xyz_gradient_accum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)


def install_backward_allreduce(group=None, enable: bool = True) -> None:
    """Autograd users (`loss.backward()` through GaussianRasterizer) cannot reach the arena: autograd copies the
    returned views into `.grad`. With this hook installed, every backward of both rasterizer packages all-reduces
    its parameter-gradient head right after the kernels, i.e. BEFORE autograd sees the gradients, so `.grad` holds
    the sum over ranks (one collective per backward call, on the current stream)."""
    from .diff_gaussian_rasterization import _C as c3
    from .diff_surfel_rasterization import _C as c2
    hook = (lambda head: allreduce_param_grads(head, group=group)) if enable else None
    c3._backward_hook[0] = hook
    c2._backward_hook[0] = hook


def render_views_sharded(render_one, view_ids: Sequence[int], P: int, M: int, surfel: bool = False, group=None):
    """Runs `render_one(view_id) -> grad_means3D` (one forward+backward of this library, returning the
    means3D gradient, which identifies the arena) for this rank's views, sums locally, reduces once.
    Returns the reduced head (split it with split_param_grads)."""
    total = None
    for v in view_ids:
        total = accumulate_param_grads(total, render_one(v), P, M, surfel)
    if total is None:       # a rank without a view still takes part in the collective
        raise ValueError("render_views_sharded needs at least one view per rank; pad the batch")
    return allreduce_param_grads(total, group=group)
