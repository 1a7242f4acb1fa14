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

import ctypes as C
import os
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


class FactoredExchange:
    """Exchange of one backward per rank (ONE view per rank per step) that moves 3 floats per Gaussian for the SH
    part instead of 3*M: the SH gradient of a view is the outer product basis(dir) (x) dL_dRGB
    (csrc/sh_exchange.cu, include/gsr_b200.h), so ranks all-gather dL_dRGB (and the camera positions) and every
    rank rebuilds sum_v basis(dir_v) (x) dL_dRGB_v itself, in rank order. The other parameter gradients
    (means3D, opacity, scales, rotations: 11 or 10 floats) are all-reduced as before; that collective runs
    underneath the rebuild kernel. Per Gaussian on the wire: 44 B all-reduced + 12 B x world gathered, against
    236 B all-reduced (M = 16); `worth_it` says when this wins.

    `rebuild(P, M, D, n_views, means3D, campos_all, drgb_all, dsh_out)` defaults to the CUDA entry point
    gsr_sh_grad_from_views; the CPU (gloo) tests of the host logic pass their own."""

    def __init__(self, group=None, rebuild=None):
        self.group, self.rebuild = group, rebuild
        self._bufs = {}
        self._pending = ()

    @staticmethod
    def worth_it(world: int, M: int) -> bool:
        """bytes received per Gaussian: gather 12*world (+ 44 reduced) vs ~2 * 4*(11+3M) for a ring all-reduce"""
        return M > 0 and 12 * world + 2 * 44 < 2 * 4 * (11 + 3 * M)

    def _buffers(self, world, P, dev):
        key = (world, P, str(dev))
        if key not in self._bufs:
            self._bufs = {key: (torch.empty((world, P, 3), dtype=torch.float32, device=dev),
                                torch.empty((world, 3), dtype=torch.float32, device=dev))}
        return self._bufs[key]

    def _world(self):
        return dist.get_world_size(self.group) if (dist.is_available() and dist.is_initialized()) else 1

    def begin(self, drgb: torch.Tensor, campos: torch.Tensor, P: int) -> None:
        """Start gathering this rank's clamped colour gradient drgb[P,3] and camera position. Called as soon as drgb
        is final (from the library's gsr_backward_early_rgb callback, i.e. while the preprocess-backward kernel is
        still to run): NCCL orders the gather after the work already enqueued on the current stream and runs it
        on its own stream underneath what is enqueued next."""
        world = self._world()
        drgb_all, campos_all = self._buffers(world, P, drgb.device)
        if world > 1:
            self._pending = (
                dist.all_gather_into_tensor(drgb_all.view(-1), drgb.reshape(-1), group=self.group, async_op=True),
                dist.all_gather_into_tensor(campos_all.view(-1), campos.reshape(-1).contiguous(), group=self.group, async_op=True))
        else:
            drgb_all[0].copy_(drgb.view(P, 3)); campos_all[0].copy_(campos.reshape(3))
            self._pending = ()

    def finish(self, head: torch.Tensor, means3D: torch.Tensor, dsh_out: torch.Tensor, P: int, M: int, D: int) -> None:
        """All-reduce the non-SH head (in place), and rebuild the SH gradient summed over ranks into dsh_out[P,M,3]
        while that collective runs."""
        world = self._world()
        drgb_all, campos_all = self._buffers(world, P, dsh_out.device)
        w_r = dist.all_reduce(head, op=dist.ReduceOp.SUM, group=self.group, async_op=True) if world > 1 else None
        for w in self._pending:
            w.wait()
        self._pending = ()
        (self.rebuild or _rebuild_cuda)(P, M, D, world, means3D, campos_all, drgb_all, dsh_out)
        if w_r is not None:
            w_r.wait()

    def __call__(self, head: torch.Tensor, drgb: torch.Tensor, campos: torch.Tensor, means3D: torch.Tensor,
                 dsh_out: torch.Tensor, P: int, M: int, D: int) -> None:
        """begin + finish in one go. head: the contiguous non-SH parameter gradients (summed in place); drgb[P,3]:
        this rank's clamped colour gradient; dsh_out[P,M,3]: receives the SH gradient summed over ranks."""
        self.begin(drgb, campos, P)
        self.finish(head, means3D, dsh_out, P, M, D)


class PeerExchange(FactoredExchange):
    """FactoredExchange whose all-gather is done by the producing kernel itself: the backward writes this rank's
    clamped colour gradient straight into every rank's gather buffer through NVLink peer pointers
    (gsr_backward_early_rgb_peers, csrc/sh_exchange.cu:clamped_rgb_push_kernel) on a side stream, underneath the
    preprocess-backward kernel; a device-side barrier over symmetric-memory signal pads replaces the collective's
    synchronisation. The gather buffers live in torch symmetric memory (torch.distributed._symmetric_memory: peer
    pointers + barrier) and are double-buffered by step parity, so a fast rank's next push cannot overwrite what a
    slow rank's rebuild kernel is still reading (the barrier bounds the skew to one step). The 11 / 10 non-SH floats
    are reduced IN THE SWITCH when the symmetric buffer has a multicast mapping (NVLS), in two phases whose per-GPU
    traffic does not grow with the number of ranks (csrc/common.cuh): the preprocess-backward kernel stores its rows
    into the local replica of a symmetric head (gsr_backward_peer_head); after a barrier every rank sums ITS 1/world
    share of the rows over all replicas with multimem.ld_reduce and broadcasts the sums with multimem.st
    (gsr_peer_reduce_shard); after a second barrier the rebuild kernel also hands the summed rows over
    (gsr_peer_finish) -- a step then contains no NCCL call at all. Without multicast support they are all-reduced by
    NCCL underneath the rebuild kernel.

    `profile = True` records CUDA events around the phases of finish(); `stage_ms()` returns their mean durations.

    Construction is collective. If symmetric memory cannot be set up (no P2P access, a backend without it), `ok` is
    False and the object behaves exactly like FactoredExchange (NCCL all-gather): both are GPU paths of this library."""

    def __init__(self, group=None, rebuild=None):
        super().__init__(group, rebuild)
        self._symm = {}
        self._parity = 0
        self._side = None
        self.ok = True
        self.why = ""
        self.profile = False
        self._events = []

    PHASES = ("wait_push_and_barrier", "head_reduce_shard", "barrier", "sh_rebuild_and_head_consume")

    def stage_ms(self):
        """Mean CUDA-event duration of each phase of finish() over the steps recorded while `profile` was set (this
        rank; the barrier phases contain the wait for the slowest rank)."""
        if not self._events:
            return None
        torch.cuda.synchronize()
        n = len(self._events)
        out = {name: sum(ev[i].elapsed_time(ev[i + 1]) for ev in self._events) / n for i, name in enumerate(self.PHASES)}
        out["steps"] = n
        self._events = []
        return out

    def _setup(self, world, P, dev):
        key = (world, P, str(dev))
        if key in self._symm:
            return self._symm[key]
        import torch.distributed._symmetric_memory as symm
        grp = self.group if self.group is not None else dist.group.WORLD
        r4 = lambda n: (n + 3) // 4 * 4
        n_rgb, n_cam, n_head = r4(world * P * 3), r4(world * 3), P * 12
        half = n_rgb + n_cam + n_head              # one parity: gathered drgb | gathered campos | switch-reduced head
        t = symm.empty(2 * half, dtype=torch.float32, device=dev)
        hdl = symm.rendezvous(t, grp)
        rank = dist.get_rank(grp)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if os.environ.get("GSR_PEER_NO_MULTICAST"):
            mc = 0
        views, dst, heads = [], [], []
        for b in range(2):
            base = b * half
            views.append((t[base: base + world * P * 3].view(world, P, 3),
                          t[base + n_rgb: base + n_rgb + world * 3].view(world, 3)))
            drgb_dst = (C.c_void_p * world)(*[p + 4 * (base + rank * P * 3) for p in ptrs])
            cam_dst = (C.c_void_p * world)(*[p + 4 * (base + n_rgb + rank * 3) for p in ptrs])
            dst.append((drgb_dst, cam_dst))
            h0 = base + n_rgb + n_cam                   # multiple of 4 floats: rows are 16-byte aligned
            heads.append((t[h0: h0 + P * 12], (mc + 4 * h0) if mc else 0))
        t.zero_()
        torch.cuda.synchronize(dev)
        dist.barrier(group=grp)
        self._symm = {key: (t, hdl, views, dst, heads)}
        self._side = torch.cuda.Stream(dev)
        self.multicast = bool(mc)
        return self._symm[key]

    def arm(self, lib, campos, P, dev):
        """Arm the next backward of this thread (called by the `_C` level in place of gsr_backward_early_rgb).
        Returns (objects to keep alive until the backward returned, list receiving an exception of the callback)."""
        from . import _lib
        world = self._world()
        errors = []
        if self.ok and world > 1:
            try:
                t, hdl, views, dst, heads = self._setup(world, P, dev)
            except Exception as ex:        # collective failure: every rank takes the same branch
                self.ok, self.why = False, repr(ex)
        if not (self.ok and world > 1):
            return None
        drgb_dst, cam_dst = dst[self._parity]
        campos = campos.reshape(3).contiguous()
        side = self._side
        use_mc = heads[self._parity][1] != 0

        def _ready(_user):
            if use_mc:
                return                             # one barrier on the main stream in finish() covers push and head
            try:
                with torch.cuda.stream(side):      # after the push kernel: all ranks' stores have landed once this passes
                    hdl.barrier(channel=0)
            except Exception as ex:                # never let an exception cross the C ABI
                errors.append(ex)

        cb = _lib.READY_FN(_ready)
        _lib.check(lib.gsr_backward_early_rgb_peers(world, drgb_dst, cam_dst, campos.data_ptr(),
                                                    C.c_void_p(side.cuda_stream), cb, None), "gsr_backward_early_rgb_peers")
        if use_mc:      # the four non-SH gradients go to the local replica of the symmetric head (phase 0 of the switch reduce)
            _lib.check(lib.gsr_backward_peer_head(C.c_void_p(heads[self._parity][0].data_ptr())), "gsr_backward_peer_head")
        return (cb, campos, drgb_dst, cam_dst), errors

    def finish(self, head, means3D, dsh_out, P, M, D):
        world = self._world()
        key = (world, P, str(dsh_out.device))
        if not (self.ok and world > 1 and key in self._symm):
            return super().finish(head, means3D, dsh_out, P, M, D)
        from . import _lib
        t, hdl, views, dst, heads = self._symm[key]
        drgb_all, campos_all = views[self._parity]
        head_sym, mc = heads[self._parity]
        self._parity ^= 1
        dev = dsh_out.device
        main = torch.cuda.current_stream(dev)
        if mc:
            lib = _lib.lib()
            st = C.c_void_p(main.cuda_stream)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)] if self.profile else None
            mark = (lambda i: ev[i].record(main)) if ev else (lambda i: None)
            mark(0)
            # push kernel (side stream) and preprocess-backward (this stream) of EVERY rank are complete after this
            main.wait_stream(self._side)
            hdl.barrier(channel=0)
            mark(1)
            with torch.cuda.device(dev):
                # phase 1: this rank's share of the head rows, summed over the replicas by the switch and broadcast
                _lib.check(lib.gsr_peer_reduce_shard(C.c_void_p(mc), P, dist.get_rank(self.group), world, st),
                           "gsr_peer_reduce_shard")
                mark(2)
                hdl.barrier(channel=1)          # every rank's share has landed in every replica
                mark(3)
                # phase 2 + SH rebuild in one launch (one thread per Gaussian does both)
                _lib.check(lib.gsr_peer_finish(P, M, D, world, means3D.data_ptr(), campos_all.data_ptr(), drgb_all.data_ptr(),
                                               dsh_out.data_ptr(), head_sym.data_ptr(), head.data_ptr(),
                                               int(head.numel() == 10 * P), st), "gsr_peer_finish")
            mark(4)
            if ev:
                self._events.append(ev)
            return
        w_r = dist.all_reduce(head, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        main.wait_stream(self._side)
        (self.rebuild or _rebuild_cuda)(P, M, D, world, means3D, campos_all, drgb_all, dsh_out)
        w_r.wait()


def _rebuild_cuda(P, M, D, n_views, means3D, campos_all, drgb_all, dsh_out):
    from . import _lib
    if not dsh_out.is_cuda:
        raise _lib.GsrError("gsr_sh_grad_from_views needs CUDA tensors: there is no CPU path")
    lib = _lib.lib()
    with torch.cuda.device(dsh_out.device):
        rc = lib.gsr_sh_grad_from_views(P, M, D, n_views, means3D.data_ptr(), campos_all.data_ptr(), drgb_all.data_ptr(),
                                        dsh_out.data_ptr(), torch.cuda.current_stream(dsh_out.device).cuda_stream)
    _lib.check(rc, "gsr_sh_grad_from_views")


def reduce_without_factoring(fx, head: torch.Tensor) -> torch.Tensor:
    """A backward the installed factored exchange cannot serve (colours precomputed: M == 0, so there is no SH outer
    product to factor) still has to return gradients summed over ranks: one all_reduce over the contiguous
    parameter-gradient head, in the exchange's group."""
    return allreduce_param_grads(head, group=getattr(fx, "group", None))


def install_factored_exchange(group=None, enable: bool = True, rebuild=None, peer: bool = False) -> None:
    """One view per rank per step: every backward of both rasterizer packages (autograd or `_C` level) returns
    gradients already summed over ranks, exchanged with FactoredExchange (peer=True: PeerExchange, the all-gather
    done by the backward's own stores into peer memory). Replaces install_backward_allreduce."""
    from .diff_gaussian_rasterization import _C as c3
    from .diff_surfel_rasterization import _C as c2
    fx = (PeerExchange(group, rebuild) if peer else FactoredExchange(group, rebuild)) if enable else None
    c3._factored[0] = fx
    c2._factored[0] = fx
    if enable:
        c3._backward_hook[0] = None
        c2._backward_hook[0] = None


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
