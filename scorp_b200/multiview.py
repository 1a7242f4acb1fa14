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
    (gsr_peer_reduce_shard); after a second barrier gsr_peer_consume hands the summed rows over. That sequence is
    pipelined with the preprocess-backward kernel row range by row range (the library runs it as CHUNKS launches and
    calls back after each), and the SH rebuild runs on the side stream right behind the push and a barrier of its own,
    so both travel underneath the preprocess-backward kernel -- a step contains no NCCL call at all. Without multicast support they are all-reduced by
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

    # row ranges the preprocess-backward kernel is split into for the pipelined head reduce (GSR_PEER_CHUNKS overrides).
    # Measured at C2 on 2 GPUs (profiles/r02_mg2_chunks.md): 1 range 1.750 ms/step, 2 ranges 1.743, 4 ranges 1.792 -- what
    # the pipelining takes off the tail mostly comes back as a slower preprocess-backward stage; at 8 GPUs 1 range 1.925,
    # 2 ranges 1.853 (profiles/r02_mg2_chunks.md, last lines). Default: 2.
    CHUNKS = int(os.environ.get("GSR_PEER_CHUNKS", "2"))

    def stage_ms(self):
        """Mean CUDA-event durations over the steps recorded while `profile` was set (this rank): `exposed_tail` is what
        the step pays after its last preprocess-backward launch (join of the reduce and side streams); per row range,
        the time from the end of its preprocess-backward launch to its summed rows sitting in the arena (two barriers,
        which include waiting for the slowest rank, the switch reduce and the hand-over); and the side stream's barrier
        and SH rebuild, which run underneath the preprocess-backward kernel."""
        if not self._events:
            return None
        torch.cuda.synchronize()
        n = len(self._events)
        out = {"exposed_tail": sum(ev[0].elapsed_time(ev[1]) for ev, _, _ in self._events) / n}
        side = [m for _, m, _ in self._events if len(m) == 3]
        if side:
            out["side:barrier_after_push"] = sum(m[0].elapsed_time(m[1]) for m in side) / len(side)
            out["side:sh_rebuild"] = sum(m[1].elapsed_time(m[2]) for m in side) / len(side)
        chunks = [c for _, _, c in self._events if c]
        if chunks:
            k = min(len(c) for c in chunks)
            out["range_reduce_ms"] = [sum(c[i][0].elapsed_time(c[i][1]) for c in chunks) / len(chunks) for i in range(k)]
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
        views, dst, heads, mc_slots = [], [], [], []
        for b in range(2):
            base = b * half
            views.append((t[base: base + world * P * 3].view(world, P, 3),
                          t[base + n_rgb: base + n_rgb + world * 3].view(world, 3)))
            drgb_dst = (C.c_void_p * world)(*[p + 4 * (base + rank * P * 3) for p in ptrs])
            cam_dst = (C.c_void_p * world)(*[p + 4 * (base + n_rgb + rank * 3) for p in ptrs])
            dst.append((drgb_dst, cam_dst))
            h0 = base + n_rgb + n_cam                   # multiple of 4 floats: rows are 16-byte aligned
            heads.append((t[h0: h0 + P * 12], (mc + 4 * h0) if mc else 0))
            mc_slots.append(((mc + 4 * (base + rank * P * 3)), (mc + 4 * (base + n_rgb + rank * 3))) if mc else None)
        t.zero_()
        torch.cuda.synchronize(dev)
        dist.barrier(group=grp)
        self._mc_slots = mc_slots
        self._symm = {key: (t, hdl, views, dst, heads)}
        self._side = torch.cuda.Stream(dev)
        # high priority: the reduce kernels and barriers of range k get SM slots as soon as blocks of the
        # preprocess-backward launch of range k+1 (main stream, normal priority, fills every SM) retire
        self._red = torch.cuda.Stream(dev, priority=-1)
        self.multicast = bool(mc)
        return self._symm[key]

    def arm(self, lib, campos, P, dev, rebuild_args=None):
        """Arm the next backward of this thread (called by the `_C` level in place of gsr_backward_early_rgb).
        rebuild_args = (means3D, dsh_out, M, D, head_dst): with a multicast mapping the SH rebuild is enqueued on the side
        stream right behind the push and its barrier, i.e. it runs underneath the preprocess-backward kernel of the main
        stream, and the switch reduce of the head rows is pipelined with that kernel row range by row range (CHUNKS
        launches; range k is reduced on the reduce stream while range k+1 is computed).
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
        use_mc = heads[self._parity][1] != 0 and rebuild_args is not None
        drgb_all, campos_all = views[self._parity]
        profile, marks = self.profile, []
        self._marks = marks
        rebuild = self.rebuild or _rebuild_cuda

        def _ready(_user):
            if use_mc:
                # the push kernel is enqueued on the side stream; behind it, on the same stream: a barrier (all ranks'
                # colour gradients have landed) and the SH rebuild, which then overlaps the preprocess-backward kernel the
                # library enqueues on the main stream right after this callback returns
                try:
                    means3D, dsh_out, M, D, _ = rebuild_args
                    with torch.cuda.stream(side):
                        if profile:
                            marks.extend(torch.cuda.Event(enable_timing=True) for _ in range(3))
                            marks[0].record(side)
                        hdl.barrier(channel=2)
                        if profile:
                            marks[1].record(side)
                        rebuild(P, M, D, world, means3D, campos_all, drgb_all, dsh_out)
                        if profile:
                            marks[2].record(side)
                except Exception as ex:
                    errors.append(ex)
                return
            try:
                with torch.cuda.stream(side):      # after the push kernel: all ranks' stores have landed once this passes
                    hdl.barrier(channel=0)
            except Exception as ex:                # never let an exception cross the C ABI
                errors.append(ex)

        cb = _lib.READY_FN(_ready)
        if use_mc and not os.environ.get("GSR_PEER_PUSH_P2P"):
            # the switch replicates the colour gradient into every rank's gather buffer (one multimem.st per 16 bytes)
            mc_drgb, mc_cam = self._mc_slots[self._parity]
            _lib.check(lib.gsr_backward_early_rgb_multicast(C.c_void_p(mc_drgb), C.c_void_p(mc_cam), campos.data_ptr(),
                                                            C.c_void_p(side.cuda_stream), cb, None), "gsr_backward_early_rgb_multicast")
        else:
            _lib.check(lib.gsr_backward_early_rgb_peers(world, drgb_dst, cam_dst, campos.data_ptr(),
                                                        C.c_void_p(side.cuda_stream), cb, None), "gsr_backward_early_rgb_peers")
        self._armed_mc = use_mc
        keep = [cb, campos, drgb_dst, cam_dst]
        if use_mc:      # the four non-SH gradients go to the local replica of the symmetric head (phase 0 of the switch reduce)
            head_sym, mc = heads[self._parity]
            head_dst = rebuild_args[4]
            surfel = int(head_dst.numel() == 10 * P)
            red, rank = self._red, dist.get_rank(self.group)
            red_st = C.c_void_p(red.cuda_stream)
            main = torch.cuda.current_stream(dev)
            cmarks = []
            self._cmarks = cmarks

            def _chunk(_user, k, row0, rows):
                # launch k of the preprocess-backward kernel (rows [row0, row0 + rows)) is enqueued on the main stream:
                # behind it, on the reduce stream, barrier -> this rank's share summed in the switch and broadcast ->
                # barrier -> the summed rows handed to the arena, while the main stream computes the next range
                try:
                    e = torch.cuda.Event(enable_timing=profile)
                    e.record(main)
                    red.wait_event(e)
                    with torch.cuda.stream(red):
                        hdl.barrier(channel=0)              # every rank has stored these rows
                        _lib.check(lib.gsr_peer_reduce_shard(C.c_void_p(mc), row0, rows, rank, world, red_st), "gsr_peer_reduce_shard")
                        hdl.barrier(channel=1)              # every rank's share has landed in every replica
                        _lib.check(lib.gsr_peer_consume(head_sym.data_ptr(), head_dst.data_ptr(), P, row0, rows, surfel, red_st),
                                   "gsr_peer_consume")
                        if profile:
                            e2 = torch.cuda.Event(enable_timing=True); e2.record(red)
                            cmarks.append((e, e2))
                except Exception as ex:                     # never let an exception cross the C ABI
                    errors.append(ex)

            ccb = _lib.CHUNK_FN(_chunk)
            keep.append(ccb)
            _lib.check(lib.gsr_backward_peer_head(C.c_void_p(head_sym.data_ptr()), self.CHUNKS, ccb, None), "gsr_backward_peer_head")
        return tuple(keep), errors

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
        if mc and getattr(self, "_armed_mc", False):
            # everything was enqueued from inside the backward call (arm()'s callbacks): the head reduce on the reduce
            # stream, row range by row range, and the SH rebuild on the side stream; the main stream only joins them
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)] if self.profile else None
            if ev:
                ev[0].record(main)
            main.wait_stream(self._red)
            main.wait_stream(self._side)
            if ev:
                ev[1].record(main)
                self._events.append((ev, list(getattr(self, "_marks", [])), list(getattr(self, "_cmarks", []))))
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
