#!/usr/bin/env python
"""bench.py -- the rasterizer hot path of BASELINE.json on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2]

A step = one forward + backward of the 3DGS rasterizer over one view of the synthetic scene
(N > 1: one view per rank, Gaussians replicated, per-Gaussian parameter gradients summed with one
NCCL allreduce -- weak scaling). Rank 0 prints ONE JSON line:
  value / ms_per_step : inputs resident in HBM, called at the reference-shaped `_C` level (C ABI)
  e2e                 : through the public API (GaussianRasterizer + autograd) with the step's host
                        inputs (camera, target-image-shaped upstream gradients) copied from pinned
                        memory and the step's scalar result read back, inside the timed region
  roofline            : dominant stage, algorithmic bytes / live CUDA-event stage time vs measured HBM peak
  cpu_baseline        : the plain-C CPU oracle port (oracle/gs_oracle.c), all host cores, one iteration
`--impl reference` times the UNMODIFIED reference CUDA build (oracle/_ref/libref_dgr.so, compiled from
/root/reference by oracle/Makefile): the reference ships no CPU rasterizer, so its own implementation
of this path is CUDA and runs on the same GPU; the CPU number beside it is the oracle port.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist

from scorp_b200 import synth

METRIC = "3dgs_rasterizer_fwd_bwd_throughput"
UNIT = "Mpix/s"
PRIME = 8       # untimed set-up iterations of each arm before its warm-up (allocator growth, lazy module init)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", default="C2", choices=list(synth.CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", choices=["auto", "dense", "factored", "peer"], default="auto",
                    help="N>1, ours: dense = one all_reduce over the gradient arena; factored = multiview.FactoredExchange "
                         "(NCCL all-gather of dL_dRGB); peer = multiview.PeerExchange (the backward stores dL_dRGB into peer memory itself)")
    ap.add_argument("--views-per-rank", type=int, default=0, help="views each rank renders per step (0: the config's default)")
    ap.add_argument("--view-loop", action="store_true",
                    help="ours, forward-only configs: one forward call per view (the reference's loop) instead of the batched gsr_rasterize_forward_views")
    return ap.parse_args()


def setup_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep NCCL's version banner off stdout: ONE JSON line
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo",
                                device_id=torch.device("cuda", local) if torch.cuda.is_available() else None)
    return rank, world, local


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi sampling DURING the timed region (the profiling recipe's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True); self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def barrier(world):
    if world > 1:
        dist.barrier()


def prime(step, world=1, min_seconds=1.0, min_steps=PRIME):
    """Untimed set-up before the W warm-ups: caching-allocator growth, lazy module loading, first-touch of the
    scratch buffers. Runs `min_steps` steps, then as many more as fill `min_seconds`; the extra count is agreed
    across ranks (MAX) because a step may contain a collective."""
    t0 = time.perf_counter()
    for _ in range(min_steps):
        step()
    torch.cuda.synchronize()
    per = max((time.perf_counter() - t0) / min_steps, 1e-5)
    extra = torch.tensor([min(int(max(0.0, min_seconds - (time.perf_counter() - t0)) / per), 2000)], device="cuda")
    if world > 1:
        dist.all_reduce(extra, op=dist.ReduceOp.MAX)
    for _ in range(int(extra.item())):
        step()
    torch.cuda.synchronize()


def time_steps(step, K, W, world, sampler=None):
    """W warm-ups, then exactly K steps between barrier+synchronize, CUDA events, max over ranks (ms).
    Also returns the per-step durations (an event after every step) so host-side jitter is visible."""
    if sampler:
        sampler.start()        # nvidia-smi start-up (NVML init) lands in the warm-up, sampling covers the timed region
    for _ in range(W):
        step()
    torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ev[0].record()
    for k in range(K):
        step()
        ev[k + 1].record()
    torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
    ms = torch.tensor([ev[0].elapsed_time(ev[K])], device="cuda", dtype=torch.float64)
    spread = {}
    if world > 1:
        # every rank's own time for the K steps: with one DIFFERENT camera per rank the slowest view sets the step
        allms = [torch.zeros_like(ms) for _ in range(world)]
        dist.all_gather(allms, ms)
        spread["rank_ms_per_step"] = [float(t) / K for t in allms]
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    per = sorted(ev[k].elapsed_time(ev[k + 1]) for k in range(K))
    clocks = sampler.stop() if sampler else None
    spread.update({"median": per[K // 2], "min": per[0], "max": per[-1]})
    return float(ms.item()), clocks, spread


def stage_bytes(P, M, Pvis, N, W, H, surfel=False):
    """Algorithmic bytes per launch group (DESIGN.md 'Kernels')."""
    if surfel:
        return {
            "preprocess_fwd": P * (40 + 12 * M) + Pvis * 87,
            "binning": 8 * P + 44 * N,
            "blend_fwd": 64 * N + 60 * W * H,
            "blend_bwd": 2 * 76 * Pvis + 68 * N + (60 + 40) * W * H,
            "preprocess_bwd": Pvis * ((76 + 36 + 40 + 12 * M) + (12 + 8 + 16 + 36 + 12 + 12 * M)),
        }
    return {
        "preprocess_fwd": P * (44 + 12 * M) + Pvis * 75,
        "binning": 8 * P + 44 * N,
        "blend_fwd": 28 * N + 24 * W * H,
        "blend_bwd": 2 * 48 * Pvis + 44 * N + 28 * W * H,
        "preprocess_bwd": Pvis * ((99 + 12 * M) + (40 + 12 * M)),
    }


STAGES = ["preprocess_fwd", "binning", "blend_fwd", "blend_bwd", "preprocess_bwd"]
STAGE_KERNEL = {"blend_bwd": "render_backward", "blend_fwd": "render_forward", "preprocess_bwd": "preprocess_backward",
                "preprocess_fwd": "preprocess", "binning": "scatter+tile_sort"}


def ncu_traffic(cfg_name, stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's dominant kernel, from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json, written by tools/ncu_summary.py); None when not captured."""
    p = ROOT / "profiles" / "ncu_traffic.json"
    try:
        return json.loads(p.read_text()).get(cfg_name, {}).get(stage)
    except Exception:
        return None


def sha_of(tensors):
    """sha256 (first 16 hex digits) over the bytes of device tensors, in order, as they sit in HBM."""
    import hashlib
    h = hashlib.sha256()
    for t in tensors:
        h.update(t.detach().contiguous().cpu().numpy().tobytes())
    return h.hexdigest()[:16]


def inputs_sha(sc, cam):
    return sha_of([sc.means3D, sc.scales, sc.rotations, sc.opacities, sc.shs, cam.viewmatrix, cam.projmatrix])


def ncu_issue(cfg_name, stage):
    """warp instructions / active lanes / issue-slot use of the stage's largest kernel from the committed `ncu --set full`
    capture (profiles/ncu_issue.json, written by tools/ncu_table.py); None when not captured."""
    p = ROOT / "profiles" / "ncu_issue.json"
    try:
        return json.loads(p.read_text()).get(cfg_name, {}).get(stage)
    except Exception:
        return None


def out_channels(surfel):
    return (3, 7) if surfel else (3, 1, 1)


def cpu_baseline(cfg_name, W, H):
    """One fwd+bwd iteration of the workload through the CPU oracle port, all host cores."""
    from oracle import gs_oracle as O
    surfel = synth.CONFIGS[cfg_name]["surfel"]
    sc = synth.make_scene(cfg_name); cam = sc.cameras[0]
    grads = synth.upstream_grads(cam.W, cam.H, 1, channels=out_channels(surfel))
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    t0 = time.perf_counter()
    fn_f, fn_b = (O.surfel_forward, O.surfel_backward) if surfel else (O.forward, O.backward)
    f = fn_f(sc.bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, cam.viewmatrix,
             cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree, cam.campos)
    fwd_only = synth.CONFIGS[cfg_name].get("forward_only", False)
    if not fwd_only:
        fn_b(f, *grads)
    dt = time.perf_counter() - t0
    src = "oracle/surfel_oracle.c" if surfel else "oracle/gs_oracle.c"
    return {"value": (W * H / 1e6) / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "ms_per_step": dt * 1e3,
            "sample": f"1 {'forward' if fwd_only else 'fwd+bwd'} iteration (one view) of the full {cfg_name} workload ({src}, OpenMP)"}


class Ours:
    """The product through its reference-shaped `_C` level (C ABI) and through the public autograd API."""
    name = "ours"

    def __init__(self, surfel):
        from scorp_b200 import _lib
        self.surfel = surfel
        if surfel:
            import scorp_b200.diff_surfel_rasterization as mod
            from scorp_b200.diff_surfel_rasterization import _C
        else:
            import scorp_b200.diff_gaussian_rasterization as mod
            from scorp_b200.diff_gaussian_rasterization import _C
        self.mod, self.C, self.lib = mod, _C, _lib.lib()
        self.E = torch.Tensor([])

    def fwd_bwd(self, sc, cam, bg, view, proj, campos, grads, fwd_only, full=False):
        C_, E = self.C, self.E
        out = C_.rasterize_gaussians(bg, sc.means3D, E, sc.opacities, sc.scales, sc.rotations, 1.0, E, view, proj,
                                     cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree, campos, False, False)
        if self.surfel:
            n, color, others, radii, geom, binning, img = out
            if fwd_only:
                return n, radii, color, None
            g = C_.rasterize_gaussians_backward(bg, sc.means3D, radii, E, sc.scales, sc.rotations, 1.0, E, view, proj,
                                                cam.tanfovx, cam.tanfovy, grads[0], grads[1], sc.shs, sc.sh_degree,
                                                campos, geom, n, binning, img, False)
        else:
            n, color, depth, alpha, radii, geom, binning, img = out
            if fwd_only:
                return n, radii, color, None
            g = C_.rasterize_gaussians_backward(bg, sc.means3D, radii, E, sc.scales, sc.rotations, 1.0, E, view, proj,
                                                cam.tanfovx, cam.tanfovy, grads[0], grads[1], grads[2], sc.shs,
                                                sc.sh_degree, campos, geom, n, binning, img, alpha, False)
        if full:        # the five parameter gradients: means3D, sh, opacity, scales, rotations
            return n, radii, color, (g[3], g[5], g[2], g[6], g[7])
        return n, radii, color, g[3]          # g[3] = dL_dmeans3D, the head of the gradient arena

    def tile_stats(self, sc, cam):
        """mean / max tile-list length and the pair count before early exit, from the ranges of one forward"""
        out = self.C.rasterize_gaussians(sc.bg, sc.means3D, self.E, sc.opacities, sc.scales, sc.rotations, 1.0, self.E,
                                         cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs,
                                         sc.sh_degree, cam.campos, False, False)
        img = out[-1]
        tiles = ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
        off = (C.c_size_t * 4)(); tot = C.c_size_t(0)
        (self.lib.gsr_surfel_image_layout if self.surfel else self.lib.gsr_image_layout)(cam.W, cam.H, off, C.byref(tot))
        rg = img[off[0]: off[0] + 8 * tiles].view(torch.int32).view(tiles, 2).long()
        ln = rg[:, 1] - rg[:, 0]
        return {"tiles": tiles, "mean_list": float(ln.float().mean()), "max_list": int(ln.max()), "pairs": int(ln.sum()) * 256}

    def binning_sha(self, sc, cam):
        """sha256 over radii and the sorted (tile | depth) keys of one forward of `cam` (bit-exactness evidence: the
        reference arm prints the same hash from its own buffers)."""
        out = self.C.rasterize_gaussians(sc.bg, sc.means3D, self.E, sc.opacities, sc.scales, sc.rotations, 1.0, self.E,
                                         cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs,
                                         sc.sh_degree, cam.campos, False, False)
        n, radii, binning = out[0], out[-4], out[-2]
        off = (C.c_size_t * 3)(); tot = C.c_size_t(0)
        self.lib.gsr_binning_layout(n, off, C.byref(tot))
        keys = binning[off[1]: off[1] + 8 * n].view(torch.int64)
        return sha_of([radii, keys]), n

    def accumulate(self, total, g_means3D, P, M):
        from scorp_b200 import multiview as mv
        return mv.accumulate_param_grads(total, g_means3D, P, M, self.surfel)

    factored = False        # multiview.install_factored_exchange active: the backward returns reduced gradients

    def reduce(self, g, P, M):
        """g: dL_dmeans3D of one backward (a view of its arena) or an already accumulated head."""
        from scorp_b200 import multiview as mv
        if self.factored:
            return
        n = P * mv.param_grad_floats(M, self.surfel)
        head = g if (g.dim() == 1 and g.numel() == n) else mv.arena_of(g)[:n]
        mv.allreduce_param_grads(head)

    def e2e_step(self, sc, cam, params, d_cam, d_gr, fwd_only, world):
        H, W, hw = cam.H, cam.W, cam.H * cam.W
        rs = self.mod.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d_cam[35:38], 1.0,
                                                    d_cam[0:16].view(4, 4), d_cam[16:32].view(4, 4), sc.sh_degree,
                                                    d_cam[32:35], False, False)
        for p in params:
            p.grad = None
        rast = self.mod.GaussianRasterizer(rs)
        with torch.set_grad_enabled(not fwd_only):
            outs = rast(means3D=params[0], means2D=torch.zeros_like(params[0], requires_grad=not fwd_only),
                        shs=params[1], opacities=params[2], scales=params[3], rotations=params[4])
        if fwd_only:
            return outs[0]
        gr = d_gr()
        if self.surfel:
            torch.autograd.backward([outs[0], outs[2]], [gr[:3 * hw].view(3, H, W), gr[3 * hw:].view(7, H, W)])
        else:
            torch.autograd.backward([outs[0], outs[2], outs[3]],
                                    [gr[:3 * hw].view(3, H, W), gr[3 * hw:4 * hw].view(1, H, W), gr[4 * hw:].view(1, H, W)])
        return outs[0]      # world > 1: the exchange runs inside backward (multiview.install_backward_allreduce)


class Reference:
    """The UNMODIFIED reference CUDA (oracle/_ref) with what its torch glue does around it: fresh outputs, three
    scratch buffers and the zero-filled gradient tensors per call."""
    name = "reference"

    def __init__(self, surfel):
        from oracle import ref_cuda
        self.surfel, self.rc = surfel, ref_cuda
        self.ok = ref_cuda.available("dsr" if surfel else "dgr")

    def fwd_bwd(self, sc, cam, bg, view, proj, campos, grads, fwd_only):
        rc = self.rc
        if self.surfel:
            f = rc.dsr_forward(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, view, proj,
                               cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree, campos, sync=False)
            self._last_fwd = f
            if fwd_only:
                return f["num_rendered"], f["radii"], f["color"], None
            g = rc.dsr_backward(f, bg, sc.means3D, None, sc.scales, sc.rotations, 1.0, None, view, proj, cam.tanfovx,
                                cam.tanfovy, grads[0], grads[1], sc.shs, sc.sh_degree, campos, sync=False)
        else:
            f = rc.dgr_forward(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, view, proj,
                               cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree, campos, sync=False)
            self._last_fwd = f
            if fwd_only:
                return f["num_rendered"], f["radii"], f["color"], None
            g = rc.dgr_backward(f, bg, sc.means3D, None, sc.scales, sc.rotations, 1.0, None, view, proj, cam.tanfovx,
                                cam.tanfovy, grads[0], grads[1], grads[2], sc.shs, sc.sh_degree, campos, sync=False)
        return f["num_rendered"], f["radii"], f["color"], g

    KEYS = ("means3D", "sh", "opacity", "scales", "rotations")

    def binning_sha(self, sc, cam):
        n, radii, color, _ = self.fwd_bwd(sc, cam, sc.bg, cam.viewmatrix, cam.projmatrix, cam.campos, None, True)
        torch.cuda.synchronize()
        f = self._last_fwd
        keys = (self.rc.dsr_intermediates(f) if self.surfel else self.rc.dgr_intermediates(f))["point_list_keys"]
        return sha_of([radii, keys]), n

    def accumulate(self, total, g, P, M):
        if total is None:
            return {k: g[k].clone() for k in self.KEYS}
        for k in self.KEYS:
            total[k].add_(g[k])
        return total

    def reduce(self, g, P, M):
        for k in self.KEYS:      # five tensors, five collectives
            dist.all_reduce(g[k])

    def e2e_step(self, sc, cam, params, d_cam, d_gr, fwd_only, world):
        H, W, hw = cam.H, cam.W, cam.H * cam.W
        gr = None
        if not fwd_only:
            g_ = d_gr()
            gr = ([g_[:3 * hw].view(3, H, W), g_[3 * hw:].view(7, H, W)] if self.surfel else
                  [g_[:3 * hw].view(3, H, W), g_[3 * hw:4 * hw].view(1, H, W), g_[4 * hw:].view(1, H, W)])
        n, radii, color, g = self.fwd_bwd(sc, cam, d_cam[35:38].contiguous(), d_cam[0:16].contiguous(),
                                          d_cam[16:32].contiguous(), d_cam[32:35].contiguous(), gr, fwd_only)
        if world > 1 and g is not None:
            self.reduce(g, sc.P, sc.shs.shape[1])
        return color


def exchange_check(arm, sc, cam, grads, world, dev):
    """Correctness of the exchange the timed steps used (N > 1, product arm): one backward through the installed
    exchange against the SAME backward with the exchange switched off followed by one plain dist.all_reduce per
    gradient tensor. Tolerance as in the single-GPU parity tests (tests/test_gpu_dgr_parity.py:_grad_ok): rtol 1e-4,
    atol = 1e-6 * max|ref| + 2 x the run-to-run spread of six runs of the plain path (float atomics in the blend backward).
    Returns {"ok", "worst_ratio", ...}; the caller exits non-zero when it fails."""
    from scorp_b200 import multiview as mv
    from scorp_b200.diff_gaussian_rasterization import _C as c3
    from scorp_b200.diff_surfel_rasterization import _C as c2
    names = ("means3D", "sh", "opacity", "scales", "rotations")
    fx = c3._factored[0]

    def run(dense_head=False):
        g = arm.fwd_bwd(sc, cam, sc.bg, cam.viewmatrix, cam.projmatrix, cam.campos, grads, False, full=True)[3]
        if dense_head:      # the un-factored product path: ONE all_reduce over the head of the gradient arena
            arm.reduce(g[0], sc.P, sc.shs.shape[1])
        return [t.clone() for t in g]
    got = run(dense_head=fx is None)                    # through the exchange that ran in the timed region
    saved = (c3._factored[0], c2._factored[0], c3._backward_hook[0], c2._backward_hook[0])
    c3._factored[0] = c2._factored[0] = None
    c3._backward_hook[0] = c2._backward_hook[0] = None
    try:
        refs = []
        for _ in range(6):                              # six plain runs: their spread is the atomics noise (three were
            # too few: the range of 3 samples of the noisiest element underestimates what a 4th can do, one run in ~10
            # then reported 1.4x on dL_drotations with every variant of the exchange)
            loc = run()
            for t in loc:
                dist.all_reduce(t)
            refs.append(loc)
    finally:
        c3._factored[0], c2._factored[0], c3._backward_hook[0], c2._backward_hook[0] = saved
    worst, worst_name = 0.0, None
    for i, name in enumerate(names):
        st = torch.stack([r[i].double().reshape(-1) for r in refs])
        mean = st.mean(0)
        spread = float((st.max(0).values - st.min(0).values).max())
        tol = 1e-6 * float(mean.abs().max()) + 2.0 * spread + 1e-4 * mean.abs()
        ratio = float(((got[i].double().reshape(-1) - mean).abs() / tol.clamp_min(1e-300)).max())
        if ratio > worst:
            worst, worst_name = ratio, name
    t = torch.tensor([worst], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"ok": bool(float(t) <= 1.0), "worst_ratio": float(t), "worst_tensor_rank0": worst_name, "rtol": 1e-4,
            "atol": "1e-6*max|ref| + 2*spread of 6 plain runs", "ranks": world,
            "path": type(fx).__name__ if fx is not None else "dense all_reduce of the arena head",
            "peer_ok": bool(getattr(fx, "ok", False)), "multicast": bool(getattr(fx, "multicast", False)),
            "against": "same backward with the exchange off + one dist.all_reduce per gradient tensor"}


def time_exchange(arm, P, M, surfel, world, dev, reps=10):
    """The exchange step alone, on buffers of the real size (ms, max over ranks) and what it moves per GPU."""
    from scorp_b200 import multiview as mv
    nh = 10 if surfel else 11
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    if arm.name == "ours" and arm.factored:
        fx = mv.FactoredExchange()
        head = torch.zeros(P * nh, device=dev); drgb = torch.zeros(P, 3, device=dev); campos = torch.zeros(3, device=dev)
        means = torch.randn(P, 3, device=dev); dsh = torch.empty(P * M * 3, device=dev)
        fn = lambda: fx(head, drgb, campos, means, dsh, P, M, int(round(M ** 0.5)) - 1)
        moved = 4 * P * (nh + 3 * (world - 1))          # reduced floats out + gathered floats in
        kind = "all_gather(dL_dRGB) + all_reduce(%d floats) + SH rebuild kernel" % nh
    elif arm.name == "ours":
        head = torch.zeros(P * mv.param_grad_floats(M, surfel), device=dev)
        fn = lambda: mv.allreduce_param_grads(head)
        moved = 4 * head.numel(); kind = "one all_reduce"
    else:
        bufs = [torch.zeros(P * w, device=dev) for w in (3, 3 * M, 1, 2 if surfel else 3, 4)]
        fn = lambda: [dist.all_reduce(b) for b in bufs]
        moved = 4 * sum(b.numel() for b in bufs); kind = "five all_reduce calls"
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    ev[0].record()
    for _ in range(reps):
        fn()
    ev[1].record(); torch.cuda.synchronize()
    t = torch.tensor([ev[0].elapsed_time(ev[1]) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    return {"kind": kind, "ms": ms, "bytes_per_gpu": moved, "gbs": moved / ms / 1e6,
            "frac_of_nvlink_900": moved / ms / 1e6 / 900.0}


def run(arm, args, rank, world, local):
    cfg = synth.CONFIGS[args.config]
    surfel, fwd_only = cfg["surfel"], cfg.get("forward_only", False)
    dev = torch.device("cuda", local)
    sweep = cfg.get("sweep_views", 0)       # C5: the whole pose sweep (128 rotations) is the job, split over the ranks
    vpr = args.views_per_rank or (max(sweep // world, 1) if sweep else cfg.get("views_per_rank", 1))
    n_views = world * vpr
    cameras = None
    if sweep:
        # the reference's own pose set: rotation_matrices/rotations_128.npz (utils/graphic.py:150-360, consumed by
        # align_3dgs_clpe_9dof.py / align_2dgs_clpe_9dof.py:298-305), committed as a fixture under tests/golden/
        import numpy as np
        rots = np.load(ROOT / "tests" / "golden" / "rotations_128.npz")["rotations"]
        cameras = [synth.orbit_camera(cfg["W"], cfg["H"], rots[v % len(rots)]) for v in range(n_views)]
    sc = synth.make_scene(args.config, n_views=n_views, cameras=cameras).to(dev)
    from scorp_b200 import multiview as mv
    my_views = mv.views_for_rank(n_views, rank, world)
    cams = [sc.cameras[v] for v in my_views]
    W, H, P, M = cams[0].W, cams[0].H, sc.P, sc.shs.shape[1]
    grads = [[g.to(dev) for g in synth.upstream_grads(W, H, 1 + v, channels=out_channels(surfel))] for v in my_views]

    batched = arm.name == "ours" and fwd_only and not surfel and len(cams) > 1 and not args.view_loop
    if batched:
        vm = torch.stack([c.viewmatrix for c in cams]); pm = torch.stack([c.projmatrix for c in cams])
        cp = torch.stack([c.campos for c in cams])

    def step_resident():
        if batched:     # the whole pose sweep of this rank in one call
            n, color, depth, alpha, radii = arm.C.rasterize_gaussians_views(
                sc.bg, sc.means3D, arm.E, sc.opacities, sc.scales, sc.rotations, 1.0, arm.E, vm, pm, cams[0].tanfovx,
                cams[0].tanfovy, H, W, sc.shs, sc.sh_degree, cp, False, False)
            return n, radii, None
        last, total = None, None
        for cam, gr in zip(cams, grads):
            n, radii, color, g = arm.fwd_bwd(sc, cam, sc.bg, cam.viewmatrix, cam.projmatrix, cam.campos, gr, fwd_only)
            last = (n, radii, g)
            if world > 1 and not fwd_only and len(cams) > 1:
                total = arm.accumulate(total, g, P, M)      # several views per rank: local sum first
        # one exchange per step
        if world > 1 and not fwd_only:
            arm.reduce(total if total is not None else last[2], P, M)
        return last

    exchange = "none"
    if world > 1 and not fwd_only:
        exchange = "one all_reduce over the parameter-gradient arena"
        want = args.exchange in ("factored", "peer") or (args.exchange == "auto" and mv.FactoredExchange.worth_it(world, M))
        if arm.name == "ours" and vpr == 1 and M > 0 and want:
            # one view per rank: gather 3 floats per Gaussian per rank and rebuild the SH gradient (multiview.FactoredExchange)
            use_peer = args.exchange in ("peer", "auto")      # auto: peer memory, NCCL gather if symmetric memory is unavailable
            mv.install_factored_exchange(peer=use_peer)
            arm.factored = True
            exchange = "factored: all_gather of dL_dRGB (12 B x world per Gaussian) + all_reduce of the 11 non-SH floats"
            if use_peer:
                exchange = "peer: dL_dRGB stored into every rank's gather buffer by the backward itself (NVLink peer pointers, symmetric memory)"
        elif arm.name == "reference":
            exchange = "five all_reduce calls (one per gradient tensor)"
    in_sha = inputs_sha(sc, cams[0])
    bin_sha, n_first = arm.binning_sha(sc, cams[0])
    n, radii, _ = step_resident()
    rank_instances = None
    if world > 1 and not batched:
        t_n = torch.tensor([int(n)], device=dev, dtype=torch.int64)
        all_n = [torch.zeros_like(t_n) for _ in range(world)]
        dist.all_gather(all_n, t_n)
        rank_instances = [int(x) for x in all_n]          # (Gaussian, tile) instances of each rank's (last) view
    Pvis = int((radii > 0).sum())          # batched: radii is [V, P] and n the total, so the stage bytes cover all views
    prime(step_resident, world)
    sampler = ClockSampler(local) if rank == 0 else None
    stages = None
    if arm.name == "ours":
        arm.lib.gsr_profile_enable(1)
    ms_total, clocks, per = time_steps(step_resident, args.steps, args.warmup, world, sampler)
    if arm.name == "ours":
        st = (C.c_float * 5)()
        arm.lib.gsr_profile_read(st)
        arm.lib.gsr_profile_enable(0)
        stages = list(st)
    ms_step = ms_total / args.steps

    if arm.name == "ours" and world > 1 and not arm.factored:
        mv.install_backward_allreduce()
    # ---- end to end: public API (ours: GaussianRasterizer + autograd), per-view host inputs from pinned memory
    params = [t.clone().requires_grad_(not fwd_only) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    h_cams = [torch.cat([c.viewmatrix.flatten(), c.projmatrix.flatten(), c.campos, sc.bg]).cpu().pin_memory() for c in cams]
    h_grads = [torch.cat([g.flatten() for g in gr]).cpu().pin_memory() for gr in grads]
    h_out = torch.empty(1, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(dev)
    d_gr_buf = torch.empty_like(h_grads[0], device=dev)

    if batched:
        h_all = torch.stack(h_cams).pin_memory()

    from scorp_b200 import capacity

    def enqueue_e2e():
        """everything of one end-to-end step that is enqueued: H2D of the step's inputs, forward, backward, the scalar and
        its D2H copy; returns without waiting (so it can also be captured in a CUDA graph)"""
        main = torch.cuda.current_stream(dev)
        acc = None
        if batched:     # one H2D of the V cameras, one forward_views call, one scalar back
            d = h_all.to(dev, non_blocking=True)
            rs = arm.mod.GaussianRasterizationSettings(H, W, cams[0].tanfovx, cams[0].tanfovy, d[0, 35:38], 1.0,
                                                       d[0, 0:16].view(4, 4), d[0, 16:32].view(4, 4), sc.sh_degree,
                                                       d[0, 32:35], False, False)
            color, _, _, _ = arm.mod.GaussianRasterizer(rs).forward_views(
                d[:, 0:16].reshape(-1, 4, 4), d[:, 16:32].reshape(-1, 4, 4), d[:, 32:35], params[0], params[2],
                shs=params[1], scales=params[3], rotations=params[4])
            h_out.copy_(color.sum().view(1), non_blocking=True)
            return
        for cam, h_cam, h_gr in zip(cams, h_cams, h_grads):
            d_cam = h_cam.to(dev, non_blocking=True)
            if not fwd_only:
                # the image-shaped upstream gradients are only needed by the backward: their H2D runs on a copy
                # stream underneath the forward
                copy_stream.wait_stream(main)
                with torch.cuda.stream(copy_stream):
                    d_gr_buf.copy_(h_gr, non_blocking=True)

            def d_gr():
                main.wait_stream(copy_stream)
                return d_gr_buf
            color = arm.e2e_step(sc, cam, params, d_cam, d_gr, fwd_only, world)
            s = color.detach().sum().view(1)      # the scalar that goes to the host (loss.item()); no graph hangs off it
            acc = s if acc is None else acc + s
        h_out.copy_(acc, non_blocking=True)

    redone = [0]

    def step_e2e():
        enqueue_e2e()
        torch.cuda.current_stream().synchronize()     # the step's result is on the host (loss.item() in train_3dgs.py)
        # no-sync mode (scorp_b200.capacity): the instance-capacity check sits at the step's own synchronisation point;
        # a step that overflowed is repeated with the raised capacity, inside the timed region
        while capacity.enabled() and capacity.overflowed(synchronize=False):
            redone[0] += 1
            enqueue_e2e()
            torch.cuda.current_stream().synchronize()

    prime(step_e2e, world)                 # setup, not warm-up: caching-allocator growth for the autograd path
    ms_e2e_total, _, per_e2e = time_steps(step_e2e, args.steps, args.warmup, world)
    e2e_modes = {"stock": {"ms_per_step": ms_e2e_total / args.steps, "spread": per_e2e,
                           "what": "drop-in API as the reference uses it: one host synchronisation inside every forward (num_rendered)"}}
    if arm.name == "ours" and not batched:
        # (1) scorp_b200.capacity.enable(): the forward no longer waits for num_rendered; the check moves to the step's sync
        capacity.reset(); capacity.enable()
        try:
            prime(step_e2e, world)
            t_ns, _, per_ns = time_steps(step_e2e, args.steps, args.warmup, world)
            e2e_modes["nosync"] = {"ms_per_step": t_ns / args.steps, "spread": per_ns, "steps_repeated_for_capacity": redone[0],
                                   "what": "same calls after scorp_b200.capacity.enable(): no host round trip in the forward, capacity checked at the step's own synchronisation"}
            # (2) the whole step (H2D copies, forward, backward, scalar D2H) captured in ONE CUDA graph and replayed
            if world == 1 and len(cams) == 1:
                try:
                    for p_ in params:
                        p_.grad = None
                    graph = torch.cuda.CUDAGraph()
                    capacity.forget_captured()
                    with torch.cuda.graph(graph):
                        enqueue_e2e()

                    def step_graph():
                        graph.replay()
                        torch.cuda.current_stream().synchronize()
                        if capacity.overflowed(synchronize=False):
                            raise RuntimeError("captured step exceeded its instance capacity (re-capture needed)")
                    prime(step_graph, world)
                    ref_val = float(h_out)
                    t_g, _, per_g = time_steps(step_graph, args.steps, args.warmup, world)
                    step_e2e()
                    e2e_modes["graph"] = {"ms_per_step": t_g / args.steps, "spread": per_g,
                                          "result_equals_eager": bool(abs(float(h_out) - ref_val) <= 1e-5 * abs(ref_val) + 1e-12),
                                          "what": "the same step captured once with torch.cuda.graph (possible because nothing in it synchronises) and replayed"}
                    del graph
                except Exception as ex:
                    e2e_modes["graph"] = {"error": repr(ex)[:300]}
        finally:
            capacity.disable(); capacity.reset(); capacity.forget_captured()
    best = min((m for m in e2e_modes if "ms_per_step" in e2e_modes[m]), key=lambda m: e2e_modes[m]["ms_per_step"])
    ms_e2e_total, per_e2e = e2e_modes[best]["ms_per_step"] * args.steps, e2e_modes[best]["spread"]
    exch, xcheck = None, None
    local_ms = None
    if world > 1 and not fwd_only and arm.name == "ours":
        # what each rank's own view costs WITHOUT any exchange: the slowest of them bounds the view-sharded step from
        # below (one different camera per rank), the rest of ms_per_step is the exchange
        from scorp_b200.diff_gaussian_rasterization import _C as _c3x
        from scorp_b200.diff_surfel_rasterization import _C as _c2x
        saved = (_c3x._factored[0], _c2x._factored[0], _c3x._backward_hook[0], _c2x._backward_hook[0])
        _c3x._factored[0] = _c2x._factored[0] = None
        _c3x._backward_hook[0] = _c2x._backward_hook[0] = None
        was_factored, arm.factored = arm.factored, True         # Ours.reduce() returns at once
        try:
            for _ in range(3):
                step_resident()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                step_resident()
            e1.record(); torch.cuda.synchronize()
            t_loc = torch.tensor([e0.elapsed_time(e1) / 10], device=dev, dtype=torch.float64)
            allt = [torch.zeros_like(t_loc) for _ in range(world)]
            dist.all_gather(allt, t_loc)
            local_ms = [float(t) for t in allt]
        finally:
            arm.factored = was_factored
            _c3x._factored[0], _c2x._factored[0], _c3x._backward_hook[0], _c2x._backward_hook[0] = saved
    if world > 1 and not fwd_only:
        exch = time_exchange(arm, P, M, surfel, world, dev)
        if arm.name == "ours":
            xcheck = exchange_check(arm, sc, cams[0], grads[0], world, dev)
    tstats = arm.tile_stats(sc, cams[0]) if arm.name == "ours" else None
    launches = 0
    if arm.name == "ours":
        # preprocess, scan, scatter, sort, blend (+ blend bwd, preprocess bwd) (+ clamped colour gradient, SH rebuild)
        per_view = (5 if fwd_only else 7) + (2 if arm.factored else 0)
        if arm.factored and args.exchange in ("peer", "auto"):
            fx_ = arm.C._factored[0]
            if getattr(fx_, "multicast", False):        # preprocess-backward in row ranges, each followed by reduce + hand-over
                per_view += (fx_.CHUNKS - 1) + 2 * fx_.CHUNKS
        launches = per_view * (1 if batched else len(cams)) * args.steps
    if arm.name == "ours" and arm.factored and args.exchange in ("peer", "auto"):
        from scorp_b200.diff_gaussian_rasterization import _C as _c3
        fx = _c3._factored[0]
        if getattr(fx, "ok", False):
            exchange += "; non-SH floats reduced in the switch in two phases (local stores, then multimem.ld_reduce + multimem.st on a 1/N share per rank)" \
                if getattr(fx, "multicast", False) else "; non-SH floats all-reduced by NCCL (no multicast mapping)"
            if exch:
                exch["note"] = "stand-alone timing of the NCCL form of this exchange, for reference; the timed steps used the peer-memory exchange (peer_phases_ms)"
                # CUDA-event times of the path that actually ran: 10 more steps with events around the phases of
                # PeerExchange.finish (rank 0; the barrier phases include waiting for the slowest rank)
                fx.profile = True
                for _ in range(10):
                    step_resident()
                fx.profile = False
                exch["peer_phases_ms"] = fx.stage_ms()
        else:
            exchange += f" [symmetric memory unavailable, ran as NCCL factored: {getattr(fx, 'why', '')}]"
    return dict(W=W, H=H, P=P, M=M, Pvis=Pvis, N=n, ms_step=ms_step, ms_e2e=ms_e2e_total / args.steps, stages=stages,
                clocks=clocks, per=per, per_e2e=per_e2e, views=n_views, vpr=vpr, surfel=surfel, fwd_only=fwd_only,
                h2d=sum(h.numel() for h in h_cams) * 4 + (0 if fwd_only else sum(h.numel() for h in h_grads) * 4),
                d2h=4, launches=launches, exchange=exchange, exch=exch, tstats=tstats, batched=batched,
                inputs_sha=in_sha, binning_sha=bin_sha, n_first=n_first, xcheck=xcheck, e2e_modes=e2e_modes, e2e_mode=best,
                rank_instances=rank_instances, local_ms=local_ms)


def main():
    args = parse()
    rank, world, local = setup_dist(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    surfel = synth.CONFIGS[args.config]["surfel"]
    arm = Ours(surfel) if args.impl == "ours" else Reference(surfel)
    if args.impl == "reference" and not arm.ok:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
        if world > 1:
            dist.destroy_process_group()
        return
    r = run(arm, args, rank, world, local)
    failed = bool(r.get("xcheck")) and not r["xcheck"]["ok"]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        if failed:
            raise SystemExit(3)
        return
    W, H = r["W"], r["H"]
    mpix = W * H * r["views"] / 1e6
    value = mpix / (r["ms_step"] * 1e-3)
    e2e_v = mpix / (r["ms_e2e"] * 1e-3)
    peak, peak_src = peaks()
    mode = "forward only" if r["fwd_only"] else "fwd+bwd"
    line = {
        "metric": METRIC if not r["fwd_only"] else "3dgs_rasterizer_fwd_throughput", "value": value, "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_step"], "ms_per_view": r["ms_step"] / r["vpr"], "ms_per_step_spread": r["per"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (seeded scene, scorp_b200/synth.py; no dataset ships with the reference)",
        "config": {"workload": synth.CONFIGS[args.config]["desc"], "id": args.config, "mode": mode, "P": r["P"],
                   "sh_coeffs": r["M"], "width": W, "height": H, "visible": r["Pvis"], "num_rendered": r["N"],
                   "views_per_step": r["views"], "views_per_rank": r["vpr"],
                   # self-verification: both arms hash the device-resident inputs and the binning result (radii + sorted
                   # tile|depth keys) of rank 0's first camera; equal hashes = identical inputs and bit-exact binning
                   "inputs_sha": r["inputs_sha"], "binning_sha": r["binning_sha"], "num_rendered_first_view": r["n_first"],
                   "host_threads": torch.get_num_threads(),
                   "parallelism": f"view-sharded dp{world}; exchange: {r['exchange']}" if world > 1 else "single GPU",
                   "l2": "per-step inputs (236 B x P parameters + 24 B x N lists) exceed the 126 MB L2; no explicit flush"
                         if r["P"] * (44 + 12 * r["M"]) > 126e6 else "working set below the 126 MB L2: numbers are L2-warm, as in the reference's own loop"},
        "e2e": {"value": e2e_v, "unit": UNIT, "ms_per_step": r["ms_e2e"], "ms_per_step_spread": r["per_e2e"], "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": r["d2h"], "mode": r["e2e_mode"], "modes": r["e2e_modes"],
                "note": "public API + autograd; per view the camera and the upstream-gradient maps are copied from pinned host memory (the gradient maps on a copy stream under the forward), scalar result read back; Gaussian parameters are model state resident in HBM as in train_3dgs.py"},
        "gpu_launches": r["launches"], "clocks": r["clocks"],
        # the contract's value/ms_per_step are MEANS over the K steps; the medians say what a typical step costs, and
        # `outlier` flags a run whose slowest step is > 3x its median (one allocator or host hiccup moves a mean)
        "value_median": mpix / (r["per"]["median"] * 1e-3), "e2e_value_median": mpix / (r["per_e2e"]["median"] * 1e-3),
        "outlier": bool(r["per"]["max"] > 3 * r["per"]["median"] or r["per_e2e"]["max"] > 3 * r["per_e2e"]["median"]),
    }
    if r["exch"]:
        line["exchange"] = r["exch"]
    if r.get("xcheck"):
        line["exchange_check"] = r["xcheck"]
    if r.get("rank_instances"):
        line["config"]["num_rendered_per_rank"] = r["rank_instances"]
    if r.get("local_ms"):
        line["view_sharding"] = {"rank_ms_without_exchange": r["local_ms"], "slowest_view_ms": max(r["local_ms"]),
                                 "exchange_overhead_ms": r["ms_step"] - max(r["local_ms"]),
                                 "note": "every rank renders a different camera: the slowest view bounds the step from below; the overhead is what the gradient exchange adds to it"}
    if r["tstats"]:
        line["config"]["tile_lists"] = r["tstats"]
    if args.impl == "ours":
        nv = r["vpr"] if r["batched"] else 1          # a batched call runs every stage once over all of the rank's views
        sb = stage_bytes(r["P"] * nv, r["M"], r["Pvis"], r["N"], W, H * nv, r["surfel"])
        line["config"]["views_per_call"] = nv
        stages = []
        for name, ms in zip(STAGES, r["stages"]):
            ms_v = ms if ms and ms > 0 else None
            gbs = sb[name] / (ms_v * 1e-3) / 1e9 if ms_v else None
            stages.append({"stage": name, "ms": ms_v, "algorithmic_bytes": sb[name], "achieved_gbs": gbs,
                           "frac": gbs / peak if gbs else None})
        dom = max(stages, key=lambda s: s["ms"] or 0)
        line["roofline"] = {"bound": "hbm", "kernel": dom["stage"], "achieved": dom["achieved_gbs"], "peak": peak,
                            "peak_source": peak_src, "unit": "GB/s", "frac": dom["frac"],
                            "traffic": ncu_traffic(args.config, dom["stage"]),
                            "note": "per view, last view of the step; blend stages are SM-issue bound (DESIGN.md 5), the HBM fraction is reported for the contract",
                            "stages": stages}
        # the roof that actually binds the blends: instruction issue, one warp instruction per clock per SM sub-partition
        # (4 x 148 per clock; measured with tools/ubench/f32x2.cu). Instruction count and lane occupancy come
        # from the committed ncu capture of the same kernel and config, the time is this run's CUDA-event stage time.
        iss = ncu_issue(args.config, dom["stage"])
        if iss and dom["ms"]:
            f_clk = ((r["clocks"] or {}).get("sm_mhz") or 1965.0) * 1e6
            peak_g = 4 * 148 * f_clk / 1e9
            ach_g = iss["warp_inst"] / (dom["ms"] * 1e-3) / 1e9
            line["roofline"]["issue"] = {"kernel": iss["kernel"], "warp_inst": iss["warp_inst"], "achieved_ginst_s": ach_g,
                                         "peak_ginst_s": peak_g, "frac": ach_g / peak_g, "lanes_active": iss["lanes_active"],
                                         "ncu_issue_active_pct": iss["issue_active_pct"], "sm_mhz": f_clk / 1e6,
                                         "source": "profiles/ncu_issue.json (ncu --set full) + live stage time"}
        if not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(args.config, W, H)
            except Exception as ex:  # the checker failing must not hide the GPU number
                line["cpu_baseline"] = {"error": repr(ex)}
    else:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                                "sample": "full workload; the reference has no CPU rasterizer, its own implementation of this path is CUDA (oracle/_ref, unmodified sources, sm_100a) on the same GPU"}
        line["e2e"]["note"] = "reference library calls + its torch glue's allocations; same host copies as the product arm"
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if failed:
        raise SystemExit("exchange_check failed: the exchanged gradients differ from the plain all_reduce (see exchange_check)")


if __name__ == "__main__":
    main()
