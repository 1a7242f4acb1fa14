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
    return ap.parse_args()


def setup_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
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
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
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


def time_steps(step, K, W, world, sampler=None):
    """W warm-ups, then exactly K steps between barrier+synchronize, CUDA events, max over ranks (ms)."""
    for _ in range(W):
        step()
    torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if sampler else None
    return float(ms.item()), clocks


def stage_bytes(P, M, Pvis, N, W, H):
    """Algorithmic bytes per launch group (DESIGN.md 'Kernels')."""
    return {
        "preprocess_fwd": P * (44 + 12 * M) + Pvis * 75,
        "binning": 8 * P + 44 * N,
        "blend_fwd": 28 * N + 24 * W * H,
        "blend_bwd": 2 * 48 * Pvis + 44 * N + 28 * W * H,
        "preprocess_bwd": Pvis * ((99 + 12 * M) + (40 + 12 * M)),
    }


STAGES = ["preprocess_fwd", "binning", "blend_fwd", "blend_bwd", "preprocess_bwd"]


def cpu_baseline(cfg_name, W, H):
    """One fwd+bwd iteration of the workload through the CPU oracle port, all host cores."""
    from oracle import gs_oracle as O
    sc = synth.make_scene(cfg_name); cam = sc.cameras[0]
    grads = synth.upstream_grads(cam.W, cam.H, 1)
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    t0 = time.perf_counter()
    f = O.forward(sc.bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, cam.viewmatrix,
                  cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree, cam.campos)
    O.backward(f, *grads)
    dt = time.perf_counter() - t0
    return {"value": (W * H / 1e6) / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "ms_per_step": dt * 1e3,
            "sample": f"1 fwd+bwd iteration of the full {cfg_name} workload (oracle/gs_oracle.c, OpenMP)"}


def run_ours(args, rank, world, local):
    from scorp_b200 import _lib
    import scorp_b200.diff_gaussian_rasterization as dgr
    from scorp_b200.diff_gaussian_rasterization import _C
    lib = _lib.lib()
    dev = torch.device("cuda", local)
    n_views = max(world, 1)
    sc = synth.make_scene(args.config, n_views=n_views).to(dev)
    cam = sc.cameras[rank % n_views]
    W, H, P, M = cam.W, cam.H, sc.P, sc.shs.shape[1]
    grads = [g.to(dev) for g in synth.upstream_grads(W, H, 1 + rank)]
    E = torch.Tensor([])
    nparam = _C.param_grad_floats(M) * P

    def step_resident():
        out = _C.rasterize_gaussians(sc.bg, sc.means3D, E, sc.opacities, sc.scales, sc.rotations, 1.0, E,
                                     cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, sc.shs,
                                     sc.sh_degree, cam.campos, False, False)
        n, color, depth, alpha, radii, geom, binning, img = out
        g = _C.rasterize_gaussians_backward(sc.bg, sc.means3D, radii, E, sc.scales, sc.rotations, 1.0, E,
                                            cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0],
                                            grads[1], grads[2], sc.shs, sc.sh_degree, cam.campos, geom, n, binning,
                                            img, alpha, False)
        if world > 1:
            dist.all_reduce(_C.grad_arena(g[3])[:nparam])
        return n, radii

    n, radii = step_resident()
    Pvis = int((radii > 0).sum())
    sampler = ClockSampler(local) if rank == 0 else None
    lib.gsr_profile_enable(1)
    ms_total, clocks = time_steps(step_resident, args.steps, args.warmup, world, sampler)
    st = (C.c_float * 5)()
    lib.gsr_profile_read(st)
    lib.gsr_profile_enable(0)
    ms_step = ms_total / args.steps

    # ---- end to end through the public API with host inputs
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    h_cam = torch.cat([cam.viewmatrix.flatten(), cam.projmatrix.flatten(), cam.campos, sc.bg]).cpu().pin_memory()
    h_grads = torch.cat([g.flatten() for g in grads]).cpu().pin_memory()
    h_out = torch.empty(1, dtype=torch.float32).pin_memory()
    hw = H * W

    copy_stream = torch.cuda.Stream(dev)
    d_gr_buf = torch.empty_like(h_grads, device=dev)

    def step_e2e():
        # camera first (the forward needs it); the image-shaped upstream gradients are only needed by
        # the backward, so their H2D runs on a copy stream underneath the forward
        main = torch.cuda.current_stream(dev)
        d_cam = h_cam.to(dev, non_blocking=True)
        copy_stream.wait_stream(main)
        with torch.cuda.stream(copy_stream):
            d_gr_buf.copy_(h_grads, non_blocking=True)
        d_gr = d_gr_buf
        rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d_cam[35:38], 1.0,
                                               d_cam[0:16].view(4, 4), d_cam[16:32].view(4, 4), sc.sh_degree,
                                               d_cam[32:35], False, False)
        for p in params:
            p.grad = None
        color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(
            means3D=params[0], means2D=torch.zeros_like(params[0], requires_grad=True), shs=params[1],
            opacities=params[2], scales=params[3], rotations=params[4])
        main.wait_stream(copy_stream)
        torch.autograd.backward([color, depth, alpha],
                                [d_gr[:3 * hw].view(3, H, W), d_gr[3 * hw:4 * hw].view(1, H, W), d_gr[4 * hw:].view(1, H, W)])
        if world > 1:
            dist.all_reduce(_C.grad_arena(params[0].grad)[:nparam])
        h_out.copy_(color.sum().view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()     # the step's result is on the host (loss.item() in train_3dgs.py)

    for _ in range(PRIME):          # setup, not warm-up: caching-allocator growth for the autograd path
        step_e2e()
    ms_e2e_total, _ = time_steps(step_e2e, args.steps, args.warmup, world)
    ms_e2e = ms_e2e_total / args.steps
    return dict(W=W, H=H, P=P, M=M, Pvis=Pvis, N=n, ms_step=ms_step, ms_e2e=ms_e2e, stages=list(st), clocks=clocks,
                h2d=(h_cam.numel() + h_grads.numel()) * 4, d2h=4, launches=7 * args.steps)


def run_reference(args, rank, world, local):
    """The UNMODIFIED reference CUDA (oracle/_ref/libref_dgr.so) with what its torch glue does around it:
    fresh outputs, three scratch buffers and ten zero-filled gradient tensors per call."""
    sys.path.insert(0, str(ROOT / "tests"))
    from oracle import ref_cuda
    if not ref_cuda.available("dgr"):
        return None
    dev = torch.device("cuda", local)
    n_views = max(world, 1)
    sc = synth.make_scene(args.config, n_views=n_views).to(dev)
    cam = sc.cameras[rank % n_views]
    W, H, P, M = cam.W, cam.H, sc.P, sc.shs.shape[1]
    grads = [g.to(dev) for g in synth.upstream_grads(W, H, 1 + rank)]

    def fwd_bwd(bg, view, proj, campos, gr):
        f = ref_cuda.dgr_forward(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, view, proj,
                                 cam.tanfovx, cam.tanfovy, H, W, sc.shs, sc.sh_degree, campos, sync=False)
        g = ref_cuda.dgr_backward(f, bg, sc.means3D, None, sc.scales, sc.rotations, 1.0, None, view, proj,
                                  cam.tanfovx, cam.tanfovy, gr[0], gr[1], gr[2], sc.shs, sc.sh_degree, campos,
                                  sync=False)
        if world > 1:
            for k in ("means3D", "sh", "opacity", "scales", "rotations"):
                dist.all_reduce(g[k])
        return f, g

    def step_resident():
        return fwd_bwd(sc.bg, cam.viewmatrix, cam.projmatrix, cam.campos, grads)

    f, _ = step_resident()
    Pvis = int((f["radii"] > 0).sum())
    sampler = ClockSampler(local) if rank == 0 else None
    ms_total, clocks = time_steps(step_resident, args.steps, args.warmup, world, sampler)
    h_cam = torch.cat([cam.viewmatrix.flatten(), cam.projmatrix.flatten(), cam.campos, sc.bg]).cpu().pin_memory()
    h_grads = torch.cat([g.flatten() for g in grads]).cpu().pin_memory()
    h_out = torch.empty(1, dtype=torch.float32).pin_memory()
    hw = H * W

    def step_e2e():
        d_cam = h_cam.to(dev, non_blocking=True)
        d_gr = h_grads.to(dev, non_blocking=True)
        gr = [d_gr[:3 * hw].view(3, H, W), d_gr[3 * hw:4 * hw].view(1, H, W), d_gr[4 * hw:].view(1, H, W)]
        f, g = fwd_bwd(d_cam[35:38].contiguous(), d_cam[0:16].contiguous(), d_cam[16:32].contiguous(),
                       d_cam[32:35].contiguous(), gr)
        h_out.copy_(f["color"].sum().view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(PRIME):
        step_e2e()
    ms_e2e_total, _ = time_steps(step_e2e, args.steps, args.warmup, world)
    return dict(W=W, H=H, P=P, M=M, Pvis=Pvis, N=f["num_rendered"], ms_step=ms_total / args.steps,
                ms_e2e=ms_e2e_total / args.steps, stages=None, clocks=clocks,
                h2d=(h_cam.numel() + h_grads.numel()) * 4, d2h=4, launches=0)


def main():
    args = parse()
    rank, world, local = setup_dist(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    r = run_ours(args, rank, world, local) if args.impl == "ours" else run_reference(args, rank, world, local)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_dgr.so not built (needs /root/reference at build time)"}))
        return
    W, H = r["W"], r["H"]
    mpix = W * H / 1e6
    value = world * mpix / (r["ms_step"] * 1e-3)
    e2e_v = world * mpix / (r["ms_e2e"] * 1e-3)
    peak, peak_src = peaks()
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (seeded scene, scorp_b200/synth.py; no dataset ships with the reference)",
        "config": {"workload": synth.CONFIGS[args.config]["desc"], "id": args.config, "P": r["P"], "sh_coeffs": r["M"],
                   "width": W, "height": H, "visible": r["Pvis"], "num_rendered": r["N"],
                   "views_per_step": world, "parallelism": f"view-sharded dp{world}" if world > 1 else "single GPU",
                   "l2": "per-step inputs (236 B x P parameters + 24 B x N lists) exceed the 126 MB L2; no explicit flush"},
        "e2e": {"value": e2e_v, "unit": UNIT, "ms_per_step": r["ms_e2e"], "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": r["d2h"],
                "note": "public API + autograd; camera and upstream-gradient maps copied from pinned host memory each step (the gradient maps on a copy stream under the forward), scalar result read back; Gaussian parameters are model state resident in HBM as in train_3dgs.py"},
        "gpu_launches": r["launches"], "clocks": r["clocks"],
    }
    if args.impl == "ours":
        sb = stage_bytes(r["P"], r["M"], r["Pvis"], r["N"], W, H)
        stages = []
        for name, ms in zip(STAGES, r["stages"]):
            gbs = sb[name] / (ms * 1e-3) / 1e9 if ms and ms > 0 else None
            stages.append({"stage": name, "ms": ms, "algorithmic_bytes": sb[name], "achieved_gbs": gbs,
                           "frac": gbs / peak if gbs else None})
        dom = max(stages, key=lambda s: s["ms"] or 0)
        line["roofline"] = {"bound": "hbm", "kernel": dom["stage"], "achieved": dom["achieved_gbs"], "peak": peak,
                            "peak_source": peak_src, "unit": "GB/s", "frac": dom["frac"], "traffic": None,
                            "note": "blend stages are SM-issue bound (DESIGN.md); HBM fraction reported for the contract",
                            "stages": stages}
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
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
