#!/usr/bin/env python
"""CUDA-event time of the SH rebuild kernel of the multi-GPU exchange (gsr_sh_grad_from_views) for V views at P
Gaussians, on one GPU: python tools/time_rebuild.py [P] [V ...]"""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from scorp_b200 import multiview as mv

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
views = [int(v) for v in sys.argv[2:]] or [2, 8]
M, D = 16, 3
torch.manual_seed(1)
means = torch.randn(P, 3, device="cuda")
for V in views:
    campos = torch.randn(V, 3, device="cuda") * 6
    drgb = torch.randn(V, P, 3, device="cuda")
    drgb[torch.rand(V, P, device="cuda") < 0.44] = 0.0           # invisible in that view, as at C2
    out = torch.empty(P * M * 3, device="cuda")
    for _ in range(3):
        mv._rebuild_cuda(P, M, D, V, means, campos, drgb, out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        mv._rebuild_cuda(P, M, D, V, means, campos, drgb, out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    mb = P * (12 * V + 12 + 192) / 1e6
    print(f"P={P} V={V}: {ms:.4f} ms, {mb:.0f} MB algorithmic = {mb / ms / 1e3:.2f} TB/s; checksum {float(out.double().abs().sum()):.6e}")
