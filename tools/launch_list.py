#!/usr/bin/env python
"""One forward + backward per config through the `_C` level, a few times: the command the ncu launch-list pass wraps
(profiles/r02_launches_*.csv). No timing here; bench.py times with CUDA events.

    python tools/launch_list.py C3 [reps]
"""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import torch
import helpers as h
from scorp_b200 import synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
surfel = synth.CONFIGS[cfg]["surfel"]
fwd_only = synth.CONFIGS[cfg].get("forward_only", False)
sc = h.make(cfg); cam = sc.cameras[0]
fwd_f, bwd_f = (h.ours_surfel_forward, h.ours_surfel_backward) if surfel else (h.ours_forward, h.ours_backward)
grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 1, channels=(3, 7) if surfel else (3, 1, 1))]
for _ in range(reps):
    fo = fwd_f(sc, cam)
    if not fwd_only:
        bwd_f(sc, cam, fo, grads)
torch.cuda.synchronize()
print(cfg, "num_rendered", fo["num_rendered"])
