#!/usr/bin/env python
"""Stage times (CUDA events inside the C ABI, gsr_profile_*) of a forward at one config, 20 repetitions: quick A/B of
binning variants selected by environment variables.   python tools/time_binning.py C3"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import torch
import helpers as h
from scorp_b200 import _lib, synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
kw = {}
if len(sys.argv) > 2:
    kw = dict(P=int(sys.argv[2]), W=int(sys.argv[3]), H=int(sys.argv[4]))
sc = h.make(cfg, **kw); cam = sc.cameras[0]
lib = _lib.lib()
for _ in range(3):
    fo = h.ours_forward(sc, cam)
lib.gsr_profile_enable(1)
for _ in range(20):
    fo = h.ours_forward(sc, cam)
st = (C.c_float * 5)()
lib.gsr_profile_read(st)
print(cfg, kw, {k: os.environ[k] for k in os.environ if k.startswith("GSR_")}, "N", fo["num_rendered"],
      "preprocess %.4f binning %.4f blend_fwd %.4f ms" % (st[0], st[1], st[2]))
