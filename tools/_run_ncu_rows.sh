#!/bin/bash
# ncu --set full of the densify and PLY kernels (one launch each), after the plain commands ran green earlier in the round
mkdir -p gpurun_out
timeout 240 ncu --set full --clock-control none --import-source on -k 'regex:densify_' --launch-skip 4 -c 4 -o gpurun_out/prof_densify_r01 -f \
    python tools/bench_next.py --rows densify --steps 2 --warmup 1 > gpurun_out/ncu_densify.log 2>&1; echo "ncu densify rc=$?"
timeout 240 ncu --set full --clock-control none --import-source on -k 'regex:ply_' -c 2 -o gpurun_out/prof_ply_r01 -f \
    python tools/bench_next.py --rows ply > gpurun_out/ncu_ply.log 2>&1; echo "ncu ply rc=$?"
ls -la gpurun_out/*.ncu-rep
