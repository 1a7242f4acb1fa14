#!/bin/bash
# densify row: tests, bench row, then (separately) the ncu launch list of the same bench command
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_densify.py -x -q 2>&1 | tail -3
python tools/bench_next.py --rows densify > gpurun_out/densify_row.json 2> gpurun_out/densify_row.err; echo "bench rc=$?"
cut -c1-700 gpurun_out/densify_row.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:densify_ -c 40 --csv \
    --log-file gpurun_out/densify_launches.csv python tools/bench_next.py --rows densify --steps 2 --warmup 1 > gpurun_out/densify_ncu.log 2>&1
echo "ncu rc=$?"; tail -12 gpurun_out/densify_launches.csv | cut -c1-260
