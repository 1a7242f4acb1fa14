python bench.py > gpurun_out/bench_ours_C2.json 2> gpurun_out/bench_ours_C2.err
python bench.py --impl reference > gpurun_out/bench_ref_C2.json 2> gpurun_out/bench_ref_C2.err
tail -c 400 gpurun_out/bench_ours_C2.json; tail -c 300 gpurun_out/bench_ref_C2.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gsr --launch-skip 40 -c 64 --csv --log-file gpurun_out/launches_C2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
