# ncu evidence: (1) --set full of the default bench (C2), one launch per kernel, (2) launch list + --set full of the
# whole-iteration loop with every 8(f) row on. Each only after the plain run exited 0.
mkdir -p gpurun_out
OURS='regex:render_|preprocess|scatter_kernel|tile_s|l1_ssim|adam_kernel|densify|sh_grad|clamped'
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c2.json 2> gpurun_out/plain_c2.err || exit 1
ncu --set full --clock-control none --import-source on -k "$OURS" --launch-skip 77 -c 9 -o gpurun_out/prof_C2_r01c -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_c2.log 2>&1; echo "ncu C2 rc=$?"
python tools/bench_next.py --rows iteration --steps 3 --warmup 2 > gpurun_out/plain_iter.json 2> gpurun_out/plain_iter.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" --launch-skip 300 -c 40 --csv --log-file gpurun_out/launches_iteration.csv python tools/bench_next.py --rows iteration --steps 3 --warmup 2 > gpurun_out/ncu_l_iter.log 2>&1; echo "ncu iter list rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:l1_ssim|adam_kernel|densify|preprocess' --launch-skip 120 -c 14 -o gpurun_out/prof_iter_r01 -f python tools/bench_next.py --rows iteration --steps 3 --warmup 2 > gpurun_out/ncu_full_iter.log 2>&1; echo "ncu iter full rc=$?"
ls -la gpurun_out/*.ncu-rep
