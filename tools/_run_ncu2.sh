mkdir -p gpurun_out
OURS='regex:render_|preprocess|scatter_kernel|tile_s|l1_ssim|adam_kernel|densify|sh_grad|clamped'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" --launch-skip 199 -c 24 --csv --log-file gpurun_out/launches_iteration.csv python tools/bench_next.py --rows iteration --steps 3 --warmup 2 > gpurun_out/ncu_l_iter.log 2>&1; echo "ncu iter list rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:l1_ssim|adam_kernel|densify|preprocess' --launch-skip 96 -c 7 -o gpurun_out/prof_iter_r01 -f python tools/bench_next.py --rows iteration --steps 3 --warmup 2 > gpurun_out/ncu_full_iter.log 2>&1; echo "ncu iter full rc=$?"
ls -la gpurun_out/*.ncu-rep
