# final verification of the round: GPU tests, smoke, every kernel once at odd sizes, default bench (both arms), launch list
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
python tools/gpu_sanitize.py > gpurun_out/all_kernels_once.log 2>&1; echo "all-kernels rc=$?"; tail -3 gpurun_out/all_kernels_once.log
python tools/bench_next.py > gpurun_out/next_rows.json 2> gpurun_out/next_rows.err; echo "next rows rc=$?"; cut -c1-420 gpurun_out/next_rows.json
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "ours rc=$?"
python - <<P
import json
for f in ("bench_ref","bench_ours"):
    try:
        d=json.loads([x for x in open(f"gpurun_out/{f}.json") if x.startswith("{")][-1])
        print(f, round(d["ms_per_step"],3), round(d["value"],1), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), d.get("roofline",{}).get("frac"), d.get("clocks"))
    except Exception as e: print(f, "failed", e)
P
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_C2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
