mkdir -p gpurun_out
python -m pytest tests/test_gpu_views.py tests/test_gpu_dgr_parity.py tests/test_gpu_api.py -q -x > gpurun_out/pytest_views.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_views.log
python bench.py --config C5 --no-cpu-baseline > gpurun_out/c5_batched.json 2> gpurun_out/c5_batched.err; echo "rc=$?"; tail -3 gpurun_out/c5_batched.err
python bench.py --config C5 --no-cpu-baseline --view-loop > gpurun_out/c5_loop.json 2> gpurun_out/c5_loop.err; echo "rc=$?"
python bench.py --config C5 --impl reference > gpurun_out/c5_ref.json 2> gpurun_out/c5_ref.err; echo "rc=$?"
python - <<P
import json
for f in ("c5_batched","c5_loop","c5_ref"):
    try:
        d=json.loads([x for x in open(f"gpurun_out/{f}.json") if x.startswith("{")][-1])
        print(f, "ms/step", round(d["ms_per_step"],3), "Mpix/s", round(d["value"],1), "e2e ms", round(d["e2e"]["ms_per_step"],3), [ (s["stage"], round(s["ms"],3)) for s in d.get("roofline",{}).get("stages",[]) if s["ms"]])
    except Exception as e: print(f, "failed", e)
P
