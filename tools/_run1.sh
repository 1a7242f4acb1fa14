timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
for c in C2 C4 C5 C1; do timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ts_$c.json 2> gpurun_out/ts_$c.err; python - <<P
import json
d=json.loads([x for x in open("gpurun_out/ts_$c.json") if x.startswith("{")][-1])
print("$c", round(d["ms_per_step"],3), [round(s["ms"],3) if s["ms"] else None for s in d["roofline"]["stages"]], "e2e", round(d["e2e"]["ms_per_step"],3))
P
done
