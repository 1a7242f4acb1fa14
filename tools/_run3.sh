for c in C2 C1; do timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/lb_$c.json 2>/dev/null; python - <<P
import json
d=json.loads([x for x in open("gpurun_out/lb_$c.json") if x.startswith("{")][-1])
print("$c", round(d["ms_per_step"],3), [round(s["ms"],3) if s["ms"] else None for s in d["roofline"]["stages"]])
P
done
