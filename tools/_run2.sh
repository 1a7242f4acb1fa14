for dm in 0 4 16 32; do
GSR_DIRECT_MAX=$dm timeout 300 python bench.py --config C2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/dm_$dm.json 2>/dev/null; python - <<P
import json
d=json.loads([x for x in open("gpurun_out/dm_$dm.json") if x.startswith("{")][-1])
print("dm=$dm", round(d["ms_per_step"],3), [round(s["ms"],3) if s["ms"] else None for s in d["roofline"]["stages"]])
P
done
