N=${1:-2}
mkdir -p gpurun_out
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --config C2 --exchange peer; }
GSR_PEER_NO_MULTICAST=1 run > gpurun_out/ab_nomc.json 2> gpurun_out/ab_nomc.err
run > gpurun_out/ab_mc.json 2> gpurun_out/ab_mc.err
GSR_PEER_NO_MULTICAST=1 run > gpurun_out/ab_nomc2.json 2> /dev/null
run > gpurun_out/ab_mc2.json 2> /dev/null
python - <<P
import json
for k in ("ab_nomc","ab_mc","ab_nomc2","ab_mc2"):
    try:
        d=json.loads([x for x in open("gpurun_out/%s.json"%k) if x.startswith("{")][-1])
        print(k, "ms", round(d["ms_per_step"],3), d["ms_per_step_spread"], "e2e ms", round(d["e2e"]["ms_per_step"],3), d["e2e"]["ms_per_step_spread"])
    except Exception as e: print(k, "failed", e)
P
