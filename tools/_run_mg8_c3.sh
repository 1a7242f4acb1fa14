# 8-GPU runs at C3 (6M Gaussians, 1080p, one view per rank): ours (factored exchange) and the reference arm
N=${1:-8}
mkdir -p gpurun_out
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline "$@"; }
run --config C3 --exchange factored > gpurun_out/mg${N}_C3_fact.json 2> gpurun_out/mg${N}_C3_fact.err
run --config C3 --impl reference    > gpurun_out/mg${N}_C3_ref.json 2> gpurun_out/mg${N}_C3_ref.err
python - <<P
import json
for k in ("C3_fact","C3_ref"):
    try:
        d=json.loads([x for x in open("gpurun_out/mg${N}_%s.json"%k) if x.startswith("{")][-1])
        print(k, "ms", round(d["ms_per_step"],3), "Mpix/s", round(d["value"],1), "e2e ms", round(d["e2e"]["ms_per_step"],3), d.get("exchange",{}).get("ms"))
    except Exception as e: print(k, "failed", e)
P
