N=${1:-2}
mkdir -p gpurun_out
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline "$@"; }
run --config C2 --exchange factored > gpurun_out/mg${N}_fact.json 2> gpurun_out/mg${N}_fact.err
run --config C2 --exchange peer     > gpurun_out/mg${N}_peer.json 2> gpurun_out/mg${N}_peer.err
python - <<P
import json
for k in ("fact","peer"):
    try:
        d=json.loads([x for x in open("gpurun_out/mg${N}_%s.json"%k) if x.startswith("{")][-1])
        print(k, "ms", round(d["ms_per_step"],3), "Mpix/s", round(d["value"],1), "e2e ms", round(d["e2e"]["ms_per_step"],3), d["ms_per_step_spread"], d["config"]["parallelism"][-90:])
    except Exception as e: print(k, "failed", e)
P
tail -3 gpurun_out/mg${N}_peer.err
