N=${1:-2}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline "$@"; }
run --config C2 --exchange dense    > gpurun_out/mg${N}_dense.json 2> gpurun_out/mg${N}_dense.err
run --config C2 --exchange factored > gpurun_out/mg${N}_fact.json 2> gpurun_out/mg${N}_fact.err
run --config C2 --impl reference    > gpurun_out/mg${N}_ref.json 2> gpurun_out/mg${N}_ref.err
python - <<P
import json
for k in ("dense","fact","ref"):
    try:
        d=json.loads([x for x in open("gpurun_out/mg${N}_%s.json"%k) if x.startswith("{")][-1])
        print(k, round(d["ms_per_step"],3), round(d["value"],1), "e2e", round(d["e2e"]["ms_per_step"],3), d["config"].get("parallelism","")[:60])
    except Exception as e: print(k, "failed", e)
P
tail -3 gpurun_out/mg${N}_fact.err
