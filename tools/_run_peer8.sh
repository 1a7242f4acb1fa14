N=${1:-8}
mkdir -p gpurun_out
run() { timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --config C2 "$@"; }
run --exchange peer > gpurun_out/mg${N}_peer.json 2> gpurun_out/mg${N}_peer.err
run --exchange factored > gpurun_out/mg${N}_fact2.json 2> gpurun_out/mg${N}_fact2.err
python - <<P
import json
for k in ("peer","fact2"):
    try:
        d=json.loads([x for x in open("gpurun_out/mg${N}_%s.json"%k) if x.startswith("{")][-1])
        print(k, "ms", round(d["ms_per_step"],3), "Mpix/s", round(d["value"],1), d["ms_per_step_spread"], "e2e ms", round(d["e2e"]["ms_per_step"],3), d["config"]["parallelism"][-60:])
    except Exception as e: print(k, "failed", e)
P
tail -4 gpurun_out/mg${N}_peer.err
