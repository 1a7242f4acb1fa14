# 8-GPU runs: C3 (6M Gaussians, 1080p, one view per rank) ours factored / dense and the reference arm; C2 ours factored
N=${1:-8}
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 15 --warmup 4 --no-cpu-baseline "$@"; }
run --config C3 --exchange factored > gpurun_out/mg${N}_C3_fact.json 2> gpurun_out/mg${N}_C3_fact.err
run --config C3 --exchange dense    > gpurun_out/mg${N}_C3_dense.json 2> gpurun_out/mg${N}_C3_dense.err
run --config C3 --impl reference    > gpurun_out/mg${N}_C3_ref.json 2> gpurun_out/mg${N}_C3_ref.err
run --config C2 --exchange factored > gpurun_out/mg${N}_C2_fact.json 2> gpurun_out/mg${N}_C2_fact.err
python - <<P
import json
for k in ("C3_fact","C3_dense","C3_ref","C2_fact"):
    try:
        d=json.loads([x for x in open("gpurun_out/mg${N}_%s.json"%k) if x.startswith("{")][-1])
        print(k, "ms", round(d["ms_per_step"],3), "Mpix/s", round(d["value"],1), "e2e ms", round(d["e2e"]["ms_per_step"],3), d.get("exchange",{}).get("ms"))
    except Exception as e: print(k, "failed", e)
P
