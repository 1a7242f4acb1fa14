python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b1.json 2> gpurun_out/b1.err || tail -5 gpurun_out/b1.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b2.json 2> gpurun_out/b2.err || tail -5 gpurun_out/b2.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --impl reference > gpurun_out/b2r.json 2> gpurun_out/b2r.err || tail -5 gpurun_out/b2r.err
python - <<P
import json
for f in ("b1","b2","b2r"):
    d=json.loads([x for x in open(f"gpurun_out/{f}.json") if x.startswith("{")][-1])
    print(f, round(d["ms_per_step"],3), d.get("exchange"), d["config"].get("tile_lists"))
P
