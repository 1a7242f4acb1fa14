"""Per-step timing diagnostic of bench.py's resident loop (ours / reference), with and without the clock sampler."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from scorp_b200 import synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
impl = sys.argv[2] if len(sys.argv) > 2 else "ours"
surfel = synth.CONFIGS[cfg]["surfel"]
arm = bench.Ours(surfel) if impl == "ours" else bench.Reference(surfel)
dev = torch.device("cuda", 0)
sc = synth.make_scene(cfg).to(dev); cam = sc.cameras[0]
grads = [g.to(dev) for g in synth.upstream_grads(cam.W, cam.H, 1, channels=bench.out_channels(surfel))]

def step():
    return arm.fwd_bwd(sc, cam, sc.bg, cam.viewmatrix, cam.projmatrix, cam.campos, grads, False)

def series(n, label, sampler=None):
    if sampler: sampler.start()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); step(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    if sampler: print(sampler.stop())
    print(f"{label:28s}", " ".join(f"{t:.2f}" for t in ts))

series(12, "cold")
series(12, "warm")
series(12, "with sampler", bench.ClockSampler(0))
series(12, "after sampler")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize(); print("20 back-to-back: ms/step", e0.elapsed_time(e1) / 20)
print(torch.cuda.memory_stats()["num_alloc_retries"], torch.cuda.memory_stats()["num_device_alloc"], torch.cuda.memory_stats()["num_device_free"])
