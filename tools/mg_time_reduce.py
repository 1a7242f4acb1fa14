#!/usr/bin/env python
"""Under torchrun: CUDA-event time of the switch reduce of the head rows (gsr_peer_reduce_shard: multimem.ld_reduce +
multimem.st on a 1/N share per rank) alone, P rows of 48 bytes, between barriers.   torchrun ... tools/mg_time_reduce.py [P]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm
from scorp_b200 import _lib

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
t = symm.empty(P * 12, dtype=torch.float32, device="cuda")
hdl = symm.rendezvous(t, dist.group.WORLD)
mc = int(hdl.multicast_ptr)
lib = _lib.lib()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
t.fill_(1.0); torch.cuda.synchronize(); dist.barrier()
for _ in range(3):
    hdl.barrier(channel=0); _lib.check(lib.gsr_peer_reduce_shard(C.c_void_p(mc), 0, P, rank, world, st), "reduce"); hdl.barrier(channel=1)
torch.cuda.synchronize(); dist.barrier()
t.fill_(1.0); torch.cuda.synchronize(); dist.barrier()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
hdl.barrier(channel=0)
ev[0].record()
_lib.check(lib.gsr_peer_reduce_shard(C.c_void_p(mc), 0, P, rank, world, st), "reduce")
ev[1].record()
hdl.barrier(channel=1)
torch.cuda.synchronize()
ok = bool((t == float(world)).all())
ms = torch.tensor([ev[0].elapsed_time(ev[1])], device="cuda"); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    share = P * 48 / world / 1e6
    print(f"world={world} P={P} blocks={os.environ.get('GSR_REDUCE_BLOCKS', '148')}: {float(ms):.4f} ms for a {share:.1f} MB share per rank "
          f"({share / float(ms) :.0f} GB/s reduced per GPU), sums correct: {ok}", flush=True)
dist.destroy_process_group()
