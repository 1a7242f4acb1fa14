"""Run under torchrun with 2+ GPUs (tests/test_gpu_peer_exchange.py launches it): the peer-memory exchange
(multiview.PeerExchange) must give every rank the same gradients as the NCCL FactoredExchange and as the dense
all-reduce of the parameter-gradient arena, over several steps (double-buffer parity) and for both rasterizers."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as h  # noqa: E402
from scorp_b200 import multiview as mv, synth  # noqa: E402


DEFAULT_CHUNKS = mv.PeerExchange.CHUNKS


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for surfel in (False, True):
        sc = synth.make_scene("C4" if surfel else "C2", P=40_000, W=320, H=208, n_views=world).to("cuda")
        fwd_f, bwd_f = (h.ours_surfel_forward, h.ours_surfel_backward) if surfel else (h.ours_forward, h.ours_backward)
        cam = sc.cameras[rank]
        grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5 + rank, channels=(3, 7) if surfel else (3, 1, 1))]
        fo = fwd_f(sc, cam)
        keys = ("means3D", "sh", "opacity", "scales", "rotations")
        # reference: plain backward + one all_reduce per tensor, four times: the mean is the comparison value and the
        # spread (float atomics in the blend backward) widens the tolerance, as tests/test_gpu_dgr_parity.py:_grad_ok does
        runs = []
        for _ in range(4):
            r = bwd_f(sc, cam, fo, grads)
            r = {k: r[k].clone() for k in keys}
            for k in keys:
                dist.all_reduce(r[k])
            runs.append(r)
        ref, tol = {}, {}
        for k in keys:
            st = torch.stack([r[k].double().reshape(-1) for r in runs])
            ref[k] = st.mean(0)
            spread = float((st.max(0).values - st.min(0).values).max())
            tol[k] = 1e-6 * float(ref[k].abs().max()) + 2.0 * spread + 1e-4 * ref[k].abs()      # north_star: rtol 1e-4
        for mode in ("factored", "peer", "peer3"):       # peer3: the head reduce pipelined over 3 row ranges
            mv.PeerExchange.CHUNKS = 3 if mode == "peer3" else DEFAULT_CHUNKS
            mv.install_factored_exchange(enable=True, peer=mode.startswith("peer"))
            try:
                for step in range(3):       # three steps: both parities of the double buffer, and reuse
                    out = bwd_f(sc, cam, fo, grads)
                    torch.cuda.synchronize()
                    for k in keys:
                        worst = float(((out[k].double().reshape(-1) - ref[k]).abs() / tol[k].clamp_min(1e-300)).max())
                        if not worst <= 1.0:
                            ok = False
                            print(f"rank {rank} surfel={surfel} {mode} step {step} {k}: worst ratio {worst}", flush=True)
                if mode.startswith("peer"):
                    from scorp_b200.diff_gaussian_rasterization import _C as c3
                    fx = c3._factored[0]
                    if rank == 0:
                        print(f"peer exchange active: {fx.ok} multicast={getattr(fx, 'multicast', None)} world={world} {fx.why}", flush=True)
                    if not fx.ok:
                        ok = False
            finally:
                mv.install_factored_exchange(enable=False)
                mv.PeerExchange.CHUNKS = DEFAULT_CHUNKS
        # precomputed colours (M == 0) under an installed factored exchange: still summed over ranks (one all_reduce)
        gcol = torch.Generator().manual_seed(11)
        colors = torch.rand(sc.P, 3, generator=gcol).cuda()
        fo_c = fwd_f(sc, cam, colors_precomp=colors)
        want_runs = []
        for _ in range(4):
            plain = bwd_f(sc, cam, fo_c, grads)
            w = {k: plain[k].clone() for k in ("means3D", "opacity", "scales", "rotations")}
            for k in w:
                dist.all_reduce(w[k])
            want_runs.append(w)
        want = want_runs[0]
        mv.install_factored_exchange(enable=True, peer=True)
        try:
            out = bwd_f(sc, cam, fo_c, grads)
            for k in want:
                good, worst = h.close_to_runs(out[k], [w[k] for w in want_runs])       # rtol 1e-4 + the runs' own spread
                if not good:
                    ok = False
                    print(f"rank {rank} surfel={surfel} precomputed colours {k}: worst ratio {worst}", flush=True)
        finally:
            mv.install_factored_exchange(enable=False)
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PEER_CHECK_OK" if int(t) == 1 else "PEER_CHECK_FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t) == 1 else 1)


if __name__ == "__main__":
    main()
