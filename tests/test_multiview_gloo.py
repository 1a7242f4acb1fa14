"""CPU-only: the N>1 host logic of the view-sharded path (scorp_b200/multiview.py) on the gloo backend,
world_size 2 and 3: view partition, one all_reduce over the parameter-gradient head of a gradient arena laid
out exactly like the library's (_C.rasterize_gaussians_backward), and the sum/sum/max reduction of the
densification statistics."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from scorp_b200 import multiview as mv


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _fake_backward(P, M, view, surfel):
    """A gradient arena with the library's layout, filled deterministically per view."""
    ns = 2 if surfel else 3
    widths = (3, 3 * M, 1, ns, 4, 3, 3, 9 if surfel else 6)
    flat = torch.empty(P * sum(widths), dtype=torch.float32)
    g = torch.Generator().manual_seed(100 + view)
    flat.copy_(torch.randn(flat.shape, generator=g))
    return flat[: P * 3].view(P, 3), flat


def _worker(rank, world, port, P, M, n_views, surfel, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = mv.views_for_rank(n_views, rank, world)
        head = mv.render_views_sharded(lambda v: _fake_backward(P, M, v, surfel)[0], mine, P, M, surfel)
        n = P * mv.param_grad_floats(M, surfel)
        expect = sum(_fake_backward(P, M, v, surfel)[1][:n].double() for v in range(n_views))
        ok = torch.allclose(head.double(), expect, rtol=1e-5, atol=1e-5)
        parts = mv.split_param_grads(head, P, M, surfel)
        ok = ok and parts[1].shape == (P, M, 3) and parts[3].shape == (P, 2 if surfel else 3) and parts[4].shape == (P, 4)
        ok = ok and parts[4].data_ptr() == head.data_ptr() + 4 * (n - 4 * P)       # views, not copies
        # densification statistics
        acc = torch.full((P, 1), float(rank + 1)); den = torch.ones(P, 1); rad = torch.full((P,), float(10 * rank))
        mv.reduce_densification_stats(acc, den, rad)
        ok = ok and float(acc[0]) == world * (world + 1) / 2 and float(den[0]) == world and float(rad[0]) == 10 * (world - 1)
        q.put((rank, bool(ok), mine))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_views,surfel", [(2, 2, False), (2, 8, True), (3, 7, False)])
def test_view_sharded_allreduce_gloo(world, n_views, surfel):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 257, 4, n_views, surfel, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    seen = sorted(v for _, _, mine in res for v in mine)
    assert seen == list(range(n_views))


def _factored_worker(rank, world, port, P, M, D, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import helpers as h
        g = torch.Generator().manual_seed(5)
        means = torch.randn(P, 3, generator=g)                                       # replicated parameters
        per_rank = []
        for r in range(world):                                                       # every rank can build all inputs
            gr = torch.Generator().manual_seed(50 + r)
            head = torch.randn(P * 11, generator=gr)
            drgb = torch.randn(P, 3, generator=gr); drgb[torch.rand(P, generator=gr) < 0.3] = 0.0
            campos = torch.randn(3, generator=gr) * 5
            per_rank.append((head, drgb, campos))
        head, drgb, campos = (t.clone() for t in per_rank[rank])
        dsh = torch.empty(P * M * 3)
        fx = mv.FactoredExchange(rebuild=h.sh_grad_from_views_torch)                  # CUDA rebuild replaced for the CPU test
        fx(head, drgb, campos, means, dsh, P, M, D)
        want_head = sum(t[0].double() for t in per_rank)
        want_dsh = torch.empty(P * M * 3)
        h.sh_grad_from_views_torch(P, M, D, world, means, torch.stack([t[2] for t in per_rank]),
                                   torch.stack([t[1] for t in per_rank]), want_dsh)
        ok = torch.allclose(head.double(), want_head, rtol=1e-5, atol=1e-5) and torch.equal(dsh, want_dsh)
        # and it is the sum of the per-rank outer products an all-reduce of dL_dsh would have produced
        mine = torch.empty(P * M * 3)
        h.sh_grad_from_views_torch(P, M, D, 1, means, campos.view(1, 3), drgb.view(1, P, 3), mine)
        dist.all_reduce(mine)
        ok = ok and torch.allclose(mine, dsh, rtol=1e-4, atol=1e-5)
        # precomputed colours (M == 0): the installed exchange cannot factor anything, the `_C` level then reduces the
        # whole head with one all_reduce (multiview.reduce_without_factoring) instead of returning rank-local gradients
        head0 = per_rank[rank][0][: P * 11].clone()
        mv.reduce_without_factoring(fx, head0)
        ok = ok and torch.allclose(head0.double(), want_head, rtol=1e-5, atol=1e-5)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_factored_exchange_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_factored_worker, args=(r, world, port, 301, 16, 3, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res)
    assert mv.FactoredExchange.worth_it(8, 16) and not mv.FactoredExchange.worth_it(32, 16) and not mv.FactoredExchange.worth_it(2, 0)


def test_views_for_rank_partition():
    for world in (1, 2, 4, 8):
        for n in (1, 8, 13, 128):
            parts = [mv.views_for_rank(n, r, world) for r in range(world)]
            assert sorted(v for p in parts for v in p) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        mv.views_for_rank(4, 2, 2)
    assert mv.param_grad_floats(16) == 59 and mv.param_grad_floats(16, surfel=True) == 58


def test_single_process_is_a_no_op():
    a, flat = _fake_backward(10, 1, 0, False)
    head = mv.accumulate_param_grads(None, a, 10, 1)
    before = head.clone()
    assert torch.equal(mv.allreduce_param_grads(head), before)
    with pytest.raises(ValueError):
        mv.arena_of(torch.zeros(3))
