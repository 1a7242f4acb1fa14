"""Factored multi-GPU exchange (scorp_b200/multiview.py:FactoredExchange, csrc/sh_exchange.cu) on ONE GPU:
with a single rank the rebuilt SH gradient must equal what the backward writes itself, bit for bit (the same
product basis_k * dL_dRGB), and every other gradient must be unaffected; the CUDA rebuild must agree with the
test-side torch restatement for several views."""
import pytest
import torch

import helpers as h
from scorp_b200 import multiview as mv, synth, _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("surfel,deg", [(False, 3), (False, 1), (True, 3), (True, 0)])
def test_single_rank_factored_equals_plain(surfel, deg):
    sc = h.make("C4" if surfel else "C2", P=30_000, W=320, H=208, sh_degree=deg); cam = sc.cameras[0]
    fwd_f, bwd_f = (h.ours_surfel_forward, h.ours_surfel_backward) if surfel else (h.ours_forward, h.ours_backward)
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5, channels=(3, 7) if surfel else (3, 1, 1))]
    fo = fwd_f(sc, cam)
    plain = bwd_f(sc, cam, fo, grads)
    plain_runs = [plain] + [bwd_f(sc, cam, fo, grads) for _ in range(3)]
    mv.install_factored_exchange(enable=True)
    try:
        fact = bwd_f(sc, cam, fo, grads)
    finally:
        mv.install_factored_exchange(enable=False)
    assert fact["sh"].shape == plain["sh"].shape
    for k in plain:                                         # two backward runs: equal up to the order of float atomics
        assert h.close_to_runs(fact[k], [r[k] for r in plain_runs])[0], k
    assert int((plain["sh"] != 0).sum()) > 0
    # the rebuild itself, fed with the colour gradient of the SAME backward run: the products basis_k * dL_dRGB the
    # backward wrote (rounding of the basis may differ by contraction, hence 1e-5 instead of bit equality)
    lib = _lib.lib()
    P, M = sc.P, sc.shs.shape[1]
    drgb = torch.empty(P, 3, device="cuda"); out = torch.empty(P * M * 3, device="cuda")
    rc = lib.gsr_clamped_color_grad(P, cam.W, cam.H, int(surfel), fo["geom"].data_ptr(), plain["colors"].data_ptr(),
                                    drgb.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    mv._rebuild_cuda(P, M, deg, 1, sc.means3D, cam.campos.view(1, 3).contiguous(), drgb.view(1, P, 3), out)
    assert torch.allclose(out.view(P, M, 3), plain["sh"], rtol=1e-5, atol=1e-7 * float(plain["sh"].abs().max()))
    # clamped channels carry no colour gradient
    assert int((drgb != 0).sum()) <= int((plain["colors"] != 0).sum())


def test_rebuild_kernel_matches_torch_for_several_views():
    torch.manual_seed(3)
    P, M, D, V = 5000, 16, 3, 4
    means = torch.randn(P, 3, device="cuda")
    campos = torch.randn(V, 3, device="cuda") * 4
    drgb = torch.randn(V, P, 3, device="cuda")
    drgb[torch.rand(V, P, device="cuda") < 0.4] = 0.0        # invisible in that view
    out = torch.empty(P * M * 3, device="cuda"); ref = torch.empty_like(out)
    mv._rebuild_cuda(P, M, D, V, means, campos, drgb, out)
    h.sh_grad_from_views_torch(P, M, D, V, means, campos, drgb, ref)
    assert torch.allclose(out, ref, rtol=1e-4, atol=1e-5)
    # degree below what M could hold: higher coefficients get zero
    out2 = torch.empty_like(out)
    mv._rebuild_cuda(P, M, 1, V, means, campos, drgb, out2)
    assert torch.equal(out2.view(P, M, 3)[:, 4:], torch.zeros(P, M - 4, 3, device="cuda"))
    with pytest.raises(_lib.GsrError):
        mv._rebuild_cuda(P, 3, 3, V, means, campos, drgb, out)   # (D+1)^2 > M


@pytest.mark.parametrize("surfel", [False, True])
def test_peer_consume_and_fused_finish_agree(surfel):
    """Phase 2 of the switch reduce on one GPU: gsr_peer_consume (head rows [P][12] -> planar head) and gsr_peer_finish
    (the same inside the SH rebuild kernel) must give identical results, and the planar head must hold the rows' fields
    in the arena order means3D | opacity | scales | rotations."""
    import ctypes as C
    torch.manual_seed(5)
    P, M, D, V = 7001, 16, 3, 3
    lib = _lib.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rows = torch.randn(P, 12, device="cuda")
    means = torch.randn(P, 3, device="cuda"); campos = torch.randn(V, 3, device="cuda") * 4
    drgb = torch.randn(V, P, 3, device="cuda")
    nh = 10 if surfel else 11
    head_a = torch.empty(P * nh, device="cuda"); head_b = torch.empty_like(head_a)
    dsh_a = torch.empty(P * M * 3, device="cuda"); dsh_b = torch.empty_like(dsh_a)
    # handed over in three row ranges, as the pipelined reduce does
    for r0, n in ((0, 3072), (3072, 3072), (6144, P - 6144)):
        assert lib.gsr_peer_consume(rows.data_ptr(), head_a.data_ptr(), P, r0, n, int(surfel), st) == 0
    mv._rebuild_cuda(P, M, D, V, means, campos, drgb, dsh_a)
    assert lib.gsr_peer_finish(P, M, D, V, means.data_ptr(), campos.data_ptr(), drgb.data_ptr(), dsh_b.data_ptr(),
                               rows.data_ptr(), head_b.data_ptr(), int(surfel), st) == 0
    assert torch.equal(head_a, head_b) and torch.equal(dsh_a, dsh_b)
    ns = 2 if surfel else 3
    want = torch.cat([rows[:, 0:3].reshape(-1), rows[:, 3], rows[:, 4:4 + ns].reshape(-1), rows[:, 4 + ns:8 + ns].reshape(-1)])
    assert torch.equal(head_a, want)
    assert lib.gsr_peer_reduce_shard(None, 0, P, 0, 1, st) != 0       # a NULL multicast address is rejected
    assert lib.gsr_peer_consume(rows.data_ptr(), head_a.data_ptr(), P, P - 5, 10, int(surfel), st) != 0      # range beyond P
