"""gsr_knn_dist2_static: the allocation-free variant of distCUDA2 (scratch in a static __device__ array inside the
library; no counterpart in the reference -- include/gsr_b200.h).

Called through the C ABI and compared with the existing entry point gsr_knn_dist2 (caller scratch), whose result is
bit-identical to the reference's (tests/test_gpu_surfel_knn_api.py): the two run the same kernels, so the tolerance is
the one those tests use -- torch.equal. Two calls in a row (a large input, then a smaller, differently distributed one:
the shared array keeps the first call's contents and they must not leak into the second), the size limit, and the
absence of device allocations by the call itself.

Lives in a file that pytest collects LAST: passes alone in a fresh process and after
the other tests."""
import ctypes as C

import pytest
import torch

from scorp_b200 import _lib

pytestmark = pytest.mark.gpu


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(lib, pts):
    """gsr_knn_dist2 with caller scratch (what scorp_b200.simple_knn._C.distCUDA2 does)."""
    P = pts.shape[0]
    out = torch.empty(P, dtype=torch.float32, device=pts.device)
    nbytes = lib.gsr_knn_workspace_bytes(P)
    tmp = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
    _lib.check(lib.gsr_knn_dist2(P, pts.data_ptr(), out.data_ptr(), tmp.data_ptr(), nbytes, _stream()), "gsr_knn_dist2")
    return out


def _static(lib, pts, out):
    _lib.check(lib.gsr_knn_dist2_static(pts.shape[0], pts.data_ptr(), out.data_ptr(), _stream()), "gsr_knn_dist2_static")
    return out


def _inputs():
    g = torch.Generator().manual_seed(41)
    a = (torch.rand(200_000, 3, generator=g) * 8.0 - 4.0)                                  # uniform cube, large
    centres = torch.randn(40, 3, generator=g) * 3.0
    b = centres[torch.randint(0, 40, (30_011,), generator=g)] + 0.05 * torch.randn(30_011, 3, generator=g)
    b[:500] = b[1500:2000]; b[500:1000] = b[1500:2000]; b[1000:1500] = b[1500:2000]         # four copies each: the three
    #                                                                                        nearest neighbours are at distance 0
    return a.cuda().contiguous(), b.cuda().contiguous()


def test_static_scratch_knn_equals_existing_entry_point_on_two_calls_in_a_row():
    lib = _lib.lib()
    a, b = _inputs()
    want_a, want_b = _existing(lib, a), _existing(lib, b)
    out_a = torch.full((a.shape[0],), -1.0, device="cuda"); out_b = torch.full((b.shape[0],), -1.0, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    free_before = torch.cuda.mem_get_info()[0]
    _static(lib, a, out_a)            # first call
    _static(lib, b, out_b)            # second call, right behind it: shares the array, which still holds call one's grid
    torch.cuda.synchronize()
    # nothing was allocated: neither through torch's allocator nor behind its back (driver-level free memory unchanged)
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before
    # (a cudaMalloc of this input's scratch would take 4.8 MB; 1 MiB of slack for anything the driver does lazily)
    assert free_before - torch.cuda.mem_get_info()[0] < (1 << 20)
    assert torch.equal(out_a, want_a), f"call 1: {int((out_a != want_a).sum())} of {a.shape[0]} differ"
    assert torch.equal(out_b, want_b), f"call 2: {int((out_b != want_b).sum())} of {b.shape[0]} differ"
    assert float(out_b[:2000].max()) == 0.0 and float(out_a.min()) > 0.0     # the inputs exercise what they claim
    # and back to the first input: same answer again
    out_a2 = torch.empty_like(out_a)
    assert torch.equal(_static(lib, a, out_a2), want_a)


def test_static_scratch_knn_limit_and_arguments():
    lib = _lib.lib()
    cap = lib.gsr_knn_static_max_points()
    assert cap == 262144
    # exactly at the limit works and still equals the existing entry point
    g = torch.Generator().manual_seed(43)
    pts = torch.randn(cap, 3, generator=g).cuda().contiguous()
    out = torch.empty(cap, device="cuda")
    assert torch.equal(_static(lib, pts, out), _existing(lib, pts))
    # one more point is rejected with an error that names the limit; nothing is launched
    big = torch.zeros(cap + 1, 3, device="cuda"); out_big = torch.full((cap + 1,), 7.0, device="cuda")
    rc = lib.gsr_knn_dist2_static(cap + 1, big.data_ptr(), out_big.data_ptr(), _stream())
    assert rc != 0
    msg = lib.gsr_last_error().decode()
    assert str(cap) in msg and "exceed" in msg, msg
    torch.cuda.synchronize()
    assert float(out_big.min()) == 7.0 and float(out_big.max()) == 7.0
    with pytest.raises(_lib.GsrError):
        _lib.check(lib.gsr_knn_dist2_static(10, None, out.data_ptr(), _stream()), "gsr_knn_dist2_static")
    assert lib.gsr_knn_dist2_static(0, None, None, _stream()) == 0          # empty input: nothing to do, as gsr_knn_dist2
    # P < 4: fewer than three other points (the existing entry point's edge case, tests/test_gpu_surfel_knn_api.py)
    tiny = torch.tensor([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0], [0.0, 2.0, 0.0]], device="cuda")
    out_t = torch.empty(3, device="cuda")
    assert torch.equal(_static(lib, tiny, out_t), _existing(lib, tiny))
