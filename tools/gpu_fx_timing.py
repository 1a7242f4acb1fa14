"""Times the pieces of the factored exchange on one GPU: the SH rebuild for n views at C2 size, and the all-gather
stand-in (device copies)."""
import torch, time
from scorp_b200 import multiview as mv

def t(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n

P, M, D = 1_000_000, 16, 3
means = torch.randn(P, 3, device="cuda")
out = torch.empty(P * M * 3, device="cuda")
for V in (1, 2, 4, 8):
    campos = torch.randn(V, 3, device="cuda") * 4
    drgb = torch.randn(V, P, 3, device="cuda"); drgb[torch.rand(V, P, device="cuda") < 0.44] = 0.0
    print(f"rebuild V={V}: {t(lambda: mv._rebuild_cuda(P, M, D, V, means, campos, drgb, out)):.3f} ms")
