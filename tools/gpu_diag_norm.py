"""GPU diagnostic: identify the float sequence of torch's F.normalize / norm(2, dim=1) on [P,4] fp32."""
import itertools
import numpy as np
import torch
import torch.nn.functional as F

g = torch.Generator().manual_seed(5)
q = (torch.randn(200_000, 4, generator=g) * (0.25 + 3.0 * torch.rand(200_000, 1, generator=g))).cuda()
n_t = q.norm(2, 1).cpu().numpy()
nrm = F.normalize(q).cpu().numpy()
x = q.cpu().numpy().astype(np.float64)
f32 = lambda v: v.astype(np.float32).astype(np.float64)
sq_r = [f32(x[:, i] * x[:, i]) for i in range(4)]       # rounded squares
sq_e = [x[:, i] * x[:, i] for i in range(4)]            # exact squares (for fma)
best = []
def report(name, s):
    for rn, r in (("sqrt_rn", np.sqrt(s).astype(np.float32)),):
        d = int((r.view(np.int32) != n_t.view(np.int32)).sum())
        best.append((d, name + " " + rn))
for perm in itertools.permutations(range(4)):
    a, b, c, d = perm
    for fz in itertools.product((0, 1), repeat=4):      # leaf fused (exact square into the add) or rounded first
        L = [sq_e[i] if fz[k] else sq_r[i] for k, i in enumerate(perm)]
        # sequential ((a+b)+c)+d ; only one operand of the first add can be fused in hardware, but enumerate loosely
        s = f32(f32(f32(L[0] + L[1]) + L[2]) + L[3]); report(f"seq {perm} fuse{fz}", s)
        s = f32(f32(L[0] + L[1]) + f32(L[2] + L[3])); report(f"pair {perm} fuse{fz}", s)
s = x[:, 0] ** 2 + x[:, 1] ** 2 + x[:, 2] ** 2 + x[:, 3] ** 2
report("double sum", f32(s))
r = np.sqrt(s).astype(np.float32); best.append((int((r.view(np.int32) != n_t.view(np.int32)).sum()), "sqrt in double"))
best.sort()
for d, name in best[:12]:
    print(d, name)
# given torch's norm, is normalize == q / max(n, eps) with IEEE division?
den = np.maximum(n_t, np.float32(1e-12)).astype(np.float32)
dv = (q.cpu().numpy() / den[:, None]).astype(np.float32)
print("normalize == q / norm (IEEE div):", int((dv.view(np.int32) != nrm.view(np.int32)).sum()), "differ of", dv.size)
rc = (q.cpu().numpy() * (np.float32(1) / den)[:, None]).astype(np.float32)
print("normalize == q * (1/norm):", int((rc.view(np.int32) != nrm.view(np.int32)).sum()))
