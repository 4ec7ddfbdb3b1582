"""Development check of the slow exact path (series_long): series the one-pass kernel must defer (almost only stale
samples), in numbers that give every CTA several of them.  python scripts/long_check.py L n"""
import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup
L = int(sys.argv[1]); n = int(sys.argv[2])
r = Rollup(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
v = torch.rand(n * L, generator=g, device="cuda", dtype=torch.float64)
mask = torch.rand(n * L, generator=g, device="cuda") < 0.99
v[mask] = float("nan")
offs = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
r.series_rollup_dev(v.data_ptr(), offs.data_ptr(), n, [0.5, 0.95, 0.99], 0, 0, 0, 0, 0, 0, q.data_ptr(),
                    torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(L, n, r.series_status(), "kernel ms", r.last_series_kernel_ms())
vv = v.view(n, L)[:4].cpu().numpy()
for i in range(4):
    x = np.sort(vv[i][~np.isnan(vv[i])])
    exp = [x[int(np.ceil(p * len(x))) - 1] for p in (0.5, 0.95, 0.99)]
    print(i, q[i].cpu().numpy().tolist(), exp)
