import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth
L, n = 40320, 2960
qs = [float(x) for x in sys.argv[1:]] if len(sys.argv) > 1 else [0.5]
r = Rollup(0)
rng = np.random.default_rng(0)
vals = np.concatenate([synth._one_series(rng, 0, L) for _ in range(n)])
offs = np.arange(n + 1, dtype=np.int64) * L
dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
q = torch.zeros((n, 4), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
for _ in range(3):
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("kernel_ms", r.last_series_kernel_ms())
