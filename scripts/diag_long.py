"""Diagnostic: device-resident speed on 30-day series (172 800 samples), which do not fit one capture pass."""
import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth
L = int(sys.argv[1]) if len(sys.argv) > 1 else 172800
n = int(sys.argv[2]) if len(sys.argv) > 2 else 592
QS = [0.5, 0.95, 0.99]
r = Rollup(0)
rng = np.random.default_rng(0)
vals = np.concatenate([synth._one_series(rng, i % 4 if i % 4 != 2 else 0, L) for i in range(n)])
offs = np.arange(n + 1, dtype=np.int64) * L
dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
for _ in range(3):
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
ms = r.last_series_kernel_ms()
print("L=%d n=%d kernel_ms=%.3f GB/s=%.0f" % (L, n, ms, n * L * 8 / ms / 1e6))
