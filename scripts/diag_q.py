import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth
L, n = 40320, 2960
r = Rollup(0)
rng = np.random.default_rng(0)
vals = np.concatenate([synth._one_series(rng, 0, L) for _ in range(n)])
offs = np.arange(n + 1, dtype=np.int64) * L
dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
q = torch.zeros((n, 4), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
for qs in ([], [0.5], [0.99], [0.95, 0.99], [0.5, 0.95], [0.5, 0.99], [0.5, 0.95, 0.99], [0.25, 0.5, 0.75]):
    for _ in range(3):
        r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr() if qs else 0,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ms = r.last_series_kernel_ms()
    out = r.series_rollup(vals[:L * 592], offs[:593], qs)
    print("qs=%-18s kernel_ms=%.3f GB/s=%.0f us/series/CTA=%.1f fallback(592)=%d" % (qs, ms, n * L * 8 / ms / 1e6, ms * 1e3 / (n / 296), r.counters()["series_fallback"]))
