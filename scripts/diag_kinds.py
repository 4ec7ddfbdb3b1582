"""Diagnostic: per series kind, time the device-resident rollup and count series needing extra passes."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth

QS = [0.5, 0.95, 0.99]
L = int(sys.argv[1]) if len(sys.argv) > 1 else 40320
n = 592
r = Rollup(0)
rng = np.random.default_rng(0)
for kind, name in ((0, "gauge"), (1, "const"), (2, "counter"), (3, "discrete")):
    vals = np.concatenate([synth._one_series(rng, kind, L) for _ in range(n)])
    offs = np.arange(n + 1, dtype=np.int64) * L
    out = r.series_rollup(vals, offs, QS)
    fb = r.counters()["series_fallback"]
    dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
    q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    s = torch.zeros(n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ms = r.last_series_kernel_ms()
    print("%-9s L=%d n=%d fallback=%d kernel_ms=%.3f GB/s=%.0f" % (name, L, n, fb, ms, n * L * 8 / ms / 1e6))
for nq, qs in ((0, []), (1, [0.5]), (1, [0.99])):
    vals = np.concatenate([synth._one_series(rng, 0, L) for _ in range(n)])
    offs = np.arange(n + 1, dtype=np.int64) * L
    dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
    q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    s = torch.zeros(n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr() if nq else 0,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ms = r.last_series_kernel_ms()
    print("gauge qs=%s kernel_ms=%.3f GB/s=%.0f" % (qs, ms, n * L * 8 / ms / 1e6))
