"""Device-resident speed of the short-window kernels (series_small / series_medium): python scripts/small_check.py"""
import sys

sys.path.insert(0, ".")
import numpy as np
import torch
from kubeoperator_b200 import Rollup

QS = [0.5, 0.95, 0.99]
r = Rollup(0)
rng = np.random.default_rng(11)
for L, n in ((20, 1000000), (32, 1000000), (60, 500000), (240, 1000000), (256, 250000), (1000, 100000)):
    values = rng.beta(2, 5, n * L)
    values[rng.random(n * L) < 1e-3] = np.nan
    dv = torch.from_numpy(values).cuda()
    do = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
    outs = [torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(5)]
    cnt = torch.zeros(n, dtype=torch.int64, device="cuda")
    qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for nq in (3, 0):
        def dev():
            r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS[:nq], outs[0].data_ptr(), outs[1].data_ptr(),
                                outs[2].data_ptr(), outs[3].data_ptr(), outs[4].data_ptr(), cnt.data_ptr(),
                                qt.data_ptr(), st)
        for _ in range(3):
            dev()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dev()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        byts = n * L * 8 + n * 8 * (6 + nq) + n * 8
        print("L=%4d n=%7d nq=%d: %.3f ms per call, %.0f GB/s (samples + offsets + outputs)" % (L, n, nq, ms, byts / ms / 1e6))
