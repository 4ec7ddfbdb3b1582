"""One short-window call for an ncu launch list: python scripts/small_one.py L n nq"""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from kubeoperator_b200 import Rollup
L, n, nq = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
r = Rollup(0)
v = torch.rand(n * L, dtype=torch.float64, device="cuda")
o = torch.arange(n + 1, dtype=torch.int64, device="cuda") * L
outs = [torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(5)]
cnt = torch.zeros(n, dtype=torch.int64, device="cuda"); qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
for _ in range(3):
    r.series_rollup_dev(v.data_ptr(), o.data_ptr(), n, [0.5, 0.95, 0.99][:nq], outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(),
                        outs[3].data_ptr(), outs[4].data_ptr(), cnt.data_ptr(), qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
