"""Small fixed workload for ncu captures of series_flow: python scripts/flow_prof.py L n_series [nq] [iters]."""
import sys

sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth

L = int(sys.argv[1]) if len(sys.argv) > 1 else 40320
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1480
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 3
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
QS = [0.5, 0.95, 0.99][:nq]
r = Rollup(0)
cfg = "c2" if L <= 40320 else "c3"
import os, numpy as np
kk = os.environ.get("KO_KINDS")     # e.g. "2" = counters only, "0,2" = gauges and counters alternating
kinds = synth.mix_kinds(cfg, n) if not kk else np.resize(np.array([int(x) for x in kk.split(",")], np.int8), n)
vals, offs, kinds = synth.make_series_torch(cfg, 1, "cuda", L=L, kinds=kinds)
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
for it in range(iters):
    r.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
gb = vals.numel() * 8 / 1e9
ms, pms = r.last_series_kernel_ms(), r.last_series_pipe_ms()
print("L=%d n=%d nq=%d: long kernels %.3f ms = %.0f GB/s; series_flow %.3f ms = %.0f GB/s; %s"
      % (L, n, nq, ms, gb / ms * 1e3, pms, gb / pms * 1e3, r.series_status()))
