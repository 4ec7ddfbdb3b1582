"""Parity at the headline length: a slice of configs[2] (30-day series of the bench's generator, every kind) rolled up
on the device and compared series by series with the C oracle - all seven outputs.
Usage: python scripts/c3_parity.py [n_series] [seed] [config: c3 | c4]"""
import sys
import time

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np
import torch
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle
import test_gpu_series
from test_gpu_series import assert_series_parity

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5920
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cfg = sys.argv[3] if len(sys.argv) > 3 else "c3"
QS = [0.5, 0.95, 0.99]
r = Rollup(0)
vals, offs, kinds = synth.make_series_torch(cfg, seed, "cuda", kinds=synth.mix_kinds(cfg, n))
outs = {k: torch.zeros(n, dtype=torch.float64, device="cuda") for k in ("sum", "mean", "min", "max", "last")}
cnt = torch.zeros(n, dtype=torch.int64, device="cuda")
qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
r.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, outs["sum"].data_ptr(), outs["mean"].data_ptr(),
                    outs["min"].data_ptr(), outs["max"].data_ptr(), outs["last"].data_ptr(), cnt.data_ptr(),
                    qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
status = r.series_status()
values, offsets = vals.cpu().numpy(), offs.cpu().numpy()
t0 = time.time()
exp = c_oracle.series_rollup(values, offsets, QS, nthreads=16)
got = {k: v.cpu().numpy() for k, v in outs.items()}
got["count"], got["quantiles"] = cnt.cpu().numpy(), qt.cpu().numpy()
test_gpu_series.SUM_RTOL = max(1e-12, 172800 * 2.0 ** -53)      # the oracle's own left-to-right error bound
assert_series_parity(got, exp, values, offsets, QS)
print(cfg + " slice: %d series x %d samples (kinds %s) equal the oracle in all seven outputs; status %s; oracle %.1f s"
      % (n, int(offsets[1] - offsets[0]), np.bincount(kinds, minlength=4).tolist(), status, time.time() - t0))
